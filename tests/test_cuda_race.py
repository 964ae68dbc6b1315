"""GPU parity tests of the race env kernel (MultiRaceAviary.step + Mellinger per substep + track tests)
through the C ABI, against the golden vectors produced by the reference's own MultiRaceAviary /
MellingerControl (tests/golden/make_golden_race.py) and against the CPU oracle.

Tolerances (BASELINE.json north_star): fp64 1e-9 relative; fp32 1e-4 m / 1e-4 rad over 1 s of closed-loop
gate tracking; gate / contact / termination flags bit-exact."""
import numpy as np
import pytest

from gym_pybullet_adrp_b200 import _abi, configs
from gym_pybullet_adrp_b200.utils.enums import Physics, RaceMode
from tests import golden_cases as gc
from tests.oracle_lib import OracleEnv

pytestmark = pytest.mark.gpu

TOL64, TOL32 = 1e-9, 1e-4
LATCH = _abi.FLAG_ELIMINATED | _abi.FLAG_FINISHED | _abi.FLAG_CRASHED | _abi.FLAG_OUT_OF_BOUNDS | _abi.FLAG_UNSTABLE


def _handle(cfg):
    from gym_pybullet_adrp_b200 import _lib
    return _lib.Handle(cfg)


def _ang_diff(a, b):
    d = np.abs(a - b)
    return np.minimum(d, 2 * np.pi - d)


@pytest.mark.parametrize("name", gc.RACE_CASES)
def test_race_golden_fp64(name):
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=_abi.F64)
    h = _handle(cfg)
    obs0 = h.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None]).cpu().numpy()
    assert gc.rel_err(obs0[0], z["obs0"]) < TOL64                   # nominal pose, randomised track
    st = h.get_state().cpu().numpy()[0]
    assert np.abs(st[:, 0:7] - z["state0"][:, 0:7]).max() < 1e-12
    worst = 0.0
    for k in range(meta["steps"]):
        noise = z["noise"][k][None] if rc.get("disturbances") else None
        obs, rew, te, tr, fl = h.step(z["actions"][k][None], noise=noise)
        o, f = obs.cpu().numpy()[0], fl.cpu().numpy()[0]
        worst = max(worst, gc.rel_err(o, z["obs"][k]))
        assert worst < TOL64, (name, k, worst)
        np.testing.assert_array_equal((f & _abi.FLAG_ELIMINATED) != 0, z["eliminated"][k])
        np.testing.assert_array_equal((f & _abi.FLAG_FINISHED) != 0, z["finished"][k])
        np.testing.assert_array_equal(o[:, 48], z["current_gate"][k])
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k])
        assert float(rew[0]) == 0
    h.close()


@pytest.mark.parametrize("name", gc.RACE_CASES)
def test_race_golden_fp32_one_second(name):
    """fp32 build vs the float64 golden trajectory: 1e-4 m / 1e-4 rad over the first second (25 control steps
    = 500 Mellinger ticks) of closed-loop tracking, and still within 1e-3 over the next second."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=_abi.F32)
    h = _handle(cfg)
    h.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(min(50, meta["steps"])):
        noise = z["noise"][k][None] if rc.get("disturbances") else None
        o = h.step(z["actions"][k][None], noise=noise)[0].cpu().numpy()[0].astype(np.float64)
        tol = TOL32 if k < 25 else 10 * TOL32
        assert np.abs(o[:, 0:3] - z["obs"][k][:, 0:3]).max() < tol, (name, k)
        assert _ang_diff(o[:, 3:6], z["obs"][k][:, 3:6]).max() < tol, (name, k)
    h.close()


@pytest.mark.parametrize("name", gc.RACE_CASES)
def test_race_flags_bit_exact_fp32(name):
    """fp32 kernel over the whole golden scenario (gate passes, crashes, eliminations); after every step the
    fp32 oracle evaluates gate progress, observation and flags on the KERNEL's state: flags, gate counters and
    in-range bits must be identical, the rest of the observation equal to rounding."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=_abi.F32)
    h, o = _handle(cfg), OracleEnv(cfg)
    draws = gc.place_draws(rc, z["draws"], meta["M"])[None]
    h.reset(draws=draws); o.reset(draws=draws)
    G, NO = 4, 4
    bits = list(range(12 + 4 * G, 12 + 5 * G)) + list(range(12 + 5 * G + 3 * NO, 12 + 5 * G + 4 * NO)) + [12 + 5 * G + 4 * NO]
    events = 0
    for k in range(meta["steps"]):
        noise = z["noise"][k][None].astype(np.float32) if rc.get("disturbances") else None
        obs, rew, te, tr, fl = h.step(z["actions"][k][None].astype(np.float32), noise=noise)
        o.set_state(h.get_state().cpu().numpy())
        obs_o, rew_o, te_o, tr_o, fl_o = o.race_post()
        og, fg = obs.cpu().numpy(), fl.cpu().numpy()
        np.testing.assert_array_equal(fg & LATCH, fl_o & LATCH, err_msg=f"{name} step {k}")
        np.testing.assert_array_equal(og[..., bits], obs_o[..., bits], err_msg=f"{name} step {k}")
        assert int(te[0]) == int(te_o[0]) and int(tr[0]) == int(tr_o[0]), (name, k)
        assert np.abs(og - obs_o).max() < 1e-5, (name, k)
        events += int((fg & LATCH).any())
    h.close(); o.close()


def _random_setpoints(rng, cfg, N, M, k):
    a = np.zeros((N, M, 4))
    for i in range(M):
        a[:, i, 0] = cfg.init_xyz[i][0] + rng.uniform(-0.4, 0.4, N)
        a[:, i, 1] = cfg.init_xyz[i][1] + rng.uniform(-0.4, 0.4, N)
        a[:, i, 2] = 0.3 + 0.02 * k + rng.uniform(0.0, 0.3, N)
        a[:, i, 3] = rng.uniform(-0.5, 0.5, N)              # non-zero yaw setpoints: exercises the firmware's libm calls
    return a


@pytest.mark.parametrize("M,racemode,physics", [(2, RaceMode.COMPETE, Physics.PYB), (3, RaceMode.COMPETE, Physics.PYB_GND_DRAG_DW),
                                                (1, RaceMode.COMPARE, Physics.PYB), (2, RaceMode.COMPARE, Physics.DYN)])
def test_race_batched_philox_fp64(M, racemode, physics):
    """Many envs, level3 (every randomisation and disturbance on), draws from the in-kernel Philox stream:
    reset state, per-substep wind / motor noise, auto-reset -- against the oracle's Philox on the CPU."""
    N = 300
    rc = configs.load_config("level3")
    if M > 2:
        # downwash on: keep the drones > 1 m apart in xy (its alpha is singular for dz -> 0+, close formations are
        # chaotic and amplify the last-bit differences between CUDA's and glibc's double-precision libm)
        for i, pos in enumerate([[0.9, 0.9, 0.05], [-0.4, 1.3, 0.05], [2.0, -0.3, 0.05]]):
            rc["init_states"][f"drone{i}"] = configs.configify({"pos": pos, "vel": [0, 0, 0], "rpy": [0, 0, 0], "pqr": [0, 0, 0]})
    rc["episode_len_sec"] = 1.0                               # truncation + auto-reset inside the run
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=physics, racemode=racemode, precision=_abi.F64,
                                auto_reset=True, seed=1337, env_offset=5)
    h, o = _handle(cfg), OracleEnv(cfg)
    np.testing.assert_allclose(h.get_state().cpu().numpy(), o.get_state(), rtol=0, atol=1e-13)
    rng = np.random.default_rng(11)
    worst, resets = 0.0, 0
    for k in range(40):
        a = _random_setpoints(rng, cfg, N, M, k % 26)
        obs, rew, te, tr, fl = h.step(a)
        obs_o, rew_o, te_o, tr_o, fl_o = o.step(a, nthreads=8)
        worst = max(worst, gc.rel_err(obs.cpu().numpy(), obs_o))
        assert worst < TOL64, (k, worst)
        np.testing.assert_array_equal(fl.cpu().numpy(), fl_o)
        np.testing.assert_array_equal(te.cpu().numpy(), te_o)
        np.testing.assert_array_equal(tr.cpu().numpy(), tr_o)
        resets += int(te_o.sum() + tr_o.sum())
        assert gc.rel_err(h.get_state().cpu().numpy(), o.get_state()) < TOL64, k
    assert resets > 0                                          # the in-kernel reset path ran
    h.close(); o.close()


@pytest.mark.parametrize("name", gc.FULLSTATE_CASES)
@pytest.mark.parametrize("precision", [_abi.F64, _abi.F32])
def test_race_fullstate_commands_golden(name, precision):
    """ADRP_ACT_MEL_FULL ((Command.FULLSTATE, args) tuples: velocity / acceleration / yaw / body-rate setpoints)
    against the reference's own run."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=precision, full_state_actions=True)
    h = _handle(cfg)
    assert h.A == 13
    h.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(meta["steps"] if precision == _abi.F64 else 25):
        obs, rew, te, tr, fl = h.step(z["actions"][k][None])
        o = obs.cpu().numpy()[0].astype(np.float64)
        if precision == _abi.F64:
            assert gc.rel_err(o, z["obs"][k]) < TOL64, (name, k)
            np.testing.assert_array_equal(o[:, 48], z["current_gate"][k])
        else:
            assert np.abs(o[:, 0:3] - z["obs"][k][:, 0:3]).max() < TOL32, (name, k)
            assert _ang_diff(o[:, 3:6], z["obs"][k][:, 3:6]).max() < TOL32, (name, k)
    h.close()


@pytest.mark.parametrize("name", gc.COMMANDER_CASES)
@pytest.mark.parametrize("precision", [_abi.F64, _abi.F32])
def test_race_commander_golden(name, precision):
    """ADRP_ACT_MEL_CMD: (Command, args) rows of every Command through the high-level commander on the device (take-off /
    land / go-to planner, command queue semantics of MellingerControl.py:32-57,:292-303,:369-374) against the reference's
    own loop -- `hardcoded`: the unmodified user_controller/HardCodedController.py flying all four gates.
    fp64: 1e-9 over the whole run; fp32: 1e-4 m / rad over the first second (25 control steps)."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=precision, command_actions=True)
    h = _handle(cfg)
    assert h.A == 16
    h.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(meta["steps"] if precision == _abi.F64 else 25):
        obs, rew, te, tr, fl = h.step(z["actions"][k][None])
        o = obs.cpu().numpy()[0].astype(np.float64)
        if precision == _abi.F64:
            assert gc.rel_err(o, z["obs"][k]) < TOL64, (name, k, gc.rel_err(o, z["obs"][k]))
            np.testing.assert_array_equal(o[:, 48], z["current_gate"][k])
            assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k])
        else:
            assert np.abs(o[:, 0:3] - z["obs"][k][:, 0:3]).max() < TOL32, (name, k)
            assert _ang_diff(o[:, 3:6], z["obs"][k][:, 3:6]).max() < TOL32, (name, k)
    h.close()


def test_race_commander_batched_and_step_n():
    """Command rows over a batch (every env its own take-off height and go-to target), single steps vs one adrp_step_n launch,
    with in-kernel auto-resets: the planner state lives in HBM between launches and is re-initialised with the env."""
    import torch
    from gym_pybullet_adrp_b200 import configs
    N, M, K = 96, 2, 40
    rng = np.random.default_rng(4)
    rc = configs.load_config("getting_started")
    rc["episode_len_sec"] = 1.0
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, precision=_abi.F64, seed=2, auto_reset=True, command_actions=True)
    rows = np.zeros((K, N, M, 16))
    rows[0, :, :, 0] = _abi.CMD_ID["tko"]
    rows[0, :, :, 1] = rng.uniform(0.3, 0.8, (N, M))
    rows[0, :, :, 2] = 1.5
    rows[12, :, :, 0] = _abi.CMD_ID["gto"]
    rows[12, :, :, 1:4] = rng.uniform(-0.2, 0.2, (N, M, 3))
    rows[12, :, :, 5], rows[12, :, :, 6] = 1.0, 1.0             # duration, relative
    rows[30] = rows[0]                                            # after the in-kernel reset at 1 s: take off again
    a, b = _handle(cfg), _handle(cfg)
    outs = [[t.clone() for t in a.step(rows[k])] for k in range(K)]
    got = b.step_n(rows, keep_outputs=True)
    for j in range(5):
        assert torch.equal(torch.stack([o[j] for o in outs]), got[j]), j
    assert torch.equal(a.get_state(), b.get_state())
    z = torch.stack([o[0] for o in outs])[:, :, :, 2]
    assert float(z[11].min()) > 0.15 and int(torch.stack([o[3] for o in outs]).sum()) >= N    # they flew; every env was truncated
    a.close(); b.close()


@pytest.mark.parametrize("name", gc.WRAPPER_CASES)
@pytest.mark.parametrize("precision", [_abi.F64, _abi.F32])
def test_race_wrappers_golden(name, precision):
    """utils/wrapper.py fused into the step (adrp_config.wrappers) against the reference's own wrapped env:
    shaped reward of drone 0, zeroed yaw command, terminated once drone 0 has passed two gates."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, precision=precision, wrappers=meta["wrappers"])
    h = _handle(cfg)
    h.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    ret = 0.0
    for k in range(meta["steps"]):
        obs, rew, te, tr, fl = h.step(z["actions"][k][None])
        o = obs.cpu().numpy()[0].astype(np.float64)
        if precision == _abi.F64:
            assert gc.rel_err(o, z["obs"][k]) < TOL64, (name, k)
            assert abs(float(rew[0]) - z["reward"][k]) < 1e-9, (name, k)
            assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k]), (name, k)
            np.testing.assert_array_equal(o[:, 48], z["current_gate"][k])
        elif k < 25:                                          # fp32: 1 s of closed-loop tracking (north_star)
            assert np.abs(o[:, 0:3] - z["obs"][k][:, 0:3]).max() < TOL32, (name, k)
            assert abs(float(rew[0]) - z["reward"][k]) < 2 * TOL32, (name, k)
        ret += float(rew[0])
    if precision == _abi.F64:
        assert bool(te[0]) and abs(ret - z["reward"].sum()) < 1e-8
    else:
        assert abs(ret - z["reward"].sum()) < 0.05           # same gates passed (2 x 5) and the same way flown
    h.close()


def test_race_wrappers_batched_vs_oracle_fp64():
    """Wrappers + randomisation + auto-reset over many envs: the device-side wrapper state (gate id, target, previous
    position) follows the oracle's through in-kernel resets."""
    N, M = 200, 2
    rc = configs.load_config("level3")
    rc["episode_len_sec"] = 0.8
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE,
                                precision=_abi.F64, auto_reset=True, seed=7, env_offset=3,
                                wrappers=_abi.WRAP_OBS | _abi.WRAP_REWARD)
    h, o = _handle(cfg), OracleEnv(cfg)
    rng = np.random.default_rng(5)
    resets = 0
    for k in range(45):
        a = _random_setpoints(rng, cfg, N, M, k % 21)
        obs, rew, te, tr, fl = h.step(a)
        obs_o, rew_o, te_o, tr_o, fl_o = o.step(a, nthreads=8)
        assert gc.rel_err(obs.cpu().numpy(), obs_o) < TOL64, k
        assert np.abs(rew.cpu().numpy() - rew_o).max() < 1e-9, k
        np.testing.assert_array_equal(te.cpu().numpy(), te_o)
        np.testing.assert_array_equal(tr.cpu().numpy(), tr_o)
        resets += int(te_o.sum() + tr_o.sum())
    assert resets > N                                          # every env went through at least one in-kernel reset
    assert np.abs(rew_o).max() > 0                             # and the shaping is live
    h.close(); o.close()


def test_race_mellinger_formation_fp32():
    """north_star target shape: 8 CF2X per env, ground effect + drag + downwash, Mellinger per substep, no track.
    Run with the Bullet free body (PYB_GND_DRAG_DW): under the reference's semantics the Mellinger loop is unstable
    on the explicit DYN model (opposite yaw-torque sign, SURVEY F5/E1 -- the reference's own run eliminates both
    drones at step 2, tests/golden/race_dyn_compete_level1.npz), so DYN + Mellinger is covered by
    test_race_batched_philox_fp64 only.  fp32 kernel vs fp64 oracle over 2 s: 1e-4 m / 1e-4 rad."""
    N, M = 64, 8
    rc = configs.load_config("getting_started")
    rc["gates"], rc["obstacles"] = [], []
    rc["bounds"] = [[-10, -10, 0], [10, 10, 5]]
    rc["init_states"] = configs.configify({f"drone{i}": {"pos": [1.2 * (i % 4), 1.2 * (i // 4), 0.05], "vel": [0, 0, 0],
                                                         "rpy": [0, 0, 0], "pqr": [0, 0, 0]} for i in range(M)})
    kw = dict(num_envs=N, num_drones=M, racemode=RaceMode.COMPARE, physics=Physics.PYB_GND_DRAG_DW, ctrl_freq=25)
    h = _handle(_abi.make_race_config(rc, precision=_abi.F32, **kw))
    o = OracleEnv(_abi.make_race_config(rc, precision=_abi.F64, **kw))
    rng = np.random.default_rng(2)
    init = np.array(h.cfg.init_xyz)[None, :M, :]
    tgt = np.zeros((N, M, 4))
    tgt[:, :, 0:2] = init[:, :, 0:2] + rng.uniform(-0.2, 0.2, (N, M, 2))
    tgt[:, :, 2] = 0.5 + 0.15 * (np.arange(M) % 3)[None, :] + rng.uniform(0, 0.05, (N, M))
    for k in range(50):
        a = tgt.copy()
        a[:, :, 0:3] = init + (tgt[:, :, 0:3] - init) * min(1.0, (k + 1) / 15)
        og = h.step(a)[0].cpu().numpy().astype(np.float64)
        oc = o.step(a, nthreads=8)[0]
        assert np.abs(og[..., 0:3] - oc[..., 0:3]).max() < TOL32, k
        assert _ang_diff(og[..., 3:6], oc[..., 3:6]).max() < TOL32, k
    assert oc[..., 2].mean() > 0.5                             # the formation took off and holds
    h.close(); o.close()


def test_race_reset_mask_and_obs():
    rc = configs.load_config("level2")
    cfg = _abi.make_race_config(rc, num_envs=37, num_drones=2, racemode=RaceMode.COMPETE, precision=_abi.F64, seed=3)
    h, o = _handle(cfg), OracleEnv(cfg)
    mask = (np.arange(37) % 3 == 0).astype(np.uint8)
    a = np.zeros((37, 2, 4)); a[..., 2] = 0.5
    for _ in range(3):
        h.step(a); o.step(a)
    og = h.reset(env_mask=mask).cpu().numpy()
    oc = o.reset(env_mask=mask)
    sel = mask.astype(bool)
    assert gc.rel_err(og[sel], oc[sel]) < TOL64
    np.testing.assert_allclose(h.get_state().cpu().numpy(), o.get_state(), rtol=0, atol=1e-12)
    h.close(); o.close()


@pytest.mark.parametrize("G,NO,N,M", [(8, 8, 21, 2), (1, 0, 1, 1), (0, 0, 3, 2), (4, 4, 5, 7)])
def test_race_track_sizes_and_ragged_batches(G, NO, N, M):
    """Maximum track (ADRP_MAX_GATES / ADRP_MAX_OBSTACLES), a single gate, no track at all (the reference returns an
    all-zero observation without gates, MultiRaceAviary.py:598-600 -- here the kinematic part is kept), one env of
    one drone, seven drones per env (padded lanes)."""
    rc = configs.load_config("level3")
    g0 = [list(map(float, g)) for g in rc["gates"]]
    o0 = [list(map(float, b)) for b in rc["obstacles"]]
    gates = [[g0[k % 4][0] + 0.3 * (k // 4), g0[k % 4][1] - 0.4 * (k // 4), g0[k % 4][2], 0, 0, g0[k % 4][5] + 0.2 * (k // 4), g0[k % 4][6]]
             for k in range(G)]
    obst = [[o0[k % 4][0] - 0.35 * (k // 4), o0[k % 4][1] + 0.45 * (k // 4), o0[k % 4][2], 0, 0, 0] for k in range(NO)]
    rc["gates"], rc["obstacles"] = gates, obst
    for i in range(M):
        rc["init_states"][f"drone{i}"] = configs.configify({"pos": [0.9 + 0.35 * (i % 4), 0.9 + 0.35 * (i // 4), 0.05],
                                                           "vel": [0, 0, 0], "rpy": [0, 0, 0], "pqr": [0, 0, 0]})
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPETE,
                                precision=_abi.F64, seed=9)
    h, o = _handle(cfg), OracleEnv(cfg)
    assert h.O == 12 + 5 * G + 4 * NO + 1 + 6 * (M - 1)
    np.testing.assert_allclose(h.reset().cpu().numpy(), o.reset(), rtol=0, atol=1e-12)
    rng = np.random.default_rng(G + NO)
    for k in range(12):
        a = _random_setpoints(rng, cfg, N, M, k)
        obs, rew, te, tr, fl = h.step(a)
        obs_o, rew_o, te_o, tr_o, fl_o = o.step(a)
        assert gc.rel_err(obs.cpu().numpy(), obs_o) < TOL64, (G, NO, k)
        np.testing.assert_array_equal(fl.cpu().numpy(), fl_o)
        np.testing.assert_array_equal(te.cpu().numpy(), te_o)
    h.close(); o.close()


def test_race_rejects_wrong_action_type():
    from gym_pybullet_adrp_b200 import _lib
    cfg = _abi.make_config(num_envs=2, action_type=_abi.ACT_MEL, env_kind=_abi.ENV_CTRL)
    with pytest.raises(_lib.AdrpError):
        _lib.Handle(cfg)
