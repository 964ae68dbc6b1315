"""CPU: pin the oracle (oracle/liboracle.so) against the golden vectors produced by the reference's
own Python code (tests/golden/make_golden.py).  fp64 tolerance 1e-9 relative (north_star)."""
import numpy as np
import pytest

from gym_pybullet_adrp_b200 import _abi
from gym_pybullet_adrp_b200.assets import drone_params
from gym_pybullet_adrp_b200.utils.enums import DroneModel
from tests import golden_cases as gc
from tests.oracle_lib import OracleEnv, dslpid_rows

TOL64 = 1e-9


@pytest.mark.parametrize("name", gc.CTRL_CASES)
def test_ctrl_open_loop_fp64(name):
    meta, z = gc.load(name)
    env = OracleEnv(gc.config_for(meta))
    obs0 = env.reset()
    assert gc.rel_err(obs0[0], z["obs0"]) < TOL64
    worst = 0.0
    for k in range(meta["steps"]):
        obs, rew, te, tr, _ = env.step(z["actions"][k][None])
        worst = max(worst, gc.rel_err(obs[0], z["obs"][k]))
        assert rew[0] == -1 and not te[0] and not tr[0]
    assert worst < TOL64, f"{name}: {worst:.3e}"


@pytest.mark.parametrize("name", gc.VEL_CASES)
def test_velocity_aviary_fp64(name):
    """VelocityAviary.step (VelocityAviary.py:129-170): velocity command -> DSLPID -> physics, as run by the reference."""
    meta, z = gc.load(name)
    env = OracleEnv(gc.config_for(meta))
    obs0 = env.reset()
    assert gc.rel_err(obs0[0], z["obs0"]) < TOL64
    worst = 0.0
    for k in range(meta["steps"]):
        obs, rew, te, tr, _ = env.step(z["actions"][k][None])
        worst = max(worst, gc.rel_err(obs[0], z["obs"][k]))
        assert rew[0] == -1 and not te[0] and not tr[0]
    assert worst < TOL64, f"{name}: {worst:.3e}"


@pytest.mark.parametrize("name", gc.RL_RESET_CASES)
def test_rl_reset_keeps_action_buffer(name):
    """reset() after `reset_at` steps: BaseRLAviary.action_buffer is created in __init__ (:66-67) and never cleared, so
    the reference's reset observation and the following steps list the previous episode's actions."""
    meta, z = gc.load(name)
    env = OracleEnv(gc.config_for(meta))
    assert gc.rel_err(env.reset()[0], z["obs0"]) < 1e-6
    for k in range(meta["steps"]):
        if k == meta["reset_at"]:
            o = env.reset()
            assert gc.rel_err(o[0, :, :12], z["obs_reset"][:, :12]) < 1e-6
            np.testing.assert_array_equal(o[0, :, 12:], z["obs_reset"][:, 12:])
            assert np.abs(o[0, :, 12:]).max() > 0                               # (it really carries old actions)
        obs, rew, te, tr, _ = env.step(z["actions"][k][None].astype(np.float64))
        assert gc.rel_err(obs[0, :, :12], z["obs"][k][:, :12]) < 1e-6, (name, k)
        np.testing.assert_array_equal(obs[0, :, 12:], z["obs"][k][:, 12:])
        assert abs(rew[0] - z["reward"][k]) < 1e-9 and bool(tr[0]) == bool(z["truncated"][k])


@pytest.mark.parametrize("name", gc.RL_CASES + gc.RL_VEL_CASES)
def test_rl_aviaries_fp64(name):
    meta, z = gc.load(name)
    env = OracleEnv(gc.config_for(meta))
    obs0 = env.reset()
    # the reference casts the 12 kinematic entries to float32 (BaseRLAviary.py:312)
    assert gc.rel_err(obs0[0], z["obs0"]) < 1e-6
    for k in range(meta["steps"]):
        obs, rew, te, tr, _ = env.step(z["actions"][k][None].astype(np.float64))
        assert gc.rel_err(obs[0, :, :12], z["obs"][k][:, :12]) < 1e-6, (name, k)
        np.testing.assert_array_equal(obs[0, :, 12:], z["obs"][k][:, 12:])      # action buffer: exact
        # ActionType.VEL: NumPy evaluates the velocity target on the float32 action (norm through BLAS sdot, whose
        # rounding depends on the BLAS build); the restatement evaluates the same expression in float64, so the
        # target differs by float32 rounding (6e-8 relative) and the reward by ~1e-9
        rtol = 1e-7 if name in gc.RL_VEL_CASES else 1e-9
        assert abs(rew[0] - z["reward"][k]) < rtol * max(1.0, abs(z["reward"][k])), (name, k)
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k]), (name, k)


@pytest.mark.parametrize("name", gc.PID_CASES)
def test_closed_loop_dslpid_fp64(name):
    """BASELINE config 1 (examples/pid.py driver loop), fused as ACT_PID_FULL."""
    meta, z = gc.load(name)
    env = OracleEnv(gc.config_for(meta))
    env.reset()
    st = np.zeros((1, 1, 20))
    o = z["obs"][0]
    st[0, 0, 0:7], st[0, 0, 7:10], st[0, 0, 10:13], st[0, 0, 16:20] = o[0:7], o[10:13], o[13:16], o[16:20]
    env.set_state(st)
    act = np.zeros((1, 1, 12))
    act[0, 0, 0:3] = meta["target"]
    worst = 0.0
    for k in range(1, meta["steps"]):
        obs, *_ = env.step(act)
        worst = max(worst, gc.rel_err(obs[0, 0], z["obs"][k]))
        assert gc.rel_err(obs[0, 0, 16:20], z["rpm"][k - 1]) < TOL64
    assert worst < TOL64, f"{name}: {worst:.3e}"
    assert abs(z["obs"][-1][2] - 1.0) < 0.1 or name != "pid_dyn_c1"    # it does reach the hover target


def test_dslpid_known_answers_fp64():
    _, z = gc.load("dslpid_kat")
    rows = z["rows"]
    for row in rows:
        ctrl = np.zeros((1, 9))
        for call in range(2):
            r = row[call * 41:(call + 1) * 41]
            cf2p, dt = int(r[0]), r[1]
            pos, quat, vel, tpos, trpy, tvel, trates = r[2:5], r[5:9], r[9:12], r[12:15], r[15:18], r[18:21], r[21:24]
            out = r[24:41]
            model = DroneModel.CF2P if cf2p else DroneModel.CF2X
            rpm, pos_e, yaw_e, ctrl = dslpid_rows(_abi.F64, _abi.MODEL_ID[model], _abi.drone_params_c(drone_params(model)),
                                                  1, dt, pos[None], quat[None], vel[None], tpos[None], trpy[None],
                                                  tvel[None], trates[None], ctrl)
            got = np.r_[rpm[0], pos_e[0], yaw_e[0], ctrl[0]]
            assert gc.rel_err(got, out) < TOL64, (cf2p, call, got - out)


@pytest.mark.parametrize("name", gc.RACE_CASES)
def test_race_env_fp64(name):
    """MultiRaceAviary + MellingerControl wrapper as run by the reference itself (tests/golden/make_golden_race.py),
    with the recorded random draws injected.  Flags, gate counters and latches must be identical."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta)
    env = OracleEnv(cfg)
    obs0 = env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    assert gc.rel_err(obs0[0], z["obs0"]) < TOL64                   # nominal pose, randomised track
    st = env.get_state()[0]
    assert np.abs(st[:, 0:7] - z["state0"][:, 0:7]).max() < 1e-12
    worst = 0.0
    for k in range(meta["steps"]):
        noise = z["noise"][k][None] if rc.get("disturbances") else None
        obs, rew, te, tr, fl = env.step(z["actions"][k][None], noise=noise)
        worst = max(worst, gc.rel_err(obs[0], z["obs"][k]))
        assert worst < TOL64, (name, k, worst)
        np.testing.assert_array_equal((fl[0] & _abi.FLAG_ELIMINATED) != 0, z["eliminated"][k])
        np.testing.assert_array_equal((fl[0] & _abi.FLAG_FINISHED) != 0, z["finished"][k])
        np.testing.assert_array_equal(obs[0][:, 48], z["current_gate"][k])
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k])
        assert rew[0] == 0


@pytest.mark.parametrize("name", gc.FULLSTATE_CASES)
def test_race_fullstate_commands_fp64(name):
    """(Command.FULLSTATE, (pos, vel, acc, yaw, rpy_rate, t)) tuples through the reference's MultiRaceAviary.step and
    MellingerControl.sendFullStateCmd: ADRP_ACT_MEL_FULL rows [pos3, vel3, acc3, yaw, rates3]."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, full_state_actions=True)
    env = OracleEnv(cfg)
    env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(meta["steps"]):
        obs, rew, te, tr, fl = env.step(z["actions"][k][None])
        assert gc.rel_err(obs[0], z["obs"][k]) < TOL64, (name, k)
        np.testing.assert_array_equal(obs[0][:, 48], z["current_gate"][k])
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k])
    assert z["current_gate"][-1].max() >= 1                      # the recorded flight does pass a gate


@pytest.mark.parametrize("name", gc.COMMANDER_CASES)
def test_race_commander_fp64(name):
    """(Command, args) tuples of every kind through the reference's own low_level_control / process_command_queue /
    _update_setpoint (MellingerControl.py:32-57,:292-303,:369-374,:547-699) over the declared planner restatement:
    `hardcoded` is user_controller/HardCodedController.py, unmodified, driven by the loop of scripts/sim.py (TAKEOFF ->
    FULLSTATE -> NOTIFY -> LAND -> NONE; both drones pass all four gates); `script` sends every other Command."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, command_actions=True)
    env = OracleEnv(cfg)
    env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(meta["steps"]):
        obs, rew, te, tr, fl = env.step(z["actions"][k][None])
        assert gc.rel_err(obs[0], z["obs"][k]) < TOL64, (name, k, gc.rel_err(obs[0], z["obs"][k]))
        np.testing.assert_array_equal(obs[0][:, 48], z["current_gate"][k])
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k])
    if "hardcoded" in name:
        assert z["current_gate"][-1].min() == 4 and z["terminated"][-1]      # the scripted flight finishes the track
    else:
        assert set(z["actions"][:, :, 0].astype(int).ravel()) == set(range(11))   # every Command occurs


@pytest.mark.parametrize("name", gc.WRAPPER_CASES)
def test_race_wrappers_fp64(name):
    """utils/wrapper.py as run by the reference on its own env: shaped reward of drone 0 (gate progress, +5 per gate),
    zeroed yaw command, terminated once drone 0 has passed two gates."""
    meta, z = gc.load(name)
    rc, cfg = gc.race_config_for(meta, wrappers=meta["wrappers"])
    env = OracleEnv(cfg)
    env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    ret = 0.0
    for k in range(meta["steps"]):
        obs, rew, te, tr, fl = env.step(z["actions"][k][None])
        assert gc.rel_err(obs[0], z["obs"][k]) < TOL64, (name, k)
        assert abs(rew[0] - z["reward"][k]) < 1e-9, (name, k, rew[0], z["reward"][k])
        assert bool(te[0]) == bool(z["terminated"][k]) and bool(tr[0]) == bool(z["truncated"][k]), (name, k)
        np.testing.assert_array_equal(obs[0][:, 48], z["current_gate"][k])
        ret += rew[0]
    assert bool(z["terminated"][-1]) and z["current_gate"][-1][0] == 2      # the recording ends on the wrapper's condition
    assert abs(ret - z["reward"].sum()) < 1e-8


def test_mellinger_schedule_known_answer():
    """the float64 sub-rate schedule of MellingerControl._step_controller (:395-411),
    first 120 ticks, computed with the reference's own expressions; the oracle's integer-stamp form
    ("run iff cur_time - last_att > 0.002") must fire on exactly the non-'1' ticks."""
    want = ("112121012121012120122210221202122022122022212022220122220221220222202221202222022220221220222202222022122022220222202212")
    last_att = last_pos = 0.0
    got = ""
    for tick in range(120):
        t = tick / 500
        if (t - last_att > 0.002) and (t - last_pos > 0.01):
            got += "0"; last_pos = last_att = t
        elif t - last_att > 0.002:
            got += "2"; last_att = t
        else:
            got += "1"
    assert got == want
    last_tick, fired = 0, ""
    for tick in range(120):
        if tick / 500.0 - last_tick / 500.0 > 0.002:
            fired += "x"; last_tick = tick
        else:
            fired += "1"
    assert all((a == "1") == (b == "1") for a, b in zip(want, fired))


def test_flop_table_is_the_mechanical_count():
    """bench.py's roofline numerators are the mechanical operation count of the oracle (oracle/flopcount.py: the same
    statements compiled with counting arithmetic types), not reading estimates."""
    import json, os, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    subprocess.run(["make", "-s", "-C", os.path.join(root, "oracle"), "liboracle_count.so"], check=True)
    out = subprocess.run([sys.executable, os.path.join(root, "oracle", "flopcount.py"), "--json"], check=True,
                         capture_output=True, text=True).stdout
    rows = json.loads(out)
    sys.path.insert(0, root)
    import bench
    key = {"c1": "C1:", "c5": "C5:", "c4": "C4:", "c2": "C2:", "c3": "C3:", "c5m": "C5M:"}
    for wl, (sub_f, sub_s, step_f, step_s) in bench.MECHANICAL_OPS.items():
        r = next(v for k, v in rows.items() if k.startswith(key[wl]))
        got = (r["per_substep_update"]["flops"], r["per_substep_update"]["sfu"], r["per_control_step"]["flops"], r["per_control_step"]["sfu"])
        for g, w in zip(got, (sub_f, sub_s, step_f, step_s)):
            assert abs(g - w) <= 0.01 * max(abs(w), 10.0), (wl, got)


KIN12 = [0, 1, 2, 7, 8, 9, 10, 11, 12, 13, 14, 15]


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
def test_obs_columns_12_is_the_kinematic_part_of_the_reference_row(precision):
    """adrp_config.obs_columns = 12 (CtrlAviary-kind; no counterpart in the reference): columns [pos3, rpy3, vel3, ang_vel3] of
    the 20-float _getDroneStateVector row (BaseAviary.py:545-565), in that order -- the oracle's two rows side by side."""
    from gym_pybullet_adrp_b200.utils.enums import Physics
    rng = np.random.default_rng(12)
    N, M = 3, 4
    xyz = [[0.8 * k, 0.3 * k, 0.4 + 0.2 * k] for k in range(M)]
    mk = lambda cols: OracleEnv(_abi.make_config(num_envs=N, num_drones=M, physics=Physics.PYB_GND_DRAG_DW, pyb_freq=240,   # noqa: E731
                                                 ctrl_freq=48, action_type=_abi.ACT_RPM_RAW, env_kind=_abi.ENV_CTRL,
                                                 precision=precision, initial_xyzs=xyz, obs_columns=cols))
    full, kin = mk(0), mk(12)
    assert full.O == 20 and kin.O == 12 and mk(20).O == 20
    assert np.array_equal(full.reset()[..., KIN12], kin.reset())
    for _ in range(12):
        a = 14468.0 * (1 + 0.05 * rng.uniform(-1, 1, (N, M, 4)))
        of, ok = full.step(a), kin.step(a)
        assert np.array_equal(of[0][..., KIN12], ok[0])
        for x, y in zip(of[1:], ok[1:]):
            assert np.array_equal(x, y)
