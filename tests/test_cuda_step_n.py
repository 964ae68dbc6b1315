"""GPU: adrp_step_n (K control steps per launch, state in registers across them) against K adrp_step calls -- for every env
family, both precisions, with in-kernel auto-resets (incl. the race env's re-randomised tracks, which the kernel must
re-stage in shared memory) and injected / Philox disturbances.

fp64 parity build (-fmad=false): bit-identical.  fp32 product build: the K-step instances are separately compiled kernels
(the single-step ones must not pay for the loop's registers), so multiply-adds are contracted differently here and there:
floats agree to rounding as the closed loops amplify it -- 1e-3 (relative above 1) over the 9 .. 40 control steps of these
cases, whose setpoints jump by decimetres every step; 2e-5 where the flight is gentle -- flags and counters exactly."""
import numpy as np
import pytest

from gym_pybullet_adrp_b200 import _abi, configs
from gym_pybullet_adrp_b200.utils.enums import DroneModel, Physics, RaceMode

pytestmark = pytest.mark.gpu
HOVER = float(np.sqrt(9.8 * 0.03454 / (4 * 3.16e-10)))


def _handle(cfg):
    from gym_pybullet_adrp_b200 import _lib
    return _lib.Handle(cfg)


def _compare(cfg, actions, noise=None, set_state=None, tol32=2e-5):
    """K single steps on one handle vs one K-step launch on its twin: every step's outputs and the final state."""
    import torch
    K = actions.shape[0]
    a, b = _handle(cfg), _handle(cfg)
    for h in (a, b):
        h.reset()
        if set_state is not None:
            h.set_state(set_state)
    outs = []
    for k in range(K):
        o = a.step(actions[k], None if noise is None else noise[k])
        outs.append([t.clone() for t in o])
    got = b.step_n(actions, noise=noise, keep_outputs=True)
    exact = cfg.precision == _abi.F64

    def same(x, y, what):
        if exact or not x.is_floating_point():
            assert torch.equal(x, y), (what, (x != y).nonzero()[:4])
        else:
            err = ((x - y).abs() / x.abs().clamp(min=1.0)).max().item()
            assert err < tol32, (what, err)

    for j in range(5):
        same(torch.stack([o[j] for o in outs]), got[j], j)
    same(a.get_state(), b.get_state(), "state")
    # ... and the handles stay interchangeable afterwards (counters, ring head, controller state were written back)
    o1, o2 = a.step(actions[0]), b.step(actions[0])
    for x, y in zip(o1, o2):
        same(x, y, "next step")
    # one slice held for all steps, outputs overwritten in place: the last step's remain
    a.reset(); b.reset()
    for k in range(3):
        last = [t.clone() for t in a.step(actions[0])]
    b.step_n(actions[0], n_steps=3)
    for x, y in zip(last, (b.obs, b.reward, b.terminated, b.truncated, b.flags)):
        same(x, y, "held action")
    a.close(); b.close()


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
@pytest.mark.parametrize("physics,M", [(Physics.DYN_GND_DRAG_DW, 8), (Physics.PYB_GND_DRAG_DW, 8), (Physics.PYB, 3),
                                       (Physics.DYN, 1), (Physics.PYB_DW, 16)])
def test_ctrl_aviary(physics, M, precision):
    N, K = 300, 9
    rng = np.random.default_rng(M)
    xyz = [[1.2 * (k % 3), 1.2 * (k // 3), 0.5 + 0.25 * (k % 4)] for k in range(M)]
    cfg = _abi.make_config(num_envs=N, num_drones=M, physics=physics, pyb_freq=240, ctrl_freq=30,
                           action_type=_abi.ACT_RPM_RAW, env_kind=_abi.ENV_CTRL, precision=precision, initial_xyzs=xyz)
    _compare(cfg, HOVER * (1 + 0.01 * rng.uniform(-1, 1, (K, N, M, 4))))


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
@pytest.mark.parametrize("kind,M,act,A", [(_abi.ENV_HOVER, 1, _abi.ACT_RPM, 4), (_abi.ENV_MULTIHOVER, 3, _abi.ACT_RPM, 4),
                                          (_abi.ENV_HOVER, 1, _abi.ACT_ONE_D_RPM, 1), (_abi.ENV_HOVER, 1, _abi.ACT_PID, 3)])
def test_rl_aviaries_with_auto_reset(kind, M, act, A, precision):
    """Episodes of 0.4 s: every env is truncated and re-initialised inside the launch several times; the action buffer
    (ring) and its head carry over from step to step and across the resets."""
    N, K = 200, 40
    rng = np.random.default_rng(7)
    cfg = _abi.make_config(num_envs=N, num_drones=M, physics=Physics.PYB, pyb_freq=240, ctrl_freq=30, action_type=act,
                           env_kind=kind, precision=precision, auto_reset=True, episode_len_sec=0.4)
    acts = rng.uniform(-1, 1, (K, N, M, A))
    if act == _abi.ACT_PID:
        acts[..., 0:2] = 0.0
        acts[..., 2] = 0.5 + 0.3 * np.abs(acts[..., 2])
    _compare(cfg, acts)


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
def test_dslpid_closed_loop_in_one_launch(precision):
    """BASELINE config 1 (examples/pid.py:101-147): 240 control steps of DSLPID hover towards a fixed target = one launch."""
    import torch
    cfg = _abi.make_config(num_envs=5, num_drones=1, physics=Physics.DYN, pyb_freq=240, ctrl_freq=30,
                           action_type=_abi.ACT_PID_FULL, env_kind=_abi.ENV_CTRL, precision=precision)
    a, b = _handle(cfg), _handle(cfg)
    act = np.zeros((5, 1, 12))
    act[:, 0, 2] = 1.0
    for _ in range(240):
        a.step(act)
    b.step_n(act, n_steps=240)
    if precision == _abi.F64:
        assert torch.equal(a.obs, b.obs) and torch.equal(a.get_state(), b.get_state())
    else:
        assert (a.obs - b.obs).abs().max().item() < 1e-4
    assert abs(float(b.obs[0, 0, 2]) - 1.0) < 0.05                 # it hovers at the target
    a.close(); b.close()


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
@pytest.mark.parametrize("level,mode,M,physics,inject", [("level3", RaceMode.COMPETE, 2, Physics.PYB, False),
                                                         ("level3", RaceMode.COMPARE, 2, Physics.PYB_GND_DRAG_DW, True),
                                                         ("getting_started", RaceMode.COMPETE, 2, Physics.PYB, False)])
def test_race_env_with_auto_reset(level, mode, M, physics, inject, precision):
    """Race env: Philox or injected disturbances per step, eliminations, in-kernel resets with re-randomised tracks."""
    N, K = 150, 30
    rng = np.random.default_rng(11)
    rc = configs.load_config(level)
    rc["episode_len_sec"] = 0.5                                    # truncation (and a fresh random track) every 13 steps
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=physics, racemode=mode, precision=precision, seed=3,
                                auto_reset=True)
    init = np.array([[cfg.init_xyz[i][k] for k in range(3)] for i in range(M)])
    acts = np.zeros((K, N, M, 4))
    acts[..., 0:3] = init[None, None] + rng.uniform(-0.3, 0.3, (K, N, M, 3))
    acts[..., 2] += 0.4
    noise = None
    if inject:
        noise = np.zeros((K, N, M, 20, 7))
        noise[..., 0:3] = rng.uniform(-0.05, 0.05, noise[..., 0:3].shape)
        noise[..., 3:7] = 1e-3 * rng.standard_normal(noise[..., 3:7].shape)
    _compare(cfg, acts, noise, tol32=1e-3)


def test_race_wrappers_and_fullstate_actions():
    """Reward shaping state (utils/wrapper.py) across the steps of one launch; 13-float FULLSTATE rows."""
    N, M, K = 64, 2, 25
    rng = np.random.default_rng(5)
    rc = configs.load_config("getting_started")
    rc["episode_len_sec"] = 0.6
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE,
                                precision=_abi.F64, seed=9, auto_reset=True, wrappers=3)
    init = np.array([[cfg.init_xyz[i][k] for k in range(3)] for i in range(M)])
    acts = np.zeros((K, N, M, 4))
    acts[..., 0:3] = init[None, None] + rng.uniform(-0.2, 0.2, (K, N, M, 3)) + [0.0, 0.0, 0.5]
    _compare(cfg, acts)
    cfg = _abi.make_race_config(rc, num_envs=N, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE,
                                precision=_abi.F32, seed=9, auto_reset=True, full_state_actions=True)
    full = np.zeros((K, N, M, 13))
    full[..., 0:3] = acts[..., 0:3]
    full[..., 3:6] = rng.uniform(-0.2, 0.2, (K, N, M, 3))
    full[..., 9] = rng.uniform(-0.3, 0.3, (K, N, M))
    _compare(cfg, full, tol32=1e-3)


def test_step_n_is_graph_capturable_and_plain_steps_too():
    """CUDA-graph replay of K adrp_step launches = the same K plain launches."""
    import torch
    N, M, K = 256, 8, 6
    rng = np.random.default_rng(2)
    xyz = [[1.2 * (k % 3), 1.2 * (k // 3), 0.5 + 0.25 * (k % 4)] for k in range(M)]
    cfg = _abi.make_config(num_envs=N, num_drones=M, physics=Physics.DYN_GND_DRAG_DW, pyb_freq=240, ctrl_freq=30,
                           action_type=_abi.ACT_RPM_RAW, env_kind=_abi.ENV_CTRL, precision=_abi.F32, initial_xyzs=xyz)
    a, b = _handle(cfg), _handle(cfg)
    acts = torch.as_tensor(HOVER * (1 + 0.01 * rng.uniform(-1, 1, (K, N, M, 4))), dtype=torch.float32, device="cuda")
    for k in range(K):
        a.step(acts[k])
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        b.step(acts[0])                                           # warm-up outside the capture (occupancy query, lazy loading)
        b.reset()
        with torch.cuda.graph(g, stream=side):
            for k in range(K):
                b.step(acts[k])
    torch.cuda.current_stream().wait_stream(side)
    b.reset()
    g.replay()
    torch.cuda.synchronize()
    assert torch.equal(a.obs, b.obs) and torch.equal(a.get_state(), b.get_state())
    a.close(); b.close()


def _random_policy_cfg(case, precision, N, env_offset=0):
    if case == "race":
        rc = configs.load_config("level3")
        rc["episode_len_sec"] = 0.5
        return _abi.make_race_config(rc, num_envs=N, num_drones=2, physics=Physics.PYB, racemode=RaceMode.COMPETE,
                                     precision=precision, seed=0x1234567812345, auto_reset=True, env_offset=env_offset), 2, 4
    kind, M, act, A = {"hover_rpm": (_abi.ENV_HOVER, 1, _abi.ACT_RPM, 4), "multihover_one_d": (_abi.ENV_MULTIHOVER, 3, _abi.ACT_ONE_D_RPM, 1),
                       "hover_pid": (_abi.ENV_HOVER, 1, _abi.ACT_PID, 3), "multihover_vel": (_abi.ENV_MULTIHOVER, 2, _abi.ACT_VEL, 4),
                       "ctrl_vel": (_abi.ENV_CTRL, 2, _abi.ACT_VEL, 4), "ctrl_rpm": (_abi.ENV_CTRL, 2, _abi.ACT_RPM, 4)}[case]
    return _abi.make_config(num_envs=N, num_drones=M, physics=Physics.PYB, pyb_freq=240, ctrl_freq=30, action_type=act, env_kind=kind,
                            precision=precision, auto_reset=True, episode_len_sec=0.4, seed=0x1234567812345, env_offset=env_offset), M, A


@pytest.mark.parametrize("precision", [_abi.F32, _abi.F64])
@pytest.mark.parametrize("case", ["hover_rpm", "multihover_one_d", "hover_pid", "multihover_vel", "race", "ctrl_vel", "ctrl_rpm"])
def test_random_policy_drawn_in_the_kernel(case, precision):
    """ADRP_ACTIONS_PHILOX (step_n(None, K)): the launch that draws its own U[-1, 1) actions equals, bit for bit in both builds
    (the same kernel instance runs both), the launch that is handed those actions -- computed here by the NumPy Philox of
    tests/philox_np.py (pinned to Random123's known answers on the CPU side); two calls in a row continue the counter, a
    shard draws the rows of the whole batch, and set_rollout_step replays.  (ctrl_rpm, fp32: a CtrlAviary-kind handle draws on
    the generic instance -- its specialised ones do not carry the draw -- so that one comparison is to rounding, 2e-5.)"""
    import torch
    from tests.philox_np import random_policy_actions
    N, K1, K2, off = 96, 7, 5, 1000
    cfg, M, A = _random_policy_cfg(case, precision, N, off)
    a, b = _handle(cfg), _handle(cfg)
    a.reset(); b.reset()
    want_actions = random_policy_actions(cfg.seed, off, N, M, A, 0, K1 + K2)
    loose = case == "ctrl_rpm" and precision == _abi.F32

    def same(x, y):
        if loose and x.is_floating_point():
            assert ((x - y).abs() / x.abs().clamp(min=1.0)).max().item() < 2e-5
        else:
            assert torch.equal(x, y), (x != y).nonzero()[:4]

    got = a.step_n(None, n_steps=K1, keep_outputs=True)
    want = b.step_n(want_actions[:K1], keep_outputs=True)
    for x, y in zip(got, want):
        same(x, y)
    got = a.step_n(None, n_steps=K2, keep_outputs=True)           # continues at rollout step K1
    want = b.step_n(want_actions[K1:], keep_outputs=True)
    for x, y in zip(got, want):
        same(x, y)
    same(a.get_state(), b.get_state())
    # a shard of the batch (env_offset = global index of its first env) reproduces those rows
    lo, hi = 32, 64
    cfg_s, _, _ = _random_policy_cfg(case, precision, hi - lo, off + lo)
    s = _handle(cfg_s)
    s.reset()
    w = _handle(cfg)                                              # (a fresh one: the RL aviaries' action buffer survives reset())
    w.reset()
    whole = w.step_n(None, n_steps=K1, keep_outputs=True)
    part = s.step_n(None, n_steps=K1, keep_outputs=True)
    for x, y in zip(whole, part):
        assert torch.equal(x[:, lo:hi], y)
    # adrp_set_rollout_step: the draws of rollout steps 5 .. 7 on fresh handles
    r1, r2 = _handle(cfg), _handle(cfg)
    r1.reset(); r2.reset()
    r1.set_rollout_step(5)
    for x, y in zip(r1.step_n(None, n_steps=3), r2.step_n(want_actions[5:8])):
        same(x, y)
    r1.step_n(None, n_steps=1)                                    # K = 1, no action buffer at all
    r1.close(); r2.close()
    for h in (a, b, s, w):
        h.close()


def test_random_policy_rejects_wide_action_rows():
    cfg = _abi.make_config(num_envs=4, num_drones=1, physics=Physics.DYN, pyb_freq=240, ctrl_freq=30,
                           action_type=_abi.ACT_PID_FULL, env_kind=_abi.ENV_CTRL, precision=_abi.F32)
    h = _handle(cfg)
    with pytest.raises(RuntimeError, match="ADRP_ACTIONS_PHILOX"):
        h.step_n(None, n_steps=2)
    h.close()
