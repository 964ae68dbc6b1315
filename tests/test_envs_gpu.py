"""GPU: the env / controller classes (the reference's interface) end to end against the golden vectors the
reference's own classes produced -- the tests read like the reference's usage (examples/pid.py, learn.py, sim.py)."""
import numpy as np
import pytest

from gym_pybullet_adrp_b200 import _abi, configs
from gym_pybullet_adrp_b200.control import DSLPIDControl
from gym_pybullet_adrp_b200.envs import CtrlAviary, HoverAviary, MultiHoverAviary, MultiRaceAviary, VelocityAviary
from gym_pybullet_adrp_b200.utils.enums import ActionType, Command, DroneModel, Physics, RaceMode
from tests import golden_cases as gc

pytestmark = pytest.mark.gpu


def test_pid_example_loop_ctrl_aviary_plus_dslpid():
    """examples/pid.py:101-147: CtrlAviary stepped with the RPMs DSLPIDControl.computeControlFromState returns
    (BASELINE config 1: one CF2X, DYN, 240/30 Hz, hover target straight above)."""
    meta, z = gc.load("pid_dyn_c1")
    env = CtrlAviary(drone_model=DroneModel.CF2X, num_drones=1, initial_xyzs=np.array(meta["init_xyz"]), physics=Physics.DYN,
                     pyb_freq=240, ctrl_freq=30, precision="fp64")
    ctrl = DSLPIDControl(drone_model=DroneModel.CF2X)
    obs, info = env.reset()
    assert obs.shape == (1, 20) and info == {"answer": 42}
    action = np.zeros((1, 4))
    worst = 0.0
    for k in range(1, meta["steps"]):
        obs, reward, terminated, truncated, info = env.step(action)
        if k == 1:                                          # the golden run starts from its first recorded state
            continue
        action[0, :], _, _ = ctrl.computeControlFromState(control_timestep=env.CTRL_TIMESTEP, state=obs[0],
                                                          target_pos=np.array(meta["target"]))
        assert reward == -1 and terminated is False and truncated is False
    # closed loop reaches the hover target like the reference's run does
    assert abs(obs[0, 2] - z["obs"][-1][2]) < 0.05 and abs(obs[0, 2] - 1.0) < 0.1
    np.testing.assert_allclose(env._getDroneStateVector(0)[0:3], obs[0, 0:3], atol=1e-12)
    env.close()


@pytest.mark.parametrize("name", ["hover_pyb_rpm", "hover_dyn_one_d_rpm", "hover_pyb_one_d_pid"])
def test_hover_aviary_matches_reference_run(name):
    meta, z = gc.load(name)
    env = HoverAviary(physics=Physics[meta["physics"]], pyb_freq=meta["pyb_freq"], ctrl_freq=meta["ctrl_freq"],
                      act=ActionType[meta["act"]], initial_xyzs=np.array(meta["init_xyz"]), precision="fp64")
    obs, _ = env.reset()
    assert obs.shape == env.observation_space.shape and obs.dtype == np.float32      # BaseRLAviary.py:312
    assert gc.rel_err(obs, z["obs0"]) < 1e-6
    for k in range(meta["steps"]):
        obs, rew, te, tr, info = env.step(z["actions"][k])
        assert gc.rel_err(obs[:, :12], z["obs"][k][:, :12]) < 1e-6
        np.testing.assert_array_equal(obs[:, 12:], z["obs"][k][:, 12:].astype(np.float32))
        assert abs(rew - z["reward"][k]) < 1e-9 * max(1, abs(z["reward"][k]))
        assert te == bool(z["terminated"][k]) and tr == bool(z["truncated"][k])
    env.close()


def test_velocity_aviary_matches_reference_run():
    """VelocityAviary.step: (vx, vy, vz, speed fraction) per drone -> DSLPID -> physics, all in one launch."""
    meta, z = gc.load("vel_pyb_m2")
    env = VelocityAviary(num_drones=meta["M"], initial_xyzs=np.array(meta["init_xyz"]), physics=Physics[meta["physics"]],
                         pyb_freq=meta["pyb_freq"], ctrl_freq=meta["ctrl_freq"], precision="fp64")
    assert abs(env.SPEED_LIMIT - meta["speed_limit"]) < 1e-15
    assert env.action_space.shape == (meta["M"], 4) and env.observation_space.shape == (meta["M"], 20)
    obs, _ = env.reset()
    assert gc.rel_err(obs, z["obs0"]) < 1e-9
    for k in range(meta["steps"]):
        obs, rew, te, tr, info = env.step(z["actions"][k])
        assert gc.rel_err(obs, z["obs"][k]) < 1e-9, k
        assert rew == -1 and te is False and tr is False and info == {"answer": 42}
    env.close()


def test_multihover_batched_envs_are_independent_copies():
    """num_envs > 1: torch tensors with a leading env axis; every env fed the golden actions reproduces the golden run."""
    import torch
    meta, z = gc.load("multihover_pyb_rpm_m2")
    N = 5
    env = MultiHoverAviary(num_drones=2, physics=Physics[meta["physics"]], pyb_freq=meta["pyb_freq"], ctrl_freq=meta["ctrl_freq"],
                           act=ActionType[meta["act"]], initial_xyzs=np.array(meta["init_xyz"]), precision="fp64", num_envs=N,
                           auto_reset=False)
    obs, _ = env.reset()
    assert isinstance(obs, torch.Tensor) and tuple(obs.shape) == (N, 2, env.observation_space.shape[1])
    for k in range(meta["steps"]):
        a = np.broadcast_to(z["actions"][k], (N,) + z["actions"][k].shape)
        obs, rew, te, tr, _ = env.step(a)
        assert tuple(rew.shape) == (N,) and te.dtype == torch.bool
        o = obs.cpu().numpy()
        assert gc.rel_err(o[:, :, :12], np.broadcast_to(z["obs"][k][:, :12], o[:, :, :12].shape)) < 1e-6
        assert np.allclose(rew.cpu().numpy(), z["reward"][k], rtol=1e-9)
    env.close()


def test_env_step_n_reproduces_the_golden_run_in_one_launch_and_draws_a_random_policy():
    """BaseAviary.step_n: the golden run's actions handed over at once = the reference's observations of every step; a held
    slice; and ``None`` = the random policy drawn in the kernel (a second env with the same seed draws the same rollout)."""
    import torch
    meta, z = gc.load("multihover_pyb_rpm_m2")
    N, K = 4, meta["steps"]
    mk = lambda: MultiHoverAviary(num_drones=2, physics=Physics[meta["physics"]], pyb_freq=meta["pyb_freq"],   # noqa: E731
                                  ctrl_freq=meta["ctrl_freq"], act=ActionType[meta["act"]], initial_xyzs=np.array(meta["init_xyz"]),
                                  precision="fp64", num_envs=N, auto_reset=False)
    env = mk()
    env.reset()
    acts = np.broadcast_to(z["actions"][:K, None], (K, N) + z["actions"].shape[1:])
    obs, rew, te, tr, _ = env.step_n(acts, keep_outputs=True)
    assert tuple(obs.shape)[:3] == (K, N, 2) and te.dtype == torch.bool
    o = obs.cpu().numpy()
    assert gc.rel_err(o[..., :12], np.broadcast_to(z["obs"][:K, None, :, :12], o[..., :12].shape)) < 1e-6
    assert np.allclose(rew.cpu().numpy(), z["reward"][:K, None], rtol=1e-9)
    assert env.step_counter == K * env.PYB_STEPS_PER_CTRL
    env.reset()
    last = env.step_n(acts[0], n_steps=3)[0].clone()
    assert tuple(last.shape) == (N, 2, o.shape[-1])
    other = mk()
    env.close()
    env = mk()
    env.reset(); other.reset()
    r1 = env.step_n(None, n_steps=6, keep_outputs=True)
    r2 = other.step_n(None, n_steps=6, keep_outputs=True)
    assert torch.equal(r1[0], r2[0]) and torch.equal(r1[1], r2[1])
    assert not torch.equal(r1[0][:, 0], r1[0][:, 1])               # ... and every env draws its own actions
    env.close(); other.close()


def test_ctrl_aviary_kinematic_columns_only():
    """CtrlAviary(obs_columns=12): (NUM_DRONES, 12) rows [pos, rpy, vel, ang_vel] = those columns of the reference's 20."""
    from gym_pybullet_adrp_b200.envs import CtrlAviary
    xyz = np.array([[0.0, 0.0, 0.5], [0.6, 0.0, 0.7]])
    full = CtrlAviary(num_drones=2, physics=Physics.PYB_GND_DRAG_DW, pyb_freq=240, ctrl_freq=48, initial_xyzs=xyz, precision="fp64")
    kin = CtrlAviary(num_drones=2, physics=Physics.PYB_GND_DRAG_DW, pyb_freq=240, ctrl_freq=48, initial_xyzs=xyz, precision="fp64",
                     obs_columns=12)
    assert kin.observation_space.shape == (2, 12) and full.observation_space.shape == (2, 20)
    keep = [0, 1, 2, 7, 8, 9, 10, 11, 12, 13, 14, 15]
    of, _ = full.reset()
    ok, _ = kin.reset()
    assert ok.shape == (2, 12) and np.array_equal(of[:, keep], ok)
    for k in range(10):
        a = np.full((2, 4), 14500.0 + 30 * k)
        of, ok = full.step(a)[0], kin.step(a)[0]
        assert np.array_equal(of[:, keep], ok)
    full.close(); kin.close()


def test_multi_race_aviary_matches_reference_run():
    """scripts/sim.py-style loop on the reference's own race run (level3, COMPARE), draws and disturbances injected."""
    meta, z = gc.load("race_pyb_compare_level3")
    rc = configs.load_config(meta["level"])
    env = MultiRaceAviary(rc, num_drones=meta["M"], physics=Physics[meta["physics"]], racemode=RaceMode[meta["racemode"]],
                          precision="fp64")
    obs, info = env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    assert obs.shape == (2, 49)
    for k in range(meta["steps"]):
        obs, reward, terminated, truncated, info = env.step(z["actions"][k], noise=z["noise"][k][None])
        assert gc.rel_err(obs, z["obs"][k]) < 1e-9, k
        np.testing.assert_array_equal(env.drones_eliminated, z["eliminated"][k])
        np.testing.assert_array_equal(env.current_gate, z["current_gate"][k])
        assert reward == 0 and terminated == bool(z["terminated"][k]) and truncated == bool(z["truncated"][k])
    obs, *_ = env.step([(Command.TAKEOFF, [0.3, 2])] * 2)          # the high-level commander runs on the device
    assert np.isfinite(obs).all()
    env.close()


def test_multi_race_aviary_command_tuples_match_reference_run():
    """HardCodedController-style actions: a list of (Command.FULLSTATE, [pos, vel, acc, yaw, rpy_rate, t]) per drone."""
    meta, z = gc.load("race_fullstate_compare_gs")
    rc = configs.load_config(meta["level"])
    env = MultiRaceAviary(rc, num_drones=meta["M"], physics=Physics.PYB, racemode=RaceMode.COMPARE, precision="fp64")
    env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    a = z["actions"]
    for k in range(60):
        cmds = [(Command.FULLSTATE, [a[k, i, 0:3], a[k, i, 3:6], a[k, i, 6:9], a[k, i, 9], a[k, i, 10:13], k / 25.0])
                for i in range(meta["M"])]
        obs, reward, terminated, truncated, info = env.step(cmds)
        assert gc.rel_err(obs, z["obs"][k]) < 1e-9, k
    obs_none, *_ = env.step([(Command.NONE, [])] * meta["M"])           # no new command: the last setpoint stays
    env.close()
    env2 = MultiRaceAviary(rc, num_drones=meta["M"], physics=Physics.PYB, racemode=RaceMode.COMPARE, precision="fp64",
                           full_state_actions=True)
    env2.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    for k in range(60):
        env2.step(a[k])                                                  # (M, 13) FULLSTATE rows, same thing
    obs_rep, *_ = env2.step(a[59])
    np.testing.assert_array_equal(obs_none, obs_rep)
    env2.close()


def _tuple_from_row(row):
    """16-float command row -> (Command, args) as user_controller/HardCodedController.py:158-188 returns them."""
    by_id = {v: k for k, v in _abi.CMD_ID.items()}
    cmd = Command(by_id[int(row[0])])
    g = row[1:]
    if cmd == Command.FULLSTATE:
        return cmd, [g[0:3].copy(), g[3:6].copy(), g[6:9].copy(), float(g[9]), g[10:13].copy(), float(g[13])]
    n = {Command.NONE: 0, Command.TAKEOFF: 2, Command.LAND: 2, Command.NOTIFY: 1, Command.STOP: 1, Command.TAKEOFFYAW: 3,
         Command.LANDYAW: 3, Command.TAKEOFFVEL: 3, Command.LANDVEL: 3}[cmd]
    return cmd, [float(v) for v in g[:n]]


def test_multi_race_aviary_runs_hardcoded_controller_commands():
    """The command stream of the UNMODIFIED user_controller/HardCodedController.py (TAKEOFF [0.3, 2] -> FULLSTATE along its
    spline -> NOTIFY -> LAND, recorded while it drove the reference's own env: tests/golden/make_golden_race.py --commander)
    as (Command, args) tuples into the env mirror, and through the batched driver of scripts/sim.py: same observations as
    the reference's run, all four gates passed by both drones, same episode time."""
    from gym_pybullet_adrp_b200.scripts.sim import simulate
    from gym_pybullet_adrp_b200.user_controller import BaseController
    meta, z = gc.load("race_commander_hardcoded_gs")
    rc = configs.load_config(meta["level"])
    M, steps = meta["M"], meta["steps"]
    env = MultiRaceAviary(rc, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE, precision="fp64")
    env.reset(draws=gc.place_draws(rc, z["draws"], M)[None])
    for k in range(steps):
        obs, reward, terminated, truncated, info = env.step([_tuple_from_row(z["actions"][k, i]) for i in range(M)])
        assert gc.rel_err(obs, z["obs"][k]) < 1e-9, k
        np.testing.assert_array_equal(env.current_gate, z["current_gate"][k])
    assert terminated and env.current_gate.min() == 4 and env.drones_finished.all()
    env.close()

    class Replay(BaseController):                              # the recorded stream, one instance per run and drone
        def predict(self, obs, reward=None, done=None, info=None, ep_time=None):
            k = min(int(round(ep_time * rc["ctrl_freq"])), steps - 1)
            return _tuple_from_row(z["actions"][k, self.drone_id])

    rc2 = configs.load_config(meta["level"])
    for key in ("random_gates_obstacles", "random_drone_state", "random_drone_inertia", "disturbances"):
        assert not rc2.get(key), "getting_started has no randomisation: the runs of the batch equal the golden run"
    stats = simulate(rc2, Replay, n_runs=3, n_drones=M, precision="fp64", return_stats=True)
    assert stats["episode_times"] == [(steps - 1) / rc["ctrl_freq"]] * 3      # sim.py:79,104: time of the last step taken


def test_wrapper_classes_match_reference_run():
    """DroneObservationWrapper(RewardWrapper(MultiRaceAviary)) -- the reference's class names, fused in the kernel."""
    from gym_pybullet_adrp_b200.utils.wrapper import DroneObservationWrapper, RewardWrapper
    meta, z = gc.load("race_wrappers_compare_gs")
    rc = configs.load_config(meta["level"])
    env = DroneObservationWrapper(RewardWrapper(MultiRaceAviary(rc, num_drones=meta["M"], physics=Physics.PYB,
                                                                racemode=RaceMode.COMPARE, precision="fp64")))
    obs, info = env.reset(draws=gc.place_draws(rc, z["draws"], meta["M"])[None])
    ret = 0.0
    for k in range(meta["steps"]):
        obs, reward, terminated, truncated, info = env.step(z["actions"][k])
        assert gc.rel_err(obs, z["obs"][k]) < 1e-9, k
        assert abs(reward - z["reward"][k]) < 1e-9 and terminated == bool(z["terminated"][k]), k
        ret += reward
    assert terminated and env.current_gate[0] == 2 and abs(ret - z["reward"].sum()) < 1e-8
    with pytest.raises(RuntimeError, match="before its first reset"):
        RewardWrapper(env.unwrapped)
    env.close()


def test_torch_vec_env_rollout():
    """SB3-style rollout loop on device tensors: shapes, done bookkeeping, and SB3's episode-end contract -- obs[dones] is
    the FIRST observation of the next episode, infos["terminal_observation"][dones] the last one of the finished episode."""
    import torch
    from gym_pybullet_adrp_b200.vec_env import TorchVecEnv
    N = 64
    with pytest.raises(ValueError, match="auto_reset=False"):
        TorchVecEnv(HoverAviary(physics=Physics.PYB, act=ActionType.RPM, num_envs=N, auto_reset=True))
    env = HoverAviary(physics=Physics.PYB, act=ActionType.RPM, num_envs=N, auto_reset=False)
    env.EPISODE_LEN_SEC = 1.0
    venv = TorchVecEnv(env)
    obs0 = venv.reset()
    assert obs0.shape == (N, int(np.prod(env.observation_space.shape))) and obs0.is_cuda
    g = torch.Generator(device="cuda").manual_seed(0)
    finished = torch.zeros(N, dtype=torch.int64, device="cuda")
    kept = []
    for k in range(80):
        actions = 0.2 * (2 * torch.rand((N,) + env.action_space.shape, device="cuda", generator=g) - 1)
        obs, rewards, dones, infos = venv.step(actions)
        kept.append(obs)
        assert rewards.shape == (N,) and dones.shape == (N,) and dones.dtype == torch.bool
        finished += dones
        assert torch.all(infos["episode_length"][dones] > 0)
        if bool(dones.any()):
            # the reset observation: initial pose at rest (columns 0:12 of the very first observation)
            assert torch.equal(obs[dones][:, :12], obs0[dones][:, :12])
            term = infos["terminal_observation"][dones]
            assert (term[:, :12] - obs0[dones][:, :12]).abs().max() > 1e-3     # the finished episode's last row is elsewhere
            # the action buffer survives the reset, like the reference's (BaseRLAviary.py:66-67)
            assert torch.equal(obs[dones][:, 12:], term[:, 12:])
        assert torch.equal(obs[~dones], infos["terminal_observation"][~dones])
    assert int(finished.min()) >= 1 and torch.isfinite(obs).all()      # 80 steps at 30 Hz > 1 s episodes: everyone reset
    assert int(venv.episode_lengths.max()) <= 31
    assert not torch.equal(kept[0], kept[1]) and kept[0].data_ptr() != kept[1].data_ptr()   # returned tensors are copies
    venv.close()


def test_torch_vec_env_race_shows_the_new_track_after_an_episode():
    """Race env through TorchVecEnv: the observation returned for a finished env is the reset observation of its NEXT
    episode -- nominal start poses seen against the re-randomised track."""
    import torch
    from gym_pybullet_adrp_b200.vec_env import TorchVecEnv
    N, M = 48, 2
    rc = configs.load_config("level3")
    rc["episode_len_sec"] = 0.3
    env = MultiRaceAviary(rc, num_drones=M, physics=Physics.PYB, racemode=RaceMode.COMPARE, num_envs=N, auto_reset=False, seed=3)
    venv = TorchVecEnv(env, flatten=False)
    obs0 = venv.reset()
    init = torch.as_tensor(np.array([rc["init_states"][f"drone{i}"]["pos"] for i in range(M)]), dtype=obs0.dtype, device=obs0.device)
    a = torch.zeros((N, M, 4), device="cuda")
    a[..., 0:3] = init + torch.tensor([0.0, 0.0, 0.4], device="cuda")
    seen = 0
    for k in range(12):
        obs, rewards, dones, infos = venv.step(a)
        if bool(dones.any()):
            seen += int(dones.sum())
            assert torch.allclose(obs[dones][:, :, 0:3], init.expand(int(dones.sum()), M, 3), atol=1e-6)
            assert (infos["terminal_observation"][dones][:, :, 2] > 0.08).all()          # they had climbed
            assert (obs[dones][:, :, 48] == 0).all()                                      # current gate of the new episode
    assert seen >= N
    venv.close()


def test_batched_eval_driver(tmp_path):
    """scripts/sim.py:18-107 over parallel runs: per-agent host controllers and the on-device predict_batch give the
    same episode times; the flight of one run lands in the reference Logger's .npy layout."""
    import glob
    from gym_pybullet_adrp_b200.scripts.sim import simulate
    from gym_pybullet_adrp_b200.user_controller import HoverController
    rc = configs.load_config("getting_started")
    rc["episode_len_sec"] = 1.2

    class HostOnlyHover(HoverController):                     # no predict_batch: forces the per-agent path
        predict_batch = None
    del HostOnlyHover.predict_batch

    class PerAgent(HoverController.__mro__[1]):
        def predict(self, obs, reward=None, done=None, info=None, ep_time=None):
            return np.array([1.5, 0.5, 1 + self.drone_id, 0])

    t_host = simulate(rc, PerAgent, n_runs=3, n_drones=2, precision="fp64", log_env=1, output_folder=str(tmp_path))
    t_dev = simulate(rc, HoverController, n_runs=3, n_drones=2, precision="fp64")
    assert t_host == t_dev and len(t_host) == 3
    # hovering towards the setpoint: the runs end by truncation, in the first env step k with k * 20 / 500 > 1.2 s
    # (k = 31); like the reference the driver counts time in the CONFIG's ctrl_freq (sim.py:79), 30 Hz here
    assert all(abs(t - 31 / rc["ctrl_freq"]) < 1e-12 for t in t_host)
    files = glob.glob(str(tmp_path / "save-flight-*.npy"))
    assert len(files) == 1
    z = np.load(files[0])
    assert z["states"].shape[0] == 2 and z["states"].shape[1] == 16 and z["states"].shape[2] >= 29
    assert z["states"][0, 2, -1] > z["states"][0, 2, 0]        # it climbed
