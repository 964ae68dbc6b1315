"""CPU: API surface of the env / controller classes (names, constructor arguments, attributes, spaces and error
behaviour of the reference) -- nothing here touches a GPU: the C-ABI handle is created lazily."""
import numpy as np
import pytest

from gym_pybullet_adrp_b200 import configs
from gym_pybullet_adrp_b200.envs import CtrlAviary, HoverAviary, MultiHoverAviary, MultiRaceAviary
from gym_pybullet_adrp_b200.utils.enums import ActionType, DroneModel, ObservationType, Physics, RaceMode


def test_enum_values_are_the_references():
    assert DroneModel("cf2x_IROS") is DroneModel.CF2X and Physics("pyb_gnd_drag_dw") is Physics.PYB_GND_DRAG_DW
    assert ActionType("one_d_pid") is ActionType.ONE_D_PID and RaceMode(1) is RaceMode.COMPETE
    with pytest.raises(ValueError):
        DroneModel("cf2x")                       # the fork renamed the value


def test_base_constants_match_survey_appendix_f():
    e = CtrlAviary()
    assert e.PYB_STEPS_PER_CTRL == 1 and e.NUM_DRONES == 1
    assert abs(e.HOVER_RPM - 16364.4219) < 1e-3 and abs(e.MAX_RPM - 24546.6328) < 1e-3
    assert abs(e.GND_EFF_H_CLIP - 0.0377637) < 1e-6 and abs(e.MAX_THRUST - 0.761607) < 1e-6
    np.testing.assert_allclose(e.INIT_XYZS, [[0, 0, 0.1125]])
    assert e.observation_space.shape == (1, 20) and e.action_space.shape == (1, 4)
    assert e.action_space.high.max() == pytest.approx(e.MAX_RPM, rel=1e-6)


def test_constructor_errors_like_reference():
    with pytest.raises(ValueError, match="not divisible"):
        CtrlAviary(pyb_freq=240, ctrl_freq=7)
    with pytest.raises(ValueError, match="initial_xyzs"):
        CtrlAviary(num_drones=2, initial_xyzs=np.zeros((3, 3)))
    with pytest.raises(NotImplementedError, match="unavailable on this box"):
        HoverAviary(gui=True)
    with pytest.raises(NotImplementedError, match="unavailable on this box"):
        HoverAviary(obs=ObservationType.RGB)


def test_velocity_aviary_surface():
    """VelocityAviary.py:87-104: speed limit and spaces, no GPU needed."""
    from gym_pybullet_adrp_b200.envs import VelocityAviary
    env = VelocityAviary(num_drones=3)
    assert abs(env.SPEED_LIMIT - 0.03 * 30 * (1000 / 3600)) < 1e-15          # cf2x max_speed_kmh = 30
    assert env.action_space.shape == (3, 4) and env.action_space.low[0].tolist() == [-1, -1, -1, 0]
    assert env.observation_space.shape == (3, 20)
    with pytest.raises(ValueError, match="DSLPIDControl"):
        VelocityAviary(drone_model=DroneModel.RACE)


def test_rl_spaces():
    e = HoverAviary()
    assert (e.CTRL_FREQ, e.PYB_FREQ, e.ACTION_BUFFER_SIZE, e.EPISODE_LEN_SEC) == (30, 240, 15, 8)
    assert e.observation_space.shape == (1, 72) and e.action_space.shape == (1, 4)
    e = HoverAviary(act=ActionType.ONE_D_RPM)
    assert e.observation_space.shape == (1, 27) and e.action_space.shape == (1, 1)
    e = MultiHoverAviary(num_drones=3, act=ActionType.PID)
    assert e.observation_space.shape == (3, 12 + 15 * 3)
    np.testing.assert_allclose(e.TARGET_POS[:, 2], e.INIT_XYZS[:, 2] + [1, 0.5, 1 / 3])


def test_race_env_surface():
    rc = configs.load_config("level3")
    e = MultiRaceAviary(rc, racemode=RaceMode.COMPETE)
    assert (e.PYB_FREQ, e.CTRL_FREQ, e.PYB_STEPS_PER_CTRL, e.NUM_DRONES) == (500, 25, 20, 2)
    assert e.observation_space.shape == (2, 55) and e.action_space.shape == (2, 4)
    assert MultiRaceAviary(rc, racemode=RaceMode.COMPARE).observation_space.shape == (2, 49)
    np.testing.assert_allclose(e.INIT_XYZS, [[0.9, 0.9, 0.05], [1.1, 1.1, 0.05]])
    cfg = e._make_config()
    assert cfg.track.num_gates == 4 and cfg.rnd.disturbances == 1 and cfg.drone.pyb_mass == 0.027   # cf2x.urdf body (E16)
    with pytest.raises(ValueError):
        MultiRaceAviary(rc, num_drones=3)._make_config()       # level3 has two init_states


def test_config_files_equal_builtin_tables():
    """The built-in level tables are the reference's YAML files (skipped where /root/reference is absent)."""
    import os
    ref = "/root/reference/config"
    if not os.path.isdir(ref):
        pytest.skip("reference not mounted")
    for name in ("getting_started", "level0", "level1", "level2", "level3", "twogates"):
        a, b = configs.load_config(os.path.join(ref, name + ".yaml")), configs.builtin_config(name)
        for key in ("gates", "obstacles", "bounds", "init_states", "episode_len_sec", "random_drone_state",
                    "random_drone_inertia", "random_gates_obstacles", "disturbances", "ctrl_freq"):
            assert a.get(key) == b.get(key), (name, key)
        for key in ("random_drone_state_info", "random_drone_inertia_info", "random_gates_obstacles_info", "disturbances_info"):
            if b.get(key.replace("_info", "")):
                assert a[key] == b[key], (name, key)


def test_wrapper_classes_surface():
    """utils/wrapper.py mirror: wrapping sets the kernel's epilogue bits; only race envs; attribute forwarding."""
    from gym_pybullet_adrp_b200 import _abi, configs
    from gym_pybullet_adrp_b200.envs import MultiRaceAviary
    from gym_pybullet_adrp_b200.utils.wrapper import DroneObservationWrapper, RewardWrapper
    base = MultiRaceAviary(configs.load_config("twogates"), num_envs=8)
    env = DroneObservationWrapper(RewardWrapper(base))
    assert env.unwrapped is base and env.NUM_DRONES == 2
    assert base._make_config().wrappers == (_abi.WRAP_OBS | _abi.WRAP_REWARD)
    with pytest.raises(TypeError, match="MultiRaceAviary"):
        RewardWrapper(HoverAviary())


def test_eval_driver_host_pieces(tmp_path):
    """scripts/sim.py helpers that need no GPU: Logger layout (utils/logger.py:100-113), load_controller, action rows."""
    import os
    from gym_pybullet_adrp_b200.scripts import sim
    from gym_pybullet_adrp_b200.utils.enums import Command
    from gym_pybullet_adrp_b200.utils.logger import Logger
    from gym_pybullet_adrp_b200.utils.utils import load_controller, map2pi
    log = Logger(logging_freq_hz=25, output_folder=str(tmp_path), num_drones=2)
    for k in range(3):
        for i in range(2):
            log.log(i, k / 25, np.arange(20.0) + 100 * i)
    z = np.load(log.save())
    assert z["timestamps"].shape == (2, 3) and z["states"].shape == (2, 16, 3) and z["controls"].shape == (2, 12, 3)
    # pos, vel, rpy, ang_vel, rpm out of the 20-float state row [pos3 quat4 rpy3 vel3 ang_vel3 rpm4]
    np.testing.assert_array_equal(z["states"][1, :, 2], 100 + np.r_[0:3, 10:13, 7:10, 13:20])
    here = os.path.dirname(os.path.abspath(sim.__file__))
    cls = load_controller(os.path.join(here, "..", "user_controller", "HoverController.py"))
    assert cls(1).predict(None).tolist() == [1.5, 0.5, 2, 0]
    # 16-float command rows (ADRP_ACT_MEL_CMD): [adrp_command, args..., 0 padding]
    row = sim._row((Command.FULLSTATE, [[1, 2, 3], [4, 5, 6], [7, 8, 9], 0.25, [0.1, 0.2, 0.3], 0.5]))
    assert row.tolist() == [1, 1, 2, 3, 4, 5, 6, 7, 8, 9, 0.25, 0.1, 0.2, 0.3, 0.5, 0]
    assert sim._row(np.array([1.0, 2.0, 3.0, 0.5]), 40).tolist() == [1, 1, 2, 3] + [0] * 6 + [0.5, 0, 0, 0, 40, 0]
    assert sim._row((Command.TAKEOFF, [0.3, 2])).tolist() == [2, 0.3, 2] + [0] * 13
    assert sim._row((Command.GOTO, [[1, 2, 3], 0.1, 2.5, True])).tolist() == [9, 1, 2, 3, 0.1, 2.5, 1] + [0] * 9
    assert sim._row((Command.NONE, [])).tolist() == [0] * 16
    assert abs(map2pi(3 * np.pi / 2) + np.pi / 2) < 1e-12
