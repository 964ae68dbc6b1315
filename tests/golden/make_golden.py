"""Generate the golden vectors under tests/golden/ by running the REFERENCE'S OWN Python code.

Run in the build container only (needs /root/reference):

    python tests/golden/make_golden.py

What runs: the unmodified reference modules (``gym_pybullet_adrp.envs.*``, ``control.DSLPIDControl``)
imported from /root/reference, with the third-party modules that are not installable here
(pybullet, pybullet_data, gymnasium, munch, transforms3d) replaced by the stand-ins in
``oracle/refstub``.  So every reference-owned line on the path (step loop order, ``_dynamics``,
``_integrateQ``, ``_physics``, ``_groundEffect``, ``_drag``, ``_downwash``, action maps, DSLPID,
rewards, flags) is executed as written, and only the PyBullet primitives are the declared
restatement of ``oracle/bullet_model.py``.  The vectors pin the oracle (tests/test_oracle_golden.py)
and, through it, the CUDA path.  Nothing here is read at GPU-test time except the .npz outputs.
"""
from __future__ import annotations

import json
import os
import sys
import warnings

warnings.filterwarnings("ignore")
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.environ.get("ADRP_GOLDEN_OUT", HERE)      # tests/test_golden_regen.py regenerates into a scratch directory
sys.path.insert(0, os.path.join(ROOT, "oracle", "refstub"))
sys.path.insert(0, os.environ.get("ADRP_REFERENCE_ROOT", "/root/reference"))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

from gym_pybullet_adrp.envs.CtrlAviary import CtrlAviary  # noqa: E402
from gym_pybullet_adrp.envs.HoverAviary import HoverAviary  # noqa: E402
from gym_pybullet_adrp.envs.MultiHoverAviary import MultiHoverAviary  # noqa: E402
from gym_pybullet_adrp.envs.VelocityAviary import VelocityAviary  # noqa: E402
from gym_pybullet_adrp.control.DSLPIDControl import DSLPIDControl  # noqa: E402
from gym_pybullet_adrp.utils.enums import ActionType, DroneModel, ObservationType, Physics  # noqa: E402


def _save(name, meta, **arrays):
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, meta=json.dumps(meta), **arrays)
    print(f"{name}: " + ", ".join(f"{k}{tuple(v.shape)}" for k, v in arrays.items()))


def ctrl_case(name, model, physics, M, pyb, ctrl, steps, seed, spread, xyz=None, rpy=None):
    """CtrlAviary open-loop RPM replay (BaseAviary.step, :262-387)."""
    rng = np.random.default_rng(seed)
    env = CtrlAviary(drone_model=model, num_drones=M, physics=physics, pyb_freq=pyb, ctrl_freq=ctrl,
                     initial_xyzs=xyz, initial_rpys=rpy, gui=False, user_debug_gui=False)
    obs0, _ = env.reset()
    actions = env.HOVER_RPM * (1.0 + spread * rng.uniform(-1, 1, size=(steps, M, 4)))
    actions[steps // 2] *= 1.6            # one step beyond MAX_RPM exercises the clip (CtrlAviary.py:140)
    obs = np.zeros((steps, M, 20))
    for k in range(steps):
        o, r, te, tr, _ = env.step(actions[k])
        obs[k] = o
    meta = dict(env="ctrl", model=model.name, physics=physics.name, M=M, pyb_freq=pyb, ctrl_freq=ctrl,
                steps=steps, init_xyz=np.asarray(env.INIT_XYZS).tolist(), init_rpy=np.asarray(env.INIT_RPYS).tolist(),
                hover_rpm=env.HOVER_RPM, max_rpm=env.MAX_RPM, gnd_eff_h_clip=env.GND_EFF_H_CLIP)
    env.close()
    _save(name, meta, obs0=np.asarray(obs0, dtype=np.float64), actions=actions, obs=obs)


def rl_case(name, cls, M, physics, act, pyb, ctrl, steps, seed, scale=1.0):
    """Hover / MultiHover aviaries with normalised actions (BaseRLAviary.py:160-239, HoverAviary.py:68-117)."""
    rng = np.random.default_rng(seed)
    kw = dict(physics=physics, pyb_freq=pyb, ctrl_freq=ctrl, gui=False, obs=ObservationType.KIN, act=act)
    env = cls(**kw) if cls is HoverAviary else cls(num_drones=M, **kw)
    obs0, _ = env.reset()
    A = env.action_space.shape[1]
    actions = (scale * rng.uniform(-1, 1, size=(steps, M, A))).astype(np.float32)
    if act == ActionType.VEL:
        actions[:, :, 2] = rng.uniform(-1, 1, size=(steps, M)).astype(np.float32)    # mostly vertical
        actions[:, :, 3] = rng.uniform(0, 1, size=(steps, M)).astype(np.float32)
    if act in (ActionType.PID,):
        # vertical targets only: the one stable DSLPID closed loop in this fork
        actions[:, :, 0:2] = np.asarray(env.INIT_XYZS)[None, :, 0:2]
        actions[:, :, 2] = 0.5 + 0.4 * np.abs(actions[:, :, 2])
    obs = np.zeros((steps,) + tuple(np.asarray(obs0).shape))
    rew, term, trunc = np.zeros(steps), np.zeros(steps, bool), np.zeros(steps, bool)
    for k in range(steps):
        o, r, te, tr, _ = env.step(actions[k])
        obs[k], rew[k], term[k], trunc[k] = o, r, te, tr
    meta = dict(env="hover" if cls is HoverAviary else "multihover", model="CF2X", physics=physics.name, M=M,
                act=act.name, pyb_freq=pyb, ctrl_freq=ctrl, steps=steps,
                init_xyz=np.asarray(env.INIT_XYZS).tolist(), init_rpy=np.asarray(env.INIT_RPYS).tolist(),
                target=np.asarray(env.TARGET_POS).tolist())
    env.close()
    _save(name, meta, obs0=np.asarray(obs0, dtype=np.float64), actions=actions, obs=obs, reward=rew,
          terminated=term, truncated=trunc)


def rl_reset_case(name, physics, act, pyb, ctrl, steps, reset_at, seed):
    """HoverAviary with a reset() in the middle of the run: BaseRLAviary fills action_buffer once in __init__
    (BaseRLAviary.py:66-67) and neither BaseAviary.reset nor _housekeeping touches it, so the observation reset()
    returns and the following ctrl_freq//2 - 1 steps still carry the previous episode's actions in obs[12:]."""
    rng = np.random.default_rng(seed)
    env = HoverAviary(physics=physics, pyb_freq=pyb, ctrl_freq=ctrl, gui=False, obs=ObservationType.KIN, act=act)
    obs0, _ = env.reset()
    A = env.action_space.shape[1]
    actions = (0.5 * rng.uniform(-1, 1, size=(steps, 1, A))).astype(np.float32)
    obs = np.zeros((steps,) + tuple(np.asarray(obs0).shape))
    rew, term, trunc = np.zeros(steps), np.zeros(steps, bool), np.zeros(steps, bool)
    obs_reset = None
    for k in range(steps):
        if k == reset_at:
            obs_reset, _ = env.reset()
        o, r, te, tr, _ = env.step(actions[k])
        obs[k], rew[k], term[k], trunc[k] = o, r, te, tr
    meta = dict(env="hover", model="CF2X", physics=physics.name, M=1, act=act.name, pyb_freq=pyb, ctrl_freq=ctrl,
                steps=steps, reset_at=reset_at, init_xyz=np.asarray(env.INIT_XYZS).tolist(),
                init_rpy=np.asarray(env.INIT_RPYS).tolist(), target=np.asarray(env.TARGET_POS).tolist())
    env.close()
    _save(name, meta, obs0=np.asarray(obs0, dtype=np.float64), obs_reset=np.asarray(obs_reset, dtype=np.float64),
          actions=actions, obs=obs, reward=rew, terminated=term, truncated=trunc)


def vel_case(name, physics, M, pyb, ctrl, steps, seed, xyz=None):
    """VelocityAviary (VelocityAviary.py:129-170): velocity commands through DSLPID inside the env step.
    Mostly vertical commands (the lateral DSLPID loop of this fork diverges), one zero-norm row."""
    rng = np.random.default_rng(seed)
    env = VelocityAviary(drone_model=DroneModel.CF2X, num_drones=M, physics=physics, pyb_freq=pyb, ctrl_freq=ctrl,
                         initial_xyzs=xyz, gui=False, user_debug_gui=False)
    obs0, _ = env.reset()
    actions = np.zeros((steps, M, 4))
    actions[:, :, 2] = rng.uniform(-1, 1, size=(steps, M))
    actions[:, :, 0:2] = 0.02 * rng.uniform(-1, 1, size=(steps, M, 2))
    actions[:, :, 3] = rng.uniform(0, 1, size=(steps, M))
    actions[steps // 3, 0, :3] = 0.0                                    # |v| == 0 branch (:148-151)
    obs = np.zeros((steps, M, 20))
    for k in range(steps):
        o, r, te, tr, _ = env.step(actions[k])
        obs[k] = o
    meta = dict(env="velocity", model="CF2X", physics=physics.name, M=M, pyb_freq=pyb, ctrl_freq=ctrl, steps=steps,
                init_xyz=np.asarray(env.INIT_XYZS).tolist(), init_rpy=np.asarray(env.INIT_RPYS).tolist(),
                speed_limit=env.SPEED_LIMIT)
    env.close()
    _save(name, meta, obs0=np.asarray(obs0, dtype=np.float64), actions=actions, obs=obs)


def pid_case(name, physics, steps):
    """BASELINE config 1: CtrlAviary + DSLPIDControl driver loop (examples/pid.py:101-147), vertical target."""
    env = CtrlAviary(drone_model=DroneModel.CF2X, num_drones=1, physics=physics, pyb_freq=240, ctrl_freq=30,
                     gui=False, user_debug_gui=False)
    ctrl = DSLPIDControl(drone_model=DroneModel.CF2X)
    obs, _ = env.reset()
    target = np.array([0.0, 0.0, 1.0])
    action = np.zeros((1, 4))
    o_log, a_log, e_log = np.zeros((steps, 20)), np.zeros((steps, 4)), np.zeros((steps, 4))
    for k in range(steps):
        obs, *_ = env.step(action)
        rpm, pos_e, yaw_e = ctrl.computeControlFromState(control_timestep=env.CTRL_TIMESTEP, state=obs[0],
                                                         target_pos=target, target_rpy=np.zeros(3))
        action[0] = rpm
        o_log[k], a_log[k], e_log[k, :3], e_log[k, 3] = obs[0], rpm, pos_e, yaw_e
    meta = dict(env="ctrl+dslpid", model="CF2X", physics=physics.name, M=1, pyb_freq=240, ctrl_freq=30,
                steps=steps, target=target.tolist(), init_xyz=np.asarray(env.INIT_XYZS).tolist())
    env.close()
    _save(name, meta, obs=o_log, rpm=a_log, err=e_log)


def dslpid_unit_case(name, seed, n):
    """DSLPIDControl.computeControl on random (not necessarily stable) states: a pure-function KAT
    of DSLPIDControl.py:149-259, two consecutive calls each so that the integral / last_rpy state is covered."""
    rng = np.random.default_rng(seed)
    import pybullet as p
    rows = []
    for model in (DroneModel.CF2X, DroneModel.CF2P):
        for _ in range(n):
            ctrl = DSLPIDControl(drone_model=model)
            rec = []
            for call in range(2):
                pos, vel = rng.uniform(-1, 1, 3), rng.uniform(-1, 1, 3)
                quat = np.array(p.getQuaternionFromEuler(rng.uniform(-0.6, 0.6, 3)))
                tpos, trpy = rng.uniform(-1, 1, 3), np.array([0, 0, rng.uniform(-1, 1)])
                tvel, trates = rng.uniform(-0.3, 0.3, 3), rng.uniform(-0.3, 0.3, 3)
                dt = 1 / 30 if call == 0 else 1 / 48
                rpm, pos_e, yaw_e = ctrl.computeControl(dt, pos, quat, vel, np.zeros(3), tpos, trpy, tvel, trates)
                rec += [np.r_[model == DroneModel.CF2P, dt, pos, quat, vel, tpos, trpy, tvel, trates],
                        np.r_[rpm, pos_e, yaw_e, ctrl.integral_pos_e, ctrl.integral_rpy_e, ctrl.last_rpy]]
            rows.append(np.concatenate(rec))
    _save(name, dict(layout="[cf2p,dt,pos3,quat4,vel3,tpos3,trpy3,tvel3,trates3 | rpm4,pos_e3,yaw_e,ipos3,irpy3,last_rpy3] x2"),
          rows=np.asarray(rows))


def main():
    X, Pm, Rc = DroneModel.CF2X, DroneModel.CF2P, DroneModel.RACE
    if len(sys.argv) > 1 and sys.argv[1] == "--new":
        # cases added after the first freeze: generate only these, the earlier .npz files stay untouched
        vel_case("vel_pyb_m2", Physics.PYB, 2, 240, 48, 60, 51, xyz=np.array([[0, 0, 0.5], [0.6, 0, 0.8]]))
        vel_case("vel_dyn_m1", Physics.DYN, 1, 240, 240, 120, 52, xyz=np.array([[0, 0, 0.4]]))
        rl_case("hover_pyb_vel", HoverAviary, 1, Physics.PYB, ActionType.VEL, 240, 30, 30, 53, scale=0.05)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "--r2":
        # round 2: BASELINE config 4's formation as bench.py flies it (examples/downwash.py:45 generalised to 8 drones on a
        # 1.2 m grid, heights 0.5 + 0.25 (k mod 4)): every downwash pair is evaluated, none sits on the formula's 1/dz^2
        # singularity, so the fp32 build can be compared with the reference's own run over the whole second
        ctrl_case("ctrl_pyb_all_s5_m8_wide", X, Physics.PYB_GND_DRAG_DW, 8, 240, 48, 48, 61, 0.005,
                  xyz=np.array([[1.2 * (k % 3), 1.2 * (k // 3), 0.5 + 0.25 * (k % 4)] for k in range(8)]))
        rl_reset_case("hover_pyb_rpm_reset", Physics.PYB, ActionType.RPM, 240, 30, 36, 12, 62)
        return
    # per-substep pins (S = 1) and 1 s flights (S = 8 / 5)
    ctrl_case("ctrl_dyn_cf2x_s1", X, Physics.DYN, 1, 240, 240, 240, 11, 0.03)
    ctrl_case("ctrl_dyn_cf2x_s8_m3", X, Physics.DYN, 3, 240, 30, 30, 12, 0.03)
    ctrl_case("ctrl_dyn_cf2p_s8", Pm, Physics.DYN, 2, 240, 30, 30, 13, 0.03)
    ctrl_case("ctrl_dyn_race_s8", Rc, Physics.DYN, 2, 240, 30, 30, 14, 0.03)
    tilt = np.array([[0.3, -0.2, 0.5], [-0.4, 0.1, -2.0]])
    ctrl_case("ctrl_dyn_cf2x_tilted", X, Physics.DYN, 2, 240, 48, 48, 15, 0.05,
              xyz=np.array([[0, 0, 1.0], [0.5, 0, 1.2]]), rpy=tilt)
    ctrl_case("ctrl_pyb_cf2x_s1", X, Physics.PYB, 1, 240, 240, 240, 21, 0.03)
    ctrl_case("ctrl_pyb_cf2x_s8_m2", X, Physics.PYB, 2, 240, 30, 30, 22, 0.03,
              xyz=np.array([[0, 0, 0.5], [0.3, 0, 0.8]]), rpy=tilt)
    ctrl_case("ctrl_pyb_cf2p_s8", Pm, Physics.PYB, 1, 240, 30, 30, 23, 0.03)
    ctrl_case("ctrl_pyb_race_s8", Rc, Physics.PYB, 1, 240, 30, 30, 24, 0.03)
    ctrl_case("ctrl_pyb_gnd_s8", X, Physics.PYB_GND, 2, 240, 30, 30, 25, 0.03,
              xyz=np.array([[0, 0, 0.06], [0.4, 0, 0.12]]))
    ctrl_case("ctrl_pyb_drag_s8", X, Physics.PYB_DRAG, 2, 240, 30, 30, 26, 0.08,
              xyz=np.array([[0, 0, 0.6], [0.4, 0, 0.9]]), rpy=tilt)
    ctrl_case("ctrl_pyb_dw_s5_m4", X, Physics.PYB_DW, 4, 240, 48, 48, 27, 0.02,
              xyz=np.array([[0, 0, 0.5], [0.05, 0, 1.5], [0.5, 0.3, 0.75], [0.5, 0.35, 1.9]]))
    ctrl_case("ctrl_pyb_all_s5_m8", X, Physics.PYB_GND_DRAG_DW, 8, 240, 48, 48, 28, 0.02,
              xyz=np.array([[0.5 * (-1) ** k, 0.3 * (k // 2), 0.5 + 0.25 * (k % 4)] for k in range(8)]))
    ctrl_case("ctrl_pyb_ground_contact", X, Physics.PYB, 2, 240, 30, 45, 29, 0.25)
    # RL aviaries
    rl_case("hover_pyb_rpm", HoverAviary, 1, Physics.PYB, ActionType.RPM, 240, 30, 40, 31)
    rl_case("hover_dyn_one_d_rpm", HoverAviary, 1, Physics.DYN, ActionType.ONE_D_RPM, 240, 30, 40, 32)
    rl_case("hover_dyn_pid", HoverAviary, 1, Physics.DYN, ActionType.PID, 240, 30, 40, 33)
    rl_case("hover_pyb_one_d_pid", HoverAviary, 1, Physics.PYB, ActionType.ONE_D_PID, 240, 30, 40, 34)
    rl_case("multihover_pyb_rpm_m2", MultiHoverAviary, 2, Physics.PYB, ActionType.RPM, 240, 30, 40, 35)
    rl_case("hover_pyb_rpm_long", HoverAviary, 1, Physics.PYB, ActionType.RPM, 240, 30, 250, 36, scale=0.3)
    # closed loop (BASELINE config 1) and DSLPID known-answer rows
    pid_case("pid_dyn_c1", Physics.DYN, 240)
    pid_case("pid_pyb_c1", Physics.PYB, 90)
    dslpid_unit_case("dslpid_kat", 41, 24)


if __name__ == "__main__":
    main()
