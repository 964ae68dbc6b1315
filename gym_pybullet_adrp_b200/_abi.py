"""ctypes mirror of include/adrp.h (struct layouts, enum values) and the config builder.

Host-side glue only: no arithmetic of the path lives here.  The same structures are handed to the
CUDA library (``_lib.py``) and, in tests, to the CPU oracle -- both take ``adrp_config`` by pointer.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import math

from .assets import DroneParams, drone_params
from .utils.enums import ActionType, DroneModel, Physics, RaceMode

ABI_VERSION = 6
MAX_DRONES, MAX_GATES, MAX_OBSTACLES, STATE_DIM = 32, 8, 8, 20

F32, F64 = 0, 1
PHY_DYN, PHY_GND, PHY_DRAG, PHY_DW, PHY_GROUND = 1, 2, 4, 8, 16
ACT_RPM_RAW, ACT_RPM, ACT_ONE_D_RPM, ACT_PID, ACT_ONE_D_PID, ACT_VEL, ACT_MEL, ACT_PID_FULL, ACT_MEL_FULL, ACT_MEL_CMD = range(10)
# enum adrp_command: Command (utils/enums.py:58-70) as the first float of an ACT_MEL_CMD row
CMD_ID = {"non": 0, "fst": 1, "tko": 2, "toy": 3, "tov": 4, "lnd": 5, "ldy": 6, "ldv": 7, "stp": 8, "gto": 9, "ntf": 10}
ENV_CTRL, ENV_HOVER, ENV_MULTIHOVER, ENV_RACE = range(4)
FLAG_ELIMINATED, FLAG_FINISHED, FLAG_CRASHED, FLAG_OUT_OF_BOUNDS, FLAG_UNSTABLE, FLAG_TUMBLE_STOP = 1, 2, 4, 8, 16, 32
WRAP_OBS, WRAP_REWARD = 1, 2                           # adrp_config.wrappers (gym_pybullet_adrp/utils/wrapper.py fused)

MODEL_ID = {DroneModel.CF2X: 0, DroneModel.CF2P: 1, DroneModel.RACE: 2}
ACTION_DIM = {ACT_RPM_RAW: 4, ACT_RPM: 4, ACT_ONE_D_RPM: 1, ACT_PID: 3, ACT_ONE_D_PID: 1, ACT_VEL: 4, ACT_MEL: 4,
              ACT_PID_FULL: 12, ACT_MEL_FULL: 13, ACT_MEL_CMD: 16}
ACTION_ID = {ActionType.RPM: ACT_RPM, ActionType.ONE_D_RPM: ACT_ONE_D_RPM, ActionType.PID: ACT_PID,
             ActionType.ONE_D_PID: ACT_ONE_D_PID, ActionType.VEL: ACT_VEL, ActionType.MEL: ACT_MEL}

#: Physics enum -> ADRP_PHY_* mask.  PYB modes get the declared plane response.
PHYSICS_MASK = {
    Physics.PYB: PHY_GROUND,
    Physics.DYN: PHY_DYN,
    Physics.PYB_GND: PHY_GND | PHY_GROUND,
    Physics.PYB_DRAG: PHY_DRAG | PHY_GROUND,
    Physics.PYB_DW: PHY_DW | PHY_GROUND,
    Physics.PYB_GND_DRAG_DW: PHY_GND | PHY_DRAG | PHY_DW | PHY_GROUND,
    Physics.DYN_GND_DRAG_DW: PHY_DYN | PHY_GND | PHY_DRAG | PHY_DW,
}

d = C.c_double
i32 = C.c_int32


class DroneParamsC(C.Structure):
    _fields_ = [(n, d) for n in (
        "mass", "arm", "kf", "km", "thrust2weight", "max_speed_kmh", "gnd_eff_coeff", "prop_radius",
        "drag_coeff_xy", "drag_coeff_z", "dw_coeff_1", "dw_coeff_2", "dw_coeff_3", "ixx", "iyy", "izz",
        "collision_r", "collision_h", "collision_z_offset")] + [
        ("prop_x", d * 4), ("prop_y", d * 4),
        ("pyb_mass", d), ("pyb_ixx", d), ("pyb_iyy", d), ("pyb_izz", d), ("lin_damping", d), ("ang_damping", d)]


class TrackC(C.Structure):
    _fields_ = [("num_gates", i32), ("num_obstacles", i32), ("gates", (d * 4) * MAX_GATES),
                ("gate_type", i32 * MAX_GATES), ("obstacles", (d * 3) * MAX_OBSTACLES), ("bounds_hi", d * 3),
                ("contact_margin", d), ("visibility_range", d)]


class RandomizationC(C.Structure):
    _fields_ = [("random_gates_obstacles", i32), ("random_drone_state", i32), ("random_drone_inertia", i32),
                ("disturbances", i32), ("gate_range", d * 2), ("obstacle_range", d * 2),
                ("pos_range", (d * 2) * 3), ("rot_range", (d * 2) * 3), ("mass_range", d * 2),
                ("ixx_range", d * 2), ("iyy_range", d * 2), ("izz_range", d * 2), ("action_noise_std", d),
                ("wind_lo", d * 3), ("wind_hi", d * 3)]


class ConfigC(C.Structure):
    _fields_ = [("abi_version", i32), ("device", i32), ("precision", i32), ("num_envs", i32),
                ("drones_per_env", i32), ("drone_model", i32), ("physics", i32), ("pyb_freq", i32),
                ("ctrl_freq", i32), ("action_type", i32), ("env_kind", i32), ("race_mode", i32),
                ("auto_reset", i32), ("dslpid_model", i32), ("bug_compat", C.c_uint32), ("wrappers", i32),
                ("g", d), ("episode_len_sec", d), ("mel_mass", d), ("seed", C.c_uint64), ("env_offset", C.c_uint64),
                ("init_xyz", (d * 3) * MAX_DRONES), ("init_rpy", (d * 3) * MAX_DRONES),
                ("init_vel", (d * 3) * MAX_DRONES), ("init_pqr", (d * 3) * MAX_DRONES),
                ("target_pos", (d * 3) * MAX_DRONES), ("drone", DroneParamsC), ("dslpid_drone", DroneParamsC),
                ("track", TrackC), ("rnd", RandomizationC), ("obs_columns", i32), ("reserved0", i32)]


def drone_params_c(p: DroneParams, body: DroneParams | None = None) -> DroneParamsC:
    """``body``: the model whose mass/inertia Bullet integrates (race: legacy cf2x)."""
    body = body or p
    out = DroneParamsC()
    for name in ("mass", "arm", "kf", "km", "thrust2weight", "max_speed_kmh", "gnd_eff_coeff", "prop_radius",
                 "drag_coeff_xy", "drag_coeff_z", "dw_coeff_1", "dw_coeff_2", "dw_coeff_3", "ixx", "iyy", "izz"):
        setattr(out, name, getattr(p, name))
    out.collision_r, out.collision_h, out.collision_z_offset = p.collision_r, p.collision_h, p.collision_z_offset
    for k in range(4):
        out.prop_x[k], out.prop_y[k] = p.prop_xy[k]
    out.pyb_mass, out.pyb_ixx, out.pyb_iyy, out.pyb_izz = body.mass, body.ixx, body.iyy, body.izz
    out.lin_damping = out.ang_damping = 0.04
    return out


def default_init_xyz(p: DroneParams, M: int):
    """BaseAviary.py:193-197."""
    z = p.collision_h / 2 - p.collision_z_offset + 0.1
    return [[i * 4 * p.arm, i * 4 * p.arm, z] for i in range(M)]


def make_config(*, num_envs, num_drones=1, drone_model=DroneModel.CF2X, physics=Physics.PYB, pyb_freq=240,
                ctrl_freq=240, action_type=ACT_RPM_RAW, env_kind=ENV_CTRL, precision=F32, device=0,
                initial_xyzs=None, initial_rpys=None, initial_vels=None, initial_pqrs=None, target_pos=None,
                episode_len_sec=8.0, auto_reset=False, race_mode=RaceMode.COMPARE, seed=0, env_offset=0,
                physics_mask=None, body_model=None, mel_mass=0.027, obs_columns=0) -> ConfigC:
    if pyb_freq % ctrl_freq != 0:
        # same check and message as BaseAviary.__init__ (:79-80)
        raise ValueError('[ERROR] in BaseAviary.__init__(), pyb_freq is not divisible by env_freq.')
    if not 1 <= num_drones <= MAX_DRONES:
        raise ValueError(f"num_drones must be in 1..{MAX_DRONES} (one env lives inside one warp)")
    p = drone_params(drone_model)
    c = ConfigC()
    c.abi_version, c.device, c.precision = ABI_VERSION, device, precision
    c.num_envs, c.drones_per_env = int(num_envs), int(num_drones)
    c.drone_model = MODEL_ID[drone_model]
    c.physics = PHYSICS_MASK[physics] if physics_mask is None else int(physics_mask)
    c.pyb_freq, c.ctrl_freq = int(pyb_freq), int(ctrl_freq)
    c.action_type, c.env_kind = int(action_type), int(env_kind)
    c.race_mode = race_mode.value if isinstance(race_mode, RaceMode) else int(race_mode)
    c.auto_reset = int(bool(auto_reset))
    c.dslpid_model = MODEL_ID[DroneModel.CF2X]           # BaseRLAviary.py:75-76 (E11)
    c.bug_compat = 0xFFFFFFFF
    c.g, c.episode_len_sec, c.mel_mass = 9.8, float(episode_len_sec), float(mel_mass)
    c.seed, c.env_offset = int(seed), int(env_offset)
    c.obs_columns = int(obs_columns)
    xyz = default_init_xyz(p, num_drones) if initial_xyzs is None else initial_xyzs
    rpy = [[0.0] * 3] * num_drones if initial_rpys is None else initial_rpys
    vel = [[0.0] * 3] * num_drones if initial_vels is None else initial_vels
    pqr = [[0.0] * 3] * num_drones if initial_pqrs is None else initial_pqrs
    for i in range(num_drones):
        for k in range(3):
            c.init_xyz[i][k], c.init_rpy[i][k] = float(xyz[i][k]), float(rpy[i][k])
            c.init_vel[i][k], c.init_pqr[i][k] = float(vel[i][k]), float(pqr[i][k])
    if target_pos is None:
        if env_kind == ENV_MULTIHOVER:                    # MultiHoverAviary.py:71
            target_pos = [[xyz[i][0], xyz[i][1], xyz[i][2] + 1.0 / (i + 1)] for i in range(num_drones)]
        else:                                             # HoverAviary.py:51
            target_pos = [[0.0, 0.0, 1.0]] * num_drones
    for i in range(num_drones):
        for k in range(3):
            c.target_pos[i][k] = float(target_pos[i][k])
    c.drone = drone_params_c(p, drone_params(body_model) if body_model else None)
    c.dslpid_drone = drone_params_c(drone_params(DroneModel.CF2X))
    c.track.contact_margin, c.track.visibility_range = 0.02, 0.45
    return c


def command_row(cmd, args):
    """``(Command, args)`` as ``BaseController.predict`` returns it -> the 16-float row of ``ACT_MEL_CMD`` (include/adrp.h):
    ``[adrp_command, flattened args, 0 ...]``.  ``cmd``: the ``Command`` enum member (or its value string)."""
    row = np.zeros(16)
    row[0] = CMD_ID[getattr(cmd, "value", cmd)]
    flat = [np.atleast_1d(np.asarray(a, dtype=float)) for a in args]
    flat = np.concatenate(flat) if flat else np.zeros(0)
    if len(flat) > 15:
        raise ValueError(f"{cmd}: too many arguments")
    row[1:1 + len(flat)] = flat
    return row


def make_race_config(race_config, *, num_envs, num_drones=2, drone_model=DroneModel.CF2X, physics=Physics.PYB,
                     pyb_freq=500, ctrl_freq=25, racemode=RaceMode.COMPARE, precision=F32, device=0, seed=None,
                     env_offset=0, auto_reset=False, physics_mask=None, mel_mass=0.027, wrappers=0,
                     full_state_actions=False, command_actions=False) -> ConfigC:
    """adrp_config for MultiRaceAviary(race_config, ...) (envs/MultiRaceAviary.py:31-123)."""
    rc = race_config
    drones = list(rc["init_states"].keys())[:num_drones]
    if len(drones) < num_drones:
        raise ValueError("race_config.init_states has fewer entries than num_drones")
    st = [rc["init_states"][d] for d in drones]
    c = make_config(num_envs=num_envs, num_drones=num_drones, drone_model=drone_model, physics=physics,
                    pyb_freq=pyb_freq, ctrl_freq=ctrl_freq, action_type=ACT_MEL_CMD if command_actions else (ACT_MEL_FULL if full_state_actions else ACT_MEL),
                    env_kind=ENV_RACE, precision=precision,
                    device=device, initial_xyzs=[s["pos"] for s in st],
                    # _drone_init uses the config rpy WITHOUT deg->rad (MultiRaceAviary.py:459)
                    initial_rpys=[s["rpy"] for s in st], initial_vels=[s["vel"] for s in st],
                    initial_pqrs=[s["pqr"] for s in st], episode_len_sec=rc["episode_len_sec"],
                    auto_reset=auto_reset, race_mode=racemode, seed=rc.get("seed", 0) if seed is None else seed,
                    env_offset=env_offset, physics_mask=physics_mask,
                    body_model="cf2x",            # Bullet mass / inertia from the legacy urdf (:409-416)
                    mel_mass=mel_mass)
    c.wrappers = int(wrappers)
    t = c.track
    gates, obst = rc["gates"], rc["obstacles"]
    if len(gates) > MAX_GATES or len(obst) > MAX_OBSTACLES:
        raise ValueError("too many gates / obstacles")
    t.num_gates, t.num_obstacles = len(gates), len(obst)
    for g, row in enumerate(gates):
        t.gates[g][0], t.gates[g][1], t.gates[g][2], t.gates[g][3] = row[0], row[1], row[2], row[5]
        if row[6] not in (0, 1):
            raise ValueError("Unknown gate type.")                      # MultiRaceAviary.py:483-484
        t.gate_type[g] = int(row[6])
    for o, row in enumerate(obst):
        for k in range(3):
            t.obstacles[o][k] = row[k]
    for k in range(3):
        t.bounds_hi[k] = rc["bounds"][1][k]
    r = c.rnd
    r.random_gates_obstacles = int(bool(rc.get("random_gates_obstacles")))
    r.random_drone_state = int(bool(rc.get("random_drone_state")))
    r.random_drone_inertia = int(bool(rc.get("random_drone_inertia")))
    r.disturbances = int(bool(rc.get("disturbances")))
    if r.random_gates_obstacles:
        info = rc["random_gates_obstacles_info"]
        r.gate_range[0], r.gate_range[1] = info["gates"]["range"]
        r.obstacle_range[0], r.obstacle_range[1] = info["obstacles"]["range"]
    if r.random_drone_state:
        info = rc["random_drone_state_info"]
        for k, ax in enumerate("xyz"):
            r.pos_range[k][0], r.pos_range[k][1] = info["pos"][ax]
        for k, ax in enumerate("rpy"):
            r.rot_range[k][0], r.rot_range[k][1] = info["rot"][ax]
    if r.random_drone_inertia:
        info = rc["random_drone_inertia_info"]
        for field, key in (("mass_range", "M"), ("ixx_range", "Ixx"), ("iyy_range", "Iyy"), ("izz_range", "Izz")):
            getattr(r, field)[0], getattr(r, field)[1] = info[key]["range"]
    if r.disturbances:
        info = rc["disturbances_info"]
        r.action_noise_std = info["action"]["std"]
        for k in range(3):
            r.wind_lo[k], r.wind_hi[k] = info["dynamics"]["low"][k], info["dynamics"]["high"][k]
    return c
