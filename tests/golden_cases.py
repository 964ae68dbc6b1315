"""Shared helpers: load a golden .npz (tests/golden/, produced by the reference's own Python) and
build the matching adrp_config.  Used by the oracle tests (CPU) and the CUDA parity tests (GPU)."""
from __future__ import annotations

import json
import os

import numpy as np

from gym_pybullet_adrp_b200 import _abi
from gym_pybullet_adrp_b200.utils.enums import DroneModel, Physics

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

CTRL_CASES = ["ctrl_dyn_cf2x_s1", "ctrl_dyn_cf2x_s8_m3", "ctrl_dyn_cf2p_s8", "ctrl_dyn_race_s8",
              "ctrl_dyn_cf2x_tilted", "ctrl_pyb_cf2x_s1", "ctrl_pyb_cf2x_s8_m2", "ctrl_pyb_cf2p_s8",
              "ctrl_pyb_race_s8", "ctrl_pyb_gnd_s8", "ctrl_pyb_drag_s8", "ctrl_pyb_dw_s5_m4",
              "ctrl_pyb_all_s5_m8", "ctrl_pyb_ground_contact", "ctrl_pyb_all_s5_m8_wide"]
RL_CASES = ["hover_pyb_rpm", "hover_dyn_one_d_rpm", "hover_dyn_pid", "hover_pyb_one_d_pid",
            "multihover_pyb_rpm_m2", "hover_pyb_rpm_long"]
RL_RESET_CASES = ["hover_pyb_rpm_reset"]                 # reset() in the middle of a run: the action buffer survives it
PID_CASES = ["pid_dyn_c1", "pid_pyb_c1"]
VEL_CASES = ["vel_pyb_m2", "vel_dyn_m1"]                 # VelocityAviary (ActionType.VEL through the env's own DSLPID)
RL_VEL_CASES = ["hover_pyb_vel"]                         # BaseRLAviary ActionType.VEL (:208-223)

ACT = {"RPM": _abi.ACT_RPM, "ONE_D_RPM": _abi.ACT_ONE_D_RPM, "PID": _abi.ACT_PID, "ONE_D_PID": _abi.ACT_ONE_D_PID,
       "VEL": _abi.ACT_VEL}


def load(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    return meta, {k: z[k] for k in z.files if k != "meta"}


def config_for(meta, precision=_abi.F64, num_envs=1, **over):
    kind = meta["env"]
    kw = dict(num_envs=num_envs, num_drones=meta["M"], drone_model=DroneModel[meta["model"]],
              physics=Physics[meta["physics"]], pyb_freq=meta["pyb_freq"], ctrl_freq=meta["ctrl_freq"],
              precision=precision, initial_xyzs=meta["init_xyz"], initial_rpys=meta.get("init_rpy"))
    if kind == "ctrl":
        kw.update(action_type=_abi.ACT_RPM_RAW, env_kind=_abi.ENV_CTRL)
    elif kind == "velocity":
        kw.update(action_type=_abi.ACT_VEL, env_kind=_abi.ENV_CTRL)
    elif kind == "ctrl+dslpid":
        kw.update(action_type=_abi.ACT_PID_FULL, env_kind=_abi.ENV_CTRL)
    else:
        kw.update(action_type=ACT[meta["act"]], env_kind=_abi.ENV_HOVER if kind == "hover" else _abi.ENV_MULTIHOVER,
                  target_pos=meta["target"] if kind == "multihover" else [meta["target"]])
    kw.update(over)
    return _abi.make_config(**kw)


def rel_err(a, b, floor=1.0):
    """max |a-b| / max(|b|, floor) -- relative where the value is large, absolute near zero."""
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), floor)))

# ---- fp32 product build vs a float64 / fp32-oracle run: BASELINE.json north_star's terms
POS_TOL32, ANG_TOL32, REST_TOL32 = 1e-4, 1e-4, 2e-4       # metres, radians (absolute); every other column relative above 1
MARGINS = {}                                              # test id -> worst (pos, ang, rest) seen: conftest writes them out


def assert_fp32_north_star(o, g, layout, what):
    """``o``: fp32 kernel observation rows (..., O); ``g``: the reference rows.  north_star: "the fp32 build must agree within
    1e-4 m and 1e-4 rad" -- absolute on the position columns and on the (wrapped) Euler angles.  The remaining columns
    (velocities, body rates, RPM, action buffer: not part of north_star's statement) keep the 2e-4 relative bound.
    layout "ctrl": 20-float CtrlAviary rows (rpy at 7:10); "kin": rows starting [pos3, rpy3, vel3, ang_v3]."""
    o, g = np.asarray(o, np.float64), np.asarray(g, np.float64)
    a0 = 7 if layout == "ctrl" else 3
    pos = float(np.abs(o[..., 0:3] - g[..., 0:3]).max())
    d = np.abs(o[..., a0:a0 + 3] - g[..., a0:a0 + 3])
    ang = float(np.minimum(d, 2 * np.pi - d).max())
    keep = np.ones(o.shape[-1], bool)
    keep[0:3] = keep[a0:a0 + 3] = False
    rest = rel_err(o[..., keep], g[..., keep]) if keep.any() else 0.0
    key = str(what[0]) if isinstance(what, tuple) else str(what)
    w = MARGINS.get(key, (0.0, 0.0, 0.0))
    MARGINS[key] = (max(w[0], pos), max(w[1], ang), max(w[2], rest))
    assert pos < POS_TOL32, (what, "position", pos)
    assert ang < ANG_TOL32, (what, "angle", ang)
    assert rest < REST_TOL32, (what, "other columns (relative)", rest)


FULLSTATE_CASES = ["race_fullstate_compare_gs"]          # (Command.FULLSTATE, args) tuples with vel / acc / yaw / rates
COMMANDER_CASES = ["race_commander_hardcoded_gs", "race_commander_script_gs"]   # every Command through the high-level commander
WRAPPER_CASES = ["race_wrappers_compare_gs"]             # DroneObservationWrapper(RewardWrapper(MultiRaceAviary))
RACE_CASES = ["race_pyb_compete_gs", "race_pyb_compare_level3", "race_dyn_compete_level1", "race_pyb_all_compete_level3",
              "race_pyb_compete_level0", "race_pyb_compare_level1", "race_pybgnd_compare_twogates"]


def race_config_for(meta, precision=_abi.F64, num_envs=1, **over):
    from gym_pybullet_adrp_b200 import configs
    from gym_pybullet_adrp_b200.utils.enums import RaceMode
    rc = configs.load_config(meta["level"])
    cfg = _abi.make_race_config(rc, num_envs=num_envs, num_drones=meta["M"], physics=Physics[meta["physics"]],
                                racemode=RaceMode[meta["racemode"]], precision=precision, **over)
    return rc, cfg


def place_draws(rc, raw, M, G=4, O=4):
    """The reference only draws for the randomisation blocks that are switched on; the C ABI's reset
    draw vector always has the full layout [gates 3G | obstacles 2O | per drone: inertia 4, pos 3, rot 3]."""
    out = np.zeros(3 * G + 2 * O + 10 * M)
    raw = list(raw)
    k = 0
    if rc.get("random_gates_obstacles"):
        n = 3 * G + 2 * O
        out[:n] = raw[k:k + n]
        k += n
    for i in range(M):
        base = 3 * G + 2 * O + 10 * i
        if rc.get("random_drone_inertia"):
            out[base:base + 4] = raw[k:k + 4]
            k += 4
        if rc.get("random_drone_state"):
            out[base + 4:base + 10] = raw[k:k + 6]
            k += 6
    assert k == len(raw), (k, len(raw))
    return out
