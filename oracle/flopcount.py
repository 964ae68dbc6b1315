#!/usr/bin/env python
"""TEST INFRASTRUCTURE -- mechanical operation count of the restated path.

Drives oracle/liboracle_count.so (the oracle's own statements compiled with counting arithmetic types,
oracle/adrp_oracle_count.cpp) over the bench workloads and prints operations per SUBSTEP-UPDATE (one drone, one
physics substep) and per CONTROL STEP (action map, observation, reward, flags), separated by running every
workload at two substep counts.  flops = add + mul + compare (a multiply-add counts 2); SFU = div, sqrt, sin/cos,
tan, atan2, asin, exp, log.  The results are recorded in BASELINE.md and used by bench.py's roofline.

    make -C oracle liboracle_count.so && python oracle/flopcount.py [--json]
"""
from __future__ import annotations

import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gym_pybullet_adrp_b200 import _abi, configs  # noqa: E402
from gym_pybullet_adrp_b200.utils.enums import Physics, RaceMode  # noqa: E402

KINDS = ["add", "mul", "cmp", "div", "sqrt", "sincos", "tan", "atan2", "asin", "exp", "log"]
FLOP, SFU = KINDS[:3], KINDS[3:]


def lib():
    L = C.CDLL(os.path.join(ROOT, "oracle", "liboracle_count.so"))
    L.cnt_create.argtypes = [C.POINTER(_abi.ConfigC), C.POINTER(C.c_void_p)]
    for f in ("cnt_destroy", "cnt_obs_dim", "cnt_action_dim"):
        getattr(L, f).argtypes = [C.c_void_p]
    L.cnt_reset.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    L.cnt_set_state.argtypes = [C.c_void_p, C.c_void_p]
    L.cnt_get_state.argtypes = [C.c_void_p, C.c_void_p]
    L.cnt_step.argtypes = [C.c_void_p] * 8
    L.cnt_counts.argtypes = [C.c_void_p, C.c_int]
    assert L.cnt_kinds() == len(KINDS)
    return L


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def run(L, cfg, actions_fn, steps, state_fn=None, skip=1):
    """Counts per drone over `steps` control steps (after `skip` uncounted steps)."""
    h = C.c_void_p()
    assert L.cnt_create(C.byref(cfg), C.byref(h)) == 0
    N, M, O, A = cfg.num_envs, cfg.drones_per_env, L.cnt_obs_dim(h), L.cnt_action_dim(h)
    obs, rew = np.zeros((N, M, O)), np.zeros(N)
    te, tr, fl = np.zeros(N, np.uint8), np.zeros(N, np.uint8), np.zeros((N, M), np.uint8)
    L.cnt_reset(h, None, ptr(obs))
    if state_fn is not None:
        st = np.zeros((N, M, 20))
        L.cnt_get_state(h, ptr(st))
        st = np.ascontiguousarray(state_fn(st))
        L.cnt_set_state(h, ptr(st))
    out = (C.c_ulonglong * len(KINDS))()
    for k in range(steps + skip):
        a = np.ascontiguousarray(actions_fn(k, N, M, A), np.float64)
        L.cnt_step(h, ptr(a), None, ptr(obs), ptr(rew), ptr(te), ptr(tr), ptr(fl))
        if k == skip - 1:
            L.cnt_counts(out, 1)                      # discard the first steps (cold controller state)
    L.cnt_counts(out, 1)
    L.cnt_destroy(h)
    assert np.isfinite(obs).all()
    return np.array(list(out), float) / (N * M)


def split(L, mk_cfg, actions_fn, state_fn=None, steps=8, pyb=240, ctrl1=30, ctrl2=15):
    """The SAME physics ticks (skipped and counted) at control rate ctrl1 = 2 ctrl2: totals T1 = 2K a + n b and
    T2 = K a + n b  ->  per control step a, per substep-update b (tick-dependent work, e.g. the Mellinger sub-rate
    schedule, is identical in both runs)."""
    assert ctrl1 == 2 * ctrl2
    n = steps * (pyb // ctrl2)
    t1 = run(L, mk_cfg(pyb, ctrl1), actions_fn, 2 * steps, state_fn, skip=2)
    t2 = run(L, mk_cfg(pyb, ctrl2), actions_fn, steps, state_fn, skip=1)
    per_step = (t1 - t2) / steps
    return (t2 - steps * per_step) / n, per_step


def fmt(v):
    d = dict(zip(KINDS, v))
    return dict(flops=round(sum(d[k] for k in FLOP), 1), sfu=round(sum(d[k] for k in SFU), 1),
                **{k: round(d[k], 2) for k in KINDS})


def main():
    L = lib()
    hover = float(np.sqrt(9.8 * 0.03454 / (4 * 3.16e-10)))
    rng = np.random.default_rng(0)

    def rpm_actions(k, N, M, A):
        return hover * (1 + 0.01 * np.random.default_rng(k).uniform(-1, 1, (N, M, A)))

    def formation(M):                                   # bench.py's formation: 3 x 3 grid, 1.2 m pitch, start heights U(0.2, 1.5)
        return [[1.2 * (i % 3), 1.2 * (i // 3), 0.2 + 1.3 * rng.uniform()] for i in range(M)]

    def ctrl_cfg(mask, M):
        xyz = formation(M)
        return lambda pyb, ctrl: _abi.make_config(num_envs=16, num_drones=M, physics=Physics.PYB, physics_mask=mask,
                                                  pyb_freq=pyb, ctrl_freq=ctrl, action_type=_abi.ACT_RPM_RAW,
                                                  env_kind=_abi.ENV_CTRL, precision=_abi.F64, initial_xyzs=xyz)

    D, G, R, W = _abi.PHY_DYN, _abi.PHY_GND, _abi.PHY_DRAG, _abi.PHY_DW
    rows = {}
    for name, mask, M in [("DYN", D, 1), ("DYN+GND", D | G, 1), ("DYN+DRAG", D | R, 1), ("DYN+DW (M=8)", D | W, 8),
                          ("C5: DYN+GND+DRAG+DW (M=8)", D | G | R | W, 8), ("PYB", 0, 1),
                          ("C4: PYB+GND+DRAG+DW (M=8)", G | R | W, 8)]:
        sub, step = split(L, ctrl_cfg(mask, M), rpm_actions)
        rows[name] = dict(per_substep_update=fmt(sub), per_control_step=fmt(step))

    # HoverAviary (C2): RPM action map, 72-float observation, reward / flags
    def hover_cfg(pyb, ctrl):
        return _abi.make_config(num_envs=16, num_drones=1, physics=Physics.PYB, pyb_freq=pyb, ctrl_freq=ctrl,
                                action_type=_abi.ACT_RPM, env_kind=_abi.ENV_HOVER, precision=_abi.F64, target_pos=[[0, 0, 1]])
    sub, step = split(L, hover_cfg, lambda k, N, M, A: 0.05 * np.random.default_rng(k).uniform(-1, 1, (N, M, A)))
    rows["C2: HoverAviary PYB, ActionType.RPM"] = dict(per_substep_update=fmt(sub), per_control_step=fmt(step))

    # DSLPID (C1): one computeControl per control step
    def pid_cfg(pyb, ctrl):
        return _abi.make_config(num_envs=16, num_drones=1, physics=Physics.DYN, pyb_freq=pyb, ctrl_freq=ctrl,
                                action_type=_abi.ACT_PID_FULL, env_kind=_abi.ENV_CTRL, precision=_abi.F64)
    def pid_actions(k, N, M, A):
        a = np.zeros((N, M, A)); a[..., 2] = 1.0
        return a
    sub, step = split(L, pid_cfg, pid_actions)
    rows["C1: DYN + DSLPID (per control step)"] = dict(per_substep_update=fmt(sub), per_control_step=fmt(step))

    # race env + Mellinger per substep: exactly bench.py's C3 and C5M workloads (track, formation, setpoints), at the
    # workload's control rate and at twice that rate (same 500 Hz physics)
    import bench
    for name, wl in [("C3: race level3 COMPETE, PYB + Mellinger (M=2)", "c3"),
                     ("C5M: PYB+GND+DRAG+DW + Mellinger (M=8)", "c5m")]:
        M = bench.WORKLOADS[wl][1]
        pool = bench.make_actions(wl, 8, M, 4, 3, 8).astype(np.float64)

        def mk(pyb, ctrl, wl=wl):
            c = bench.make_cfg(wl, 8, precision=_abi.F64)
            c.pyb_freq, c.ctrl_freq, c.auto_reset = pyb, ctrl, 0
            c.rnd.disturbances = 0                       # Philox draws are integer work; the Box-Muller is listed apart
            return c
        sub, step = split(L, mk, lambda k, N, M_, A, pool=pool: pool[k % len(pool)], pyb=500, ctrl1=50, ctrl2=25, steps=6)
        rows[name] = dict(per_substep_update=fmt(sub), per_control_step=fmt(step))

    if "--json" in sys.argv:
        print(json.dumps(rows, indent=1))
        return
    print(f"{'configuration':52s} | per substep-update: flops  SFU | per control step: flops  SFU")
    for name, r in rows.items():
        a, b = r["per_substep_update"], r["per_control_step"]
        print(f"{name:52s} | {a['flops']:24.1f} {a['sfu']:5.1f} | {b['flops']:23.1f} {b['sfu']:5.1f}")
    print("\nSFU detail per substep-update:")
    for name, r in rows.items():
        a = r["per_substep_update"]
        print(f"  {name:50s} " + " ".join(f"{k}={a[k]:g}" for k in SFU if a[k]))


if __name__ == "__main__":
    main()
