#!/usr/bin/env python
"""bench.py -- drone-steps/s of the fused control-step kernel (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c1|c2|c3|c4|c5|c5m] [--scaling weak|strong]
                    [--impl reference] [--no-workloads] [--no-cpu-baseline]

One "step" = one control step (S fused physics substeps) of every drone of the workload.  Default
workload C5 (BASELINE.json configs[4], the scaling sweep the metric is quoted on): 2^20 drones per GPU,
131072 envs x 8 CF2X drones, DYN + ground effect + drag + downwash, 240/30 Hz (S = 8), random RPM
actions, CtrlAviary observation.  Weak scaling: every rank owns its own env range, no collective on the
step path.  See DESIGN.md section "Measurement" for the roofline byte count.

The timed region is `repeats` back-to-back blocks of exactly --steps control steps, each block between two CUDA events on the
launching stream, the whole region between barrier + synchronize; `value` comes from the MEDIAN block (max over ranks per
block), so that the driver's 20-step run is not a 2 ms sample.  At N = 1 the default run also measures every other BASELINE
config (C1 .. C4) and the north_star's target configuration (C5M) into "workloads", each with an untimed parity spot check
of that workload's handle against the CPU oracle (64 envs x 4 control steps) before its timing.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from gym_pybullet_adrp_b200 import _abi  # noqa: E402
from gym_pybullet_adrp_b200.utils.enums import DroneModel, Physics  # noqa: E402

HOVER_RPM = float(np.sqrt(9.8 * 0.03454 / (4 * 3.16e-10)))
WORKLOADS = {
    # name: (envs per GPU, drones per env, physics, pyb, ctrl, env kind, action type, description)
    "c1": (1, 1, Physics.DYN, 240, 30, _abi.ENV_CTRL, _abi.ACT_PID_FULL,
           "C1: 1 env x 1 CF2X, DYN, 240/30 Hz, DSLPIDControl towards [0, 0, 1] (examples/pid.py:101-147), obs (1,1,20); "
           "launch-latency-bound on a GPU: adrp_step_n is the number to read"),
    "c5": (131072, 8, Physics.DYN_GND_DRAG_DW, 240, 30, _abi.ENV_CTRL, _abi.ACT_RPM_RAW,
           "C5: 2^20 CF2X drones/GPU (131072 envs x 8), DYN+GND+DRAG+DW, 240/30 Hz, random RPM, obs (N,8,20)"),
    "c4": (32768, 8, Physics.PYB_GND_DRAG_DW, 240, 48, _abi.ENV_CTRL, _abi.ACT_RPM_RAW,
           "C4: 32768 envs x 8 CF2X, PYB_GND_DRAG_DW, 240/48 Hz, open-loop RPM, obs (N,8,20)"),
    "c2": (65536, 1, Physics.PYB, 240, 30, _abi.ENV_HOVER, _abi.ACT_RPM,
           "C2: HoverAviary 65536 envs x 1 CF2X, PYB, 240/30 Hz, RPM actions in [-1,1], obs (N,1,72), auto-reset"),
    "c3": (16384, 2, Physics.PYB, 500, 25, _abi.ENV_RACE, _abi.ACT_MEL,
           "C3: race env level3 COMPETE, 16384 envs x 2 CF2X, PYB, 500/25 Hz, Mellinger per substep, gates + obstacles + "
           "drone-drone contacts, all randomisation and disturbances (Philox), obs (N,2,55), auto-reset"),
    "c5m": (131072, 8, Physics.PYB_GND_DRAG_DW, 500, 25, _abi.ENV_RACE, _abi.ACT_MEL,
            "C5M: 2^20 CF2X drones/GPU (131072 envs x 8), Bullet free body + GND+DRAG+DW, 500/25 Hz, Mellinger per substep, "
            "formation hover setpoints, no track, obs (N,8,13)"),
}


def formation_xyz(M, z0=None):
    """3 x 3 grid, 1.2 m pitch: the downwash term is evaluated for every pair but stays physical (its alpha is
    singular for dz -> 0+ at small lateral distance, BaseAviary.py:808-810)."""
    return [[1.2 * (k % 3), 1.2 * (k // 3), (0.5 + 0.25 * (k % 4)) if z0 is None else z0] for k in range(M)]


def race_rc(wl):
    from gym_pybullet_adrp_b200 import configs
    M = WORKLOADS[wl][1]
    if wl == "c3":
        return configs.load_config("level3")
    rc = configs.load_config("getting_started")
    rc["gates"], rc["obstacles"] = [], []
    rc["bounds"] = [[-10, -10, 0], [10, 10, 5]]
    rc["episode_len_sec"] = 1e9
    rc["init_states"] = configs.configify({f"drone{i}": {"pos": p, "vel": [0, 0, 0], "rpy": [0, 0, 0], "pqr": [0, 0, 0]}
                                           for i, p in enumerate(formation_xyz(M, 0.05))})
    return rc


def make_cfg(wl, n_envs, device=0, env_offset=0, precision=_abi.F32, obs_columns=0):
    N, M, phy, pyb, ctrl, kind, act, _ = WORKLOADS[wl]
    if kind == _abi.ENV_RACE:
        from gym_pybullet_adrp_b200.utils.enums import RaceMode
        return _abi.make_race_config(race_rc(wl), num_envs=n_envs, num_drones=M, physics=phy, pyb_freq=pyb, ctrl_freq=ctrl,
                                     racemode=RaceMode.COMPETE if wl == "c3" else RaceMode.COMPARE, precision=precision,
                                     device=device, seed=1337, env_offset=env_offset, auto_reset=(wl == "c3"))
    xyz = formation_xyz(M) if M > 1 else None
    return _abi.make_config(num_envs=n_envs, num_drones=M, drone_model=DroneModel.CF2X, physics=phy, pyb_freq=pyb,
                            ctrl_freq=ctrl, action_type=act, env_kind=kind, precision=precision, device=device,
                            initial_xyzs=xyz, auto_reset=(kind != _abi.ENV_CTRL), seed=5, env_offset=env_offset,
                            obs_columns=obs_columns)


def make_actions(wl, n_envs, M, A, seed, count):
    rng = np.random.default_rng(seed)
    kind = WORKLOADS[wl][5]
    if kind == _abi.ENV_RACE:
        # full-state setpoints [x, y, z, yaw]: c3 climbs and heads for the first gate, c5m holds a formation
        cfg = make_cfg(wl, 1)
        init = np.array([[cfg.init_xyz[i][k] for k in range(3)] for i in range(M)], np.float32)
        # one hover target per drone (formation slot + U(-0.2, 0.2) m, staggered heights), re-sent every control step
        # with centimetre jitter: setpoint JUMPS of decimetres every 40 ms would tumble some of the 10^6 drones, and a
        # tumbling drone that crosses a neighbour's height meets the downwash singularity
        a = np.zeros((count, n_envs, M, 4), np.float32)
        xy = init[None, :, 0:2] + rng.uniform(-0.2, 0.2, (n_envs, M, 2))
        z = (0.4 if wl == "c3" else 0.6) + 0.15 * (np.arange(M) % 3)[None, :] + rng.uniform(0, 0.05, (n_envs, M))
        for c in range(count):
            a[c, :, :, 0:2] = xy + rng.uniform(-0.01, 0.01, (n_envs, M, 2))
            a[c, :, :, 2] = z + rng.uniform(-0.01, 0.01, (n_envs, M))
        return a
    if WORKLOADS[wl][6] == _abi.ACT_PID_FULL:              # examples/pid.py:138-147: a fixed target, rpy / vel / rates zero
        a = np.zeros((count, n_envs, M, A), np.float32)
        a[..., 2] = 1.0
        return a
    if WORKLOADS[wl][6] == _abi.ACT_RPM_RAW:
        # rpm = HOVER_RPM (1 + 0.05 a) (BaseRLAviary.py:192) with a = collective U(-1,1) + 1 % per-motor differential:
        # open-loop quadrotors tumble within a second under fully independent motor commands, and tumbling drones
        # fly into the singularity of the reference's downwash formula (alpha ~ 1/dz^2, BaseAviary.py:808) -- the
        # arithmetic per step is the same, but the rollout stays a flight instead of an overflow
        u = rng.uniform(-1, 1, (count, n_envs, M, 1)) + 0.01 * rng.uniform(-1, 1, (count, n_envs, M, A))
        u -= u.mean(axis=0, keepdims=True)           # the pool is replayed cyclically: no net thrust / torque bias per cycle
        return (HOVER_RPM * (1 + 0.05 * u)).astype(np.float32)
    return rng.uniform(-1, 1, (count, n_envs, M, A)).astype(np.float32)


def initial_state(wl, n_envs, M, seed):
    """C5 / C4: the formation grid with start heights U(0.2, 1.5)."""
    rng = np.random.default_rng(seed)
    st = np.zeros((n_envs, M, 20), np.float32)
    xyz = np.array(formation_xyz(M), np.float32)
    st[:, :, 0] = xyz[None, :, 0]
    st[:, :, 1] = xyz[None, :, 1]
    st[:, :, 2] = rng.uniform(0.2, 1.5, (n_envs, M))
    st[:, :, 6] = 1.0
    return st


def algorithmic_bytes_per_drone_step(wl, O, A, M):
    """fp32 bytes that MUST cross HBM per drone per control step (DESIGN.md 'Measurement')."""
    phy = _abi.PHYSICS_MASK[WORKLOADS[wl][2]]
    kind = WORKLOADS[wl][5]
    state_rd = 17                                        # pos3 quat4 vel3 (rates3 | ang_v3) last_rpm4
    state_wr = 20 if phy & _abi.PHY_DYN else 17          # DYN also stores world ang_v next to the body rates
    ring, extra = 0, 0
    if kind == _abi.ENV_RACE:
        # Mellinger wrapper + firmware state (10 T + 18 f32 + 5 i32) and gate / latch word, read and written;
        # Bullet body mass / inertia read; the env's track instance (6 floats per gate, 3 per obstacle) read once per env
        extra = 2 * (10 + 18 + 5 + 1) + 4 + (6 * 4 + 3 * 4 if wl == "c3" else 0) / M
    elif kind != _abi.ENV_CTRL:
        ring = (O - 12) + A                               # read the action buffer once, write the newest slot
    if WORKLOADS[wl][6] == _abi.ACT_PID_FULL:
        extra = 2 * 9                                     # DSLPID integrators / last_rpy, read and written
    per_drone = 4 * (state_rd + A + state_wr + O + ring + extra) + 1
    per_env = 4 + 2 + 4 + 4                               # reward, two flag bytes, step counter read+write
    return per_drone + per_env / M


# algorithmic work of the restated path.  MECHANICAL COUNT of the frozen oracle (oracle/flopcount.py:
# the oracle's own statements compiled with counting arithmetic types, run on these workloads):
#   (flops per substep-update, SFU ops per substep-update, flops per control step, SFU ops per control step)
# flops = add + sub + mul + compare (a multiply-add counts 2); SFU = div, sqrt, sin/cos, atan2, asin, exp, log.
# tests/test_oracle_golden.py::test_flop_table_is_the_mechanical_count keeps this table equal to the counter's output.
MECHANICAL_OPS = {
    "c1": (216.0, 11.0, 295.0, 32.0),
    "c5": (480.0, 46.5, 32.0, 3.0),
    "c4": (570.0, 54.5, 32.0, 3.0),
    "c2": (428.0, 25.0, 80.0, 7.0),
    "c3": (825.1, 89.8, 2782.9, 306.2),
    "c5m": (1000.3, 118.5, 125.6, 38.1),
}


def algorithmic_flops_per_substep(wl):
    """flops per substep-update including the per-control-step part spread over the S substeps."""
    S = WORKLOADS[wl][3] // WORKLOADS[wl][4]
    sub, _, step, _ = MECHANICAL_OPS[wl]
    return sub + step / S


def reading_estimate_flops_per_substep(wl):
    """SURVEY 8(d)'s reading estimate (+-20 %), kept beside the mechanical count for the record."""
    N, M, physics, pyb, ctrl, kind, act, _ = WORKLOADS[wl]
    phy = _abi.PHYSICS_MASK[physics]
    fl = 160 if phy & _abi.PHY_DYN else 200
    fl += 20 if (phy & _abi.PHY_GND or kind == _abi.ENV_RACE) else 0
    fl += 45 if phy & _abi.PHY_GND else 0
    fl += 25 if phy & _abi.PHY_DRAG else 0
    fl += 17 * (M - 1) if phy & _abi.PHY_DW else 0
    fl += 450 if kind == _abi.ENV_RACE else 0
    fl += 330 / (pyb // ctrl) if act == _abi.ACT_PID_FULL else 0
    return fl


def executed_flops_per_substep(wl, n_envs):
    """FP32 flops the kernel actually EXECUTES per substep-update: 32 lanes x (2 FFMA + FMUL + FADD, packed forms twice) warp
    instructions of one launch, from the committed SASS opcode histogram of this workload's ncu capture (the newest
    profiles/r*_sass_hist_<wl>.txt), / (drones x substeps).  None without a capture at this size."""
    import glob
    import re
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", f"r*_sass_hist_{wl}.txt")))
    if n_envs != WORKLOADS[wl][0] or not found:
        return None
    w = {"FFMA": 2, "FMUL": 1, "FADD": 1, "FFMA2": 4, "FMUL2": 2, "FADD2": 2}
    total = 0.0
    with open(found[-1]) as f:
        for line in f:
            m = re.match(r"^([A-Z0-9_.]+)\s+(\d+)", line)
            if m and m.group(1).split(".")[0] in w:
                total += 32.0 * w[m.group(1).split(".")[0]] * int(m.group(2))
    N, M, S = WORKLOADS[wl][0], WORKLOADS[wl][1], WORKLOADS[wl][3] // WORKLOADS[wl][4]
    return total / (N * M * S) if total else None


def ncu_traffic_bytes(wl, n_envs):
    """dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the step kernel, from the committed
    `ncu --set full` capture of this workload (the newest profiles/r*_ncu_<wl>_metrics.csv); None if there is none
    or the capture was taken at another size."""
    import glob
    found = sorted(glob.glob(os.path.join(ROOT, "profiles", f"r*_ncu_{wl}_metrics.csv")))
    if n_envs != WORKLOADS[wl][0] or not found:
        return None
    path = found[-1]
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total = 0.0
    with open(path) as f:
        for name, unit, value in csv.reader(f):
            if name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                total += float(value) * scale.get(unit, 1.0)
    return total or None


class ClockSampler(threading.Thread):
    """Polls NVML (SM clock + throttle reasons) for the duration of the timed region.  NVML is initialised in the
    constructor (before the region starts) and polled every millisecond, so that even a 20 ms region is sampled."""

    NAMES = {"nvmlClocksThrottleReasonHwSlowdown": "hw_slowdown",
             "nvmlClocksThrottleReasonHwThermalSlowdown": "hw_thermal_slowdown",
             "nvmlClocksThrottleReasonSwThermalSlowdown": "sw_thermal_slowdown",
             "nvmlClocksThrottleReasonSwPowerCap": "sw_power_cap"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.stop_flag, self.samples, self.reasons, self.max_mhz = False, [], set(), None
        self.nv = self.h = None
        try:
            import pynvml as nv
            nv.nvmlInit()
            # LOCAL_RANK indexes CUDA_VISIBLE_DEVICES; NVML enumerates the whole box
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = vis.split(",")[index] if vis else str(index)
            self.h = nv.nvmlDeviceGetHandleByIndex(int(phys)) if phys.isdigit() else nv.nvmlDeviceGetHandleByUUID(phys)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
            self.names = {getattr(nv, k): v for k, v in self.NAMES.items()}
            self.nv = nv
        except Exception as e:  # noqa: BLE001
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def poll(self):
        nv, h = self.nv, self.h
        self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
        r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        for bit, name in self.names.items():
            if r & bit:
                self.reasons.add(name)

    def run(self):
        if self.nv is None:
            return
        try:
            while not self.stop_flag:
                self.poll()
                time.sleep(0.001)
        except Exception as e:  # noqa: BLE001
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def result(self):
        return {"sm_mhz": float(np.median(self.samples)) if self.samples else None, "sm_max_mhz": self.max_mhz,
                "samples": len(self.samples), "reasons": sorted(self.reasons)}


def cpu_port_rate(wl, n_envs_sample, steps, threads):
    """Oracle (CPU port of the reference arithmetic) on a bounded sample of the same workload."""
    from tests.oracle_lib import OracleEnv
    N, M = n_envs_sample, WORKLOADS[wl][1]
    threads = max(1, min(threads, N))                   # (envs are what the threads split: one env = one thread)
    cfg = make_cfg(wl, N, precision=_abi.F64)           # the reference computes in numpy float64
    env = OracleEnv(cfg)
    if WORKLOADS[wl][5] == _abi.ENV_CTRL and M > 1:
        env.set_state(initial_state(wl, N, M, 5))
    acts = make_actions(wl, N, M, env.A, 5, 2)
    env.step(acts[0], nthreads=threads)
    t0 = time.perf_counter()
    for k in range(steps):
        env.step(acts[k % 2], nthreads=threads)
    dt = time.perf_counter() - t0
    env.close()
    return N * M * steps / dt, dt


def cpu_sample_envs(wl):
    M, S = WORKLOADS[wl][1], WORKLOADS[wl][3] // WORKLOADS[wl][4]
    return max(1, min(WORKLOADS[wl][0], max(256, min(65536, (1 << 19) // (M * S)))))    # ~0.5 M substep-updates per control step


def cpu_envs_for_budget(wl, steps, threads, budget_s):
    """The FULL workload when `steps` control steps of it fit the time budget on this host (then the CPU arm runs the same
    config as the GPU arm), else the bounded sample.  Decided from a short run of the sample."""
    n_small = cpu_sample_envs(wl)
    cpu_port_rate(wl, n_small, 1, threads)
    rate, _ = cpu_port_rate(wl, n_small, 2, threads)
    N_full, M = WORKLOADS[wl][0], WORKLOADS[wl][1]
    est = N_full * M * (steps + 1) / rate
    return (N_full, True) if est <= budget_s else (n_small, False)


def run_reference(args):
    """--impl reference: the reference's own CPU arithmetic (oracle port; the reference itself needs
    PyBullet, which is not installable here or on the GPU box) on all host cores: the full workload when K steps of it
    fit ~30 s, else a bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(os.sched_getaffinity(0))
    wl = args.workload
    M, S = WORKLOADS[wl][1], WORKLOADS[wl][3] // WORKLOADS[wl][4]
    n_sample, full = cpu_envs_for_budget(wl, args.steps + max(args.warmup, 1), threads, 30.0)
    for _ in range(max(min(args.warmup, 2), 1)):
        cpu_port_rate(wl, n_sample, 1, threads)
    rate, dt = cpu_port_rate(wl, n_sample, args.steps, threads)
    sample = (f"{'the full workload: ' if full else ''}{n_sample} envs x {M} drones x {args.steps} control steps, fp64, "
              f"OpenMP threads of one process over envs")
    print(json.dumps({
        "impl": "reference", "metric": "drone_steps_per_s", "value": rate, "unit": "drone-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS[wl][7], "sample": sample, "same_config": full},
        "cpu_baseline": {"value": rate, "unit": "drone-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "drone-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "substep_updates_per_s": rate * S}))


def parity_spot_check(wl, device, envs=64, steps=4):
    """Untimed: `envs` envs x `steps` control steps of this workload on the fp32 product build against the CPU oracle's fp32
    build on the same inputs (the oracle as the checker, never the thing measured).  Kinematic columns of the observation
    (positions, angles, velocities) to 2e-4 (relative above 1), reward to 1e-3, terminated / truncated equal."""
    from gym_pybullet_adrp_b200 import _lib
    from tests.oracle_lib import OracleEnv
    M = WORKLOADS[wl][1]
    envs = min(envs, WORKLOADS[wl][0])
    cfg = make_cfg(wl, envs, device=device)
    gpu, cpu = _lib.Handle(cfg), OracleEnv(cfg)
    race, ctrl = WORKLOADS[wl][5] == _abi.ENV_RACE, WORKLOADS[wl][5] == _abi.ENV_CTRL
    if race:
        gpu.reset(); cpu.reset()                          # both draw the randomisation from Philox(seed; env, episode)
    if ctrl and M > 1:
        st = initial_state(wl, envs, M, 99)
        gpu.set_state(st); cpu.set_state(st)
    acts = make_actions(wl, envs, M, gpu.A, 99, steps)
    kin = 12 if not ctrl else 16
    worst, rew_err, flags_equal = 0.0, 0.0, True
    for k in range(steps):
        og, rg, teg, trg, _ = gpu.step(acts[k])
        oc, rc, tec, trc, _ = cpu.step(acts[k])
        og = og.cpu().numpy()[..., :kin].astype(np.float64)
        oc = oc[..., :kin].astype(np.float64)
        worst = max(worst, float(np.max(np.abs(og - oc) / np.maximum(np.abs(oc), 1.0))))
        rew_err = max(rew_err, float(np.max(np.abs(rg.cpu().numpy().astype(np.float64) - rc))))
        flags_equal &= bool(np.array_equal(teg.cpu().numpy(), tec) and np.array_equal(trg.cpu().numpy(), trc))
    gpu.close(); cpu.close()
    tol = 2e-4
    return {"envs": envs, "steps": steps, "max_rel_err": worst, "tol": tol, "reward_abs_err": rew_err,
            "flags_equal": flags_equal, "ok": bool(worst < tol and rew_err < 1e-3 and flags_equal),
            "against": "oracle/liboracle.so, fp32 build"}


def measure(wl, N, args, *, rank=0, world=1, local=0, dist=None, steps=None, min_time_s=0.3, max_repeats=400, e2e=True,
            step_n_k=16, peaks=None, fp32_peak=None, fp32_nominal=None):
    """Device-resident throughput (median of repeated blocks of exactly `steps` control steps), adrp_step_n throughput,
    end-to-end throughput through adrp_step_host, and the rooflines of one workload.  Returns the dict of its bench line."""
    import torch
    from gym_pybullet_adrp_b200 import _lib
    M = WORKLOADS[wl][1]
    S = WORKLOADS[wl][3] // WORKLOADS[wl][4]
    K = steps or args.steps
    ctrl = WORKLOADS[wl][5] == _abi.ENV_CTRL
    obs_columns = getattr(args, "obs_columns", 0) if ctrl else 0
    h = _lib.Handle(make_cfg(wl, N, device=local, env_offset=rank * N, obs_columns=obs_columns))
    pool = 4
    acts_np = make_actions(wl, N, M, h.A, 5 + rank, pool)
    acts_dev = [torch.as_tensor(a, device=h.device) for a in acts_np]
    state0 = torch.as_tensor(initial_state(wl, N, M, 5 + rank), device=h.device) if (ctrl and M > 1) else None

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def restart():
        """Open-loop workloads: put the formation back (untimed) so that every block measures flight, not the late
        tumbling of an uncontrolled quadrotor."""
        if state0 is not None:
            h.set_state(state0)
        elif ctrl:
            h.reset()

    restart()
    # ---------------- warm-up, and the block count for >= min_time_s of timed work
    for k in range(max(args.warmup, 3)):
        h.step(acts_dev[k % pool])
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(K):
        h.step(acts_dev[k % pool])
    e1.record()
    torch.cuda.synchronize()
    est = torch.tensor([max(e0.elapsed_time(e1), 1e-3)], device=h.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(est, op=dist.ReduceOp.MAX)        # every rank must time the SAME number of blocks (the block times
    est_ms = float(est.item())                            # are all-reduced element by element below)
    repeats = int(min(max_repeats, max(25, np.ceil(min_time_s * 1e3 / est_ms))))
    restart_every = max(1, 200 // K)                      # blocks between two re-seatings of the formation
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(repeats)]
    restart()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    l0 = h.launch_count
    restarts = 0
    for r in range(repeats):
        if r and r % restart_every == 0:
            restart()                                     # outside the block's event pair
            restarts += 1
        ev[r][0].record()
        for k in range(K):
            h.step(acts_dev[k % pool])
        ev[r][1].record()
    if sampler.nv is not None and not sampler.samples:
        sampler.poll()                                    # the queue is still draining: a sample under load
    barrier()
    sampler.stop_flag = True
    launches = h.launch_count - l0 - restarts
    block_ms = torch.tensor([a.elapsed_time(b) for a, b in ev], device=h.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(block_ms, op=dist.ReduceOp.MAX)   # per block: the slowest rank
    med = float(block_ms.median().item())
    value = world * N * M * K / (med * 1e-3)
    kernel_ms = med / K                                   # K back-to-back launches and nothing else between the two events

    # ---------------- the same through adrp_step_n: step_n_k control steps per launch (state in registers across them)
    acts_n = torch.stack([acts_dev[k % pool] for k in range(step_n_k)])
    restart()
    h.step_n(acts_n)
    reps_n = int(min(max_repeats, max(10, np.ceil(min_time_s * 1e3 / (kernel_ms * step_n_k)))))
    evn = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps_n)]
    restart()
    barrier()
    for r in range(reps_n):
        if r and (r * step_n_k) % 208 < step_n_k:
            restart()
        evn[r][0].record()
        h.step_n(acts_n)
        evn[r][1].record()
    barrier()
    n_ms = torch.tensor([a.elapsed_time(b) for a, b in evn], device=h.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(n_ms, op=dist.ReduceOp.MAX)
    n_med = float(n_ms.median().item())
    del acts_n
    # ... and, for the RL aviaries, with the random policy's U[-1, 1) actions drawn in the kernel (ADRP_ACTIONS_PHILOX: SURVEY 8d's
    # own definition of C2's actions; no action tensor is read)
    rnd_med = None
    if WORKLOADS[wl][5] in (_abi.ENV_HOVER, _abi.ENV_MULTIHOVER):
        evr = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(reps_n)]
        restart()
        h.step_n(None, n_steps=step_n_k)
        barrier()
        for r in range(reps_n):
            if r and (r * step_n_k) % 208 < step_n_k:
                restart()
            evr[r][0].record()
            h.step_n(None, n_steps=step_n_k)
            evr[r][1].record()
        barrier()
        r_ms = torch.tensor([a.elapsed_time(b) for a, b in evr], device=h.device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(r_ms, op=dist.ReduceOp.MAX)
        rnd_med = float(r_ms.median().item())

    # ---------------- kernel time with one event pair around every launch (kept for the record: it adds ~7 us per launch)
    restart()
    h.set_profiling(True)
    for k in range(min(K, 50)):
        h.step(acts_dev[k % pool])
    torch.cuda.synchronize()
    kernel_ms_events = h.step_kernel_ms()
    h.set_profiling(False)

    # ---------------- end to end through the host-buffer call (H2D + kernel + D2H every step)
    e2e_out = None
    if e2e:
        e2e_steps = max(3, min(K, 20))
        restart()
        acts_pin = []
        for a in acts_np:                                 # page-locked host copies of the action pool
            buf = h.pinned_empty(a.shape)
            buf[...] = a
            acts_pin.append(buf)

        def host_rate(hh):
            for k in range(3):
                hh.step_host(acts_pin[k % pool])
            barrier()
            t0 = time.perf_counter()
            for k in range(e2e_steps):
                hh.step_host(acts_pin[k % pool])
            barrier()
            t_e2e = torch.tensor([time.perf_counter() - t0], device=hh.device, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
            return world * N * M * e2e_steps / float(t_e2e.item())

        isz = 4
        d2h = lambda hh: N * M * hh.O * isz + N * isz + 2 * N + N * M   # noqa: E731
        e2e_out = {"value": host_rate(h), "unit": "drone-steps/s",
                   "h2d_bytes_per_step": N * M * h.A * isz, "d2h_bytes_per_step": d2h(h), "steps": e2e_steps}
        if ctrl and h.A == 4:
            # the same call for a caller who reads only the kinematic columns (adrp_config.obs_columns = 12; NOT the
            # reference's 20-float row, so not the headline): 40 % less D2H on a link-bound path
            h12 = _lib.Handle(make_cfg(wl, N, device=local, env_offset=rank * N, obs_columns=12))
            if state0 is not None:
                h12.set_state(state0)
            e2e_out["kinematic_columns_only"] = {"value": host_rate(h12), "obs_columns": 12, "d2h_bytes_per_step": d2h(h12)}
            h12.close()
            restart()

    stats = torch.stack([h.obs[:, :, 2].double().mean(), h.reward.double().mean()])
    O, A = h.O, h.A
    h.close()

    peak = float((peaks or {}).get("hbm_gbs", 6650.0))
    bpd = algorithmic_bytes_per_drone_step(wl, O, A, M)
    achieved = bpd * N * M / (kernel_ms * 1e-3) / 1e9
    fps = algorithmic_flops_per_substep(wl)
    fps_survey = reading_estimate_flops_per_substep(wl)
    fps_exec = executed_flops_per_substep(wl, N)
    per_s = S * N * M / (kernel_ms * 1e-3) / 1e12         # substep-updates per second / 1e12
    fp32 = {"achieved": fps * per_s, "peak": fp32_peak, "unit": "TFLOP/s", "frac": fps * per_s / fp32_peak if fp32_peak else None,
            "algorithmic_flops_per_substep_update": fps,
            "sfu_ops_per_substep_update": MECHANICAL_OPS[wl][1] + MECHANICAL_OPS[wl][3] / S,
            # all three numerators and both denominators, so that every reading of "fraction of the FP32 roofline" can be
            # recomputed from the line: mechanical = every arithmetic statement of the oracle (oracle/flopcount.py); survey =
            # SURVEY 8(d)'s reading estimate; executed = FFMA x 2 + FMUL + FADD of the kernel's own SASS (committed histogram)
            "numerators": {"flops_mechanical": fps, "flops_survey": fps_survey, "flops_executed": fps_exec},
            "denominators": {"peak_measured_fma_chain": fp32_peak, "peak_nominal_sms_x_128_x_2_x_clock": fp32_nominal},
            "frac_by": {f"{n}/{d}": (v * per_s / p if (v and p) else None)
                        for n, v in (("mechanical", fps), ("survey", fps_survey), ("executed", fps_exec))
                        for d, p in (("measured", fp32_peak), ("nominal", fp32_nominal))},
            "count": "mechanical (oracle/flopcount.py)",
            "peak_source": "measured live: adrp_fp32_peak_tflops (16 FMA chains per thread)"}
    return {
        "value": value, "ms_per_step": med / K, "steps": K, "repeats": repeats, "timed_region_ms": float(block_ms.sum().item()),
        "block_ms": {"median": med, "min": float(block_ms.min().item()), "max": float(block_ms.max().item())},
        "substep_updates_per_s": value * S,
        "step_n": {"value": world * N * M * step_n_k / (n_med * 1e-3), "control_steps_per_launch": step_n_k,
                   "ms_per_step": n_med / step_n_k, "launches": reps_n, "substep_updates_per_s": world * N * M * step_n_k * S / (n_med * 1e-3),
                   "random_policy_value": None if rnd_med is None else world * N * M * step_n_k / (rnd_med * 1e-3)},
        "config": {"workload": WORKLOADS[wl][7] + (" [obs_columns=12: NOT the reference's row]" if O == 12 and ctrl else ""),
                   "envs_per_gpu": N, "drones_per_env": M, "substeps_per_step": S,
                   "l2": "working set per step (state+action+obs) exceeds the 126 MB L2" if N * M * bpd > 126e6
                   else "working set fits L2 (small workload)", "action_pool": pool},
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": ncu_traffic_bytes(wl, N), "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback",
                     "kernel_ms": kernel_ms, "kernel_ms_event_pairs": kernel_ms_events, "algorithmic_bytes_per_drone_step": bpd},
        "fp32_roofline": fp32, "e2e": e2e_out, "gpu_launches": launches, "clocks": sampler.result(),
        "episode_stats_local": [float(stats[0]), float(stats[1])],
    }


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default=os.environ.get("ADRP_BENCH_WORKLOAD", "c5"), choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="weak: --envs-per-gpu (default: the workload's) on every GPU; strong: --total-drones split over the GPUs")
    ap.add_argument("--total-drones", type=int, default=1 << 20, help="strong scaling: drones of the whole job (SURVEY 8d: 2^20, 2^23)")
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--obs-columns", type=int, default=0, choices=[0, 12, 20],
                    help="diagnostic: CtrlAviary-kind workloads with adrp_config.obs_columns (12 = kinematic columns only; not the "
                         "reference's row, so not a headline figure)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-workloads", action="store_true", help="skip the other configs (C1 .. C5M) of the default N = 1 run")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from gym_pybullet_adrp_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import datetime
        dist.init_process_group("nccl", device_id=torch.device("cuda", local), timeout=datetime.timedelta(seconds=180))
    wl = args.workload
    M = WORKLOADS[wl][1]
    if args.scaling == "strong":
        N = max(1, args.total_drones // (M * world))      # total work fixed: every GPU gets 1 / world of the envs
    else:
        N = args.envs_per_gpu or WORKLOADS[wl][0]
    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    fp32_peak = _lib.fp32_peak_tflops(local)
    prop = torch.cuda.get_device_properties(local)
    sm_max_mhz = ClockSampler(local).max_mhz or peaks.get("sm_max_mhz")
    fp32_nominal = prop.multi_processor_count * 128 * 2 * float(sm_max_mhz) * 1e6 / 1e12 if sm_max_mhz else None
    common = dict(rank=rank, world=world, local=local, dist=dist, peaks=peaks, fp32_peak=fp32_peak, fp32_nominal=fp32_nominal)

    parity = parity_spot_check(wl, local) if rank == 0 else None
    res = measure(wl, N, args, **common)

    # ---------------- optional episode statistics reduce (the only collective; off the step path)
    stats = torch.tensor(res.pop("episode_stats_local"), device=torch.device("cuda", local), dtype=torch.float64)
    if world > 1:
        dist.all_reduce(stats)
        stats /= world

    if rank == 0:
        S = WORKLOADS[wl][3] // WORKLOADS[wl][4]
        line = {
            "metric": "drone_steps_per_s", "value": res["value"], "unit": "drone-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": res["ms_per_step"], "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": dict(res["config"], total_drones=world * N * M,
                           timing="median of `repeats` back-to-back blocks of exactly `steps` control steps, one CUDA event pair per "
                                  "block on the launching stream, max over ranks per block"),
            "repeats": res["repeats"], "timed_region_ms": res["timed_region_ms"], "block_ms": res["block_ms"],
            "substep_updates_per_s": res["value"] * S, "step_n": res["step_n"],
            "roofline": res["roofline"], "fp32_roofline": res["fp32_roofline"], "e2e": res["e2e"],
            "gpu_launches": res["gpu_launches"], "clocks": res["clocks"], "parity": parity,
            "episode_stats": {"mean_z": float(stats[0]), "mean_reward": float(stats[1])},
        }
        if world == 1 and not args.no_workloads and wl == "c5" and args.scaling == "weak" and not args.envs_per_gpu:
            # every other BASELINE config and the north_star's target configuration, each >= 0.3 s of timed work
            others = {}
            for w in ("c1", "c2", "c3", "c4", "c5m"):
                p = parity_spot_check(w, local)
                r = measure(w, WORKLOADS[w][0], args, steps=max(args.steps, 50), **common)
                r.pop("episode_stats_local")
                r["parity"] = p
                if not args.no_cpu_baseline:               # the CPU port beside it: a bounded sample, ~1 s of all host cores
                    threads = len(os.sched_getaffinity(0))
                    n_s = cpu_sample_envs(w)
                    _, d1 = cpu_port_rate(w, n_s, 2, threads)
                    k_cpu = int(max(2, min(20000, 1.0 / max(d1 / 2, 1e-6))))
                    rate, dt = cpu_port_rate(w, n_s, k_cpu, threads)
                    r["cpu_baseline"] = {"value": rate, "unit": "drone-steps/s", "cores": min(threads, n_s), "kind": "port",
                                         "sample": f"{n_s} envs x {WORKLOADS[w][1]} drones x {k_cpu} control steps ({dt:.1f} s), fp64, "
                                                   f"OpenMP threads of one process over envs"}
                others[w] = r
            line["workloads"] = others
        if world == 1 and not args.no_cpu_baseline:
            threads = len(os.sched_getaffinity(0))
            n_sample, full = cpu_envs_for_budget(wl, 8, threads, 20.0)
            r1, d1 = cpu_port_rate(wl, n_sample, 2, threads)
            steps_cpu = int(max(2, min(400, 12.0 / (d1 / 2))))           # ~12 s of CPU work
            rate, dt = cpu_port_rate(wl, n_sample, steps_cpu, threads)
            line["cpu_baseline"] = {"value": rate, "unit": "drone-steps/s", "cores": threads, "kind": "port",
                                    "sample": f"{'the full workload: ' if full else ''}{n_sample} envs x {M} drones x {steps_cpu} "
                                              f"control steps ({dt:.1f} s), fp64, OpenMP threads of one process over envs"}
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
