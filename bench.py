#!/usr/bin/env python
"""bench.py -- drone-steps/s of the fused control-step kernel (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c5|c2|c4] [--impl reference]

One "step" = one control step (S fused physics substeps) of every drone of the workload.  Default
workload C5 (BASELINE.json configs[4], the scaling sweep the metric is quoted on): 2^20 drones per GPU,
131072 envs x 8 CF2X drones, DYN + ground effect + drag + downwash, 240/30 Hz (S = 8), random RPM
actions, CtrlAviary observation.  Weak scaling: every rank owns its own env range, no collective on the
step path.  See DESIGN.md section "Measurement" for the roofline byte count.
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


def make_cfg(wl, n_envs, device=0, env_offset=0, precision=_abi.F32):
    N, M, phy, pyb, ctrl, kind, act, _ = WORKLOADS[wl]
    if kind == _abi.ENV_RACE:
        from gym_pybullet_adrp_b200.utils.enums import RaceMode
        return _abi.make_race_config(race_rc(wl), num_envs=n_envs, num_drones=M, physics=phy, pyb_freq=pyb, ctrl_freq=ctrl,
                                     racemode=RaceMode.COMPETE if wl == "c3" else RaceMode.COMPARE, precision=precision,
                                     device=device, seed=1337, env_offset=env_offset, auto_reset=(wl == "c3"))
    xyz = formation_xyz(M) if M > 1 else None
    return _abi.make_config(num_envs=n_envs, num_drones=M, drone_model=DroneModel.CF2X, physics=phy, pyb_freq=pyb,
                            ctrl_freq=ctrl, action_type=act, env_kind=kind, precision=precision, device=device,
                            initial_xyzs=xyz, auto_reset=(kind != _abi.ENV_CTRL), seed=5, env_offset=env_offset)


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
    per_drone = 4 * (state_rd + A + state_wr + O + ring + extra) + 1
    per_env = 4 + 2 + 4 + 4                               # reward, two flag bytes, step counter read+write
    return per_drone + per_env / M


# algorithmic work of the restated path.  MECHANICAL COUNT of the frozen oracle (oracle/flopcount.py:
# the oracle's own statements compiled with counting arithmetic types, run on these workloads):
#   (flops per substep-update, SFU ops per substep-update, flops per control step, SFU ops per control step)
# flops = add + sub + mul + compare (a multiply-add counts 2); SFU = div, sqrt, sin/cos, atan2, asin, exp, log.
# tests/test_oracle_golden.py::test_flop_table_is_the_mechanical_count keeps this table equal to the counter's output.
MECHANICAL_OPS = {
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
    return fl


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
    cfg = make_cfg(wl, N, precision=_abi.F64)           # the reference computes in numpy float64
    env = OracleEnv(cfg)
    if WORKLOADS[wl][5] == _abi.ENV_CTRL:
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
    return max(256, min(65536, (1 << 19) // (M * S)))    # ~0.5 M substep-updates per control step of the sample


def run_reference(args):
    """--impl reference: the reference's own CPU arithmetic (oracle port; the reference itself needs
    PyBullet, which is not installable here or on the GPU box) on all host cores, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = len(os.sched_getaffinity(0))
    wl = args.workload
    M, S = WORKLOADS[wl][1], WORKLOADS[wl][3] // WORKLOADS[wl][4]
    n_sample = cpu_sample_envs(wl)
    for _ in range(max(args.warmup, 1)):
        cpu_port_rate(wl, n_sample, 1, threads)
    rate, dt = cpu_port_rate(wl, n_sample, args.steps, threads)
    sample = f"{n_sample} envs x {M} drones x {args.steps} control steps of the workload, OpenMP over envs"
    print(json.dumps({
        "impl": "reference", "metric": "drone_steps_per_s", "value": rate, "unit": "drone-steps/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOADS[wl][7], "sample": sample},
        "cpu_baseline": {"value": rate, "unit": "drone-steps/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": rate, "unit": "drone-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "substep_updates_per_s": rate * S}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--workload", default=os.environ.get("ADRP_BENCH_WORKLOAD", "c5"), choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--envs-per-gpu", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = args.workload
    N = args.envs_per_gpu or WORKLOADS[wl][0]
    M = WORKLOADS[wl][1]
    S = WORKLOADS[wl][3] // WORKLOADS[wl][4]
    h = _lib.Handle(make_cfg(wl, N, device=local, env_offset=rank * N))
    if WORKLOADS[wl][5] == _abi.ENV_CTRL:
        h.set_state(initial_state(wl, N, M, 5 + rank))
    pool = 4
    acts_np = make_actions(wl, N, M, h.A, 5 + rank, pool)
    acts_dev = [torch.as_tensor(a, device=h.device) for a in acts_np]

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def restart():
        """Open-loop workloads: put the formation back (untimed) so that every phase measures flight, not the
        late tumbling of an uncontrolled quadrotor."""
        if WORKLOADS[wl][5] == _abi.ENV_CTRL:
            h.set_state(initial_state(wl, N, M, 5 + rank))

    # ---------------- device-resident throughput ("value")
    for k in range(max(args.warmup, 3)):
        h.step(acts_dev[k % pool])
    restart()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    l0 = h.launch_count
    e0.record()
    for k in range(args.steps):
        h.step(acts_dev[k % pool])
    e1.record()
    if sampler.nv is not None and not sampler.samples:
        sampler.poll()                                   # the queue is still draining: a sample under load
    barrier()
    sampler.stop_flag = True
    launches = h.launch_count - l0
    ms = torch.tensor([e0.elapsed_time(e1)], device=h.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms_total = float(ms.item())
    value = world * N * M * args.steps / (ms_total * 1e-3)

    # ---------------- kernel-only time for the roofline (CUDA events on the launching stream)
    restart()
    h.set_profiling(True)
    for k in range(min(args.steps, 50)):
        h.step(acts_dev[k % pool])
    torch.cuda.synchronize()
    kernel_ms_events = h.step_kernel_ms()
    h.set_profiling(False)
    # The timed region above is `launches` back-to-back launches of the step kernel and nothing else on the stream, bracketed
    # by two CUDA events: its mean is the kernel's average launch duration under the conditions the throughput is quoted on
    # (launch gaps included, so it is an upper bound of the pure kernel time).  The per-launch event pairs of the profiling
    # pass put an event record between consecutive launches, which adds ~7 us per launch and defeats the programmatic
    # dependent launch; that figure is kept as `kernel_ms_event_pairs`.
    kernel_ms = ms_total / launches if launches == args.steps else kernel_ms_events

    # ---------------- end to end through the host-buffer call (H2D + kernel + D2H every step)
    e2e_steps = max(3, min(args.steps, 20))
    restart()
    acts_pin = []
    for a in acts_np:                                    # page-locked host copies of the action pool
        buf = h.pinned_empty(a.shape)
        buf[...] = a
        acts_pin.append(buf)
    for k in range(3):
        h.step_host(acts_pin[k % pool])
    barrier()
    t0 = time.perf_counter()
    for k in range(e2e_steps):
        out = h.step_host(acts_pin[k % pool])
    barrier()
    t_e2e = torch.tensor([time.perf_counter() - t0], device=h.device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t_e2e, op=dist.ReduceOp.MAX)
    e2e_value = world * N * M * e2e_steps / float(t_e2e.item())
    isz = 4
    h2d = N * M * h.A * isz
    d2h = N * M * h.O * isz + N * isz + 2 * N + N * M

    # ---------------- optional episode statistics reduce (the only collective; off the step path)
    stats = torch.stack([h.obs[:, :, 2].double().mean(), h.reward.double().mean()])
    if world > 1:
        dist.all_reduce(stats)
        stats /= world

    if rank == 0:
        peaks = {}
        try:
            with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        bpd = algorithmic_bytes_per_drone_step(wl, h.O, h.A, M)
        achieved = bpd * N * M / (kernel_ms * 1e-3) / 1e9 if kernel_ms > 0 else None
        fps = algorithmic_flops_per_substep(wl)
        fp32_peak = _lib.fp32_peak_tflops(local)
        fp32_ach = fps * S * N * M / (kernel_ms * 1e-3) / 1e12 if kernel_ms > 0 else None
        line = {
            "metric": "drone_steps_per_s", "value": value, "unit": "drone-steps/s", "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_total / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOADS[wl][7], "envs_per_gpu": N, "drones_per_env": M, "substeps_per_step": S,
                       "l2": "working set per step (state+action+obs) exceeds the 126 MB L2" if N * M * bpd > 126e6
                       else "working set fits L2 (small workload)", "action_pool": pool},
            "substep_updates_per_s": value * S,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                         "frac": achieved / peak if achieved else None, "traffic": ncu_traffic_bytes(wl, N),
                         "peak_source": "measured (MEASURED_PEAKS.json)" if peaks else "fallback",
                         "kernel_ms": kernel_ms, "kernel_ms_event_pairs": kernel_ms_events,
                         "algorithmic_bytes_per_drone_step": bpd},
            "fp32_roofline": {"achieved": fp32_ach, "peak": fp32_peak, "unit": "TFLOP/s", "frac": fp32_ach / fp32_peak if fp32_ach else None,
                              "algorithmic_flops_per_substep_update": fps,
                              "sfu_ops_per_substep_update": MECHANICAL_OPS[wl][1] + MECHANICAL_OPS[wl][3] / S,
                              "count": "mechanical (oracle/flopcount.py); SURVEY reading estimate: %d" % reading_estimate_flops_per_substep(wl),
                              "peak_source": "measured live: adrp_fp32_peak_tflops (16 FMA chains per thread)"},
            "e2e": {"value": e2e_value, "unit": "drone-steps/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps},
            "gpu_launches": launches,
            "clocks": sampler.result(),
            "episode_stats": {"mean_z": float(stats[0]), "mean_reward": float(stats[1])},
        }
        if world == 1 and not args.no_cpu_baseline:
            threads = len(os.sched_getaffinity(0))
            n_sample = cpu_sample_envs(wl)
            cpu_port_rate(wl, n_sample, 1, threads)
            # size the sample for ~10-20 s of CPU work
            r1, d1 = cpu_port_rate(wl, n_sample, 4, threads)
            steps_cpu = int(max(4, min(400, 12.0 / (d1 / 4))))
            rate, dt = cpu_port_rate(wl, n_sample, steps_cpu, threads)
            line["cpu_baseline"] = {"value": rate, "unit": "drone-steps/s", "cores": threads, "kind": "port",
                                    "sample": f"{n_sample} envs x {M} drones x {steps_cpu} control steps ({dt:.1f} s), fp64, OpenMP over envs"}
        print(json.dumps(line))
    h.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
