#!/usr/bin/env python
"""Raw host<->device copy ceiling for the byte counts one adrp_step_host call of a bench workload moves -- no kernel, no
library: page-locked host buffers, one cudaMemcpyAsync per copy (torch ``copy_(non_blocking=True)``), every rank at once.

    python tools/pcie_probe.py [--workload c5] [--iters 30]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 tools/pcie_probe.py

Per rank and per step: H2D of the action tensor, D2H of observation + reward + flags (bench.py's h2d / d2h_bytes_per_step).
Two schedules: "serial" (H2D then D2H on one stream: what a naive step does) and "duplex" (the step cut into chunks over
two streams so that chunk c+1's upload overlaps chunk c's download: what adrp_step_host does).  Wall clock between
barriers, max over ranks; rank 0 prints ONE JSON line with the aggregate GB/s and the drone-steps/s the copies alone allow.
bench.py's e2e divided by `duplex.drone_steps_per_s_ceiling` of the same rank count is the fraction of the link that
adrp_step_host reaches (profiles/README.md)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c5")
    ap.add_argument("--iters", type=int, default=30)
    ap.add_argument("--chunks", type=int, default=8)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    import bench
    from gym_pybullet_adrp_b200 import _abi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    wl = args.workload
    N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
    A = _abi.ACTION_DIM[bench.WORKLOADS[wl][6]]
    cfg = bench.make_cfg(wl, 2)
    O = 20 if bench.WORKLOADS[wl][5] == _abi.ENV_CTRL else None
    if O is None:                                          # RL / race observation widths come from the library
        from gym_pybullet_adrp_b200 import _lib
        h = _lib.Handle(cfg)
        O = h.O
        h.close()
    h2d = N * M * A * 4
    d2h = N * M * O * 4 + N * 4 + 2 * N + N * M
    dev = torch.device("cuda", local)
    hin = torch.empty(h2d, dtype=torch.uint8, pin_memory=True)
    hout = torch.empty(d2h, dtype=torch.uint8, pin_memory=True)
    hin.fill_(1); hout.fill_(0)                            # first touch on this rank's cores
    din = torch.empty(h2d, dtype=torch.uint8, device=dev)
    dout = torch.ones(d2h, dtype=torch.uint8, device=dev)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def serial():
        with torch.cuda.stream(s1):
            din.copy_(hin, non_blocking=True)
            hout.copy_(dout, non_blocking=True)
        s1.synchronize()

    def duplex():
        c = args.chunks
        for k in range(c):
            a0, a1 = k * h2d // c, (k + 1) * h2d // c
            b0, b1 = k * d2h // c, (k + 1) * d2h // c
            with torch.cuda.stream(s1 if k % 2 == 0 else s2):
                din[a0:a1].copy_(hin[a0:a1], non_blocking=True)
                hout[b0:b1].copy_(dout[b0:b1], non_blocking=True)
        s1.synchronize(); s2.synchronize()

    out = {}
    for name, fn in (("serial", serial), ("duplex", duplex)):
        for _ in range(3):
            fn()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.iters):
            fn()
        barrier()
        t = torch.tensor([time.perf_counter() - t0], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        sec = float(t.item()) / args.iters
        out[name] = {"ms_per_step": 1e3 * sec, "gbs_per_rank": (h2d + d2h) / sec / 1e9,
                     "gbs_aggregate": world * (h2d + d2h) / sec / 1e9, "drone_steps_per_s_ceiling": world * N * M / sec}
    if rank == 0:
        print(json.dumps({"probe": "pinned cudaMemcpyAsync, no kernel", "workload": wl, "n_gpus": world,
                          "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "iters": args.iters, "chunks": args.chunks,
                          "host_cores": len(os.sched_getaffinity(0)), **out}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
