#!/usr/bin/env python
"""Single-step launches vs adrp_step_n (K control steps per launch) on the bench workloads.
    python tools/bench_step_n.py [c2 c3 ...] [--k 16]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from gym_pybullet_adrp_b200 import _abi, _lib  # noqa: E402


def timed(fn, reps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    wls = [a for a in sys.argv[1:] if not a.startswith("--") and not a.isdigit()] or ["c2", "c3", "c4", "c5", "c5m"]
    K = int(sys.argv[sys.argv.index("--k") + 1]) if "--k" in sys.argv else 16
    for wl in wls:
        N, M = bench.WORKLOADS[wl][0], bench.WORKLOADS[wl][1]
        h = _lib.Handle(bench.make_cfg(wl, N))
        ctrl = bench.WORKLOADS[wl][5] == _abi.ENV_CTRL
        def restart():
            if ctrl:
                h.set_state(bench.initial_state(wl, N, M, 5))
        acts = torch.as_tensor(bench.make_actions(wl, N, M, h.A, 5, K), device=h.device)
        restart()
        for k in range(K):
            h.step(acts[k])
        restart()
        single = timed(lambda: [h.step(acts[k]) for k in range(K)], 10) / K
        restart()
        h.step_n(acts)
        restart()
        multi = timed(lambda: h.step_n(acts), 10) / K
        restart()
        big = (torch.empty((K, N, M, h.O), device=h.device), torch.empty((K, N), device=h.device),
               torch.empty((K, N), dtype=torch.uint8, device=h.device), torch.empty((K, N), dtype=torch.uint8, device=h.device),
               torch.empty((K, N, M), dtype=torch.uint8, device=h.device))
        def keep():
            h._check(h.lib.adrp_step_n(h._h, K, 1, 1, acts.data_ptr(), None, *[t.data_ptr() for t in big], h._stream()), "step_n")
        keep()
        restart()
        multi_keep = timed(keep, 10) / K
        print(json.dumps({"workload": wl, "K": K, "ms_per_step_single": single, "ms_per_step_n": multi,
                          "ms_per_step_n_keep_outputs": multi_keep, "drone_steps_per_s_single": N * M / single * 1e3,
                          "drone_steps_per_s_n": N * M / multi * 1e3, "drone_steps_per_s_n_keep": N * M / multi_keep * 1e3}))
        h.close()


if __name__ == "__main__":
    main()
