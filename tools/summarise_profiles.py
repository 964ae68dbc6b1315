#!/usr/bin/env python
"""Turns the captures of tools/collect_profiles.sh (gpurun_out/TAG_*) into the summaries committed under profiles/:

    TAG_bench_<wl>.json, TAG_launches_<wl>.csv      copied
    TAG_ncu_<wl>_metrics.csv                        selected `--set full` metrics of the step kernel's launch
    TAG_sass_hist_<wl>.txt                          warp-level SASS instructions executed, by opcode
    TAG_line_hist_<wl>.txt                          ... per source line (tools/line_hist.py, leaf frames + inclusive)

    python tools/summarise_profiles.py TAG [--obj gym_pybullet_adrp_b200/build]
"""
from __future__ import annotations

import argparse
import collections
import csv
import glob
import io
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
METRICS = [
    "gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "sm__inst_issued.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "lts__t_sectors_srcunit_tex_op_read.sum", "lts__t_sectors_srcunit_tex_op_write.sum",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__occupancy_limit_registers",
    "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio",
    "smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_st.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_st.sum",
    "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum",
]


def raw_metrics(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, vals = rows[0], rows[1], rows[2]
    col = {k: i for i, k in enumerate(head)}
    return vals[col["Kernel Name"]], {k: (units[i], vals[i]) for k, i in col.items()}


def sass_hist(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True, check=True).stdout
    hist, header = collections.Counter(), None
    for rec in csv.reader(io.StringIO(out)):
        if rec and rec[0] == "Address":
            header = {k: i for i, k in enumerate(rec)}
        elif header and rec and rec[0].startswith("0x"):
            toks = rec[header["Source"]].split()
            op = toks[1] if toks[0].startswith("@") else toks[0]
            hist[".".join(op.split(".")[:2]) if op.startswith("MUFU") else op.split(".")[0]] += int(rec[header["Instructions Executed"]] or 0)
    return hist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("tag")
    ap.add_argument("--obj", default=os.path.join(ROOT, "gym_pybullet_adrp_b200", "build"), help="object file or directory of *.o")
    a = ap.parse_args()
    src, dst = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
    for f in glob.glob(os.path.join(src, f"{a.tag}_bench_*.json")) + glob.glob(os.path.join(src, f"{a.tag}_launches_*.csv")) + glob.glob(os.path.join(src, f"{a.tag}_sweep_*.csv")):
        if os.path.getsize(f):  # (empty: the run failed)
            shutil.copy(f, dst)
    for rep in sorted(glob.glob(os.path.join(src, f"{a.tag}_ncu_*.ncu-rep"))):
        wl = os.path.basename(rep)[len(a.tag) + 5:-8]
        kernel, m = raw_metrics(rep)
        with open(os.path.join(dst, f"{a.tag}_ncu_{wl}_metrics.csv"), "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(["metric", "unit", "value"])
            w.writerow(["kernel", "", kernel])
            for k in METRICS:
                if k in m:
                    w.writerow([k, m[k][0], m[k][1]])
        hist = sass_hist(rep)
        total = sum(hist.values())
        with open(os.path.join(dst, f"{a.tag}_sass_hist_{wl}.txt"), "w") as f:
            f.write(f"# warp-level SASS instructions executed by one launch, by opcode (gpurun_out/{os.path.basename(rep)}, ncu source page)\n# total {total}\n")
            for op, c in hist.most_common():
                if c:
                    f.write(f"{op:<16}{c:>12}  {100.0 * c / total:4.1f}%\n")
        base = kernel.split("(")[0].split()[-1]           # adrp::step_kernel<float, ...>
        name = kernel[kernel.index(base.split("<")[0].split("::")[-1]):].split(">(")[0] + ">"
        with open(os.path.join(dst, f"{a.tag}_line_hist_{wl}.txt"), "w") as f:
            for extra in ([], ["--inclusive"]):
                r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "line_hist.py"), rep, a.obj, name, "--top", "60"] + extra,
                                   capture_output=True, text=True)
                f.write(r.stdout if r.returncode == 0 else f"# line_hist failed: {r.stderr}\n")
                f.write("\n")
        print(wl, kernel[:60], "instructions", total)


if __name__ == "__main__":
    main()
