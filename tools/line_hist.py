#!/usr/bin/env python
"""Warp instructions executed per SOURCE line of one kernel launch.

    python tools/line_hist.py REPORT.ncu-rep OBJECT.o KERNEL_SUBSTRING [--top N] [--inclusive]

Joins (a) the per-SASS-instruction "Instructions Executed" column of an `ncu --set full --import-source on`
report (`ncu -i REPORT --page source --csv --print-source sass`) with (b) the line table of the built object
(`cuobjdump -xelf` + `nvdisasm -gi`, which needs -lineinfo), instruction by instruction in program order; the
opcodes of the two listings are compared as a check.  `--inclusive` charges every frame of the inline chain
(leaf first, callers after), i.e. a line of the kernel body is charged with everything inlined into it.
"""
from __future__ import annotations

import argparse
import collections
import csv
import glob
import io
import os
import re
import subprocess
import tempfile


def _norm(name):
    """ncu prints template arguments differently per page ("(int)8" / "8", with or without the namespace)."""
    return re.sub(r"\((?:int|bool)\)|adrp::|\s", "", name)


def sass_counts(report, kernel_sub):
    out = subprocess.run(["ncu", "-i", report, "--page", "source", "--csv", "--print-source", "sass"],
                         capture_output=True, text=True, check=True).stdout
    rows, name, header = [], None, None
    for rec in csv.reader(io.StringIO(out)):
        if not rec:
            continue
        if rec[0] == "Kernel Name":
            name, header = rec[1], None
            continue
        if rec[0] == "Address":
            header = {k: i for i, k in enumerate(rec)}
            continue
        if header and name and _norm(kernel_sub) in _norm(name):
            rows.append((rec[header["Source"]].strip(), int(rec[header["Instructions Executed"]] or 0),
                         {k: int(rec[i] or 0) for k, i in header.items() if k.startswith("stall_") and "Not Issued" not in k}))
    return name, rows


def line_table(obj, kernel_sub):
    """obj: one object file, or a directory whose *.o are all searched (one translation unit per kernel family)."""
    objs = sorted(glob.glob(os.path.join(obj, "*.o"))) if os.path.isdir(obj) else [obj]
    txt = ""
    for o in objs:
        tmp = tempfile.mkdtemp()
        subprocess.run(["cuobjdump", "-xelf", "all", os.path.abspath(o)], cwd=tmp, capture_output=True)
        for cubin in glob.glob(os.path.join(tmp, "*.cubin")):
            txt += subprocess.run(["nvdisasm", "-gi", "-c", cubin], capture_output=True, text=True, check=True).stdout
    insts, frames, pending, inside, sections = [], [], [], False, {}
    for line in txt.splitlines():
        if line.startswith(".text."):
            inside = kernel_sub in line and "$" not in line
            frames, pending = [], []
            insts = sections.setdefault(line, []) if inside else []
            continue
        if line.startswith("//-----"):
            inside = False
            continue
        if not inside:
            continue
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', line)
        if m:
            pending.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(.*?) ;", line)
        if m:
            if pending:
                frames, pending = pending, []
            insts.append((m.group(1).strip(), list(frames)))
    return sections


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("report"), ap.add_argument("obj"), ap.add_argument("kernel")
    ap.add_argument("--top", type=int, default=60)
    ap.add_argument("--inclusive", action="store_true")
    ap.add_argument("--stalls", action="store_true", help="also print the sampled stall reasons per line")
    a = ap.parse_args()
    name, rows = sass_counts(a.report, a.kernel)
    insts = line_table(a.obj, re.sub(r"[^A-Za-z0-9_]", "", a.kernel.split("<")[0].split("::")[-1]))
    # template instances share the base name: take the section whose opcode sequence is the report's
    ops = [s.split()[0] if not s.startswith("@") else s.split()[1] for s, _, _ in rows]
    # (the report's listing ends with the section's padding: a trailing BRA + NOPs the disassembly does not print)
    match = [v for v in insts.values() if 0 <= len(rows) - len(v) < 32 and
             [(d.split()[1] if d.startswith("@") else d.split()[0]) for d, _ in v] == ops[:len(v)]]
    if not match:
        raise SystemExit(f"no section of {a.obj} has the report's {len(rows)} instructions: sizes {[len(v) for v in insts.values()]} "
                         "(object and report from different builds?)")
    insts = match[0]
    mism = sum(1 for (s, _, _), (d, _) in zip(rows, insts) if s.split()[0:2] != d.split()[0:2] and s.split()[0] != d.split()[0])
    total = sum(c for _, c, _ in rows)
    hist, files, stalls = collections.Counter(), collections.Counter(), collections.defaultdict(collections.Counter)
    for (_, c, st), (_, fr) in zip(rows, insts):
        fr = fr or [("?", 0)]
        for f in (fr if a.inclusive else fr[:1]):
            hist[f] += c
            for k, v in st.items():
                stalls[f][k] += v
        files[fr[0][0]] += c
    src = {}
    print("# warp instructions executed per source line (nvdisasm -gi line table joined with the ncu source page%s)"
          % ("; inclusive over the inline chain" if a.inclusive else ""))
    print(f"# {name}")
    print(f"# total warp instructions {total}  (opcode mismatches: {mism})")
    for (f, n), c in hist.most_common(a.top):
        if f not in src:
            base = os.path.abspath(a.obj) if os.path.isdir(a.obj) else os.path.dirname(os.path.abspath(a.obj))
            cand = glob.glob(os.path.join(base, "..", "csrc", f))
            src[f] = open(cand[0]).read().splitlines() if cand else []
        text = src[f][n - 1].strip()[:100] if 0 < n <= len(src[f]) else ""
        extra = ""
        if a.stalls:
            top = [f"{k[6:]}={v}" for k, v in stalls[(f, n)].most_common(3) if v]
            extra = "  [" + " ".join(top) + "]"
        print(f"{100.0 * c / total:6.2f}%  {f}:{n}   {text}{extra}")
    print("# by file (leaf frame)")
    for f, c in files.most_common():
        print(f"{100.0 * c / total:6.2f}%  {f}")


if __name__ == "__main__":
    main()
