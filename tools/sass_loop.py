#!/usr/bin/env python
"""Static opcode histogram of the innermost backward-branch loop(s) of one kernel's SASS (no GPU needed).
    tools/sass_loop.py <object file> <mangled-name substring> [--all]
Prints every backward branch (loop) with its instruction count and opcode mix; conditional regions inside are included."""
import collections
import re
import subprocess
import sys

obj, pat = sys.argv[1], sys.argv[2]
names = subprocess.run(["cuobjdump", "-sass", obj], capture_output=True, text=True).stdout
funcs = re.split(r"\n\s*Function : ", names)
for fn in funcs[1:]:
    name = fn.split("\n", 1)[0].strip()
    dem = subprocess.run(["c++filt", name], capture_output=True, text=True).stdout.strip()
    if pat not in dem:
        continue
    ins = [(int(m.group(1), 16), m.group(2).strip()) for m in re.finditer(r"/\*([0-9a-f]{4,5})\*/\s+(.*?);", fn)]
    print(dem[:110], "| total", len(ins))
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            lo = int(m.group(1), 16)
            body = [re.sub(r"^@!?U?P\d\s+", "", x) for b, x in ins if lo <= b <= a]
            c = collections.Counter(x.split()[0].split(".")[0] for x in body)
            print(f"  loop 0x{lo:x}..0x{a:x}: {len(body)} instructions:", ", ".join(f"{k} {v}" for k, v in c.most_common(14)))
