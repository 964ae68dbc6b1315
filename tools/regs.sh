#!/bin/bash
# Register / spill report of the product-build kernel instances (ptxas -v), one line per kernel.
#   tools/regs.sh [translation units without .cu ...]
cd "$(dirname "$0")/../gym_pybullet_adrp_b200/csrc" || exit 1
TUS=${@:-adrp_step_f32_forces adrp_step_f32_plain adrp_race_f32_phy16 adrp_race_f32_phy30}
for f in $TUS; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xptxas -v $ADRP_EXTRA -c -o /tmp/regs_$$.o $f.cu 2>&1 | c++filt | awk '
    /Compiling entry function/ { name=$0; sub(/.*Compiling entry function ./,"",name); sub(/\(adrp::KParams.*/,"",name) }
    /bytes stack frame/ { spill=$0; sub(/^ */,"",spill) }
    /Used [0-9]+ registers/ { match($0,/Used [0-9]+ registers/); r=substr($0,RSTART+5,RLENGTH-15); print r, name, "|", spill }'
done
rm -f /tmp/regs_$$.o
