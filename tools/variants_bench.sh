#!/bin/bash
# Runs ON THE GPU BOX: the race kernel's two register budgets (ADRP_RACE_DENSE=0: 168 registers, =1: 128) on C3 at its own
# batch and at 8x, and on C5M.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
for dense in 0 1; do
  for spec in c3:16384 c3:131072 c5m:131072 c5m:16384; do
    wl=${spec%%:*}; n=${spec##*:}
    ADRP_RACE_DENSE=$dense python bench.py --workload $wl --envs-per-gpu $n --steps 50 --warmup 5 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('dense=$dense $wl envs=$n %.4e step_n %.4e'%(d['value'],d['step_n']['value']))"
  done
done
