#!/bin/bash
# Runs ON THE GPU BOX: one build, an environment switch off / on, over the workloads that fit one wave (and C5 as the control).
#   gpurun -- tools/ab_env.sh ADRP_PDL_ONEWAVE
cd "${GRAFT_REPO_ROOT:-/root/repo}"
VAR=${1:-ADRP_PDL_ONEWAVE}
for rep in 1 2; do
for val in 0 1; do
  for wl in c2 c3 c1 c5; do
    env $VAR=$val python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$VAR=$val $wl %.4e step_n %.4e'%(d['value'],d['step_n']['value']))"
  done
done
done
