#!/bin/bash
# Runs ON THE GPU BOX: product build vs a variant library (tools/variant_build.sh) on the step-kernel workloads.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
for lib in "" gym_pybullet_adrp_b200/variants/${1:-step_minb6}/libadrp_b200.so; do
  for wl in c4 c2 c5; do
    ADRP_LIB_PATH=$lib python bench.py --workload $wl --steps 50 --warmup 5 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${lib:-product} $wl %.4e step_n %.4e'%(d['value'],d['step_n']['value']))"
  done
done
