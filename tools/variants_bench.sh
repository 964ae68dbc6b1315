#!/bin/bash
# Runs ON THE GPU BOX: product build vs a variant library (tools/variant_build.sh NAME flags) on the race workloads with disturbance draws.
#   gpurun -- tools/variants_bench.sh [NAME]      (no NAME: the product build only)
cd "${GRAFT_REPO_ROOT:-/root/repo}"
libs=("")
[ -n "$1" ] && libs+=("gym_pybullet_adrp_b200/variants/$1/libadrp_b200.so")
for lib in "${libs[@]}"; do
  for spec in c3:16384 c3:131072; do
    wl=${spec%%:*}; n=${spec##*:}
    ADRP_LIB_PATH=$lib python bench.py --workload $wl --envs-per-gpu $n --steps 50 --warmup 5 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${lib:-product} $wl envs=$n %.4e step_n %.4e'%(d['value'],d['step_n']['value']))"
  done
done
