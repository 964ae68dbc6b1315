#!/bin/bash
# Runs ON THE GPU BOX: the product build against a variant library (tools/variant_build.sh NAME flags) on the given workloads, twice.
#   gpurun -- tools/ab_variant.sh NAME [workloads ...]
cd "${GRAFT_REPO_ROOT:-/root/repo}"
NAME=$1; shift
WLS=${@:-c5 c4}
for rep in 1 2; do
for lib in "" gym_pybullet_adrp_b200/variants/$NAME/libadrp_b200.so; do
  for wl in $WLS; do
    ADRP_LIB_PATH=$lib python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('${lib:-product} $wl %.4e step_n %.4e parity %s %.2e'%(d['value'],d['step_n']['value'],d['parity']['ok'],d['parity']['max_rel_err']))"
  done
done
done
