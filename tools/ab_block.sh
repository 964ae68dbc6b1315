#!/bin/bash
# Runs ON THE GPU BOX: the block size of the multi-wave workloads (ADRP_BLOCK: default = 128 for grids longer than one wave).
cd "${GRAFT_REPO_ROOT:-/root/repo}"
for blk in 0 64; do
  for wl in ${@:-c5 c4}; do
    ADRP_BLOCK=$blk python bench.py --workload $wl --steps 100 --warmup 10 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('ADRP_BLOCK=$blk $wl %.4e step_n %.4e'%(d['value'],d['step_n']['value']))"
  done
done
