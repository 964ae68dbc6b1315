#!/bin/bash
# Runs ON THE GPU BOX (gpurun -- tools/quick_check.sh c5 c3 ...): the -m gpu tests, then a short bench line per workload.
cd "${GRAFT_REPO_ROOT:-/root/repo}"
python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for wl in "$@"; do python bench.py --workload $wl --steps 200 --warmup 20 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
print('$wl', 'value %.4g' % d['value'], 'ms %.4f' % d['ms_per_step'], 'fp32 %.3f' % d['fp32_roofline']['frac'], 'hbm %.3f' % d['roofline']['frac'], 'e2e %.3g' % d['e2e']['value'])
"; done
