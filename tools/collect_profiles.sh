#!/bin/bash
# Runs ON THE GPU BOX (gpurun -- tools/collect_profiles.sh TAG): bench lines, ncu launch lists and one
# `ncu --set full` capture of the step kernel per workload, all into gpurun_out/TAG_*.  Numbers printed under
# ncu are never bench values: the bench lines come from the plain runs.  tools/summarise_profiles.py TAG then
# turns the captures into the text summaries committed under profiles/.
TAG=${1:-rXX}
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
for wl in c5 c5m c3 c4 c2; do
  steps=200; [ $wl = c5m ] && steps=100
  python bench.py --workload $wl --steps $steps --warmup 20 > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err
done
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/${TAG}_bench_reference_c5.json 2>/dev/null
for wl in c5 c5m c3; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv \
      python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
  ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 4 -c 1 -f -o gpurun_out/${TAG}_ncu_$wl \
      python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline > /dev/null 2>&1
done
ls -la gpurun_out | grep ${TAG}_
