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
  python bench.py --workload $wl --steps $steps --warmup 20 --no-workloads > gpurun_out/${TAG}_bench_$wl.json 2> gpurun_out/${TAG}_bench_$wl.err
done
python bench.py --impl reference --steps 20 --warmup 2 > gpurun_out/${TAG}_bench_reference_c5.json 2>/dev/null
for wl in ${NCU_WORKLOADS:-c5 c5m c3}; do
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/${TAG}_launches_$wl.csv \
      python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline --no-workloads > /dev/null 2>&1
  # (-s 12: past the 4 launches of the 64-env parity spot check, the warm-up and the first timed steps of the real handle)
  ncu --set full --clock-control none --import-source on -k regex:step_kernel -s 12 -c 1 -f -o gpurun_out/${TAG}_ncu_$wl \
      python bench.py --workload $wl --steps 6 --warmup 3 --no-cpu-baseline --no-workloads > /dev/null 2>&1
done
ls -la gpurun_out | grep ${TAG}_
# size sweep of the headline workload (D = 1, 2, 4, 8 x 2^20 drones on one GPU) and C3 at 8x its batch
echo "workload,envs_per_gpu,drones,drone_steps_per_s,ms_per_step,kernel_ms,fp32_frac,hbm_frac" > gpurun_out/${TAG}_sweep_size.csv
for spec in c5:16384 c5:32768 c5:65536 c5:131072 c5:262144 c5:524288 c5:1048576 c5m:32768 c5m:131072 c5m:524288 c3:16384 c3:131072; do
  wl=${spec%%:*}; n=${spec##*:}
  python bench.py --workload $wl --envs-per-gpu $n --steps 60 --warmup 6 --no-cpu-baseline --no-workloads 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read())
c=d['config']
print('$wl,$n,%d,%.4g,%.4f,%.4f,%.3f,%.3f' % (c['envs_per_gpu']*c['drones_per_env'], d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['fp32_roofline']['frac'], d['roofline']['frac']))
" >> gpurun_out/${TAG}_sweep_size.csv
done
cat gpurun_out/${TAG}_sweep_size.csv
