#!/bin/bash
# Runs ON THE GPU BOX (gpurun --gpus N -- tools/scale_run.sh TAG N): weak- and strong-scaling bench lines of C5 at N GPUs
# and the raw pinned-copy ceiling (tools/pcie_probe.py) at the same rank count, into gpurun_out/TAG_*.
TAG=${1:-rXX}; N=${2:-1}
cd "${GRAFT_REPO_ROOT:-/root/repo}"
mkdir -p gpurun_out
run() {  # run <output stem> <script> <args...>
  local out=$1; shift
  if [ "$N" -gt 1 ]; then
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 "$@" > gpurun_out/${TAG}_${out}_${N}gpu.json 2> gpurun_out/${TAG}_${out}_${N}gpu.err
  else
    timeout 300 python "$@" > gpurun_out/${TAG}_${out}_${N}gpu.json 2> gpurun_out/${TAG}_${out}_${N}gpu.err
  fi
}
run scale_weak_c5 bench.py --gpus $N --steps 20 --warmup 5 --no-workloads --no-cpu-baseline
run scale_strong8m_c5 bench.py --gpus $N --steps 20 --warmup 5 --no-workloads --no-cpu-baseline --scaling strong --total-drones 8388608
[ "$N" -eq 8 ] && run scale_strong1m_c5 bench.py --gpus $N --steps 20 --warmup 5 --no-workloads --no-cpu-baseline --scaling strong --total-drones 1048576
[ "$N" -eq 8 ] && run scale_weak_c5m bench.py --gpus $N --steps 20 --warmup 5 --no-workloads --no-cpu-baseline --workload c5m
run pcie_probe_c5 tools/pcie_probe.py
for f in gpurun_out/${TAG}_*_${N}gpu.json; do echo "$f: $(head -c 160 $f)"; done
