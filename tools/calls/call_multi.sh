#!/bin/bash
# Multi-GPU call (round 2): N = number of GPUs of the box.  configs[4] gather check (N ranks vs 1 GPU, bit-identical) + full-map bench
# (strong scaling) + the default weak-scaling bench.   usage: bash tools/calls/call_multi.sh N tag
N=${1:-2}; tag=${2:-r02d}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then
  timeout 600 $TR --master-port 29511 tools/fullmap_check.py 4 30 8 > gpurun_out/${tag}_fullmap_check_${N}gpu.txt 2>&1; echo "fullmap_check rc=$?"
  grep "\[fullmap\]" gpurun_out/${tag}_fullmap_check_${N}gpu.txt | cut -c1-400
fi
timeout 600 $TR --master-port 29512 bench.py --gpus $N --workload fullmap --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_bench_fullmap_${N}gpu.json 2> gpurun_out/${tag}_bench_fullmap_${N}gpu.err; echo "fullmap bench rc=$?"
head -c 260 gpurun_out/${tag}_bench_fullmap_${N}gpu.json; echo
timeout 400 $TR --master-port 29513 bench.py --gpus $N --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/${tag}_bench_${N}gpu.json 2> gpurun_out/${tag}_bench_${N}gpu.err; echo "bench rc=$?"
head -c 260 gpurun_out/${tag}_bench_${N}gpu.json; echo
