mkdir -p gpurun_out
N=${NGPU:-8}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_${N}gpu.json 2> gpurun_out/bench_${N}gpu.err; echo "rc=$?"; cut -c1-260 gpurun_out/bench_${N}gpu.json; tail -2 gpurun_out/bench_${N}gpu.err
