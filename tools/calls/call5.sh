#!/bin/bash
# GPU call 5 (round 2): split-row self-attention (8 softmax warps per stream)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py -q -s -p no:cacheprovider > gpurun_out/r02f_attn_tests.txt 2>&1; rc=$?
tail -3 gpurun_out/r02f_attn_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02f_attn_tests.txt | head
(for cfg in 0 32 33 34 22 23; do echo "== TOPO_ATTN_SPLIT=$cfg"; TOPO_ATTN_SPLIT=$cfg timeout 120 python tools/attn_bench.py 38 | grep "N4096"; done) > gpurun_out/r02f_attn_bench.txt 2>&1; cat gpurun_out/r02f_attn_bench.txt
if [ $rc -ne 0 ]; then export TOPO_ATTN_SPLIT=0; echo "SPLIT DISABLED"; fi
timeout 500 python -m pytest tests/test_gpu_e2e.py tests/test_gpu_fullsize.py -q -s -p no:cacheprovider > gpurun_out/r02f_e2e_tests.txt 2>&1; tail -2 gpurun_out/r02f_e2e_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02f_e2e_tests.txt | head
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02f_bench_1gpu.json 2> gpurun_out/r02f_bench_1gpu.err; tail -c 300 gpurun_out/r02f_bench_1gpu.json; echo
TOPO_ATTN_SPLIT=0 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02f_bench_1gpu_split0.json 2> gpurun_out/r02f_bench_1gpu_split0.err; head -c 200 gpurun_out/r02f_bench_1gpu_split0.json; echo
