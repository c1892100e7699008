#!/bin/bash
# GPU call 1 (round 2): attention tests (new K3 kernel) -> xattn A/B -> full gpu test suite -> ablation -> bench -> shape breakdown
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py -q -s -p no:cacheprovider > gpurun_out/r02a_attn_tests.txt 2>&1; rc=$?
tail -3 gpurun_out/r02a_attn_tests.txt
if [ $rc -ne 0 ]; then export TOPO_XATTN=0; echo "XATTN DISABLED"; grep -E "FAILED|Error|error" gpurun_out/r02a_attn_tests.txt | head -20
else (python tools/xattn_bench.py; TOPO_XATTN=0 python tools/xattn_bench.py) > gpurun_out/r02a_xattn_bench.txt 2>&1; cat gpurun_out/r02a_xattn_bench.txt; fi
(timeout 900 python -m pytest tests -m gpu -q -s -p no:cacheprovider --deselect tests/test_gpu_attention.py > gpurun_out/r02a_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02a_gpu_tests.txt)
tail -5 gpurun_out/r02a_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02a_gpu_tests.txt | head
timeout 300 python tools/parity_ablation.py 64 > gpurun_out/r02a_ablation.txt 2>&1; tail -12 gpurun_out/r02a_ablation.txt
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02a_bench_1gpu.json 2> gpurun_out/r02a_bench_1gpu.err; tail -c 400 gpurun_out/r02a_bench_1gpu.json
timeout 200 python tools/shape_breakdown.py 38 > gpurun_out/r02a_shape_breakdown.txt 2>&1; head -5 gpurun_out/r02a_shape_breakdown.txt
