#!/bin/bash
# GPU call 16 (round 2): warp-uniform producer / issuer warps in the attention kernels: parity + micro-bench + step breakdown
mkdir -p gpurun_out
tag=${1:-r02o}
(timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_gemm.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_attn_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_attn_tests.txt)
tail -4 gpurun_out/${tag}_attn_tests.txt; grep -E "^FAILED|^ERROR|Error" gpurun_out/${tag}_attn_tests.txt | head
timeout 300 python tools/attn_bench.py > gpurun_out/${tag}_attn_bench.txt 2>&1; cat gpurun_out/${tag}_attn_bench.txt
timeout 300 python tools/gemm_bench.py attn >> gpurun_out/${tag}_attn_bench.txt 2>&1; tail -5 gpurun_out/${tag}_attn_bench.txt
timeout 300 python tools/xattn_bench.py > gpurun_out/${tag}_xattn_bench.txt 2>&1; cat gpurun_out/${tag}_xattn_bench.txt
timeout 300 python tools/shape_breakdown.py > gpurun_out/${tag}_shape_breakdown.txt 2>&1; head -12 gpurun_out/${tag}_shape_breakdown.txt
