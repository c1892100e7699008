#!/bin/bash
# GPU call 14 (round 2): 16-warp row-per-thread epilogue (bf16 staging, 16-byte stores): parity, micro-bench, step breakdown
mkdir -p gpurun_out
tag=${1:-r02l}
(timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_gemm_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gemm_tests.txt)
tail -4 gpurun_out/${tag}_gemm_tests.txt; grep -E "^FAILED|^ERROR|Error" gpurun_out/${tag}_gemm_tests.txt | head
timeout 300 python tools/gemm_bench.py > gpurun_out/${tag}_gemm_bench.txt 2>&1; cat gpurun_out/${tag}_gemm_bench.txt
timeout 300 python tools/shape_breakdown.py > gpurun_out/${tag}_shape_breakdown.txt 2>&1; head -45 gpurun_out/${tag}_shape_breakdown.txt
