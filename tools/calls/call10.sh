#!/bin/bash
# GPU call 10 (round 2): LoRA pre-pass walk tests; split attention with 0 / 1 of 8 pairs on the polynomial
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_attention.py -q -s -p no:cacheprovider > gpurun_out/r02i_kernel_tests.txt 2>&1; tail -3 gpurun_out/r02i_kernel_tests.txt; grep -E "^FAILED|^ERROR|LoRA pre-pass" gpurun_out/r02i_kernel_tests.txt | cut -c1-160 | head -12
(for cfg in 30 31 32 33; do echo "== TOPO_ATTN_SPLIT=$cfg"; TOPO_ATTN_SPLIT=$cfg timeout 120 python tools/attn_bench.py 38 | grep "l_col= 40"; done) > gpurun_out/r02i_attn_bench.txt 2>&1; cat gpurun_out/r02i_attn_bench.txt
