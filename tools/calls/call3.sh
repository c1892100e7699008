#!/bin/bash
# GPU call 3 (round 2): cross-attention v3 (TMA store), packed GEGLU epilogue, 192-wide fused qkv; golden tests vs the reference's own outputs
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_attention.py tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py tests/test_gpu_models.py tests/test_gpu_rows.py -q -s -p no:cacheprovider > gpurun_out/r02c_kernel_tests.txt 2>&1
tail -3 gpurun_out/r02c_kernel_tests.txt; grep -E "^FAILED|^ERROR|vs reference" gpurun_out/r02c_kernel_tests.txt | cut -c1-200 | head -30
(echo "== TMA store"; timeout 120 python tools/xattn_bench.py; echo "== TOPO_XATTN_TMAST=0"; TOPO_XATTN_TMAST=0 timeout 120 python tools/xattn_bench.py) > gpurun_out/r02c_xattn_bench.txt 2>&1; cat gpurun_out/r02c_xattn_bench.txt
ITERS=7 timeout 200 python tools/gemm_bench.py gemm > gpurun_out/r02c_gemm_bench.txt 2>&1; cat gpurun_out/r02c_gemm_bench.txt
timeout 200 python tools/lora_bench.py 2>&1 | grep " r0 " > gpurun_out/r02c_qkv_bench.txt; cat gpurun_out/r02c_qkv_bench.txt
timeout 400 python -m pytest tests/test_gpu_e2e.py -q -s -p no:cacheprovider > gpurun_out/r02c_e2e_tests.txt 2>&1; tail -2 gpurun_out/r02c_e2e_tests.txt
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02c_bench_1gpu.json 2> gpurun_out/r02c_bench_1gpu.err; tail -c 300 gpurun_out/r02c_bench_1gpu.json
timeout 200 python tools/shape_breakdown.py 38 > gpurun_out/r02c_shape_breakdown.txt 2>&1; head -3 gpurun_out/r02c_shape_breakdown.txt
