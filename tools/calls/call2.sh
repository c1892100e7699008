#!/bin/bash
# GPU call 2 (round 2): LoRA pre-pass kernel + cross-attention v2: tests, micro-benches with A/B switches, bench (default + other workloads)
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py -q -s -p no:cacheprovider > gpurun_out/r02b_kernel_tests.txt 2>&1; rc=$?
tail -3 gpurun_out/r02b_kernel_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02b_kernel_tests.txt | head
(for cfg in 23 20 22 40 42 43; do echo "== TOPO_XATTN_CFG=$cfg"; TOPO_XATTN_CFG=$cfg timeout 120 python tools/xattn_bench.py; done) > gpurun_out/r02b_xattn_bench.txt 2>&1; cat gpurun_out/r02b_xattn_bench.txt
(echo "== default tile widths"; timeout 200 python tools/lora_bench.py; echo "== TOPO_LORA_BN=128"; TOPO_LORA_BN=128 timeout 200 python tools/lora_bench.py) > gpurun_out/r02b_lora_bench.txt 2>&1; cat gpurun_out/r02b_lora_bench.txt
(timeout 900 python -m pytest tests -m gpu -q -s -p no:cacheprovider --deselect tests/test_gpu_attention.py --deselect tests/test_gpu_gemm.py --deselect tests/test_gpu_pair_misc.py > gpurun_out/r02b_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_gpu_tests.txt)
tail -4 gpurun_out/r02b_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02b_gpu_tests.txt | head
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02b_bench_1gpu.json 2> gpurun_out/r02b_bench_1gpu.err; tail -c 300 gpurun_out/r02b_bench_1gpu.json
for w in addlrSD qkvo; do timeout 300 python bench.py --workload $w --steps 2 --warmup 2 --no-cpu-baseline > gpurun_out/r02b_bench_$w.json 2> gpurun_out/r02b_bench_$w.err; head -c 300 gpurun_out/r02b_bench_$w.json; echo; done
timeout 200 python tools/shape_breakdown.py 38 > gpurun_out/r02b_shape_breakdown.txt 2>&1; head -3 gpurun_out/r02b_shape_breakdown.txt
timeout 300 python bench.py --workload fullmap --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02b_bench_fullmap_1gpu.json 2> gpurun_out/r02b_bench_fullmap_1gpu.err; head -c 300 gpurun_out/r02b_bench_fullmap_1gpu.json; tail -3 gpurun_out/r02b_bench_fullmap_1gpu.err
