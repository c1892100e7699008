#!/bin/bash
# GPU call 4 (round 2): ones-column row sum in flash attention; source-level ncu capture of the N=4096 d=40 kernel
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attention.py tests/test_gpu_models.py -q -s -p no:cacheprovider > gpurun_out/r02e_kernel_tests.txt 2>&1
tail -3 gpurun_out/r02e_kernel_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02e_kernel_tests.txt | head
timeout 200 python tools/attn_bench.py 38 > gpurun_out/r02e_attn_bench.txt 2>&1; cat gpurun_out/r02e_attn_bench.txt
timeout 500 python -m pytest tests/test_gpu_e2e.py tests/test_gpu_fullsize.py -q -s -p no:cacheprovider > gpurun_out/r02e_e2e_tests.txt 2>&1; tail -2 gpurun_out/r02e_e2e_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02e_e2e_tests.txt | head
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r02e_bench_1gpu.json 2> gpurun_out/r02e_bench_1gpu.err; tail -c 300 gpurun_out/r02e_bench_1gpu.json; echo
TOPO_ATTN_LCOL=0 timeout 400 python bench.py --no-cpu-baseline --steps 2 --warmup 2 > gpurun_out/r02e_bench_1gpu_lcol0.json 2> gpurun_out/r02e_bench_1gpu_lcol0.err; head -c 200 gpurun_out/r02e_bench_1gpu_lcol0.json; echo
timeout 120 python tools/attn_bench.py 8 > gpurun_out/r02e_attn_plain.log 2>&1 && timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 4 -c 1 -f -o gpurun_out/r02e_attn_src python tools/attn_bench.py 8 > gpurun_out/r02e_attn_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r02e_attn_ncu.log
