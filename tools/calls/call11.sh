#!/bin/bash
# GPU call 11 (round 2): FINAL build (split attention default <3,1>): test_zz_no_alloc alone in a fresh process, then the full -m gpu suite (it runs last there), smoke, default bench line, ncu of the attention family
mkdir -p gpurun_out
(timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -p no:cacheprovider > gpurun_out/r02y_no_alloc_alone.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02y_no_alloc_alone.txt)
tail -4 gpurun_out/r02y_no_alloc_alone.txt
(timeout 900 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r02y_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02y_gpu_tests.txt)
tail -3 gpurun_out/r02y_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02y_gpu_tests.txt | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/r02y_bench_1gpu.json 2> gpurun_out/r02y_bench_1gpu.err; echo "bench rc=$?"; head -c 300 gpurun_out/r02y_bench_1gpu.json; echo
ncu --set full --clock-control none --profile-from-start off --kernel-name-base demangled -k "regex:attention_split_kernel" -s 1 -c 3 -f -o gpurun_out/r02y_attn python tools/profile_step.py 38 > gpurun_out/r02y_attn.log 2>&1; echo "ncu rc=$?"
ncu -i gpurun_out/r02y_attn.ncu-rep --page raw --csv --print-units base > gpurun_out/r02y_attn.csv 2>/dev/null && rm -f gpurun_out/r02y_attn.ncu-rep
