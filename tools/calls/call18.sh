#!/bin/bash
# GPU call 18 (round 2): full -m gpu suite, smoke and the default bench line on the current build
mkdir -p gpurun_out
tag=${1:-r02w}
(timeout 1200 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/${tag}_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gpu_tests.txt)
tail -3 gpurun_out/${tag}_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/${tag}_gpu_tests.txt | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/${tag}_bench_1gpu.json 2> gpurun_out/${tag}_bench_1gpu.err; echo "bench rc=$?"; head -c 400 gpurun_out/${tag}_bench_1gpu.json; echo
