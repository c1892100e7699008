#!/bin/bash
# GPU call 21 (round 2): FINAL build, 1 GPU: full -m gpu suite, smoke, default bench line as the driver runs it, the other BASELINE
# configs, step breakdown, launch list
mkdir -p gpurun_out
tag=${1:-r03a}
(timeout 1200 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/${tag}_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gpu_tests.txt)
tail -3 gpurun_out/${tag}_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/${tag}_gpu_tests.txt | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/${tag}_bench_1gpu.json 2> gpurun_out/${tag}_bench_1gpu.err; echo "bench rc=$?"; head -c 300 gpurun_out/${tag}_bench_1gpu.json; echo
for w in addlrSD qkvo; do timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/${tag}_bench_$w.json 2> gpurun_out/${tag}_bench_$w.err; head -c 200 gpurun_out/${tag}_bench_$w.json; echo; done
timeout 400 python bench.py --workload fullmap --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/${tag}_bench_fullmap_1gpu.json 2> gpurun_out/${tag}_bench_fullmap_1gpu.err; head -c 200 gpurun_out/${tag}_bench_fullmap_1gpu.json; echo
timeout 200 python tools/shape_breakdown.py 38 > gpurun_out/${tag}_shape_breakdown.txt 2>&1; head -3 gpurun_out/${tag}_shape_breakdown.txt
timeout 200 python tools/vae_breakdown.py > gpurun_out/${tag}_vae_breakdown.txt 2>&1; head -2 gpurun_out/${tag}_vae_breakdown.txt
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/${tag}_launches.csv python tools/profile_step.py 38 > gpurun_out/${tag}_launches.log 2>&1; echo "launch list rc=$?"
