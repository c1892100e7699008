#!/bin/bash
# GPU call 9 (round 2): final 1-GPU bench lines on the final build: default workload as the driver runs it (with the CPU arm),
# the reference arm, and the other BASELINE configs
mkdir -p gpurun_out
timeout 900 python bench.py > gpurun_out/r02z_bench_1gpu.json 2> gpurun_out/r02z_bench_1gpu.err; echo "bench rc=$?"; head -c 400 gpurun_out/r02z_bench_1gpu.json; echo
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02z_bench_reference_arm.json 2> gpurun_out/r02z_bench_reference_arm.err; echo "reference rc=$?"; head -c 300 gpurun_out/r02z_bench_reference_arm.json; echo
for w in addlrSD qkvo; do timeout 400 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r02z_bench_$w.json 2> gpurun_out/r02z_bench_$w.err; head -c 200 gpurun_out/r02z_bench_$w.json; echo; done
timeout 400 python bench.py --workload fullmap --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r02z_bench_fullmap_1gpu.json 2> gpurun_out/r02z_bench_fullmap_1gpu.err; head -c 200 gpurun_out/r02z_bench_fullmap_1gpu.json; echo
timeout 200 python tools/shape_breakdown.py 38 > gpurun_out/r02z_shape_breakdown.txt 2>&1; head -3 gpurun_out/r02z_shape_breakdown.txt
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02z_launches.csv python tools/profile_step.py 38 > gpurun_out/r02z_launches.log 2>&1; echo "launch list rc=$?"
