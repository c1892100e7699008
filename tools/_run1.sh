mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cut -c1-200 gpurun_out/bench.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:gemm_kernel -c 80 -o /tmp/gemm_full -f python tools/profile_step.py > gpurun_out/ncu3.log 2>&1; echo "ncu gemm rc=$?"
ncu -i /tmp/gemm_full.ncu-rep --page raw --csv > gpurun_out/gemm_full_raw.csv
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:"attention_kernel|gn_|layernorm" -c 16 -o /tmp/misc_full -f python tools/profile_step.py > gpurun_out/ncu5.log 2>&1; echo "ncu misc rc=$?"
ncu -i /tmp/misc_full.ncu-rep --page raw --csv > gpurun_out/misc_full_raw.csv
du -sh gpurun_out
