mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/launches.csv python tools/profile_step.py > gpurun_out/ncu1.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:attention_kernel -c 4 -o gpurun_out/attn_full -f python tools/profile_step.py > gpurun_out/ncu2.log 2>&1; echo "ncu attn rc=$?"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:gemm_kernel -c 70 -o /tmp/gemm_full -f python tools/profile_step.py > gpurun_out/ncu3.log 2>&1; echo "ncu gemm rc=$?"
ncu -i /tmp/gemm_full.ncu-rep --page raw --csv > gpurun_out/gemm_full_raw.csv
timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:"gemm_kernel<256, 0, 6, 1, 1>" -c 2 -o gpurun_out/pairconv_full -f python tools/profile_step.py > gpurun_out/ncu4.log 2>&1; echo "ncu pair rc=$?"
timeout 900 ncu --set full --clock-control none --profile-from-start off -k regex:"gn_|layernorm" -c 12 -o /tmp/norm_full -f python tools/profile_step.py > gpurun_out/ncu5.log 2>&1; echo "ncu norm rc=$?"
ncu -i /tmp/norm_full.ncu-rep --page raw --csv > gpurun_out/norm_full_raw.csv
du -sh gpurun_out
