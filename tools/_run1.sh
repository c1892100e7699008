mkdir -p gpurun_out
timeout 120 python tools/pair_check.py 2>&1 | tail -2
python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; echo "stage rc=$?"; grep -E "tiles=" gpurun_out/stage.log
