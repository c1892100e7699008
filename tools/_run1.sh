mkdir -p gpurun_out
python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; echo "stage rc=$?"; grep -E "tiles=|conv_direct" gpurun_out/stage.log
timeout 900 python -m pytest tests/test_gpu_e2e.py -m gpu -x -q > gpurun_out/e2e.log 2>&1; echo "e2e rc=$?"; tail -3 gpurun_out/e2e.log
