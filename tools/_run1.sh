mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_norm_elem.py -m gpu -q -x > gpurun_out/norm.log 2>&1; echo "norm rc=$?"; tail -3 gpurun_out/norm.log
timeout 600 python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; echo "stage rc=$?"; grep -E "tiles=|groupnorm" gpurun_out/stage.log
