mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_norm_elem.py -x -q -m gpu 2>&1 | tail -1
XREP=2 python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; grep -E "tiles=" gpurun_out/stage.log; grep -E "groupnorm" gpurun_out/stage.log
