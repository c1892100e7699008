mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_models.py -m gpu -q -s -x > gpurun_out/models.log 2>&1; echo "models rc=$?"; grep -E "parity|Error|error|assert|passed|failed" gpurun_out/models.log | tail -40
