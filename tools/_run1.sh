mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_metrics.py -x -q -s -m gpu > gpurun_out/metrics.log 2>&1; echo "rc=$?"; grep -E "parity|passed|failed|Error|error|assert" gpurun_out/metrics.log | head -20
timeout 900 python -m pytest tests/test_gpu_e2e.py tests/test_gpu_norm_elem.py -x -q -m gpu 2>&1 | tail -3
