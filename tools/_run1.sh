timeout 900 python -m pytest tests/test_gpu_e2e.py -k eager_pytorch -x -q -s -m gpu 2>&1 | grep -E "timing|passed|failed|Error" | head
