timeout 900 python -m pytest tests/test_gpu_textenc.py -x -q -s -m gpu 2>&1 | grep -E "parity|passed|failed|Error" | head
