mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_textenc.py tests/test_gpu_attention.py tests/test_gpu_gemm.py -m gpu -q -s > gpurun_out/textenc.log 2>&1; echo "rc=$?"; grep -E "parity\] (BERT|encode|adapter)|Error|error|passed|failed" gpurun_out/textenc.log | tail -30
