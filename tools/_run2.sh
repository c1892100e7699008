mkdir -p gpurun_out
timeout 120 python tools/pair_check.py 2>&1 | tail -12; echo "check rc=$?"
timeout 300 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x 2>&1 | tail -2
ITERS=5 timeout 300 python tools/gemm_bench.py 2>&1 | grep -E "gemm|conv" 
