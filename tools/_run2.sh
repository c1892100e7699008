mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x 2>&1 | tail -2
ITERS=5 python tools/gemm_bench.py 2>&1 | grep -E "gemm|conv" 
