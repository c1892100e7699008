mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py -x -q -m gpu 2>&1 | tail -1
ITERS=7 timeout 300 python tools/gemm_bench.py "attn" 2>&1 | grep attn
for i in 1 2; do XREP=2 python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; grep -E "tiles=" gpurun_out/stage.log; done
