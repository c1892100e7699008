mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py tests/test_gpu_models.py -x -q -m gpu 2>&1 | tail -3
ITERS=7 timeout 300 python tools/gemm_bench.py "gemm" 2>&1 | grep gemm
python tools/stage_breakdown.py > gpurun_out/stage.log 2>&1; echo "stage rc=$?"; grep -E "tiles=" gpurun_out/stage.log
