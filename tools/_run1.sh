mkdir -p gpurun_out
python -m pytest tests/test_gpu_attention.py -m gpu -x -q -s > gpurun_out/attn_tests.log 2>&1; echo "attn tests rc=$?" 
for ns in 1 2 3 4; do for np in 0 2; do echo "NS=$ns NPOLY=$np"; TOPO_ATTN_NS=$ns TOPO_ATTN_NPOLY=$np ITERS=7 python tools/gemm_bench.py attn 2>&1 | tail -4; done; done > gpurun_out/attn_sweep.log 2>&1
cat gpurun_out/attn_sweep.log; tail -5 gpurun_out/attn_tests.log
