mkdir -p gpurun_out
ITERS=7 python tools/gemm_bench.py Nk77 2>&1 | tail -1
TOPO_ATTN_NOALIAS=1 ITERS=7 python tools/gemm_bench.py Nk77 2>&1 | tail -1
