mkdir -p gpurun_out
for np in 0 2; do
TOPO_ATTN_NPOLY=$np ITERS=1 timeout 600 ncu --set full --import-source on --clock-control none -k regex:attention_kernel -s 2 -c 1 -f -o gpurun_out/attn_np$np python tools/gemm_bench.py Nk4096 > gpurun_out/ncu_attn_np$np.log 2>&1; echo "ncu np=$np rc=$?"
done
ls -la gpurun_out/
