#!/bin/bash
# GPU call 6 (round 2): GroupNorm L2-sized sample chunks A/B; source-level ncu capture of attention_split_kernel<3,3>
mkdir -p gpurun_out
timeout 200 python -m pytest tests/test_gpu_norm_elem.py -q -p no:cacheprovider 2>&1 | tail -2
(for mb in 0 24 40 64; do echo "== TOPO_GN_CHUNK_MB=$mb"; TOPO_GN_CHUNK_MB=$mb timeout 200 python tools/gn_bench.py 38 4096x320,4096x640,4096x960,1024x640,1024x1280,1024x1920; done) > gpurun_out/r02g_gn_bench.txt 2>&1; cat gpurun_out/r02g_gn_bench.txt
python tools/batch_invariance.py 2>&1 | tail -3
for mb in 0 40; do TOPO_GN_CHUNK_MB=$mb timeout 200 python tools/shape_breakdown.py 38 2>&1 | grep -E "^total|groupnorm" | head -4; done
TOPO_ATTN_SPLIT=33 timeout 120 python tools/attn_bench.py 8 > gpurun_out/r02g_attn_plain.log 2>&1 && TOPO_ATTN_SPLIT=33 timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_split_kernel -s 4 -c 1 -f -o gpurun_out/r02g_attn_split_src python tools/attn_bench.py 8 > gpurun_out/r02g_attn_ncu.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/r02g_attn_ncu.log
