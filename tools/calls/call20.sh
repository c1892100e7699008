#!/bin/bash
# GPU call 20 (round 2): cross-attention with warp-uniform producer / issuer warps, same-box A/B
mkdir -p gpurun_out
tag=${1:-r02ad}
PK=$(ls -d interactive*_b200)
(timeout 600 python -m pytest tests/test_gpu_attention.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3)
for v in A B A B A B; do
  if [ $v = A ]; then export TOPO_LIB=$PWD/$PK/ab/libtopo_A.so; else unset TOPO_LIB; fi
  echo "== $v" >> gpurun_out/${tag}_xattn_ab.txt
  timeout 300 python tools/xattn_bench.py >> gpurun_out/${tag}_xattn_ab.txt 2>&1
done
cat gpurun_out/${tag}_xattn_ab.txt
