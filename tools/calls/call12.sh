#!/bin/bash
# GPU call 12 (round 2): resident-operand short-K projections (gemm_resident_kernel): parity, A/B micro-bench, step breakdown
mkdir -p gpurun_out
(timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x -s -p no:cacheprovider > gpurun_out/r02j_gemm_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02j_gemm_tests.txt)
tail -4 gpurun_out/r02j_gemm_tests.txt; grep -E "^FAILED|^ERROR|Error" gpurun_out/r02j_gemm_tests.txt | head
for m in 0 1 2 3; do
  echo "== TOPO_RESIDENT=$m" >> gpurun_out/r02j_gemm_ab.txt
  TOPO_RESIDENT=$m timeout 300 python tools/gemm_bench.py "K3" >> gpurun_out/r02j_gemm_ab.txt 2>&1
done
cat gpurun_out/r02j_gemm_ab.txt
TOPO_RESIDENT=0 timeout 300 python tools/shape_breakdown.py > gpurun_out/r02j_shape_breakdown_off.txt 2>&1; head -1 gpurun_out/r02j_shape_breakdown_off.txt
timeout 300 python tools/shape_breakdown.py > gpurun_out/r02j_shape_breakdown.txt 2>&1; head -1 gpurun_out/r02j_shape_breakdown.txt
grep "resident" gpurun_out/r02j_shape_breakdown.txt
