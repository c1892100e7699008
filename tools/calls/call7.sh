#!/bin/bash
# GPU call 7 (round 2): pair kernels at K = 640 (A/B), 160-wide tiles at N = 640 (A/B), then the full -m gpu suite on the candidate final build
mkdir -p gpurun_out
(echo "== default"; ITERS=7 timeout 200 python tools/gemm_bench.py "K640 "; ITERS=7 timeout 100 python tools/gemm_bench.py "N640 K2560"
 echo "== TOPO_PAIR_MINK=640"; TOPO_PAIR_MINK=640 ITERS=7 timeout 200 python tools/gemm_bench.py "K640 "
 echo "== TOPO_BN160_WIDE=1"; TOPO_BN160_WIDE=1 ITERS=7 timeout 200 python tools/gemm_bench.py "K640 ") > gpurun_out/r02h_gemm_ab.txt 2>&1; cat gpurun_out/r02h_gemm_ab.txt
(timeout 900 python -m pytest tests -m gpu -q -s -p no:cacheprovider > gpurun_out/r02h_gpu_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02h_gpu_tests.txt)
tail -4 gpurun_out/r02h_gpu_tests.txt; grep -E "^FAILED|^ERROR" gpurun_out/r02h_gpu_tests.txt | head
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
