mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_e2e.py -x -q -m gpu 2>&1 | tail -2
for v in 0 1 1; do echo "TWO_STREAMS=$v"; TOPO_TWO_STREAMS=$v XREP=2 python tools/stage_breakdown.py > gpurun_out/stage$v.log 2>&1; grep -E "tiles=" gpurun_out/stage$v.log; done
