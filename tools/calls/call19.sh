#!/bin/bash
# GPU call 19 (round 2): re-test of the dispatch switches on the new epilogue (same box, alternating)
mkdir -p gpurun_out
tag=${1:-r02aa}
(timeout 600 python -m pytest tests/test_gpu_gemm.py -m gpu -q -x -p no:cacheprovider 2>&1 | tail -3)
for r in 1 2; do
for e in "X=0" "TOPO_BN320=2" "TOPO_PAIR_MINK=640" "TOPO_BN160_WIDE=1" "TOPO_PAIR_MINK=320"; do
  echo "== $e" >> gpurun_out/${tag}_switches.txt
  env $e timeout 300 python tools/gemm_bench.py "gemm M" 2>&1 | grep -E "N320 K1280|K640|K2560|N1280 K1280|N2560 K320" >> gpurun_out/${tag}_switches.txt
done; done
python - <<PY
import re,collections
rows=collections.OrderedDict(); cur=None
for l in open("gpurun_out/${tag}_switches.txt"):
    if l.startswith("== "): cur=l.split()[1]; continue
    m=re.match(r"(.{52})\s+([0-9.]+) us.*?(\S+<.*>)\s*$",l)
    if m: rows.setdefault(m.group(1).strip(),collections.OrderedDict()).setdefault(cur,[]).append((float(m.group(2)),m.group(3)))
for k,v in rows.items():
    base=min(x[0] for x in v["X=0"])
    print(k)
    for e,xs in v.items():
        t=min(x[0] for x in xs); print(f"    {e:22s} {t:8.1f} us {100*(t/base-1):+6.1f}%  {xs[0][1]}")
PY
