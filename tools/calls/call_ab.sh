#!/bin/bash
# Same-box A/B: library A = <pkg>/ab/libtopo_A.so (an earlier build), B = the in-tree build.  usage: call_ab.sh <tag> [gemm_bench filter]
mkdir -p gpurun_out
tag=${1:-ab}; filt=${2:-}
PK=$(ls -d interactive*_b200)
(timeout 900 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_pair_misc.py -m gpu -q -x -p no:cacheprovider > gpurun_out/${tag}_gemm_tests.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/${tag}_gemm_tests.txt)
tail -3 gpurun_out/${tag}_gemm_tests.txt; grep -E "^FAILED|^ERROR|Error" gpurun_out/${tag}_gemm_tests.txt | head
for v in A B A B; do
  if [ $v = A ]; then export TOPO_LIB=$PWD/$PK/ab/libtopo_A.so; else unset TOPO_LIB; fi
  echo "== $v" >> gpurun_out/${tag}_gemm_bench.txt
  timeout 300 python tools/gemm_bench.py $filt 2>&1 | grep -v "^attn" >> gpurun_out/${tag}_gemm_bench.txt
done
unset TOPO_LIB
python - <<PY
import re,collections
rows=collections.OrderedDict()
cur=None
for l in open("gpurun_out/${tag}_gemm_bench.txt"):
    if l.startswith("== "): cur=l.split()[1]; continue
    m=re.match(r"(.{52})\s+([0-9.]+) us",l)
    if m: rows.setdefault(m.group(1).strip(),{}).setdefault(cur,[]).append(float(m.group(2)))
for k,v in rows.items():
    a=min(v.get("A",[0])); b=min(v.get("B",[0]))
    print(f"{k:52s} A {a:8.1f}  B {b:8.1f}  {100*(b/a-1) if a else 0:+6.1f}%")
PY
for r in 1 2; do for v in A B; do
  if [ $v = A ]; then export TOPO_LIB=$PWD/$PK/ab/libtopo_A.so; else unset TOPO_LIB; fi
  timeout 300 python tools/shape_breakdown.py > gpurun_out/${tag}_shape_breakdown_$v$r.txt 2>&1; echo "$v$r $(head -1 gpurun_out/${tag}_shape_breakdown_$v$r.txt)"
done; done
unset TOPO_LIB
python tools/ab_breakdown.py gpurun_out/${tag}_shape_breakdown_A1.txt gpurun_out/${tag}_shape_breakdown_B1.txt gpurun_out/${tag}_shape_breakdown_A2.txt gpurun_out/${tag}_shape_breakdown_B2.txt | tee gpurun_out/${tag}_ab_breakdown.txt
