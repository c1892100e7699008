#!/bin/bash
# ncu --set full captures of the north star's named kernels on the CURRENT build, taken from a real apply_model (tools/profile_step.py,
# Be = 38, ctrlora r=16): one ncu pass per kernel family (-k filter, a few launches each), reports under gpurun_out/<tag>_*.ncu-rep.
# Run under gpurun AFTER the plain command exited 0:   bash tools/ncu_named.sh r02a
# (the raw pages are exported to CSV on the box, the reports are dropped: gpurun returns at most 64 MiB)
# Here (no GPU):  python tools/ncu_summary.py json profiles/r02_ncu.json "<note>" gpurun_out/<tag>_*.csv
set -u
tag=${1:-r02}
variant=${2:-ctrlora}
mkdir -p gpurun_out
run() {  # name regex(on the demangled name) skip count
  ncu --set full --clock-control none --profile-from-start off --kernel-name-base demangled -k "regex:$2" -s $3 -c $4 \
      -f -o gpurun_out/${tag}_$1 python tools/profile_step.py 38 $variant > gpurun_out/${tag}_$1.log 2>&1
  echo "$1 rc=$? $(grep -c '==PROF== Profiling' gpurun_out/${tag}_$1.log) launches profiled"
  # gpurun brings back at most 64 MiB: keep the raw page as CSV (base units), drop the report
  ncu -i gpurun_out/${tag}_$1.ncu-rep --page raw --csv --print-units base > gpurun_out/${tag}_$1.csv 2>/dev/null && rm -f gpurun_out/${tag}_$1.ncu-rep
  ls -la gpurun_out/${tag}_$1.csv | awk '{print $5, $9}'
}
run lora    'gemm_kernel<\(int\)[0-9]+, \(int\)(16|32),' 0 6
run attn    'attention_split_kernel|attention_kernel<\(int\)(80|160), \(int\)2' 1 6
run xattn   'cross_attention_kernel' 0 6
run gn      'gn_(stats|apply|small)_kernel' 0 8
run ln      'layernorm' 0 4
run conv    'gemm_kernel<\(int\)(192|256|320), \(int\)0, \(int\)[0-9], \(bool\)1, \(bool\)1>' 2 6
run proj    'gemm_resident_kernel|gemm_kernel<\(int\)(128|160|192|256), \(int\)0, \(int\)[0-9], \(bool\)0, \(bool\)[01]>' 8 16
