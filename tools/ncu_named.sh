#!/bin/bash
# ncu --set full captures of the north star's named kernels on the CURRENT build, taken from a real apply_model (tools/profile_step.py,
# Be = 38, ctrlora r=16): one ncu pass per kernel family (-k filter, a few launches each), reports under gpurun_out/<tag>_*.ncu-rep.
# Run under gpurun AFTER the plain command exited 0:   bash tools/ncu_named.sh r02a
# Here (no GPU):  for f in gpurun_out/r02a_*.ncu-rep; do ncu -i $f --page raw --csv --print-units base > ${f%.ncu-rep}.csv; done
#                 python tools/ncu_summary.py json profiles/r02_ncu.json "<note>" gpurun_out/r02a_*.csv
set -u
tag=${1:-r02}
variant=${2:-ctrlora}
mkdir -p gpurun_out
python tools/profile_step.py 38 $variant > gpurun_out/${tag}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
run() {  # name regex(on the demangled name) skip count
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled -k "regex:$2" -s $3 -c $4 \
      -f -o gpurun_out/${tag}_$1 python tools/profile_step.py 38 $variant > gpurun_out/${tag}_$1.log 2>&1
  echo "$1 rc=$? $(grep -c '==PROF== Profiling' gpurun_out/${tag}_$1.log) launches profiled"
}
run lora    'gemm_kernel<\(int\)[0-9]+, \(int\)(16|32),' 0 6
run attn    'attention_split_kernel|attention_kernel<\(int\)(80|160), \(int\)2' 1 6
run xattn   'cross_attention_kernel' 0 6
run gn      'gn_(stats|apply|small)_kernel' 0 8
run ln      'layernorm' 0 4
run conv    'gemm_kernel<\(int\)(256|320), \(int\)0, \(int\)[0-9], \(bool\)1, \(bool\)1>' 2 4
run proj    'gemm_kernel<\(int\)(128|160|192|256), \(int\)0, \(int\)[0-9], \(bool\)0, \(bool\)[01]>' 8 10
