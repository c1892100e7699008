#!/bin/bash
# GPU call 22 (round 2): ncu --set full of the kernel instances the family passes of tools/ncu_named.sh did not reach
mkdir -p gpurun_out
tag=r03n
run() {
  ncu --set full --clock-control none --profile-from-start off --kernel-name-base demangled -k "regex:$2" -s $3 -c $4 \
      -f -o gpurun_out/${tag}_$1 python tools/profile_step.py 38 > gpurun_out/${tag}_$1.log 2>&1
  echo "$1 rc=$? $(grep -c '==PROF== Profiling' gpurun_out/${tag}_$1.log) launches profiled"
  ncu -i gpurun_out/${tag}_$1.ncu-rep --page raw --csv --print-units base > gpurun_out/${tag}_$1.csv 2>/dev/null && rm -f gpurun_out/${tag}_$1.ncu-rep
}
run proj4   'gemm_kernel<\(int\)256, \(int\)0, \(int\)4, \(bool\)0, \(bool\)0>' 4 4
run conv192 'gemm_kernel<\(int\)192, \(int\)0, \(int\)6, \(bool\)1, \(bool\)1>' 2 3
run res     'gemm_resident_kernel<\(int\)(160|192), \(int\)4' 0 6
run proj160 'gemm_kernel<\(int\)160, \(int\)0, \(int\)5, \(bool\)0, \(bool\)0>' 0 3
