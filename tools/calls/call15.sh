#!/bin/bash
# GPU call 15 (round 2): GEMM parity on the 16-warp epilogue + source-level ncu of three short-K cases
mkdir -p gpurun_out
tag=${1:-r02m}
true
true
prof() {  # name filter
  ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k "regex:gemm_resident_kernel|gemm_kernel" -s 2 -c 1 \
      -f -o gpurun_out/${tag}_$1 python tools/gemm_bench.py "$2" > gpurun_out/${tag}_$1.log 2>&1
  echo "$1 ncu rc=$?"
  ncu -i gpurun_out/${tag}_$1.ncu-rep --page source --csv > gpurun_out/${tag}_$1_source.csv 2>/dev/null
  ncu -i gpurun_out/${tag}_$1.ncu-rep --page raw --csv --print-units base > gpurun_out/${tag}_$1_raw.csv 2>/dev/null
  rm -f gpurun_out/${tag}_$1.ncu-rep
}
prof res0 "N320 K320 res0"
prof res1 "N320 K320 res1"
prof geglu "N2560 K320"
ls -la gpurun_out/${tag}_*
