#!/bin/bash
# GPU call 13 (round 2): source-level ncu of the short-K projection kernels (where do the epilogue warps stall?)
mkdir -p gpurun_out
prof() {  # tag filter
  ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k "regex:gemm_resident_kernel" -s 2 -c 1 \
      -f -o gpurun_out/r02k_$1 python tools/gemm_bench.py "$2" > gpurun_out/r02k_$1.log 2>&1
  echo "$1 ncu rc=$?"
  ncu -i gpurun_out/r02k_$1.ncu-rep --page source --csv > gpurun_out/r02k_$1_source.csv 2>/dev/null
  ncu -i gpurun_out/r02k_$1.ncu-rep --page raw --csv --print-units base > gpurun_out/r02k_$1_raw.csv 2>/dev/null
  ls -la gpurun_out/r02k_$1*
  rm -f gpurun_out/r02k_$1.ncu-rep
}
timeout 200 python tools/gemm_bench.py "N320 K320 res1" || exit 1
prof res1 "N320 K320 res1"
prof geglu "N2560 K320"
