#!/bin/bash
# GPU call 17 (round 2): source-level ncu of attention_split_kernel<3, 1> (B = 8 to keep the report small)
mkdir -p gpurun_out
tag=${1:-r02v}
timeout 300 python tools/attn_bench.py 8 || exit 1
ncu --set full --import-source on --clock-control none --kernel-name-base demangled -k "regex:attention_split_kernel" -s 2 -c 1 \
    -f -o gpurun_out/${tag}_split python tools/attn_bench.py 8 > gpurun_out/${tag}_split.log 2>&1
echo "ncu rc=$?"
ncu -i gpurun_out/${tag}_split.ncu-rep --page source --csv > gpurun_out/${tag}_split_source.csv 2>/dev/null
ncu -i gpurun_out/${tag}_split.ncu-rep --page raw --csv --print-units base > gpurun_out/${tag}_split_raw.csv 2>/dev/null
rm -f gpurun_out/${tag}_split.ncu-rep; ls -la gpurun_out/${tag}_*
