#!/bin/bash
# Run every -m gpu test id in its own process (a device trap kills the CUDA context of the process
# that hit it), with a per-test timeout; summary in gpurun_out/isolated.txt.
mkdir -p gpurun_out
FILES="${@:-tests}"
python -m pytest $FILES -m gpu --collect-only -q 2>/dev/null | grep "::" > gpurun_out/ids.txt
: > gpurun_out/isolated.txt
while read -r id; do
  timeout 120 python -m pytest "$id" -x -q -s > gpurun_out/one.log 2>&1
  rc=$?
  echo "rc=$rc $id" >> gpurun_out/isolated.txt
  grep -E "\[parity\]|Error|error|assert" gpurun_out/one.log | head -8 >> gpurun_out/isolated.txt
done < gpurun_out/ids.txt
cat gpurun_out/isolated.txt
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
