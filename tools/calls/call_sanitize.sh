#!/bin/bash
# usage: bash tools/calls/call_sanitize.sh memcheck|racecheck|synccheck   (ONE tool per gpurun call)
tool=${1:-memcheck}
mkdir -p gpurun_out
timeout 300 python tools/sanitize_small.py > gpurun_out/r02_sanitize_plain.txt 2>&1; rc=$?; tail -3 gpurun_out/r02_sanitize_plain.txt
if [ $rc -ne 0 ]; then echo "plain run failed (rc=$rc): not running the sanitizer"; exit 1; fi
timeout 1500 compute-sanitizer --tool $tool --print-limit 50 python tools/sanitize_small.py > gpurun_out/r02_sanitize_$tool.txt 2>&1; echo "sanitizer rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|MISMATCHES" gpurun_out/r02_sanitize_$tool.txt | head -20
tail -3 gpurun_out/r02_sanitize_$tool.txt
