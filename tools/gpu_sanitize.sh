#!/bin/bash
# one compute-sanitizer tool per gpurun call: bash tools/gpu_sanitize.sh memcheck|racecheck|initcheck
T=$1; O=gpurun_out; mkdir -p $O
python tools/sanitize_case.py > $O/sanitize_plain.log 2>&1 || { tail -5 $O/sanitize_plain.log; exit 1; }
timeout 1200 compute-sanitizer --tool $T --print-limit 50 python tools/sanitize_case.py > $O/r2_sanitize_$T.log 2>&1
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize_case ok|Error|hazard" $O/r2_sanitize_$T.log | head -20
