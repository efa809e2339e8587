#!/bin/bash
# one compute-sanitizer tool per gpurun call (B200_PROFILING.md); plain run first
mkdir -p gpurun_out
tool=${1:-memcheck}
python tools/sanitize_small.py > gpurun_out/sanitize_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitize_plain.log; exit 1; }
tail -3 gpurun_out/sanitize_plain.log
timeout 1500 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_small.py > gpurun_out/sanitize_$tool.log 2>&1
echo "sanitizer $tool rc=$?"
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|SYNCCHECK SUMMARY|hazard|Invalid|error" gpurun_out/sanitize_$tool.log | head -20
tail -4 gpurun_out/sanitize_$tool.log
