#!/bin/bash
# compute-sanitizer over every entry point at small sizes (tools/sanitize_run.py).
set -u
mkdir -p gpurun_out
timeout 300 python tools/sanitize_run.py || exit 1
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --log-file gpurun_out/sanitizer_$tool.log python tools/sanitize_run.py > gpurun_out/sanitizer_$tool.out 2>&1
  echo "$tool rc=$?"; tail -2 gpurun_out/sanitizer_$tool.out; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|Error|hazard" gpurun_out/sanitizer_$tool.log | sort | uniq -c | head -10
done
