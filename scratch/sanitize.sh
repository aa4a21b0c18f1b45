#!/bin/bash
# compute-sanitizer over scratch/sanitize_driver.py; logs -> gpurun_out/sanitizer_<tool>.log
mkdir -p gpurun_out
for tool in memcheck racecheck synccheck; do
  timeout 900 compute-sanitizer --tool $tool --print-limit 20 python scratch/sanitize_driver.py > gpurun_out/sanitizer_$tool.log 2>&1
  echo "== $tool: rc $?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|all ok|Error|hazard" gpurun_out/sanitizer_$tool.log | sort | uniq -c | head -8
done
