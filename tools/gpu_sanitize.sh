#!/bin/bash
# compute-sanitizer memcheck + racecheck over tools/sanitize_case.py (run on the GPU box through gpurun).
# Summaries land in gpurun_out/sanitize_{memcheck,racecheck}.txt; copy them to profiles/ when they are to be kept.
mkdir -p gpurun_out
for tool in memcheck racecheck; do
  timeout 1200 compute-sanitizer --tool $tool --print-limit 20 python tools/sanitize_case.py > gpurun_out/sanitize_$tool.full.txt 2>&1
  echo "exit $?" >> gpurun_out/sanitize_$tool.full.txt
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|hazard|Invalid|exit |sanitize_case done| ok " gpurun_out/sanitize_$tool.full.txt | head -60 > gpurun_out/sanitize_$tool.txt
done
tail -5 gpurun_out/sanitize_memcheck.txt gpurun_out/sanitize_racecheck.txt
