#!/bin/bash
# ncu launch list (gpu__time_duration per launch, cold-cache, serialised) of one bench step: bash tools/gpu_launches.sh <tag> [bench args]
tag=${1:-x}; shift
mkdir -p gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$tag.csv \
  python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-per-function "$@" > gpurun_out/launches_$tag.log 2>&1
python tools/launch_summary.py gpurun_out/launches_$tag.csv
