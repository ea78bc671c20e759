#!/bin/bash
# ncu --set full with source counters on the two accumulation kernels (one launch each) at cfg2; the report and the
# per-line / raw CSV pages land in gpurun_out/.  Usage: bash tools/gpu_profile_acc.sh <tag> [D P n]
tag=${1:-acc}; D=${2:-2}; P=${3:-1000000}; n=${4:-4096}
mkdir -p gpurun_out
python tools/gpu_quick.py $D $P $n > gpurun_out/quick_$tag.jsonl 2> gpurun_out/quick_$tag.err || { tail -5 gpurun_out/quick_$tag.err; exit 1; }
cat gpurun_out/quick_$tag.jsonl
timeout 900 ncu --set full --import-source on --clock-control none -k regex:'dgs_forward_kernel|dgs_backward_kernel' --launch-skip 2 -c 2 \
  -o gpurun_out/ncu_$tag -f python tools/gpu_quick.py $D $P $n > gpurun_out/ncu_$tag.log 2>&1
ncu -i gpurun_out/ncu_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/ncu_$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_${tag}_source_sass.csv 2>/dev/null
ls -la gpurun_out/ncu_$tag*
