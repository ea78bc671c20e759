#!/bin/bash
# ncu --set full + source counters on the preprocess kernels (one launch each) at cfg2: bash tools/gpu_profile_pre.sh <tag> [D P n]
tag=${1:-pre}; D=${2:-2}; P=${3:-1000000}; n=${4:-4096}
mkdir -p gpurun_out
timeout 900 ncu --set full --import-source on --clock-control none \
  -k regex:'dgs_l1_kernel|dgs_l1_sort|dgs_l1_mask|dgs_sample_keys|dgs_fine_scatter|radix_scatter' --launch-skip 8 -c 9 \
  -o gpurun_out/ncu_$tag -f python tools/gpu_quick.py $D $P $n > gpurun_out/ncu_$tag.log 2>&1
ncu -i gpurun_out/ncu_$tag.ncu-rep --page raw --csv > gpurun_out/ncu_${tag}_raw.csv 2>/dev/null
ncu -i gpurun_out/ncu_$tag.ncu-rep --page source --csv --print-source sass > gpurun_out/ncu_${tag}_source_sass.csv 2>/dev/null
ls -la gpurun_out/ncu_$tag*
