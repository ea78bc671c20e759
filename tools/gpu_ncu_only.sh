#!/usr/bin/env bash
# The two ncu passes of tools/gpu_measure.sh alone (launch list of one bench step; --set full of the accumulation kernels).
set -u
tag=${1:-x}; O=gpurun_out; mkdir -p $O
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-per-function"
timeout 600 $CMD > $O/plain_$tag.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_$tag.csv $CMD > $O/ncu_launches_$tag.log 2>&1
echo "launch list rc=$?"; python tools/launch_summary.py $O/launches_$tag.csv > $O/launches_${tag}_one_step.txt 2>&1; tail -1 $O/launches_${tag}_one_step.txt
timeout 600 $CMD > $O/plain2_$tag.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'dgs_(forward|backward)_kernel' -s 6 -c 2 -f -o $O/ncu_full_$tag $CMD > $O/ncu_full_$tag.log 2>&1
echo "ncu full rc=$?"
ncu -i $O/ncu_full_$tag.ncu-rep --page source --csv --print-source sass > $O/ncu_full_${tag}_source_sass.csv 2>/dev/null
python tools/ncu_blocks.py $O/ncu_full_${tag}_source_sass.csv 10 > $O/ncu_full_${tag}_blocks.txt 2>&1
