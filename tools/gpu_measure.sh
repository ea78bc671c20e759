#!/usr/bin/env bash
# Round measurement at N = 1: default bench line, reference arm, cfg1/cfg3/cfg4 lines, ncu launch list of one bench step and
# ncu --set full of the two accumulation kernels (each ncu pass only after the same command has exited 0 without ncu).
# Usage: bash tools/gpu_measure.sh <tag>
set -u
tag=${1:-x}; O=gpurun_out; mkdir -p $O
timeout 900 python bench.py > $O/bench_cfg2_$tag.json 2> $O/bench_cfg2_$tag.err; echo "bench cfg2 rc=$?"; tail -c 600 $O/bench_cfg2_$tag.json | head -c 300; echo
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $O/bench_reference_$tag.json 2> $O/bench_reference_$tag.err; echo "reference rc=$?"
for w in cfg1 cfg3 cfg4; do
  timeout 900 python bench.py --workload $w --steps 5 --warmup 3 > $O/bench_${w}_$tag.json 2> $O/bench_${w}_$tag.err; echo "bench $w rc=$?"
done
timeout 600 python bench.py --workload cfg1 --impl reference --steps 5 --warmup 3 > $O/bench_cfg1_reference_$tag.json 2> $O/bench_cfg1_reference_$tag.err; echo "cfg1 reference rc=$?"
CMD="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e --no-per-function"
timeout 600 $CMD > $O/plain_$tag.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/launches_$tag.csv $CMD > $O/ncu_launches_$tag.log 2>&1
echo "launch list rc=$?"; python tools/launch_summary.py $O/launches_$tag.csv > $O/launches_${tag}_one_step.txt 2>&1; tail -1 $O/launches_${tag}_one_step.txt
timeout 600 $CMD > $O/plain2_$tag.log 2>&1 && timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'dgs_(forward|backward)_kernel' -s 6 -c 2 -f -o $O/ncu_full_$tag $CMD > $O/ncu_full_$tag.log 2>&1
echo "ncu full rc=$?"
python tools/ncu_summary.py $O/ncu_full_$tag.ncu-rep $O/ncu_full_$tag.json > $O/ncu_summary_$tag.log 2>&1; echo "summary rc=$?"
ncu -i $O/ncu_full_$tag.ncu-rep --page source --csv --print-source sass > $O/ncu_full_${tag}_source_sass.csv 2>/dev/null
python tools/ncu_blocks.py $O/ncu_full_${tag}_source_sass.csv 10 > $O/ncu_full_${tag}_blocks.txt 2>&1
