#!/usr/bin/env bash
# One gpurun call: GPU parity tests, quick probes, the bench (both arms), the ncu launch list.
# Every step has its own timeout; logs land in gpurun_out/.
set -u
mkdir -p gpurun_out
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/smi.txt 2>&1
echo "== pytest gpu"; timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "rc=$?"; tail -5 $O/pytest_gpu.log
echo "== quick cfg1"; timeout 300 python tools/gpu_quick.py 2 10000 256 > $O/quick_cfg1.log 2>&1; echo "rc=$?"; tail -2 $O/quick_cfg1.log
echo "== quick cfg2"; timeout 600 python tools/gpu_quick.py 2 1000000 4096 > $O/quick_cfg2.log 2>&1; echo "rc=$?"; tail -2 $O/quick_cfg2.log
echo "== quick cfg3"; timeout 600 python tools/gpu_quick.py 3 1000000 256 > $O/quick_cfg3.log 2>&1; echo "rc=$?"; tail -2 $O/quick_cfg3.log
echo "== bench ours"; timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_ours.log 2>&1; echo "rc=$?"; tail -2 $O/bench_ours.log
echo "== bench reference"; timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > $O/bench_ref.log 2>&1; echo "rc=$?"; tail -2 $O/bench_ref.log
echo "== ncu launch list"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $O/ncu_launches.log 2>&1; echo "rc=$?"
