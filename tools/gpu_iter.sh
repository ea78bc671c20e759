#!/usr/bin/env bash
# One gpurun call per iteration: parity tests, probes, bench, ncu launch list + full capture of the
# two accumulation kernels.  Logs in gpurun_out/.
set -u
mkdir -p gpurun_out
O=gpurun_out
echo "== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; rc=$?; echo "rc=$rc"; tail -15 $O/pytest_gpu.log
echo "== quick cfg1"; timeout 300 python tools/gpu_quick.py 2 10000 256 > $O/quick_cfg1.log 2>&1; echo "rc=$?"; tail -1 $O/quick_cfg1.log
echo "== quick cfg2"; timeout 600 python tools/gpu_quick.py 2 1000000 4096 > $O/quick_cfg2.log 2>&1; echo "rc=$?"; tail -1 $O/quick_cfg2.log
echo "== quick cfg3"; timeout 600 python tools/gpu_quick.py 3 1000000 256 > $O/quick_cfg3.log 2>&1; echo "rc=$?"; tail -1 $O/quick_cfg3.log
if [ "${SKIP_BENCH:-0}" = "0" ]; then
echo "== bench ours"; timeout 900 python bench.py --steps 5 --warmup 3 > $O/bench_ours.log 2>&1; echo "rc=$?"; tail -1 $O/bench_ours.log
fi
if [ "${WITH_NCU:-0}" = "1" ]; then
echo "== ncu"
CMD="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --no-per-function"
timeout 600 $CMD > $O/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches.csv $CMD > $O/ncu_launches.log 2>&1
echo "rc=$?"
timeout 600 $CMD > $O/plain2.log 2>&1 &&
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:'dgs_(forward|backward)_kernel' -s 6 -c 2 -o $O/prof_fwd_bwd $CMD > $O/ncu_full.log 2>&1
echo "rc=$?"; tail -3 $O/ncu_full.log
fi
