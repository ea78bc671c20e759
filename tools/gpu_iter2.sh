#!/usr/bin/env bash
# One gpurun call per iteration: GPU parity tests (optionally a subset: PYTEST_ARGS), then the quick probes with in-library
# per-kernel timing.  Logs in gpurun_out/.  Usage: bash tools/gpu_iter2.sh <tag>
set -u
tag=${1:-x}
mkdir -p gpurun_out
O=gpurun_out
if [ "${SKIP_TESTS:-0}" = "0" ]; then
echo "== pytest gpu"; timeout 1500 python -m pytest tests -m gpu -q -x ${PYTEST_ARGS:-} > $O/pytest_gpu_$tag.log 2>&1; echo "rc=$?"; tail -8 $O/pytest_gpu_$tag.log
fi
echo "== quick cfg2"; DGS_QUICK_PROFILE=1 timeout 600 python tools/gpu_quick.py 2 1000000 4096 > $O/quick_cfg2_$tag.json 2> $O/quick_cfg2_$tag.err; echo "rc=$?"; tail -1 $O/quick_cfg2_$tag.json
echo "== quick cfg3"; DGS_QUICK_PROFILE=1 timeout 600 python tools/gpu_quick.py 3 1000000 256 > $O/quick_cfg3_$tag.json 2> $O/quick_cfg3_$tag.err; echo "rc=$?"; tail -1 $O/quick_cfg3_$tag.json
if [ "${WITH_CFG1:-0}" = "1" ]; then
echo "== quick cfg1"; DGS_QUICK_PROFILE=1 timeout 300 python tools/gpu_quick.py 2 10000 256 > $O/quick_cfg1_$tag.json 2> $O/quick_cfg1_$tag.err; echo "rc=$?"; tail -1 $O/quick_cfg1_$tag.json
fi
