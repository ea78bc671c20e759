#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/sweep3.jsonl
: > $O
echo "== pytest gpu"; timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for f in 4 3 2; do
  DGS_FPC=$f timeout 300 python tools/gpu_quick.py 2 1000000 4096 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"fpc=$f\", /" >> $O
done
timeout 300 python tools/gpu_quick.py 3 1000000 256 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"default\", /" >> $O
timeout 300 python tools/gpu_quick.py 2 10000 256 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"default\", /" >> $O
for fn in 1 2 3; do
timeout 300 python tools/gpu_quick.py 2 1000000 4096 1 $fn 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"fn=$fn\", /" >> $O
done
python - <<'PY'
import json
for l in open("gpurun_out/sweep3.jsonl"):
    d = json.loads(l)
    print(d['D'], d['P'], d['variant'], 'T1', d['T1'], 'fpc', d['fpc'], 'pre %.2f fwd %.2f bwd %.2f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']), 'pairs/pt %.0f' % d['pairs_per_pt'], 'n_l1', d['n_l1'])
PY
