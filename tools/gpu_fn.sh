#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/fn.jsonl
: > $O
for cfg in "2 1000000 4096 1 0" "2 1000000 4096 1 1" "2 1000000 4096 1 2" "2 1000000 4096 1 3" "3 1000000 256 1 0" "3 1000000 256 1 1" "3 1000000 256 1 2" "2 1000000 4096 4 0" "2 10000 256 1 0"; do
timeout 300 python tools/gpu_quick.py $cfg 2>&1 | grep '^{' >> $O
done
python - <<'PY'
import json
for l in open("gpurun_out/fn.jsonl"):
    d = json.loads(l)
    print(d['D'], d['P'], 'C', d['C'], 'fn', d['fn'], 'T1', d['T1'], 'fpc', d['fpc'], 'pre %.2f fwd %.2f bwd %.2f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']), 'pairs/pt %.0f' % d['pairs_per_pt'])
PY
