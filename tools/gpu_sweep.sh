#!/usr/bin/env bash
# Sweep of the tile-size knob (samples per fine tile) on cfg2 and cfg3; one JSON line per setting.
set -u
mkdir -p gpurun_out
O=gpurun_out/sweep.jsonl
: > $O
for t in 96 128 160 192 256; do
  echo "cfg2 target $t"; DGS_TILE_TARGET=$t timeout 300 python tools/gpu_quick.py 2 1000000 4096 2>&1 | grep '^{' | sed "s/^{/{\"target\": $t, /" >> $O
done
for t in 64 96 128 176 256; do
  echo "cfg3 target $t"; DGS_TILE_TARGET=$t timeout 300 python tools/gpu_quick.py 3 1000000 256 2>&1 | grep '^{' | sed "s/^{/{\"target\": $t, /" >> $O
done
cat $O | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l)
    print(d['D'], d['target'], 'T1', d['T1'], 'fpc', d['fpc'], 'pre %.2f fwd %.2f bwd %.2f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']), 'pairs/pt %.0f' % d['pairs_per_pt'], 'n_l1', d['n_l1'])
"
