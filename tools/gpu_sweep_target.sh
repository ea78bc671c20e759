#!/bin/bash
# Sweep of the samples-per-fine-tile target (DGS_TILE_TARGET) with today's kernels.  bash tools/gpu_sweep_target.sh <tag>
tag=${1:-x}; O=gpurun_out/sweep_target_$tag.jsonl; : > $O
for t in 128 160 176 192 208 224 256; do DGS_TILE_TARGET=$t timeout 300 python tools/gpu_quick.py 2 1000000 4096 2>/dev/null | tail -1 | sed "s/^{/{\"target\": $t, /" >> $O; done
for t in 80 96 110 128 150 176 220; do DGS_TILE_TARGET=$t timeout 300 python tools/gpu_quick.py 3 1000000 256 2>/dev/null | tail -1 | sed "s/^{/{\"target\": $t, /" >> $O; done
python - <<PY
import json
for l in open("$O"):
    d=json.loads(l)
    print(d['D'], 'target', d['target'], 'T1',d['T1'],'fpc',d['fpc'], 'pre %.3f fwd %.3f bwd %.3f total %.3f'%(d['t_pre_ms'],d['t_fwd_ms'],d['t_bwd_ms'],d['t_pre_ms']+d['t_fwd_ms']+d['t_bwd_ms']), 'ppp %.1f'%d['pairs_per_pt'])
PY
