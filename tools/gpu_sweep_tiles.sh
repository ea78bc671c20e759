#!/bin/bash
# Tile-hierarchy sweep (coarse tiles per cell axis T1, fine tiles per coarse-tile axis fpc) at cfg2 and cfg3: total of
# preprocess + forward + backward per setting.  bash tools/gpu_sweep_tiles.sh <tag>
tag=${1:-x}; O=gpurun_out/sweep_tiles_$tag.jsonl; : > $O
for cfg in "2 1000000 4096 19 4" "2 1000000 4096 25 3" "2 1000000 4096 15 5" "2 1000000 4096 13 6" "2 1000000 4096 38 2" "2 1000000 4096 22 4" "2 1000000 4096 17 4" \
           "3 1000000 256 5 3" "3 1000000 256 7 2" "3 1000000 256 4 4" "3 1000000 256 8 2" "3 1000000 256 6 3" "3 1000000 256 4 3" "3 1000000 256 6 2"; do
  set -- $cfg
  DGS_T1=$4 DGS_FPC=$5 timeout 300 python tools/gpu_quick.py $1 $2 $3 >> $O 2>/dev/null || echo "{\"failed\": \"$cfg\"}" >> $O
done
python - <<PY
import json
for l in open("$O"):
    d=json.loads(l)
    if 'failed' in d: print(d); continue
    print(d['D'], 'T1',d['T1'],'fpc',d['fpc'], 'pre %.3f fwd %.3f bwd %.3f total %.3f'%(d['t_pre_ms'],d['t_fwd_ms'],d['t_bwd_ms'],d['t_pre_ms']+d['t_fwd_ms']+d['t_bwd_ms']), 'n_l1',d['n_l1'],'ppp %.1f'%d['pairs_per_pt'])
PY
