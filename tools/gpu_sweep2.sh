#!/usr/bin/env bash
# occupancy variants (launch-bounds min blocks) and forced T1 in 3D
set -u
mkdir -p gpurun_out
O=gpurun_out/sweep2.jsonl
: > $O
P=$PWD/diff_gaussian_sampling_b200
for v in libdgs_b200 libdgs_v_f4 libdgs_v_f6 libdgs_v_f8; do
  for cfg in "2 1000000 4096" "3 1000000 256"; do
    echo "$v $cfg"; DGS_LIB_PATH=$P/$v.so timeout 300 python tools/gpu_quick.py $cfg 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"$v\", /" >> $O
  done
done
for t1 in 4 5 6; do
  echo "cfg3 T1 $t1"; DGS_T1=$t1 timeout 300 python tools/gpu_quick.py 3 1000000 256 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"T1=$t1\", /" >> $O
done
for t1 in 4 5; do
  echo "cfg3 T1 $t1 fpc3"; DGS_T1=$t1 DGS_TILE_TARGET=100 timeout 300 python tools/gpu_quick.py 3 1000000 256 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"T1=$t1,target100\", /" >> $O
done
python - <<'PY'
import json
for l in open("gpurun_out/sweep2.jsonl"):
    d = json.loads(l)
    print(d['D'], d['variant'], 'T1', d['T1'], 'fpc', d['fpc'], 'pre %.2f fwd %.2f bwd %.2f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']), 'pairs/pt %.0f' % d['pairs_per_pt'], 'n_l1', d['n_l1'])
PY
