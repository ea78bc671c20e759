#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/sweep5.jsonl
: > $O
P=$PWD/diff_gaussian_sampling_b200
for v in libdgs_b200 libdgs_v_p0 libdgs_v_u4 libdgs_v_u4b8 libdgs_v_u2b8 libdgs_v_u4b6; do
  for cfg in "2 1000000 4096" "3 1000000 256"; do
  DGS_LIB_PATH=$P/$v.so timeout 300 python tools/gpu_quick.py $cfg 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"$v\", /" >> $O
  done
done
python - <<'PY'
import json
for l in open("gpurun_out/sweep5.jsonl"):
    d = json.loads(l)
    print(d['D'], d['variant'], 'pre %.2f fwd %.3f bwd %.3f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']))
PY
