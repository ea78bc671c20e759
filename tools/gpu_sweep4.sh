#!/usr/bin/env bash
set -u
mkdir -p gpurun_out
O=gpurun_out/sweep4.jsonl
: > $O
P=$PWD/diff_gaussian_sampling_b200
echo "== covariance + aggregate tests"; timeout 600 python -m pytest tests/test_covariance.py tests/test_aggregate.py -m gpu -q 2>&1 | tail -4
for v in libdgs_b200 libdgs_v_b7 libdgs_v_b9 libdgs_v_b10 libdgs_v_f5 libdgs_v_f7 libdgs_b200; do
  DGS_LIB_PATH=$P/$v.so timeout 300 python tools/gpu_quick.py 2 1000000 4096 2>&1 | grep '^{' | sed "s/^{/{\"variant\": \"$v\", /" >> $O
done
python - <<'PY'
import json
for l in open("gpurun_out/sweep4.jsonl"):
    d = json.loads(l)
    print(d['D'], d['variant'], 'pre %.2f fwd %.3f bwd %.3f' % (d['t_pre_ms'], d['t_fwd_ms'], d['t_bwd_ms']))
PY
echo "== aggregate timing"
timeout 300 python tools/gpu_aggregate.py 3 1000000 > gpurun_out/agg_cfg4.json 2> gpurun_out/agg_cfg4.err; tail -1 gpurun_out/agg_cfg4.json; tail -2 gpurun_out/agg_cfg4.err
timeout 300 python tools/gpu_aggregate.py 2 1000000 > gpurun_out/agg_2d_1m.json 2>> gpurun_out/agg_cfg4.err; tail -1 gpurun_out/agg_2d_1m.json
timeout 600 python tools/gpu_aggregate.py 2 30000 16 16 4 --ref > gpurun_out/agg_vs_ref.json 2>> gpurun_out/agg_cfg4.err; tail -1 gpurun_out/agg_vs_ref.json
echo "== cfg5 per-rank stress (P=8M, 16.7M points, 3D)"
timeout 600 python tools/gpu_quick.py 3 8000000 256 > gpurun_out/quick_cfg5rank.log 2>&1; tail -1 gpurun_out/quick_cfg5rank.log
