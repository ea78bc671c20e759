#!/bin/bash
python -m pytest tests/test_gpu_parity.py tests/test_jet.py -m gpu -x -q 2>&1 | tail -3
: > gpurun_out/exp_mc.jsonl
for fn in 1 2 4; do
python tools/gpu_quick.py 2 1000000 4096 3 $fn | sed 's/^{/{"lib": "moments", /' >> gpurun_out/exp_mc.jsonl
DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_m0.so python tools/gpu_quick.py 2 1000000 4096 3 $fn | sed 's/^{/{"lib": "per-pair", /' >> gpurun_out/exp_mc.jsonl
done
python tools/gpu_quick.py 3 1000000 256 4 1 | sed 's/^{/{"lib": "moments", /' >> gpurun_out/exp_mc.jsonl
DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_m0.so python tools/gpu_quick.py 3 1000000 256 4 1 | sed 's/^{/{"lib": "per-pair", /' >> gpurun_out/exp_mc.jsonl
python -c "
import json
for l in open('gpurun_out/exp_mc.jsonl'):
    d = json.loads(l); print(d['lib'], 'D', d['D'], 'C', d['C'], 'fn', d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
