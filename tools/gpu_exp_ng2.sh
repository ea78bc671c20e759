#!/bin/bash
: > gpurun_out/exp_ng2.jsonl
for fn in 1 2 3 4; do DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_ng2.so python tools/gpu_quick.py 2 1000000 4096 1 $fn | sed 's/^{/{"lib": "ng2", /' >> gpurun_out/exp_ng2.jsonl; done
DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_ng2.so python tools/gpu_quick.py 1 1000000 4194304 1 1 | sed 's/^{/{"lib": "ng2", /' >> gpurun_out/exp_ng2.jsonl
python tools/gpu_quick.py 1 1000000 4194304 1 1 | sed 's/^{/{"lib": "default", /' >> gpurun_out/exp_ng2.jsonl
python -c "
import json
for l in open('gpurun_out/exp_ng2.jsonl'):
    d = json.loads(l); print(d['lib'], d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
