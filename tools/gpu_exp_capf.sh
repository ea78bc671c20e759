#!/bin/bash
: > gpurun_out/exp_capf.jsonl
python tools/gpu_quick.py 2 1000000 4096 1 0 | sed 's/^{/{"lib": "default", /' >> gpurun_out/exp_capf.jsonl
for v in 96 128; do DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_c$v.so python tools/gpu_quick.py 2 1000000 4096 1 0 | sed "s/^{/{\"lib\": \"c$v\", /" >> gpurun_out/exp_capf.jsonl; done
for v in 96 128; do DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_c$v.so python tools/gpu_quick.py 3 1000000 256 1 0 | sed "s/^{/{\"lib\": \"c$v\", /" >> gpurun_out/exp_capf.jsonl; done
python -c "
import json
for l in open('gpurun_out/exp_capf.jsonl'):
    d = json.loads(l); print(d['lib'], d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],3), round(d['t_bwd_ms'],2))"
