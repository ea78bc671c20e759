#!/bin/bash
# occupancy sweep of the 3D gaussian kernels (cfg3) + N=2 bench with pipelined e2e is run separately
: > gpurun_out/exp3d.jsonl
python tools/gpu_quick.py 3 1000000 256 1 0 | sed 's/^{/{"lib": "default", /' >> gpurun_out/exp3d.jsonl
for v in 4 6 7; do DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_t$v.so python tools/gpu_quick.py 3 1000000 256 1 0 | sed "s/^{/{\"lib\": \"t$v\", /" >> gpurun_out/exp3d.jsonl; done
python -c "
import json
for l in open('gpurun_out/exp3d.jsonl'):
    d = json.loads(l); print(d['lib'], d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
