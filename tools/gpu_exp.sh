#!/bin/bash
# experiment: branch-free mask kernel (default lib) and two Gaussians per lane for derivative orders (ng2 variant)
python -m pytest tests/test_gpu_parity.py tests/test_jet.py -m gpu -x -q 2>&1 | tail -3
: > gpurun_out/exp.jsonl
for fn in 0 1 2 4; do python tools/gpu_quick.py 2 1000000 4096 1 $fn | sed 's/^{/{"lib": "default", /' >> gpurun_out/exp.jsonl; done
python tools/gpu_quick.py 3 1000000 256 1 0 | sed 's/^{/{"lib": "default", /' >> gpurun_out/exp.jsonl
for fn in 1 2 4; do DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_ng2.so python tools/gpu_quick.py 2 1000000 4096 1 $fn | sed 's/^{/{"lib": "ng2", /' >> gpurun_out/exp.jsonl; done
DGS_LIB_PATH=$PWD/diff_gaussian_sampling_b200/libdgs_b200_ng2.so python -m pytest tests/test_jet.py -m gpu -x -q 2>&1 | tail -3
python -c "
import sys, json
for l in open('gpurun_out/exp.jsonl'):
    d = json.loads(l); print(d['lib'], d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
