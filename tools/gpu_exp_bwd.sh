#!/bin/bash
# moments-form backward of the derivative orders: parity tests + timings (2D cfg2 fn 1,2,3,4; 3D cfg3 fn 1,2,4)
python -m pytest tests/test_gpu_parity.py tests/test_jet.py -m gpu -x -q 2>&1 | tail -5
: > gpurun_out/exp_bwd.jsonl
for fn in 1 2 3 4; do python tools/gpu_quick.py 2 1000000 4096 1 $fn >> gpurun_out/exp_bwd.jsonl; done
for fn in 1 2 4; do python tools/gpu_quick.py 3 1000000 256 1 $fn >> gpurun_out/exp_bwd.jsonl; done
python -c "
import json
for l in open('gpurun_out/exp_bwd.jsonl'):
    d = json.loads(l); print(d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
