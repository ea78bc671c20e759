#!/bin/bash
python -m pytest tests/test_gpu_parity.py tests/test_jet.py -m gpu -x -q 2>&1 | tail -3
: > gpurun_out/exp_cb3.jsonl
for fn in 0 1 2 4; do python tools/gpu_quick.py 2 1000000 4096 3 $fn >> gpurun_out/exp_cb3.jsonl; done
python tools/gpu_quick.py 3 1000000 256 3 0 >> gpurun_out/exp_cb3.jsonl
python tools/gpu_quick.py 3 1000000 256 3 1 >> gpurun_out/exp_cb3.jsonl
python tools/gpu_quick.py 2 1000000 4096 1 0 >> gpurun_out/exp_cb3.jsonl
python -c "
import json
for l in open('gpurun_out/exp_cb3.jsonl'):
    d = json.loads(l); print('D', d['D'], 'C', d['C'], 'fn', d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2))"
