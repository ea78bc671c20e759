#!/bin/bash
# jet (fn 4) against the three separate traversals (fn 0, 1, 2) at cfg2 / cfg3; results -> gpurun_out/jet.jsonl
set -x
python -m pytest tests/test_jet.py -m gpu -x -q 2>&1 | tail -15
: > gpurun_out/jet.jsonl
for fn in 0 1 2 4; do python tools/gpu_quick.py 2 1000000 4096 1 $fn >> gpurun_out/jet.jsonl; done
for fn in 0 1 2 4; do python tools/gpu_quick.py 3 1000000 256 1 $fn >> gpurun_out/jet.jsonl; done
cat gpurun_out/jet.jsonl | python -c "
import sys, json
for l in sys.stdin:
    d = json.loads(l); print(d['D'], d['fn'], round(d['t_pre_ms'],2), round(d['t_fwd_ms'],2), round(d['t_bwd_ms'],2), round(d['pairs_per_pt'],1))"
