#!/bin/bash
# N-rank strong-scaling check of the bench (cfg2 and cfg3) with the untimed multi-GPU parity self-check: bash tools/gpu_n2.sh <N> <tag>
N=${1:-2}; tag=${2:-x}
mkdir -p gpurun_out
for w in cfg2 cfg3; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus $N --steps 5 --warmup 3 --workload $w --no-cpu-baseline --no-per-function > gpurun_out/bench_${w}_n${N}_$tag.log 2>&1
  echo "rc=$? $w"; tail -1 gpurun_out/bench_${w}_n${N}_$tag.log | cut -c1-3000
done
