#!/bin/bash
# Like gpu_variants.sh, but prints the per-kernel preprocess times (in-library profiling).  bash tools/gpu_variants_pre.sh <tag>
tag=${1:-x}; O=gpurun_out/variants_pre_$tag.jsonl; : > $O
for so in "" diff_gaussian_sampling_b200/_variants/libdgs_v*.so; do
  for cfg in "2 1000000 4096" "3 1000000 256"; do
    if [ -z "$so" ]; then r=$(DGS_QUICK_PROFILE=1 timeout 300 python tools/gpu_quick.py $cfg 2>/dev/null | tail -1); name=default
    else r=$(DGS_QUICK_PROFILE=1 DGS_LIB_PATH=$PWD/$so timeout 300 python tools/gpu_quick.py $cfg 2>/dev/null | tail -1); name=$(basename $so .so); fi
    echo "$r" | sed "s/^{/{\"lib\": \"$name\", /" >> $O
  done
done
python - <<PY
import json
for l in open("$O"):
    try: d=json.loads(l)
    except Exception: print('bad', l[:80]); continue
    k=d['kernel_ms']
    print(d['lib'], d['D'], 'pre %.3f'%d['t_pre_ms'], 'l1_count %.3f l1_emit %.3f l1_sort %.3f masks %.3f'%(k['l1_count'],k['l1_emit'],k['l1_sort'],k['l1_masks']))
PY
