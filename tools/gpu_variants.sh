#!/bin/bash
# Times tuning builds of the library (diff_gaussian_sampling_b200/_variants/libdgs_v*.so, selected through DGS_LIB_PATH) with
# the quick probe at cfg2 and cfg3.  bash tools/gpu_variants.sh <tag>
tag=${1:-x}; O=gpurun_out/variants_$tag.jsonl; : > $O
for so in "" diff_gaussian_sampling_b200/_variants/libdgs_v*.so; do
  for cfg in "2 1000000 4096" "3 1000000 256"; do
    if [ -z "$so" ]; then r=$(timeout 300 python tools/gpu_quick.py $cfg 2>/dev/null | tail -1); name=default
    else r=$(DGS_LIB_PATH=$PWD/$so timeout 300 python tools/gpu_quick.py $cfg 2>/dev/null | tail -1); name=$(basename $so .so); fi
    echo "$r" | sed "s/^{/{\"lib\": \"$name\", /" >> $O
  done
done
python - <<PY
import json
for l in open("$O"):
    try: d=json.loads(l)
    except Exception: print('bad', l[:80]); continue
    print(d['lib'], d['D'], 'pre %.3f fwd %.3f bwd %.3f'%(d['t_pre_ms'],d['t_fwd_ms'],d['t_bwd_ms']))
PY
