#!/bin/bash
# A/B of tuning options on the headline bench.  gpurun -- 'bash tools/gpu_variants.sh tag "opt1=v opt2=v" "opt3=v" ...'
set -u
tag=${1:-var}; shift
out=gpurun_out; mkdir -p $out
i=0
for v in "" "$@"; do
  args=""
  for kv in $v; do args="$args --opt $kv"; done
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 10 $args > $out/${tag}_$i.json 2> $out/${tag}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_$i.json"))
    print("[$v]",'ms/step',round(d['ms_per_step'],3),{k:round(x,3) for k,x in d['phases_ms'].items()},'frac',round(d['roofline']['frac'],3),'| e2e ms',round(d['e2e']['ms_per_step'],3),{k:round(x,3) for k,x in d['e2e'].get('phases_ms',{}).items()})
except Exception as e: print("[$v] failed",e)
PY
  i=$((i+1))
done
