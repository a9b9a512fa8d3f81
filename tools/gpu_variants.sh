#!/bin/bash
# A/B of tuning options on the headline bench.  gpurun -- 'bash tools/gpu_variants.sh tag "opt1=v opt2=v" "opt3=v" ...'
set -u
tag=${1:-var}; shift
out=gpurun_out; mkdir -p $out
i=0
for v in "" "$@"; do
  args=""
  for kv in $v; do args="$args --opt $kv"; done
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 3 $args > $out/${tag}_$i.json 2> $out/${tag}_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_$i.json"))
    print("[$v]",'ms/step',round(d['ms_per_step'],3),{k:round(x,3) for k,x in d['phases_ms'].items()},'frac',round(d['roofline']['frac'],3))
except Exception as e: print("[$v] failed",e)
PY
  i=$((i+1))
done
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "c2_perovskite_batch_uncapped or c1_shipped" > $out/${tag}_pytest.log 2>&1; tail -n 2 $out/${tag}_pytest.log
