#!/bin/bash
# same-box A/B of two builds of the library: default vs $1 (path of the alternative .so), alternating, headline bench.
set -u
alt=$1; tag=${2:-ab}
out=gpurun_out; mkdir -p $out
for i in 1 2 3; do
  for v in base alt; do
    if [ $v = alt ]; then export DGNN_LIB=$alt; else unset DGNN_LIB; fi
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs --no-c5 --e2e-steps 2 > $out/${tag}_${v}_$i.json 2> $out/${tag}_${v}_$i.err
    python - <<PY
import json
try:
    d=json.load(open("$out/${tag}_${v}_$i.json"))
    print("$v $i",'ms/step',round(d['ms_per_step'],3),{k:round(x,3) for k,x in d['phases_ms'].items() if k in ('count_ms','fill_ms')})
except Exception as e: print("$v failed",e)
PY
  done
done
