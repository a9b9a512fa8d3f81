#!/bin/bash
# same-box A/B of several builds of the library on the headline bench: gpu_ab_libs.sh tag rounds lib1.so lib2.so ...
# ("-" = the in-tree library), alternating; prints ms/step and the count / fill phases.
set -u
tag=$1; rounds=$2; shift 2
out=gpurun_out; mkdir -p $out
for i in $(seq 1 $rounds); do
  for lib in "$@"; do
    if [ "$lib" = "-" ]; then unset DGNN_LIB; name=base; else export DGNN_LIB=$PWD/$lib; name=$(basename $lib .so); fi
    timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-configs --no-c5 --e2e-steps 2 > $out/${tag}_${name}_$i.json 2> $out/${tag}_${name}_$i.err
    python - "$name" $i $out/${tag}_${name}_$i.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[3]))
    print(sys.argv[1], sys.argv[2], 'ms/step', round(d['ms_per_step'],3), {k:round(x,3) for k,x in d['phases_ms'].items() if k in ('count_ms','fill_ms')})
except Exception as e: print(sys.argv[1], "failed", e)
PY
  done
done
