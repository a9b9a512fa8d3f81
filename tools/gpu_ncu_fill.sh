#!/bin/bash
# ncu --set full of the dominant kernels of a warmed-up bench run.  usage: gpu_ncu_fill.sh tag [kernel-regex] [skip] [count]
set -u
tag=${1:-run}; rx=${2:-'^k_direct$|^k_fill_hits$'}; skip=${3:-6}; cnt=${4:-2}
out=gpurun_out
mkdir -p $out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$CMD > $out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$rx" -s $skip -c $cnt -f -o $out/${tag}_prof $CMD \
    > $out/${tag}_ncu_full.log 2>&1
echo "ncu full exit $?"
