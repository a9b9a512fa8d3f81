#!/bin/bash
# ncu --set full of selected kernels of one tools/bench_configs.py config.  usage: gpu_ncu_cfg.sh tag config regex skip count
set -u
tag=$1; cfg=$2; rx=$3; skip=${4:-0}; cnt=${5:-2}
out=gpurun_out; mkdir -p $out
CMD="python tools/bench_configs.py --reps 1 --only $cfg"
$CMD > $out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:$rx" -s $skip -c $cnt -f -o $out/${tag}_prof $CMD > $out/${tag}_ncu.log 2>&1
echo "ncu exit $?"; tail -2 $out/${tag}_ncu.log
