#!/bin/bash
# `ncu --set full` capture of kernels of the bench.py command (C2 headline), summarised on the box.
# usage: gpu_ncu_bench.sh tag regex skip count linefilter ["--opt a=b --opt c=d"]
set -u
tag=$1; rx=$2; skip=${3:-0}; cnt=${4:-1}; filt=${5:-}; opts=${6:-}
out=gpurun_out; mkdir -p $out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1 $opts"
timeout 300 $CMD > $out/${tag}_plain.log 2>&1 || { echo "plain failed"; tail -3 $out/${tag}_plain.log; exit 1; }
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$rx" -s $skip -c $cnt -f -o /tmp/${tag} $CMD > $out/${tag}_ncu.log 2>&1
echo "ncu exit $?"; tail -1 $out/${tag}_ncu.log
python tools/ncu_summary.py /tmp/${tag}.ncu-rep > $out/${tag}_full.txt 2>&1
[ -n "$filt" ] && python tools/ncu_by_line.py /tmp/${tag}.ncu-rep "$filt" 60 > $out/${tag}_lines.txt 2>&1
[ -n "$filt" ] && python tools/ncu_sass_hot.py /tmp/${tag}.ncu-rep "$filt" 40 > $out/${tag}_sass.txt 2>&1
rm -f /tmp/${tag}.ncu-rep
