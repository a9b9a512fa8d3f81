#!/bin/bash
# ncu --set full of the dominant kernel only (k_direct, one count + one fill launch of a warmed-up bench run).
set -u
tag=${1:-run}
out=gpurun_out
mkdir -p $out
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$CMD > $out/${tag}_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k 'regex:^k_direct$' -s 6 -c 2 -f -o $out/${tag}_prof $CMD \
    > $out/${tag}_ncu_full.log 2>&1
echo "ncu full exit $?"
bash tools/sweep.sh "fuse_rbf=1" "fuse_rbf=0" > $out/${tag}_sweep.log 2>&1
cat $out/${tag}_sweep.log
