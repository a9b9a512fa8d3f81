#!/bin/bash
# One gpurun call: GPU parity tests, the bench line, the ncu launch list of the same bench command and one
# `ncu --set full` capture of the dominant kernel (k_direct, count + fill).  Outputs land in gpurun_out/.
#   gpurun --timeout 1500 -- 'bash tools/gpu_profile.sh <tag>'
set -u
tag=${1:-run}
out=gpurun_out
mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1
echo "pytest exit $?" | tee -a $out/${tag}_pytest.log
tail -3 $out/${tag}_pytest.log
python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err
echo "bench exit $?"
cat $out/${tag}_bench.json
python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err
cat $out/${tag}_bench_ref.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --e2e-steps 1"
$CMD > $out/${tag}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $out/${tag}_launches.csv $CMD \
    > $out/${tag}_ncu_launches.log 2>&1
echo "ncu launches exit $?"
$CMD > $out/${tag}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k "regex:^k_direct$|^k_direct_count_sym$|^k_fill_hits$" -s 6 -c 2 -f -o $out/${tag}_prof $CMD \
    > $out/${tag}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $out
