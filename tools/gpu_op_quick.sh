#!/bin/bash
# quick single-pass check: the single-pass parity test + headline bench both ways.  gpurun -- 'bash tools/gpu_op_quick.sh tag'
set -u
tag=${1:-opq}
out=gpurun_out; mkdir -p $out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "single_pass or c2_perovskite_batch_uncapped" > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $out/${tag}_pytest.log
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 5 > $out/${tag}_bench_onepass.json 2> $out/${tag}_bench.err; echo "bench exit $?"
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 5 --opt onepass=0 > $out/${tag}_bench_twopass.json 2>> $out/${tag}_bench.err; echo "bench exit $?"
python - <<PY
import json
for n in ("onepass","twopass"):
    try:
        d=json.load(open("$out/${tag}_bench_%s.json"%n))
        print(n,'ms/step',round(d['ms_per_step'],3),'Gedges/s',round(d['value']/1e9,2),'e2e',round(d['e2e']['value']/1e9,2),{k:round(v,3) for k,v in d['phases_ms'].items()},'frac',round(d['roofline']['frac'],3),d.get('build'))
    except Exception as e: print(n,'failed',e)
PY
tail -3 $out/${tag}_bench.err
