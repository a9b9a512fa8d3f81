#!/bin/bash
# single-pass build vs count + fill: parity tests, headline bench both ways, secondary configs both ways.
#   gpurun --timeout 1200 -- 'bash tools/gpu_onepass.sh tag'
set -u
tag=${1:-op}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $out/${tag}_pytest.log
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
timeout 400 python tools/bench_configs.py --only C1,C2cap,C3 > $out/${tag}_cfg_onepass.jsonl 2> $out/${tag}_cfg.err; echo "cfg exit $?"
timeout 400 python tools/bench_configs.py --only C1,C2cap,C3 --opt onepass=0 > $out/${tag}_cfg_twopass.jsonl 2>> $out/${tag}_cfg.err; echo "cfg exit $?"
python - <<PY
import json
for n in ("onepass","twopass"):
    for l in open("$out/${tag}_cfg_%s.jsonl"%n):
        d=json.loads(l); print(n, d['config'][:40], 'ms',round(d['ms_per_build'],3), {k:round(v,3) for k,v in d['phases_ms'].items()}, 'onepass',d.get('onepass'),'fb',d.get('onepass_fallback_rows'),'hit',d.get('hit_cap'))
PY
