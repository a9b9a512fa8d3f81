#!/bin/bash
# full evidence pass: all GPU tests, headline bench (+ reference arm), secondary configs.  gpurun -- 'bash tools/gpu_round.sh tag'
set -u
tag=${1:-rnd}
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 $out/${tag}_pytest.log
timeout 400 python bench.py --steps 10 --warmup 3 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"
timeout 400 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err
timeout 600 python tools/bench_configs.py > $out/${tag}_cfg.jsonl 2> $out/${tag}_cfg.err; echo "cfg exit $?"
python - <<PY
import json
d=json.load(open("$out/${tag}_bench.json"))
print('ms/step',round(d['ms_per_step'],3),'Gedges/s',round(d['value']/1e9,2),'e2e',round(d['e2e']['value']/1e9,2),{k:round(v,3) for k,v in d['phases_ms'].items()},'frac',round(d['roofline']['frac'],3),d.get('build'))
for l in open("$out/${tag}_cfg.jsonl"):
    d=json.loads(l)
    if 'device_ms' in d: print(d['config'][:60],'ms',round(d['device_ms'],3),'GB/s',round(d['achieved_gbs'],1))
    elif 'ms_per_call' in d: print(d['config'][:50],'onepass',d['onepass'],'ms/call',round(d['ms_per_call'],3))
    else: print(d['config'][:50],'ms',round(d['ms_per_build'],3),{k:round(v,3) for k,v in d['phases_ms'].items()},'onepass',d.get('onepass'),'hit',d.get('hit_cap'))
PY
tail -n 3 $out/${tag}_bench.err; tail -n 3 $out/${tag}_cfg.err; true
