#!/bin/bash
# quick GPU check: parity tests + one bench line (no CPU baseline leg).  gpurun --timeout 600 -- 'bash tools/gpu_quick.sh tag [bench args]'
set -u
tag=${1:-q}; shift || true
out=gpurun_out; mkdir -p $out
python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -5 $out/${tag}_pytest.log
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 3 "$@" > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"
python -c "
import json,sys
d=json.load(open('$out/${tag}_bench.json'))
print('ms/step', d['ms_per_step'], 'Gedges/s', d['value']/1e9, 'e2e', d['e2e']['value']/1e9, d['phases_ms'], 'frac', d['roofline']['frac'])
" || tail -5 $out/${tag}_bench.err
