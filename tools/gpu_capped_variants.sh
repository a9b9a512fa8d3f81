#!/bin/bash
# capped plan (C2 cap 20, 10 k structures) under different build plans.  gpurun -- bash tools/gpu_capped_variants.sh
for o in "" "--opt onepass=1" "--opt hit_cap=0" "--opt onepass=0 --opt hit_cap=0 --opt sym_count=0"; do
  echo "== [$o]"
  python tools/bench_configs.py --only C2cap $o 2>&1 | python -c '
import sys, json
for l in sys.stdin:
    try:
        d = json.loads(l)
    except Exception:
        print(l.strip()[:200]); continue
    print(d["config"][:40], round(d["ms_per_build"], 3), d["phases_ms"], "onepass", d["onepass"], "hit_cap", d["hit_cap"], round(d["frac_of_measured_peak"], 3))
'
done
