for o in "$@"; do echo "== $o"; python tools/bench_configs.py --reps 3 --only C4 $o | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(round(d['ms_per_build'],3), d['phases_ms'])
"; done
