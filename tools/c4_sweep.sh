# usage: c4_sweep.sh "opts variant" ...   (each variant: space-separated name=value dgnn options)
for v in "$@"; do
  o=""; for kv in $v; do [ "$kv" != "-" ] && o="$o --opt $kv"; done
  echo "== $v"; python tools/bench_configs.py --only C4 --reps 3 $o 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print(round(d['ms_per_build'],3), {k:round(v,3) for k,v in d['phases_ms'].items()}, 'frac',round(d['frac_of_measured_peak'],3))"
done
