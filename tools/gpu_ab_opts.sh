#!/bin/bash
# same-box A/B of bench.py under different dgnn_set_option settings.  usage: gpu_ab_opts.sh tag "name=v name2=v" "..." ...
# ("-" = defaults).  One JSON line per variant in gpurun_out/<tag>_ab.jsonl + a summary on stdout.
set -u
tag=$1; shift
out=gpurun_out; mkdir -p $out
: > $out/${tag}_ab.jsonl
for v in "$@"; do
  opts=""
  [ "$v" != "-" ] && for kv in $v; do opts="$opts --opt $kv"; done
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --e2e-steps 3 $opts > $out/${tag}_ab_tmp.json 2>> $out/${tag}_ab.err || { echo "variant [$v] failed"; tail -3 $out/${tag}_ab.err; continue; }
  python - "$v" $out/${tag}_ab_tmp.json >> $out/${tag}_ab.jsonl <<'PY'
import json,sys
d=json.load(open(sys.argv[2])); d["variant"]=sys.argv[1]; print(json.dumps(d))
PY
done
python - $out/${tag}_ab.jsonl <<'PY'
import json,sys
for l in open(sys.argv[1]):
    d=json.loads(l)
    print(f"[{d['variant']:40s}] ms/step {d['ms_per_step']:.3f}  G edges/s {d['value']/1e9:.2f}  e2e {d['e2e']['value']/1e9:.2f}  " + " ".join(f"{k}={v:.3f}" for k,v in d['phases_ms'].items()) + f"  frac {d['roofline']['frac']:.3f}")
PY
