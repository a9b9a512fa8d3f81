#!/bin/bash
# ncu launch list (gpu__time_duration per launch) of one tools/bench_configs.py config.  usage: gpu_launches_cfg.sh tag cfg [opts]
set -u
tag=$1; cfg=$2; shift 2
out=gpurun_out; mkdir -p $out
CMD="python tools/bench_configs.py --reps 1 --only $cfg $*"
timeout 600 $CMD > $out/${tag}_plain.log 2>&1 || { echo plain failed; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file $out/${tag}_launches.csv $CMD > $out/${tag}_ncu.log 2>&1
echo "ncu exit $?"
python - $out/${tag}_launches.csv <<'PY'
import csv,sys,collections
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>5]
hdr=rows[0]; ki=hdr.index("Kernel Name"); vi=hdr.index("Metric Value"); ui=hdr.index("Metric Unit")
agg=collections.OrderedDict(); 
for r in rows[1:]:
    v=float(r[vi].replace(',','')); u=r[ui]
    v = v/1e3 if u in ('ns','nsecond') else (v if u in ('us','usecond') else v*1e3)
    k=r[ki].split('(')[0][:60]
    a=agg.setdefault(k,[0,0.0]); a[0]+=1; a[1]+=v
tot=sum(a[1] for a in agg.values())
for k,a in sorted(agg.items(), key=lambda kv:-kv[1][1])[:25]: print(f"{k:62s} n={a[0]:4d} total {a[1]:10.1f} us  {100*a[1]/tot:5.1f}%  avg {a[1]/a[0]:9.1f} us")
PY
