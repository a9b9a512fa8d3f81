#!/bin/bash
# Several `ncu --set full` captures in one gpurun call.  usage: gpu_ncu_multi.sh tag "cfg|regex|skip|count|linefilter" ...
# cfg is a tools/bench_configs.py --only name; each capture runs only after the plain command exited 0.  The reports
# are summarised ON THE BOX (tools/ncu_summary.py, tools/ncu_by_line.py) and then deleted: gpurun_out/ is capped at
# 64 MiB and an --import-source report is 5-30 MB.
set -u
tag=$1; shift
out=gpurun_out; mkdir -p $out
for spec in "$@"; do
  IFS='|' read -r cfg rx skip cnt filt <<< "$spec"
  name=$(echo "${cfg}_${rx}" | tr -c 'A-Za-z0-9_\n' '_')
  CMD="python tools/bench_configs.py --reps 1 --only $cfg ${DGNN_CFG_OPTS:-}"
  timeout 600 $CMD > $out/${tag}_${cfg}_plain.log 2>&1 || { echo "plain $cfg failed"; tail -3 $out/${tag}_${cfg}_plain.log; continue; }
  timeout 900 ncu --set full --clock-control none --import-source on -k "regex:$rx" -s ${skip:-0} -c ${cnt:-1} -f \
      -o /tmp/${tag}_${name} $CMD > $out/${tag}_${name}_ncu.log 2>&1
  echo "ncu $cfg $rx exit $?"; tail -1 $out/${tag}_${name}_ncu.log
  python tools/ncu_summary.py /tmp/${tag}_${name}.ncu-rep > $out/${tag}_${name}_full.txt 2>&1
  [ -n "${filt:-}" ] && python tools/ncu_by_line.py /tmp/${tag}_${name}.ncu-rep "$filt" 40 > $out/${tag}_${name}_lines.txt 2>&1
  rm -f /tmp/${tag}_${name}.ncu-rep
done
ls -la $out | tail -30
