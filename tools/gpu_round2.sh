#!/bin/bash
# Round-2 evidence pass on one B200: all GPU tests, the bench line (own arm + reference arm), the ncu launch list of the
# bench command and `ncu --set full` summaries of every kernel DESIGN.md quotes.  Outputs: gpurun_out/<tag>_*.
set -u
tag=${1:-r02fin}
out=gpurun_out; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -2 $out/${tag}_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > $out/${tag}_bench.json 2> $out/${tag}_bench.err; echo "bench exit $?"
( time timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/${tag}_bench_ref.json 2>> $out/${tag}_bench.err ) 2> $out/${tag}_ref_time.txt; tail -3 $out/${tag}_ref_time.txt
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-configs --no-c5 --e2e-steps 1"
timeout 300 $CMD > $out/${tag}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $out/${tag}_launches.csv $CMD > $out/${tag}_ncu_launches.log 2>&1
echo "launch list exit $?"
bash tools/gpu_ncu_bench.sh ${tag}_fill "^k_direct$" 4 1 "k_direct<" ""
bash tools/gpu_ncu_bench.sh ${tag}_count "k_direct_count_sym" 4 1 "k_direct_count_sym" ""
bash tools/gpu_ncu_bench.sh ${tag}_rbf "k_rbf_f32_vec4" 3 1 "k_rbf_f32_vec4" "--opt fuse_rbf=0"
bash tools/gpu_ncu_multi.sh ${tag} "C4|k_cell32|4|2|k_cell32" "C4|k_cloud_scatter|2|1|k_cloud_scatter" "C2cap|^k_direct$|2|1|k_direct" "C2cap|^k_fill_hits|2|1|k_fill_hits" "F1|k_collate_edges|2|1|k_collate_edges" "C3|^k_direct$|3|1|k_direct" > $out/${tag}_multi.log 2>&1
grep "ncu " $out/${tag}_multi.log
python tools/ncu_traffic_txt.py $out/${tag}_fill_full.txt > $out/${tag}_traffic.json 2>/dev/null; cat $out/${tag}_traffic.json
ls $out | wc -l
