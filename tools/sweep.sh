# usage: bash tools/sweep.sh "opt1=v" "opt2=v --opt opt3=v" ...   (each argument is one bench configuration)
for o in "$@"; do
  echo "== $o"; python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 --opt $o | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['phases_ms'], d['roofline']['frac'])"
done
