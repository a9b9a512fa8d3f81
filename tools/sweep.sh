for o in "fuse_rbf=1" "fuse_rbf=0" "fuse_rbf=1 --opt direct_min_blocks=3"; do
  echo "== $o"; python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 --opt $o | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['phases_ms'], d['roofline']['frac'])"
done
