#!/bin/bash
# build + bench compile-time variants on the GPU box: bash tools/variants.sh "<bench opts>" "-DX=1" "-DY=2 -DZ=3" ...
opts=$1; shift
for v in "$@"; do
  make -C defect-gnn-cpp_b200/csrc clean > /dev/null; make -C defect-gnn-cpp_b200/csrc EXTRA="$v" > /dev/null 2>&1 || echo BUILD FAILED
  echo "== [$v] $opts"
  python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-steps 1 $opts | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['ms_per_step'], d['phases_ms'])"
done
