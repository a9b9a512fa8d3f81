#!/bin/bash
# GPU tests only.  gpurun -- 'bash tools/gpu_tests.sh tag [pytest args]'
set -u
tag=${1:-t}; shift || true
out=gpurun_out; mkdir -p $out
timeout 900 python -m pytest tests -m gpu -q "$@" > $out/${tag}_pytest.log 2>&1; echo "pytest exit $?"; tail -n 15 $out/${tag}_pytest.log
