#!/bin/bash
# Quick GPU check: a pytest selection (-k "$2"), then optionally the bench.  Usage: bash scripts/gpu_quick.sh <tag> <k-expr> [bench]
TAG=${1:-q}
OUT=gpurun_out
mkdir -p $OUT
timeout 1200 python -m pytest tests -m gpu -x -q -k "$2" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/${TAG}_pytest.log
if [ "$3" == "bench" ]; then
  timeout 600 python bench.py --no-cpu-baseline > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cat $OUT/${TAG}_bench.json; tail -5 $OUT/${TAG}_bench.err
fi
