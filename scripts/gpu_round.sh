#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), ncu launch list of one train step, ncu --set full of the
# SpMM kernels.  Everything lands in gpurun_out/<tag>_*.  Usage: bash scripts/gpu_round.sh <tag> [skip-tests]
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
if [ "$2" != "skip-tests" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
  tail -5 $OUT/${TAG}_pytest.log
  timeout 300 python __graft_entry__.py smoke > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
fi
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cat $OUT/${TAG}_bench.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_ref.json 2>> $OUT/${TAG}_bench.err; cat $OUT/${TAG}_bench_ref.json
timeout 300 python scripts/spmm_bench.py --modes=-2,-1,116,100 > $OUT/${TAG}_spmm.json 2>&1; cat $OUT/${TAG}_spmm.json
# launch list of ONE train step (bench has exited 0 above without ncu)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
  --log-file $OUT/${TAG}_launches_step.csv python bench.py --steps 2 --warmup 3 --no-e2e --no-roofline --no-cpu-baseline --profile-step \
  > $OUT/${TAG}_ncu_step.log 2>&1; echo "ncu launches rc=$?"
# full capture of the forward and transposed SpMM
timeout 900 ncu --set full --clock-control none --import-source on -k regex:spmm -c 6 -f -o $OUT/${TAG}_spmm_full \
  python scripts/spmm_bench.py --modes=100 --iters 1 > $OUT/${TAG}_ncu_spmm.log 2>&1; echo "ncu spmm rc=$?"
ls -la $OUT | tail -20
