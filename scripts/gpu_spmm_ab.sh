#!/bin/bash
# A/B of the forward SpMM kernels + the kernel-agreement parity test.  Usage: bash scripts/gpu_spmm_ab.sh <tag>
TAG=${1:-ab}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -k "spmm" > $OUT/${TAG}_pytest_spmm.log 2>&1; echo "pytest rc=$?"; tail -15 $OUT/${TAG}_pytest_spmm.log
timeout 300 python scripts/spmm_bench.py --modes=-2,16,116,108,100 > $OUT/${TAG}_spmm_f32.json 2>&1; cat $OUT/${TAG}_spmm_f32.json
timeout 300 python scripts/spmm_bench.py --B 128 --N 1000 --F 512 --dtype bf16 --modes=-2,-1,8,108,116,100 > $OUT/${TAG}_spmm_bf16.json 2>&1; cat $OUT/${TAG}_spmm_bf16.json
timeout 300 python scripts/spmm_bench.py --B 1024 --N 200 --F 256 --modes=-2,16,116,108,132,100 > $OUT/${TAG}_spmm_n200.json 2>&1; cat $OUT/${TAG}_spmm_n200.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_slab2 -c 1 -f -o $OUT/${TAG}_slab2 \
  python scripts/spmm_bench.py --modes=100 --iters 1 > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
