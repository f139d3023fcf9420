#!/bin/bash
TAG=${1:-probe}
OUT=gpurun_out
mkdir -p $OUT
timeout 120 scripts/microbench/membw > $OUT/${TAG}_membw.txt 2>&1; cat $OUT/${TAG}_membw.txt | grep -E "grid  2368|grid  4736|Memset|Memcpy|status"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_slab2 -c 2 -f -o $OUT/${TAG}_slab2_116 \
  python scripts/spmm_bench.py --modes=116 --iters 1 > $OUT/${TAG}_ncu1.log 2>&1; echo "ncu rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_slab2 -c 2 -f -o $OUT/${TAG}_slab2_132_n200 \
  python scripts/spmm_bench.py --B 1024 --N 200 --modes=132 --iters 1 > $OUT/${TAG}_ncu2.log 2>&1; echo "ncu rc=$?"
