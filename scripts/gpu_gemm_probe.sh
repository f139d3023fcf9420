#!/bin/bash
TAG=${1:-gp}
OUT=gpurun_out
mkdir -p $OUT
for promo in 1 2 4 8; do
  echo "promo=$promo persistent"; python scripts/gemm_bench.py --x3-only --promo $promo --only "gcn1 dX,gcn1 proj,w_rand" 2>&1 | cut -c1-160
  echo "promo=$promo one-cta-per-tile"; python scripts/gemm_bench.py --x3-only --promo $promo --one-cta-per-tile --only "gcn1 dX,gcn1 proj,w_rand" 2>&1 | cut -c1-160
done > $OUT/${TAG}_probe.txt 2>&1
cat $OUT/${TAG}_probe.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_x3 -c 1 --launch-skip 3 -f -o $OUT/${TAG}_x3_dx \
  python scripts/gemm_bench.py --x3-only --only "gcn1 dX" > $OUT/${TAG}_ncu.log 2>&1; echo "ncu rc=$?"
