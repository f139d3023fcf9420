#!/bin/bash
TAG=${1:-g}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q -k "gemm or model or linear or train" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -6 $OUT/${TAG}_pytest.log
timeout 300 python scripts/gemm_bench.py > $OUT/${TAG}_gemm_bench.jsonl 2>&1; python - $TAG <<'PY'
import json,sys
for l in open("gpurun_out/%s_gemm_bench.jsonl" % sys.argv[1] if len(sys.argv)>1 else "g"):
    try: d=json.loads(l)
    except Exception: print(l.strip()); continue
    print(d["gemm"], d["N"], d["K"], "x3", d["x3_ms"], d["x3_useful_TFs"], "tf32", d["tf32_ms"], "cublas_tf32", d["cublas_tf32_ms"])
PY
timeout 600 python bench.py --no-cpu-baseline --no-e2e > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-400 $OUT/${TAG}_bench.json; tail -3 $OUT/${TAG}_bench.err
