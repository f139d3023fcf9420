#!/usr/bin/env python
"""Throughput of the tcgen05 GEMM modes on the model's contraction shapes (cfg2: M = 128 000), next to cuBLAS.
Usage (GPU box): python scripts/gemm_bench.py [--M 128000]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import ops  # noqa: E402


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--M", type=int, default=128000)
    ap.add_argument("--promo", type=int, default=2, help="TF32X3 promotion interval (K-blocks per TMEM chunk)")
    ap.add_argument("--one-cta-per-tile", action="store_true", help="TF32X3: non-persistent launch (A/B)")
    ap.add_argument("--only", default="", help="comma-separated GEMM names to run")
    ap.add_argument("--x3-only", action="store_true")
    args = ap.parse_args()
    from graph_representation_learning_b200 import _lib
    lib = _lib.load()
    lib.grl_gemm_x3_set_promo(args.promo)
    lib.grl_gemm_x3_set_promo(-1 if args.one_cta_per_tile else -2)
    dev = torch.device("cuda:0")
    M = args.M
    shapes = [("emb1", 256, 4368), ("gcn1 proj", 256, 1792), ("gcn3 proj", 256, 3584), ("gcn1 dX", 1792, 256),
              ("gcn3 dX", 3584, 256), ("w_rand", 1280, 128), ("w_rand dX", 128, 1280), ("emb2", 128, 512)]
    out = []
    for name, N, K in shapes:
        if args.only and name not in args.only.split(","):
            continue
        a = torch.randn(M, K, device=dev)
        b = torch.randn(N, K, device=dev) / K ** 0.5
        ab, bb = a.bfloat16(), b.bfloat16()
        gf = 2.0 * M * N * K / 1e9
        row = {"gemm": name, "M": M, "N": N, "K": K, "GF": round(gf, 1)}
        ops.GemmConfig.fp32_mode = "tf32x3"
        t = timeit(lambda: ops.gemm_nt([a], [b]))
        row["x3_ms"], row["x3_useful_TFs"] = round(t, 3), round(gf / t, 1)
        if args.x3_only:
            print(json.dumps(row), flush=True)
            continue
        ops.GemmConfig.fp32_mode = "tf32"
        t = timeit(lambda: ops.gemm_nt([a], [b]))
        row["tf32_ms"], row["tf32_TFs"] = round(t, 3), round(gf / t, 1)
        ops.GemmConfig.fp32_mode = "tf32x3"
        t = timeit(lambda: ops.gemm_nt([ab], [bb]))
        row["bf16_ms"], row["bf16_TFs"] = round(t, 3), round(gf / t, 1)
        torch.backends.cuda.matmul.allow_tf32 = True
        t = timeit(lambda: torch.mm(a, b.t()))
        row["cublas_tf32_ms"], row["cublas_tf32_TFs"] = round(t, 3), round(gf / t, 1)
        torch.backends.cuda.matmul.allow_tf32 = False
        t = timeit(lambda: torch.mm(a, b.t()), n=3)
        row["cublas_fp32_ms"], row["cublas_fp32_TFs"] = round(t, 3), round(gf / t, 1)
        t = timeit(lambda: torch.mm(ab, bb.t()))
        row["cublas_bf16_ms"], row["cublas_bf16_TFs"] = round(t, 3), round(gf / t, 1)
        out.append(row)
        print(json.dumps(row), flush=True)
        del a, b, ab, bb


if __name__ == "__main__":
    main()
