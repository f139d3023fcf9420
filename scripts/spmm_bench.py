#!/usr/bin/env python
"""A/B timing of the forward / transposed SpMM kernels (grl_spmm_set_tile_mode) on one workload shape.
Usage: python scripts/spmm_bench.py [--B 256 --N 500 --F 256 --dtype f32 --modes -2,-1,116,100]
Every mode is checked against the legacy row kernel (bit-for-bit for fp32: same per-row summation order)."""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--B", type=int, default=256)
    ap.add_argument("--N", type=int, default=500)
    ap.add_argument("--F", type=int, default=256)
    ap.add_argument("--degree", type=float, default=8.0)
    ap.add_argument("--dtype", default="f32")
    ap.add_argument("--modes", default="-2,-1,116,100")
    ap.add_argument("--iters", type=int, default=20)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    dt = torch.float32 if args.dtype == "f32" else torch.bfloat16
    wl = synthetic.Workload("bench", args.B, args.N, input_dim=8, degree=args.degree)
    _, A, _ = synthetic.make_batch(wl, 7, dev)
    g = PackedGraphBatch.from_dense(A)
    del A
    X = torch.randn(g.num_nodes, args.F, device=dev, dtype=dt)
    w = g.edge_weights(0.3, seed=1, stream_id=1)
    es = X.element_size()
    alg = 4 * (g.num_rows + 1) + 8 * g.nnz + X.numel() * es + g.num_rows * args.F * es
    lib = _lib.load()
    ref = None
    out = {}
    for mode in [int(m) for m in args.modes.split(",")]:
        ops._spmm_stream_enabled[0] = mode >= 100         # 100 / 108 / 116 / 132: node-stream slab kernel
        _lib.check(lib.grl_spmm_set_tile_mode(0 if mode == 100 else mode), "set_tile_mode")
        Y = torch.full((g.num_rows, args.F), float("nan"), device=dev, dtype=dt)
        ops.spmm_fwd_(g, w, X, Y)
        torch.cuda.synchronize()
        if ref is None:
            ref = Y.clone()
        err = float((Y.float() - ref.float()).abs().max())
        for _ in range(3):
            ops.spmm_fwd_(g, w, X, Y)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(args.iters):
            ops.spmm_fwd_(g, w, X, Y)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / args.iters
        out[mode] = {"ms": round(ms, 4), "GBps": round(alg / ms / 1e6, 1), "max_abs_diff_vs_first": err}
    lib.grl_spmm_set_tile_mode(0)
    ops._spmm_stream_enabled[0] = True
    # the per-layer DropEdge + node-stream pack that feeds the stream kernel
    for _ in range(3):
        g.edge_weights(0.3, seed=1, stream_id=1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.iters):
        g.edge_weights(0.3, seed=1, stream_id=1)
    e1.record()
    torch.cuda.synchronize()
    out["dropedge_stream_ms"] = round(e0.elapsed_time(e1) / args.iters, 4)
    # transposed (gradient) kernel: dX = A'^T dY over CSC + perm
    dY = torch.randn(g.num_rows, args.F, device=dev, dtype=dt)
    dX = torch.empty(g.num_nodes, args.F, device=dev, dtype=dt)
    for _ in range(3):
        ops.spmm_bwd_(g, w, dY, dX)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(args.iters):
        ops.spmm_bwd_(g, w, dY, dX)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / args.iters
    alg_bwd = 4 * (g.num_nodes + 1) + 12 * g.nnz + dY.numel() * es + dX.numel() * es
    bwd = {"ms": round(ms, 4), "GBps": round(alg_bwd / ms / 1e6, 1), "alg_bytes": alg_bwd}
    print(json.dumps({"shape": vars(args), "nnz": g.nnz, "alg_bytes": alg, "fwd": out, "bwd": bwd}))


if __name__ == "__main__":
    main()
