#!/usr/bin/env python
"""A/B of one GraphConv forward layer: unfused (grl_spmm_fwd_stream + grl_gemm_nt, new_V through HBM) against the fused
kernel (grl_graphconv_fused_fwd; slab and direct producers).  Prints one JSON line per shape.
Usage: python scripts/fused_bench.py [--shapes 256x500x256x256,64x1000x256x256] [--iters 20]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops, synthetic  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shapes", default="256x500x256x256,256x500x512x256,64x500x256x256,64x1000x256x256,64x2000x256x256")
    ap.add_argument("--degree", type=float, default=8.0)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--model", default="", help="comma list of BxN: time the whole eval forward (S=256, in_dim 4369) fused vs unfused")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    lib = _lib.load()
    if args.model:
        from graph_representation_learning_b200 import GraphCNNDropEdge
        torch.manual_seed(0)
        model = GraphCNNDropEdge(4369, 53, 6, net_size=256).to(dev).eval()
        for shp in args.model.split(","):
            B, N = (int(v) for v in shp.split("x"))
            wl = synthetic.Workload("bench", B, N, input_dim=4369, degree=args.degree)
            V, edges, _ = synthetic.make_edge_batch(wl, 7, dev)
            res = {"model_eval": {"B": B, "N": N}}
            outs = {}
            for on in (False, True):
                ops.FusedGraphConv.enabled = on

                def fwd():
                    with torch.no_grad():
                        return model([V, PackedGraphBatch.from_edges(edges, B, N, 6, with_transpose=False)])
                outs[on] = fwd()
                res["fused_ms" if on else "unfused_ms"] = round(timed(fwd, args.iters), 4)
            ops.FusedGraphConv.enabled = "auto"
            res["logits_rel_diff"] = float((outs[True].double() - outs[False].double()).norm() / outs[False].double().norm())
            print(json.dumps(res), flush=True)
        return
    for shp in args.shapes.split(","):
        B, N, F, C = (int(v) for v in shp.split("x"))
        wl = synthetic.Workload("bench", B, N, input_dim=8, degree=args.degree)
        _, edges, _ = synthetic.make_edge_batch(wl, 7, dev, features=False)
        g = PackedGraphBatch.from_edges(edges, B, N, wl.num_edges)
        L = g.L
        if F > 256:      # gcn3: virtual concat of two inputs
            xs = [torch.randn(g.num_nodes, F // 2, device=dev), torch.randn(g.num_nodes, F // 2, device=dev)]
        else:
            xs = [torch.randn(g.num_nodes, F, device=dev)]
        K = (L + 1) * F
        h = torch.randn(K, C, device=dev) / K ** 0.5
        bias = torch.randn(C, device=dev)
        w = g.edge_weights(0.0)
        res = {"shape": {"B": B, "N": N, "F": F, "C": C, "nnz": g.nnz}}
        with torch.no_grad():
            ref = ops.linear_act([ops.spmm_aggregate(g, w, xs)], h, bias, True, weight_is_kn=True)
            res["unfused_ms"] = round(timed(lambda: ops.linear_act([ops.spmm_aggregate(g, w, xs)], h, bias, True, weight_is_kn=True), args.iters), 4)
            res["spmm_ms"] = round(timed(lambda: ops.spmm_aggregate(g, w, xs), args.iters), 4)
            for name, mode in (("fused_auto", 0), ("fused_direct", 2)):
                _lib.check(lib.grl_graphconv_fused_set_mode(mode), "set_mode")
                out = ops.graph_conv_fused(xs, g, w, h, bias, True)
                if out is None:
                    res[name + "_ms"] = None
                    continue
                torch.cuda.synchronize()
                res[name + "_err_vs_unfused"] = float((out.double() - ref.double()).norm() / ref.double().norm())
                res[name + "_ms"] = round(timed(lambda: ops.graph_conv_fused(xs, g, w, h, bias, True), args.iters), 4)
            lib.grl_graphconv_fused_set_mode(0)
        newv_bytes = g.num_rows * F * 4
        res["new_v_GB_saved_write_plus_read"] = round(2 * newv_bytes / 1e9, 3)
        print(json.dumps(res), flush=True)
        del g, xs, h, ref


if __name__ == "__main__":
    main()
