#!/usr/bin/env python
"""cfg4 (one graph, 1M nodes, degree 16 per type): ONE launch of the forward aggregation over the packed graph and ONE
over its typed transpose (what the backward runs since r02), for an `ncu --set full -k regex:spmm_nodes` capture of
their DRAM traffic (profiles/spmm_traffic_cfg4.json).  Usage: python scripts/cfg4_spmm.py [--nodes 1000000]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import PackedGraphBatch, ops, synthetic  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=1_000_000)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    wl = synthetic.CONFIGS["cfg4"]
    _, edges, _ = synthetic.make_edge_batch(wl, 1234, dev, N=args.nodes, features=False)
    g = PackedGraphBatch.from_edges(edges, 1, args.nodes, wl.num_edges)
    del edges
    S = wl.net_size
    X = torch.randn(args.nodes, S, device=dev)
    Y = torch.empty(g.num_rows, S, device=dev)
    w = g.edge_weights(0.3, seed=1, stream_id=1)
    gT, wT = g.typed_transpose(), g.transposed_weights(w)
    torch.cuda.synchronize()
    ops.spmm_fwd_(g, w, X, Y)
    ops.spmm_fwd_(gT, wT, X, Y)
    torch.cuda.synchronize()
    kept = int((w != 0).sum())
    print({"nodes": args.nodes, "nnz": g.nnz, "kept": kept, "bytes_gather": 4 * (g.num_rows + 1) + 8 * g.nnz + kept * S * 4 + Y.numel() * 4})


if __name__ == "__main__":
    main()
