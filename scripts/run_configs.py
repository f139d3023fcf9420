#!/usr/bin/env python
"""The BASELINE.json configurations that are not the bench line, on one B200: cfg3 (per-GPU share, bf16 S=512),
cfg4 (one 1M-node graph, fwd+bwd), cfg5 (inference sweep over nodes per graph at batch 64).  Prints one JSON line per
measurement.  Usage: python scripts/run_configs.py [--only cfg3,cfg4,cfg5] [--cfg4-nodes 1000000]"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch, synthetic  # noqa: E402
from graph_representation_learning_b200.trainer import KVTrainStep  # noqa: E402


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def latencies(fn, steps, warmup):
    """per-call device latency (one event pair per call, a synchronize between calls as a serving loop would see)"""
    for _ in range(warmup):
        fn()
    out = []
    for _ in range(steps):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        out.append(e0.elapsed_time(e1))
    out.sort()
    return out


def run_train(wl, dev, steps, warmup, edges_mode):
    torch.manual_seed(0)
    model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, wl.use_attention).to(dev)
    model.set_compute_dtype(wl.dtype)
    model.rng_seed = 1000
    step = KVTrainStep(model)
    fdt = torch.float32 if wl.dtype == torch.float32 else torch.bfloat16
    if edges_mode:
        V, edges, y = synthetic.make_edge_batch(wl, 1234, dev, feature_dtype=fdt)
        g = PackedGraphBatch.from_edges(edges, wl.B, wl.N, wl.num_edges)
        nnz = g.nnz
        batch = {"textline_encoding": V, "adjacency_matrix": g, "node_label": y}
    else:
        V, A, y = synthetic.make_batch(wl, 1234, dev, feature_dtype=fdt)
        batch = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}
        nnz = None
    losses = []
    ms = timed(lambda: losses.append(step.run_train_step(batch)), steps, warmup)
    loss = [float(x) for x in losses]
    assert all(l == l and abs(l) < 1e6 for l in loss), f"non-finite loss: {loss}"
    return {"config": wl.name, "ms_per_step": round(ms, 3), "graphs_per_s": round(wl.B / ms * 1e3, 2),
            "nodes_per_s": round(wl.B * wl.N / ms * 1e3, 1), "nnz": nnz, "first_loss": loss[0], "last_loss": loss[-1],
            "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 1e9, 2),
            "step": "fwd+bwd+clip+adam, device-resident" + (" packed graph" if edges_mode else " dense fp16 A packed in the step")}


def run_infer_sweep(dev, sizes, B):
    wl = synthetic.Workload("cfg5", B, 0)
    torch.manual_seed(0)
    model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size, True).to(dev).eval()
    out = []
    for N in sizes:
        torch.cuda.empty_cache()
        V, edges, _ = synthetic.make_edge_batch(wl, 77 + N, dev, B=B, N=N)
        def fwd():
            with torch.no_grad():
                g = PackedGraphBatch.from_edges(edges, B, N, wl.num_edges, with_transpose=False)
                return model([V, g])
        ms = timed(fwd, 5 if N >= 5000 else 10, 2)
        lat = latencies(fwd, 20 if N >= 5000 else 50, 2)
        logits = fwd()
        assert torch.isfinite(logits).all()
        out.append({"config": f"cfg5: eval forward, batch {B} x {N} nodes (edge-list pack + model)", "N": N,
                    "latency_ms": round(ms, 3), "latency_p50_ms": round(lat[len(lat) // 2], 3),
                    "latency_p99_ms": round(lat[min(len(lat) - 1, int(0.99 * len(lat)))], 3), "graphs_per_s": round(B / ms * 1e3, 1),
                    "peak_mem_GB": round(torch.cuda.max_memory_allocated() / 1e9, 2)})
        print(json.dumps(out[-1]), flush=True)
        del V, edges, logits
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="cfg3,cfg4,cfg5")
    ap.add_argument("--cfg4-nodes", type=int, default=1_000_000)
    ap.add_argument("--sizes", default="100,200,500,1000,2000,5000,10000")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    only = args.only.split(",")
    if "cfg3" in only:
        print(json.dumps(run_train(synthetic.CONFIGS["cfg3"], dev, 5, 2, edges_mode=False)), flush=True)
        torch.cuda.empty_cache(); torch.cuda.reset_peak_memory_stats()
    if "cfg5" in only:
        run_infer_sweep(dev, [int(x) for x in args.sizes.split(",")], 64)
        torch.cuda.empty_cache(); torch.cuda.reset_peak_memory_stats()
    if "cfg4" in only:
        wl = synthetic.CONFIGS["cfg4"]
        if args.cfg4_nodes != wl.N:
            wl = synthetic.Workload(wl.name.replace("1M", str(args.cfg4_nodes)), 1, args.cfg4_nodes, degree=16.0,
                                    use_attention=False)
        print(json.dumps(run_train(wl, dev, 3, 1, edges_mode=True)), flush=True)


if __name__ == "__main__":
    main()
