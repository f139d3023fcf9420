#!/usr/bin/env python
"""Train-step latency at small batches: eager KVTrainStep against trainer.GraphedTrainStep (forward + CE + backward in one
CUDA graph).  python scripts/train_graph_latency.py [--shapes 8x200,16x100,64x100]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import GraphCNNDropEdge, synthetic  # noqa: E402
from graph_representation_learning_b200.trainer import GraphedTrainStep, KVTrainStep  # noqa: E402


def timed(fn, iters=20):
    for _ in range(5):
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
    ap.add_argument("--shapes", default="8x200,16x100,64x100,32x200")
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    for shp in args.shapes.split(","):
        B, N = (int(v) for v in shp.split("x"))
        wl = synthetic.Workload("lat", B, N)
        V, A, y = synthetic.make_batch(wl, 3, dev)
        batch = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}
        res = {"B": B, "N": N}
        for graphed in (False, True):
            torch.manual_seed(0)
            model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size).to(dev)
            step = KVTrainStep(model)
            runner = GraphedTrainStep(step, B, N, wl.input_dim, wl.num_edges) if graphed else step
            res["graphed_ms" if graphed else "eager_ms"] = round(timed(lambda: runner.run_train_step(batch)), 4)
            del runner, step, model
            torch.cuda.empty_cache()
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
