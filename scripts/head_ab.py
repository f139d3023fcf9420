"""Train step at cfg2 with the w_rand -> classifier head as one autograd node (gated GEMM epilogue) vs two layers."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch, synthetic
from graph_representation_learning_b200.trainer import KVTrainStep
dev = torch.device("cuda:0")
wl = synthetic.CONFIGS["cfg2"]
V, A, y = synthetic.make_batch(wl, 1234, dev)
for rep in range(2):
    for fuse in (False, True):
        torch.manual_seed(0)
        model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size).to(dev)
        model.fuse_head = fuse
        model.rng_seed = 1
        step = KVTrainStep(model)
        batch = {"textline_encoding": V, "adjacency_matrix": A, "node_label": y}
        for _ in range(4):
            step.run_train_step(batch)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(15):
            step.run_train_step(batch)
        e1.record(); torch.cuda.synchronize()
        print(json.dumps({"fuse_head": fuse, "ms_per_step": round(e0.elapsed_time(e1) / 15, 3)}), flush=True)
        del model, step
