"""cfg5, small documents: eval forward latency eager vs CUDA-graph replay (batch 64, dense fp16 adjacency input)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from graph_representation_learning_b200 import GraphCNNDropEdge, synthetic
from graph_representation_learning_b200.procedures import GraphedEvalForward
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = GraphCNNDropEdge(4369, 53, 6, 256).to(dev).eval()
def timeit(fn, n=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
for N in (100, 200, 500):
    wl = synthetic.Workload("g", 64, N)
    V, A, _ = synthetic.make_batch(wl, 5, dev)
    with torch.no_grad():
        t_eager = timeit(lambda: model([V, A]))
    gf = GraphedEvalForward(model, 64, N, 4369, 6)
    t_graph = timeit(lambda: gf(V, A))
    with torch.no_grad():
        same = bool(torch.equal(gf(V, A), model([V, A])))
    print(json.dumps({"config": f"cfg5: eval forward, batch 64 x {N} nodes (dense fp16 adjacency pack + model)", "N": N,
                      "eager_ms": round(t_eager, 3), "graph_ms": round(t_graph, 3), "bit_identical": same,
                      "graphs_per_s_graph": round(64 / t_graph * 1e3, 1)}))
    del gf
