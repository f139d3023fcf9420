"""Dense-adjacency CSR pack (cfg2 shape) timing: PackedGraphBatch.from_dense with / without the transpose."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from graph_representation_learning_b200 import PackedGraphBatch, synthetic
dev = torch.device("cuda:0")
wl = synthetic.Workload("b", 256, 500, input_dim=8)
_, A, _ = synthetic.make_batch(wl, 7, dev)
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
t0 = timeit(lambda: PackedGraphBatch.from_dense(A, with_transpose=False))
t1 = timeit(lambda: PackedGraphBatch.from_dense(A, with_transpose=True))
print(f"from_dense (CSR only) {t0:.3f} ms  ({A.numel()*2/t0/1e6:.0f} GB/s over one read of A);  with CSC {t1:.3f} ms")
