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

# the two passes on their own (C ABI called directly)
from graph_representation_learning_b200 import _lib
lib = _lib.load()
B, N, L = A.shape[0], A.shape[1], A.shape[2]
R = B * N * (L + 1)
st = _lib.stream_ptr(dev)
code = _lib.dtype_code(A)
rowptr = torch.zeros(R + 1, dtype=torch.int32, device=dev)
bitmap = torch.empty(lib.grl_csr_bitmap_bytes(B, N, L), dtype=torch.uint8, device=dev)
g = PackedGraphBatch.from_dense(A, with_transpose=False)
col, val = torch.empty_like(g.col), torch.empty_like(g.val)
tc = timeit(lambda: lib.grl_csr_count_bitmap(A.data_ptr(), code, B, N, L, 1, rowptr.data_ptr(), bitmap.data_ptr(), st))
tf = timeit(lambda: lib.grl_csr_fill_bitmap(A.data_ptr(), code, bitmap.data_ptr(), B, N, L, 1, g.rowptr.data_ptr(),
                                            col.data_ptr(), val.data_ptr(), st))
assert torch.equal(col, g.col) and torch.equal(val, g.val)
print(f"count+bitmap {tc:.3f} ms ({A.numel()*2/tc/1e6:.0f} GB/s);  fill from bitmap {tf:.3f} ms;  nnz {g.nnz}")
A2 = A.clone(); A2[0, 0, 0, 1] = 0.5                    # one stored value that is not 1: the fill pass fetches values
g2 = PackedGraphBatch.from_dense(A2, with_transpose=False)
col2, val2 = torch.empty_like(g2.col), torch.empty_like(g2.val)
lib.grl_csr_count_bitmap(A2.data_ptr(), code, B, N, L, 1, rowptr.data_ptr(), bitmap.data_ptr(), st)
tw = timeit(lambda: lib.grl_csr_fill_bitmap(A2.data_ptr(), code, bitmap.data_ptr(), B, N, L, 1, g2.rowptr.data_ptr(),
                                            col2.data_ptr(), val2.data_ptr(), st))
assert torch.equal(col2, g2.col) and torch.equal(val2, g2.val)
print(f"fill from bitmap, weighted adjacency {tw:.3f} ms")

# typed transpose of the batch: one CTA per graph (csrc/typed_transpose.cu) against CSC + regroup by type
def typed(direct):
    h = PackedGraphBatch(g.B, g.N, g.L, g.nnz, g.rowptr, g.col, g.val)
    h.direct_typed_transpose = direct
    return h.typed_transpose()
td, tc2 = timeit(lambda: typed(True)), timeit(lambda: typed(False))
a, b = typed(True), typed(False)
assert torch.equal(a.rowptr, b.rowptr) and torch.equal(a.col, b.col) and torch.equal(a.t_perm, b.t_perm)
print(f"typed transpose: direct {td:.3f} ms, through the CSC {tc2:.3f} ms")
