"""Backward of GraphConv as aggregate-then-project on the typed transpose of the batch (ops._GraphConv,
csrc/csr_pack.cu csc_typed_kernel, grl_stream_from_weights): the autograd of
gnn/models/networks/robust_gcn.py:45-51 in the association dX = sum_t (A_t^T dZ) W_t^T.
Index structures bit-exact against a dense transpose; gradients against fp64 autograd of the dense formula (2e-6, the
fp32-grade GEMMs' bar) and against the project-then-aggregate path (same mathematics, other association)."""
import pytest
import torch

from _golden import rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


@pytest.mark.parametrize("B,N,density,weighted", [(3, 37, 0.12, False), (2, 260, 0.03, True), (1, 8, 0.0, False)])
def test_typed_transpose_is_the_dense_transpose(dev, B, N, density, weighted):
    from graph_representation_learning_b200 import PackedGraphBatch
    gen = torch.Generator().manual_seed(N)
    L = 6
    A = (torch.rand(B, N, L, N, generator=gen) < density).float()
    if weighted:
        A = A * (torch.randint(1, 200, (B, N, L, N), generator=gen).float() / 64)     # exact in fp16
    g = PackedGraphBatch.from_dense(A.half().to(dev))
    w = g.edge_weights(0.3, seed=11, stream_id=2)
    gT = g.typed_transpose()
    wT = g.transposed_weights(w)
    nnz = int(g.rowptr[-1])
    assert int(gT.rowptr[-1]) == nnz and int(gT.rowptr[0]) == 0
    dense = g.to_dense(w).view(B, N, L + 1, N)                      # [b, i, t, j]
    denseT = gT.to_dense(wT).view(B, N, L + 1, N)                   # [b, j, t, i]
    assert torch.equal(denseT, dense.permute(0, 3, 2, 1))
    # source nodes ascend inside every typed row (deterministic summation order), weights are a permutation of w
    rp, col = gT.rowptr.cpu().long(), gT.col.cpu()[:nnz]
    rows = torch.repeat_interleave(torch.arange(rp.numel() - 1), rp[1:] - rp[:-1])
    same_row = rows[1:] == rows[:-1]
    assert bool((col[1:][same_row] > col[:-1][same_row]).all())
    assert torch.equal(wT[:nnz], w[gT.t_perm.long()[:nnz]])
    assert getattr(wT, "_grl_stream", None) is not None


def _both_builds(g):
    """(t_rowptr, t_col, t_perm) from the one-CTA-per-graph kernel and from the CSC route"""
    from graph_representation_learning_b200 import PackedGraphBatch
    outs = []
    for direct in (True, False):
        h = PackedGraphBatch(g.B, g.N, g.L, g.nnz, g.rowptr, g.col, g.val)
        h.direct_typed_transpose = direct
        gT = h.typed_transpose()
        assert (h.colptr is None) == direct, "the wrong build ran"
        outs.append((gT.rowptr, gT.col, gT.t_perm))
    return outs


@pytest.mark.parametrize("case", ["sparse", "mid_rows", "hub", "many_hubs", "dense_small", "empty", "two_waves", "edges_dup"])
def test_direct_typed_transpose_equals_the_csc_route_bit_for_bit(dev, case):
    """csrc/typed_transpose.cu against grl_csr_transpose + grl_csc_typed_regroup: rows sorted by their owner thread
    (<= 32 entries), by a warp through shared memory (<= 256), hub rows rebuilt by the CTA (listed, and more than the list
    holds), more graphs than SMs, graphs without edges, duplicate edges from an edge list."""
    from graph_representation_learning_b200 import PackedGraphBatch
    gen = torch.Generator().manual_seed(len(case))
    L = 6
    if case == "edges_dup":
        B, N = 2, 50
        e = torch.stack([torch.randint(0, B, (4000,), generator=gen), torch.randint(0, N, (4000,), generator=gen),
                         torch.randint(0, L, (4000,), generator=gen), torch.randint(0, 3, (4000,), generator=gen)], 1).int()
        g = PackedGraphBatch.from_edges(e.to(dev), B, N, L, with_transpose=False)     # ~ 4 copies of every (i, l, j < 3)
    else:
        B, N, dens = {"sparse": (3, 300, 0.02), "mid_rows": (2, 400, 0.3), "hub": (2, 600, 0.01), "many_hubs": (1, 300, 0.95),
                      "dense_small": (4, 40, 1.0), "empty": (3, 64, 0.0), "two_waves": (310, 24, 0.2)}[case]
        A = (torch.rand(B, N, L, N, generator=gen) < dens)
        if case == "hub":
            A[:, :, 2, 7] = True                                     # every node points at node 7 through type 2: d = N
            A[1, :, 5, 599] = True
        g = PackedGraphBatch.from_dense(A.half().to(dev), with_transpose=False)
    (ra, ca, pa), (rb, cb, pb) = _both_builds(g)
    nnz = int(g.rowptr[-1])
    assert torch.equal(ra, rb)
    assert torch.equal(pa[:nnz], pb[:nnz])
    assert torch.equal(ca[:nnz], cb[:nnz])


def test_direct_typed_transpose_declines_graphs_whose_counters_do_not_fit(dev):
    from graph_representation_learning_b200 import PackedGraphBatch, _lib
    N, L = 5000, 6                                                   # 35 000 typed rows per graph > 34 816
    e = torch.tensor([[0, 1, 2, 3], [0, 4999, 5, 0]], dtype=torch.int32, device=dev)
    g = PackedGraphBatch.from_edges(e, 1, N, L, with_transpose=False)
    t = [torch.empty(g.num_rows + 1, dtype=torch.int32, device=dev), torch.empty(g.nnz, dtype=torch.int32, device=dev),
         torch.empty(g.nnz, dtype=torch.int32, device=dev)]
    rc = _lib.load().grl_typed_transpose(g.rowptr.data_ptr(), g.col.data_ptr(), 1, N, L, t[0].data_ptr(), t[1].data_ptr(),
                                         t[2].data_ptr(), _lib.stream_ptr(dev))
    assert rc == _lib.GRL_E_UNSUP
    gT = g.typed_transpose()                                         # the public path still answers, through the CSC
    assert g.colptr is not None and int(gT.rowptr[-1]) == int(g.rowptr[-1])


@pytest.mark.parametrize("dtype,widths,C,N", [(torch.float32, (64,), 64, 130), (torch.float32, (128, 128), 128, 300),
                                              (torch.bfloat16, (128,), 128, 257)])
def test_graphconv_gradients_aggregate_first(dev, dtype, widths, C, N):
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(C + N)
    B, L = 2, 6
    A = (torch.rand(B, N, L, N, generator=gen) < 0.04).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    K = (L + 1) * sum(widths)
    xs0 = [torch.randn(B * N, f, generator=gen).to(dev) for f in widths]
    h0 = (torch.randn(K, C, generator=gen) / K ** 0.5).to(dev)
    b0 = torch.randn(C, generator=gen).to(dev)
    dOut = torch.randn(B * N, C, generator=gen).to(dev)
    w = g.edge_weights(0.3, seed=4, stream_id=9)

    def run(agg_first):
        ops.GraphConvBackward.aggregate_first = agg_first
        try:
            xs = [x.detach().clone().to(dtype).requires_grad_(True) for x in xs0]
            h, b = h0.clone().requires_grad_(True), b0.clone().requires_grad_(True)
            out = ops.graph_conv(xs, g, w, h, b, relu=True, p=0.5, seed=21, stream_id=3)
            out.backward(dOut.to(dtype))
            return out.detach(), [x.grad for x in xs], h.grad, b.grad
        finally:
            ops.GraphConvBackward.aggregate_first = True

    out1, dx1, dh1, db1 = run(True)
    out0, dx0, dh0, db0 = run(False)
    assert torch.equal(out1, out0)                                 # same forward kernels
    assert torch.equal(dh1, dh0) and torch.equal(db1, db0)         # same dZ / dW / db kernels
    # fp64 autograd of the dense formula with the SAME dropout pattern (read off the output zeros is not enough: use the mask)
    Ad = g.to_dense(w).double()
    xs64 = [x.detach().clone().to(dtype).double().requires_grad_(True) for x in xs0]
    h64 = h0.double().requires_grad_(True)
    X = torch.cat(xs64, 1).view(B, N, -1)
    pre = torch.bmm(Ad, X).view(B, N, L + 1, -1).reshape(B * N, -1) @ h64 + b0.double()
    # the activation / dropout pattern is the kernel's (read off its output: kept and positive, scale 1 / (1 - 0.5)): a
    # pre-activation whose sign differs between bf16 and fp64 must not count as a gradient error of the backward kernels
    keep = (out1 != 0).double() * 2.0
    (pre * keep).backward(dOut.to(dtype).double())
    tol, tol_paths = (2e-6, 2e-6) if dtype == torch.float32 else (1e-2, 1e-2)
    for k in range(len(widths)):
        assert rel_err(dx1[k], xs64[k].grad) < tol, (k, rel_err(dx1[k], xs64[k].grad))
        assert rel_err(dx1[k], dx0[k]) < tol_paths


@pytest.mark.parametrize("M,C,p", [(1000, 512, 0.5), (777, 64, 0.0), (4097, 256, 0.3)])
def test_bf16_dz_and_bias_gradient_in_one_pass(dev, M, C, p):
    """grl_dz_colsum_bf16 (the bf16 backward of F.relu + nn.Dropout + the bias gradient in one pass) against the two
    kernels it replaces: dZ bit for bit, the bias gradient to fp32 summation order -- and deterministic."""
    from graph_representation_learning_b200 import _lib, ops
    gen = torch.Generator().manual_seed(M + C)
    dY = torch.randn(M, C, generator=gen).to(dev).bfloat16()
    Y = (torch.randn(M, C, generator=gen).clamp_min(0) * (torch.rand(M, C, generator=gen) < 0.6)).to(dev).bfloat16()
    ref_dz = torch.empty_like(dY)
    ops.epilogue_bwd_(dY, Y, ref_dz, True, p)
    ref_db = ops.colsum(ref_dz)
    lib = _lib.load()
    outs = []
    for _ in range(2):
        dZ = torch.empty_like(dY)
        db = torch.empty(C, dtype=torch.float32, device=dev)
        ws = torch.empty(lib.grl_dz_split_workspace_bytes(M, C), dtype=torch.uint8, device=dev)
        flags = _lib.EPI_RELU | (_lib.EPI_DROPOUT if p > 0.0 else 0)
        _lib.check(lib.grl_dz_colsum_bf16(dY.data_ptr(), dY.stride(0), Y.data_ptr(), Y.stride(0), dZ.data_ptr(), dZ.stride(0), M, C,
                                          flags, float(p), db.data_ptr(), ws.data_ptr(), ws.numel(), _lib.stream_ptr(dev)),
                   "grl_dz_colsum_bf16")
        outs.append((dZ, db))
    assert torch.equal(outs[0][0], ref_dz)
    assert torch.equal(outs[0][1], outs[1][1])
    assert rel_err(outs[0][1], ref_dz.double().sum(0)) < 1e-6 and rel_err(ref_db, ref_dz.double().sum(0)) < 1e-5
