"""Out-of-bounds guards of our own (compute-sanitizer is closed on this GPU pool, profiles/r02_s40_sanitizer_closed.txt):
every kernel that writes through a caller-supplied pointer writes into a view cut out of a larger sentinel-filled buffer
-- guard rows before / after and guard columns right of every row -- and the guards must come back untouched, on shapes
with ragged tails (rows not a multiple of the tile, graphs not a multiple of the TMA box, columns below the pitch)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SENTINEL = -12345.0


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


class Guarded:
    """[rows, cols] view in the middle of a sentinel buffer: 32 guard rows on both sides, >= 8 guard columns"""

    def __init__(self, rows, cols, dtype, dev, pad_cols=8):
        self.rows, self.cols = rows, cols
        ld = (cols + pad_cols + 7) // 8 * 8
        self.big = torch.full((rows + 64, ld), SENTINEL, dtype=dtype, device=dev)
        self.view = self.big[32:32 + rows, :cols]

    def check(self, what):
        torch.cuda.synchronize()
        s = torch.tensor(SENTINEL, dtype=self.big.dtype).item()
        assert bool((self.big[:32] == s).all()) and bool((self.big[32 + self.rows:] == s).all()), f"{what}: guard rows written"
        assert bool((self.big[32:32 + self.rows, self.cols:] == s).all()), f"{what}: guard columns written"
        assert not bool((self.view == s).any()), f"{what}: part of the output was not written"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,N,F", [(3, 37, 64), (2, 263, 128), (1, 700, 256)])
def test_spmm_kernels_stay_inside_their_output(dev, dtype, B, N, F):
    from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops
    gen = torch.Generator().manual_seed(N)
    g = PackedGraphBatch.from_dense((torch.rand(B, N, 6, N, generator=gen) < 0.05).half().to(dev))
    X = torch.randn(B * N, F, generator=gen).to(dev, dtype)
    w = g.edge_weights(0.3, seed=1, stream_id=1)
    lib = _lib.load()
    try:
        for mode, stream in ((0, True), (-1, False), (-2, False)):      # node-stream slab kernel, the two gather kernels
            ops._spmm_stream_enabled[0] = stream
            _lib.check(lib.grl_spmm_set_tile_mode(mode), "grl_spmm_set_tile_mode")
            Y = Guarded(g.num_rows, F, dtype, dev)
            ops.spmm_fwd_(g, w, X, Y.view)
            Y.check(f"spmm_fwd mode {mode}")
    finally:
        lib.grl_spmm_set_tile_mode(0)
        ops._spmm_stream_enabled[0] = True
    dY = torch.randn(g.num_rows, F, generator=gen).to(dev, dtype)
    dX = Guarded(g.num_nodes, F, dtype, dev)
    ops.spmm_bwd_(g, w, dY, dX.view)
    dX.check("spmm_bwd")
    gT, wT = g.typed_transpose(), g.transposed_weights(w)               # backward aggregation over the typed transpose
    U = Guarded(g.num_rows, F, dtype, dev)
    ops.spmm_fwd_(gT, wT, X, U.view)
    U.check("spmm_fwd over the typed transpose")


@pytest.mark.parametrize("B,N,widths,C,mode", [(2, 130, (32,), 64, 0), (1, 77, (96,), 192, 0), (2, 300, (64, 64), 512, 0),
                                               (2, 263, (64,), 256, 2)])
def test_fused_graphconv_stays_inside_its_output(dev, B, N, widths, C, mode):
    from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops
    gen = torch.Generator().manual_seed(N + C)
    L = 6
    g = PackedGraphBatch.from_dense((torch.rand(B, N, L, N, generator=gen) < 0.05).half().to(dev))
    xs = [torch.randn(B * N, f, generator=gen).to(dev) for f in widths]
    K = (L + 1) * sum(widths)
    h = (torch.randn(K, C, generator=gen) / K ** 0.5).to(dev)
    hi, lo = ops._prep_weight(h.t(), torch.float32, _lib.GEMM_TF32X3)
    bias = torch.randn(C, generator=gen).to(dev)
    w = g.edge_weights(0.0)
    Y = Guarded(g.num_nodes, C, torch.float32, dev)
    x1 = xs[1] if len(xs) == 2 else None
    lib = _lib.load()
    try:
        _lib.check(lib.grl_graphconv_fused_set_mode(mode), "grl_graphconv_fused_set_mode")
        _lib.check(lib.grl_graphconv_fused_fwd(
            g.rowptr.data_ptr(), g.col.data_ptr(), w.data_ptr(), xs[0].data_ptr(), xs[0].stride(0), xs[0].shape[1],
            _lib.ptr(x1), x1.stride(0) if x1 is not None else 0, x1.shape[1] if x1 is not None else 0,
            hi.data_ptr(), lo.data_ptr(), hi.stride(0), bias.data_ptr(), Y.view.data_ptr(), Y.view.stride(0), g.B, g.N, g.L, C,
            _lib.EPI_BIAS, 0.0, 0, 0, 0, _lib.stream_ptr(dev)), "grl_graphconv_fused_fwd")
    finally:
        lib.grl_graphconv_fused_set_mode(0)
    Y.check("grl_graphconv_fused_fwd")
    with torch.no_grad():
        ref = ops.linear_act([ops.spmm_aggregate(g, w, xs)], h, bias, False, weight_is_kn=True)
    assert float((Y.view - ref).norm() / ref.norm()) < 1e-6


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_epilogue_kernels_stay_inside_their_output(dev, dtype):
    from graph_representation_learning_b200 import ops
    gen = torch.Generator().manual_seed(4)
    M, C = 1001, 72
    X = torch.randn(M, C, generator=gen).to(dev, dtype)
    bias = torch.randn(C, generator=gen).to(dev)
    Y = Guarded(M, C, dtype, dev)
    ops.epilogue_fwd_(X, bias, Y.view, True, 0.5, None, 3, 1)
    torch.cuda.synchronize()
    s = torch.tensor(SENTINEL, dtype=dtype).item()
    assert bool((Y.big[:32] == s).all()) and bool((Y.big[32 + M:] == s).all()) and bool((Y.big[32:32 + M, C:] == s).all())
    dX = Guarded(M, C, dtype, dev)
    ops.epilogue_bwd_(X, Y.view, dX.view, True, 0.5)
    torch.cuda.synchronize()
    assert bool((dX.big[:32] == s).all()) and bool((dX.big[32 + M:] == s).all()) and bool((dX.big[32:32 + M, C:] == s).all())
