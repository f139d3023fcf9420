"""tcgen05 GEMM (csrc/gemm_tcgen05.cu) against an fp64 torch matmul of the same operands.

Tolerances: the fp32 parity modes TF32X3 (three tf32 products) and TF32_BF16C (tf32 + bf16 correction product, activation
operand through TMEM; the default) 2e-6 norm-wise; single-pass TF32 2e-3; BF16 operands are quantised
first, so only the fp32 accumulation differs from the fp64 product: 1e-5 (2e-2 once the OUTPUT is bf16)."""
import pytest
import torch

from _golden import rel_err

pytestmark = pytest.mark.gpu

SHAPES = [(128, 256, 64), (128, 128, 32), (300, 256, 1000), (1000, 128, 512), (257, 53, 1280), (100, 16, 128),
          (640, 1280, 128), (129, 256, 4369), (5, 8, 7)]


def _mk(M, N, K, gen, dev, dtype):
    a = torch.randn(M, K, generator=gen)
    b = torch.randn(N, K, generator=gen) / K ** 0.5
    if dtype == torch.bfloat16:
        a, b = a.bfloat16().float(), b.bfloat16().float()
    return a, b


@pytest.mark.parametrize("M,N,K", SHAPES)
@pytest.mark.parametrize("mode", ["tf32bc", "tf32x3", "tf32", "bf16"])
def test_gemm_nt_plain(M, N, K, mode):
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M * 7 + N * 3 + K)
    dtype = torch.bfloat16 if mode == "bf16" else torch.float32
    a, b = _mk(M, N, K, gen, dev, dtype)
    ref = a.double() @ b.double().t()
    prev = ops.GemmConfig.fp32_mode
    ops.GemmConfig.fp32_mode = mode if mode != "bf16" else prev
    try:
        out = ops.gemm_nt([a.to(dev, dtype)], [b.to(dev, dtype)], out_dtype=torch.float32)
    finally:
        ops.GemmConfig.fp32_mode = prev
    tol = {"tf32bc": 2e-6, "tf32x3": 2e-6, "tf32": 2e-3, "bf16": 1e-5}[mode]
    assert out.shape == (M, N)
    assert rel_err(out.cpu(), ref) < tol


@pytest.mark.parametrize("mode", ["tf32bc", "tf32x3", "bf16"])
def test_gemm_nt_two_segments_strided_views_and_epilogue(mode):
    """virtual concat [x1 | x2] @ W^T with W an [N, K1+K2] nn.Linear weight, x1/x2 column slices of wider buffers,
    bias + ReLU + injected dropout mask fused."""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(42)
    dtype = torch.bfloat16 if mode == "bf16" else torch.float32
    M, N, K1, K2 = 777, 128, 256, 192
    big1 = torch.randn(M, K1 + 64, generator=gen).to(dtype)
    big2 = torch.randn(M, K2 + 32, generator=gen).to(dtype)
    W = (torch.randn(N, K1 + K2, generator=gen) / 20).to(dtype)
    bias = torch.randn(N, generator=gen)
    keep = torch.rand(M, N, generator=gen) < 0.5
    x1, x2 = big1[:, 64:], big2[:, :K2]
    ref = torch.relu(torch.cat([x1, x2], 1).double() @ W.double().t() + bias.double()) * keep / 0.5
    d1, d2, dW = big1.to(dev), big2.to(dev), W.to(dev)
    prev = ops.GemmConfig.fp32_mode
    ops.GemmConfig.fp32_mode = mode if mode != "bf16" else prev
    try:
        out = ops.gemm_nt([d1[:, 64:], d2[:, :K2]], [dW[:, :K1], dW[:, K1:]], bias=bias.to(dev), relu=True, p=0.5,
                          mask=keep.to(dev, torch.uint8).contiguous())
    finally:
        ops.GemmConfig.fp32_mode = prev
    assert out.dtype == dtype
    assert rel_err(out.cpu(), ref) < (2e-6 if mode != "bf16" else 2e-2)


def test_gemm_nt_kn_weight_and_philox_matches_unfused_epilogue():
    """h_weights-style [K, N] weight through a transposed view; fused Philox dropout == the standalone epilogue kernel."""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(1)
    M, K, N = 515, 448, 256
    x = torch.randn(M, K, generator=gen).to(dev)
    h = (torch.randn(K, N, generator=gen) / 20).to(dev)
    b = torch.randn(N, generator=gen).to(dev)
    fused = ops.gemm_nt([x], [h.t()], bias=b, relu=True, p=0.5, seed=77, stream_id=5)
    plain = ops.gemm_nt([x], [h.t()])
    assert rel_err(plain.cpu(), x.double().cpu() @ h.double().cpu()) < 2e-6
    unfused = torch.empty_like(plain)
    ops.epilogue_fwd_(plain, b, unfused, True, 0.5, None, 77, 5)
    assert torch.equal(fused, unfused)
    frac = float((fused != 0).float().mean())
    assert 0.2 < frac < 0.3          # relu keeps ~half, dropout keeps half of those


def test_gemm_large_k_tail_and_m_tail_fp32_grade():
    """emb1-like: K = 4369 (not a multiple of the 32-element K block), leading dimension padded to 4372."""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(3)
    M, N, K = 1000, 256, 4369
    buf = torch.rand(M, 4372, generator=gen)
    W = torch.randn(N, K, generator=gen) / 60
    ref = buf[:, :K].double() @ W.double().t()
    out = ops.gemm_nt([buf.to(dev)[:, :K]], [W.to(dev)])
    assert rel_err(out.cpu(), ref) < 2e-6
    # unaligned input (ld = 4369): re-laid internally, same result
    out2 = ops.gemm_nt([buf[:, :K].contiguous().to(dev)], [W.to(dev)])
    assert torch.equal(out, out2)


@pytest.mark.parametrize("R,Ma,Nb", [(1000, 256, 128), (128, 128, 128), (33, 40, 24), (5000, 1792, 256), (777, 53, 1280),
                                     (20000, 300, 256), (4100, 16, 128), (1000, 129, 7)])
@pytest.mark.parametrize("mode", ["tf32bc", "tf32x3"])
def test_gemm_tn_weight_gradient(R, Ma, Nb, mode):
    """dW = A^T B over the node dimension (split-K, deterministic reduce), fp32-grade, both parity modes; covers the
    wide-operand swap + transposed store of the TMEM kernel (Nb > Ma), reduction tails (R % 32 != 0) and narrow operands
    whose width is not a multiple of 8."""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(R + Ma + Nb)
    a = torch.randn(R, Ma, generator=gen)
    b = torch.randn(R, Nb, generator=gen)
    ref = a.double().t() @ b.double()
    out = ops.gemm_tn(a.to(dev), b.to(dev), fp32_mode=mode)
    again = ops.gemm_tn(a.to(dev), b.to(dev), fp32_mode=mode)
    assert out.shape == (Ma, Nb) and out.dtype == torch.float32
    assert rel_err(out.cpu(), ref) < 2e-6
    assert torch.equal(out, again)          # split-K partials are reduced in a fixed order


def test_gemm_bc_small_magnitudes_keep_fp32_grade():
    """gradient-sized operands (1e-6): the bf16 correction terms share fp32's exponent range, nothing underflows"""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(5)
    a = torch.randn(2000, 256, generator=gen) * 1e-6
    b = torch.randn(384, 256, generator=gen) * 1e-3
    out = ops.gemm_nt([a.to(dev)], [b.to(dev)], fp32_mode="tf32bc")
    assert rel_err(out.cpu(), a.double() @ b.double().t()) < 2e-6
    dz = torch.randn(2000, 128, generator=gen) * 1e-7
    dw = ops.gemm_tn(a.to(dev), dz.to(dev), fp32_mode="tf32bc")
    assert rel_err(dw.cpu(), a.double().t() @ dz.double()) < 2e-6


def test_gemm_tn_strided_operands_emb1_like():
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(8)
    R, K, N = 3000, 4369, 256
    buf = torch.rand(R, 4376, generator=gen)
    dz = torch.randn(R, N, generator=gen)
    ref = dz.double().t() @ buf[:, :K].double()
    out = ops.gemm_tn(dz.to(dev), buf.to(dev)[:, :K])          # nn.Linear layout [out, in]
    assert rel_err(out.cpu(), ref) < 2e-6


@pytest.mark.parametrize("R,Ma,Nb", [(1000, 256, 128), (4096, 512, 512), (130, 72, 40), (20000, 1792, 512), (777, 56, 1280)])
def test_gemm_tn_bf16_weight_gradient(R, Ma, Nb):
    """bf16 activations, MN-major kind::f16 kernel, fp32 accumulation/output: only the accumulation order differs from
    the fp64 product of the same bf16 values (1e-5); deterministic."""
    from graph_representation_learning_b200 import ops
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(R + Ma + Nb)
    a = torch.randn(R, Ma, generator=gen).bfloat16()
    b = torch.randn(R, Nb, generator=gen).bfloat16()
    ref = a.double().t() @ b.double()
    out = ops.gemm_tn(a.to(dev), b.to(dev))
    assert out.shape == (Ma, Nb) and out.dtype == torch.float32
    assert rel_err(out.cpu(), ref) < 1e-5
    assert torch.equal(out, ops.gemm_tn(a.to(dev), b.to(dev)))


@pytest.mark.parametrize("M,N,K", [(19000, 300, 100), (19100, 128, 64), (40000, 53, 1280)])
def test_tf32bc_weight_multicast_clusters_identical(M, N, K):
    """grl_gemm_bc_set_multicast(1): CTA pairs (thread-block clusters) that share their weight tiles through TMA
    multicast -- odd row-tile counts (the last pair has one tile past M), N tails past the second weight half, short K --
    must reproduce the single-CTA kernel bit for bit."""
    from graph_representation_learning_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=gen).to(dev)
    b = (torch.randn(N, K, generator=gen) / K ** 0.5).to(dev)
    bias = torch.randn(N, generator=gen).to(dev)
    outs = []
    try:
        lib.grl_gemm_bc_set_pairs(0)                 # the cta_group::2 kernel (default) would take both runs
        for on in (0, 1):
            lib.grl_gemm_bc_set_multicast(on)
            outs.append(ops.gemm_nt([a], [b], bias=bias, relu=True, fp32_mode="tf32bc"))
            torch.cuda.synchronize()
    finally:
        lib.grl_gemm_bc_set_multicast(0)
        lib.grl_gemm_bc_set_pairs(1)
    assert torch.equal(outs[0], outs[1])
    ref = torch.relu(a.double() @ b.double().t() + bias.double())
    assert float((outs[1].double() - ref).norm() / ref.norm()) < 2e-6


@pytest.mark.parametrize("M,N,K,mode,two,p_drop", [(19000, 300, 100, "tf32bc", False, 0.0), (19100, 128, 64, "tf32x3", False, 0.5),
                                                   (40000, 53, 1280, "tf32x3", False, 0.0), (9577, 256, 160, "tf32x3", True, 0.0),
                                                   (19001, 200, 4369, "tf32bc", False, 0.3)])
def test_cta_pair_kernel_identical_to_single_cta(M, N, K, mode, two, p_drop):
    """grl_gemm_bc_set_pairs(1) (default): the fp32-grade NT products on CTA pairs -- tcgen05 cta_group::2, one MMA of
    M = 256 over two CTAs that each hold half of every weight tile -- must reproduce the single-CTA kernel bit for bit:
    odd row-tile counts (the pair's second tile past M), N tails inside / past the second weight half, K tails, two K
    segments (virtual concat), bias + ReLU + Philox dropout epilogue."""
    from graph_representation_learning_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M + N + K)
    Kp = (K + 3) // 4 * 4
    a = torch.randn(M, Kp, generator=gen).to(dev)[:, :K]
    b = (torch.randn(N, K, generator=gen) / K ** 0.5).to(dev)
    bias = torch.randn(N, generator=gen).to(dev)
    a_list, b_list = ([a[:, :K // 8 * 4], a[:, K // 8 * 4:]], [b[:, :K // 8 * 4], b[:, K // 8 * 4:]]) if two else ([a], [b])
    outs = []
    try:
        for on in (0, 1):
            lib.grl_gemm_bc_set_pairs(on)
            n0 = _lib.launches()
            outs.append(ops.gemm_nt(a_list, b_list, bias=bias, relu=True, p=p_drop, seed=5, stream_id=2, fp32_mode=mode))
            torch.cuda.synchronize()
    finally:
        lib.grl_gemm_bc_set_pairs(1)
    assert torch.equal(outs[0], outs[1])
    if p_drop == 0.0:
        ref = torch.relu(a.double() @ b.double().t() + bias.double())
        assert float((outs[1].double() - ref).norm() / ref.norm()) < 2e-6


@pytest.mark.parametrize("M,N,K,two", [(19001, 300, 100, False), (19001, 515, 1000, False), (38000, 256, 192, True)])
def test_bf16_cta_pair_kernel_identical_to_single_cta(M, N, K, two):
    """bf16 NT products with N > 128 run on CTA pairs by default (gemm_nt_bf16_2sm_kernel: one 256 x 256 MMA over two CTAs,
    each loading its rows of A and half of the weight tile); same bits as the single-CTA persistent kernel, with odd
    row-tile counts, N tails inside either weight half and two K segments; bias + ReLU + Philox dropout epilogue."""
    from graph_representation_learning_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M + N + K)
    a = torch.randn(M, K, generator=gen).to(dev).bfloat16()
    b = (torch.randn(N, K, generator=gen) / K ** 0.5).to(dev)
    bias = torch.randn(N, generator=gen).to(dev)
    a_list, b_list = ([a[:, :K // 16 * 8], a[:, K // 16 * 8:]], [b[:, :K // 16 * 8], b[:, K // 16 * 8:]]) if two else ([a], [b])
    outs = []
    try:
        for on in (0, 1):
            lib.grl_gemm_bc_set_pairs(on)
            outs.append(ops.gemm_nt(a_list, b_list, bias=bias, relu=True, p=0.5, seed=9, stream_id=1))
            torch.cuda.synchronize()
    finally:
        lib.grl_gemm_bc_set_pairs(1)
    assert torch.equal(outs[0], outs[1])
    plain = ops.gemm_nt(a_list, b_list, out_dtype=torch.float32)
    ref = a.double() @ b.bfloat16().double().t()
    assert float((plain.double() - ref).norm() / ref.norm()) < 1e-5


@pytest.mark.parametrize("M,N,K,two", [(1000, 256, 512, False), (777, 53, 100, False), (4096, 300, 37, False),
                                       (2000, 128, 256, True)])
def test_tf32x3_tmem_and_smem_operand_kernels_identical(M, N, K, two):
    """The three-product arithmetic runs by default on the TMEM-operand kernel (gemm_bc_kernel<NT, x3>); the
    shared-memory-operand kernel (grl_gemm_x3_set_promo(-4)) must give the same bits: same split, same products, same
    accumulation order.  One and two K segments, N and K tails, fused bias + ReLU."""
    from graph_representation_learning_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M + N + K)
    a = [torch.randn(M, K, generator=gen).to(dev)] + ([torch.rand(M, 96, generator=gen).to(dev)] if two else [])
    b = [(torch.randn(N, x.shape[1], generator=gen) / K ** 0.5).to(dev) for x in a]
    bias = torch.randn(N, generator=gen).to(dev)
    outs = []
    try:
        for sw in (-4, -3):
            lib.grl_gemm_x3_set_promo(sw)
            outs.append(ops.gemm_nt(a, b, bias=bias, relu=True, fp32_mode="tf32x3"))
            torch.cuda.synchronize()
    finally:
        lib.grl_gemm_x3_set_promo(-3)
    assert torch.equal(outs[0], outs[1])
    ref = torch.relu(sum(x.double() @ w.double().t() for x, w in zip(a, b)) + bias.double())
    assert float((outs[1].double() - ref).norm() / ref.norm()) < 2e-6
