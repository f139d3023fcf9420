"""Fused GraphConv forward (csrc/graphconv_fused.cu, grl_graphconv_fused_fwd) against the unfused pair
(grl_spmm_fwd_stream + grl_gemm_nt) and against an fp64 restatement of GraphConv.forward
(reference gnn/models/networks/robust_gcn.py:45-51).  Tolerances: 2e-6 norm-wise against fp64 (the three-product GEMM's
own bar, tests/test_gpu_gemm.py), 1e-6 against the unfused path (same arithmetic, different K-block order)."""
import pytest
import torch

from _golden import MODEL_CASES, load_model_case, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


def _fp64_graphconv(g, w, xs, h_weights, bias, relu):
    """new_V = A' V per typed row, then ONE projection over K = (L+1) F (robust_gcn.py:45-51), in fp64 on the device"""
    A = g.to_dense(w).double()                                     # [B, (L+1) N, N]
    X = torch.cat([x.double() for x in xs], 1).view(g.B, g.N, -1)
    new_v = torch.bmm(A, X).view(g.B, g.N, g.L + 1, -1).reshape(g.B * g.N, -1)
    out = new_v @ h_weights.double()
    if bias is not None:
        out = out + bias.double()
    return torch.relu(out) if relu else out


CASES = [
    # B, N,   widths,      C,   density, mode  (0 auto, 1 slab, 2 direct)
    (3, 200, (64,), 128, 0.04, 0),             # one CTA per tile
    (2, 500, (256,), 256, 0.016, 0),           # cfg2's gcn1: clusters of 2, slab
    (2, 300, (128, 128), 512, 0.03, 0),        # virtual concat of two inputs, clusters of 4
    (2, 130, (32,), 64, 0.05, 0),              # C below one tile, N just over one tile
    (1, 77, (96,), 192, 0.1, 0),               # C not a multiple of 128 (second CTA half empty), tiny graph
    (2, 500, (256,), 256, 0.016, 2),           # the same through the direct (global / L2) producer
    (1, 1500, (64,), 256, 0.006, 0),           # slab does not fit -> direct producer chosen
    (2, 257, (64, 32), 128, 0.0, 0),           # no edges at all except nothing: every typed row but the self loop is empty
]


@pytest.mark.parametrize("B,N,widths,C,density,mode", CASES)
def test_fused_graphconv_matches_unfused_and_fp64(dev, B, N, widths, C, density, mode):
    from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops
    gen = torch.Generator().manual_seed(B * 1000 + N + C)
    L = 6
    A = (torch.rand(B, N, L, N, generator=gen) < density).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    xs = [torch.randn(B * N, f, generator=gen).to(dev) for f in widths]
    K = (L + 1) * sum(widths)
    h = (torch.randn(K, C, generator=gen) / K ** 0.5).to(dev)
    bias = torch.randn(C, generator=gen).to(dev)
    lib = _lib.load()
    for p_drop, relu in ((0.0, True), (0.3, False)):
        w = g.edge_weights(p_drop, seed=3, stream_id=1)
        ref = _fp64_graphconv(g, w, xs, h, bias, relu)
        with torch.no_grad():
            unfused = ops.linear_act([ops.spmm_aggregate(g, w, xs)], h, bias, relu, weight_is_kn=True)
            try:
                _lib.check(lib.grl_graphconv_fused_set_mode(mode), "grl_graphconv_fused_set_mode")
                fused = ops.graph_conv_fused(xs, g, w, h, bias, relu)
            finally:
                lib.grl_graphconv_fused_set_mode(0)
        assert fused is not None, "shape should be supported"
        torch.cuda.synchronize()
        assert rel_err(fused, ref) < 2e-6, (p_drop, rel_err(fused, ref))
        assert rel_err(fused, unfused) < 1e-6, (p_drop, rel_err(fused, unfused))


def test_fused_dropout_epilogue_uses_the_same_philox_stream(dev):
    """bias + ReLU + Philox dropout in the fused epilogue drop exactly the elements the unfused GEMM epilogue drops"""
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(5)
    B, N, L, F, C = 2, 260, 6, 64, 256
    A = (torch.rand(B, N, L, N, generator=gen) < 0.03).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    x = torch.randn(B * N, F, generator=gen).to(dev)
    h = (torch.randn((L + 1) * F, C, generator=gen) / 20).to(dev)
    bias = torch.randn(C, generator=gen).to(dev)
    w = g.edge_weights(0.0)
    with torch.no_grad():
        a = ops.linear_act([ops.spmm_aggregate(g, w, [x])], h, bias, True, 0.5, None, 77, 4, weight_is_kn=True)
        b = ops.graph_conv_fused([x], g, w, h, bias, True, 0.5, None, 77, 4)
    assert torch.equal(a == 0, b == 0)
    assert rel_err(b, a) < 1e-6


def test_unsupported_shapes_fall_back(dev):
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(6)
    B, N, L = 1, 40, 6
    g = PackedGraphBatch.from_dense((torch.rand(B, N, L, N, generator=gen) < 0.1).half().to(dev))
    x = torch.randn(B * N, 24, generator=gen).to(dev)             # width not a multiple of 32
    h = torch.randn((L + 1) * 24, 16, generator=gen).to(dev)
    w = g.edge_weights(0.0)
    assert ops.graph_conv_fused([x], g, w, h, None) is None
    with torch.no_grad():
        out = ops.graph_conv([x], g, w, h, None)                  # the public op still answers (unfused pair)
    assert rel_err(out, _fp64_graphconv(g, w, [x], h, None, False)) < 2e-6


@pytest.mark.parametrize("name", MODEL_CASES)
def test_eval_forward_fused_equals_unfused_and_reference(dev, name):
    """whole model, eval: fused GraphConv layers vs the unfused pair (1e-5) and vs the reference's logits"""
    from graph_representation_learning_b200 import GraphCNNDropEdge, _lib, ops
    c = load_model_case(name)
    m = c["meta"]
    model = GraphCNNDropEdge(m["in_dim"], m["out_dim"], m["L"], net_size=m["S"], use_attention=m.get("use_attention", True))
    model.load_state_dict(c["sd0"], strict=True)
    model = model.to(dev).eval()
    V, A = c["V"].to(dev), c["A"].to(dev)
    outs = {}
    for on in (True, False):
        ops.FusedGraphConv.enabled = on
        try:
            n0 = _lib.launches()
            with torch.no_grad():
                outs[on] = model([V, A])
            outs[("launches", on)] = _lib.launches() - n0
        finally:
            ops.FusedGraphConv.enabled = "auto"
    assert outs[("launches", True)] < outs[("launches", False)], "the fused path did not run"
    assert rel_err(outs[True], outs[False]) < 1e-5
    assert rel_err(outs[True].cpu(), c["logits_eval"]) < 1e-5
