"""GPU parity tests: the CUDA path (through the C ABI) against the oracle and the committed reference vectors.

Tolerances (BASELINE.json north_star): integer CSR/CSC work bit-exact; fp32 within 1e-5 norm-wise for outputs
(gradients: max(1e-5, 2 x the reference's own fp32 noise) -- SURVEY.md 8c measured 3-5e-5 on attention-projection
gradients, so 1e-4 is used for gradients); bf16 within 1e-2.
"""
import os

import numpy as np
import pytest
import torch

from _golden import (GOLDEN, MODEL_CASES, PRODUCTION_CASES, TAME_CASES, allowed, load_model_case, load_production_case,
                     production_model, rel_err)
from oracle import reference_port as rp

pytestmark = pytest.mark.gpu

TOL_OUT = 1e-5
TOL_GRAD = 1e-4
TOL_BF16 = 1e-2


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


def _scipy_like_csr(Ap):
    B, RN, N = Ap.shape
    csr = Ap.reshape(B * RN, N).to_sparse_csr()
    return (csr.crow_indices().to(torch.int32), csr.col_indices().to(torch.int32), csr.values())


# ---------------------------------------------------------------------------------------- integer packing
def test_csr_pack_matches_reference_vector(dev):
    from graph_representation_learning_b200 import PackedGraphBatch
    z = np.load(os.path.join(GOLDEN, "preprocess_adj.npz"))
    for dtype in (torch.float16, torch.float32):
        g = PackedGraphBatch.from_dense(torch.from_numpy(z["A"]).to(dev, dtype))
        assert np.array_equal(g.rowptr.cpu().numpy(), z["rowptr"])
        assert np.array_equal(g.col.cpu().numpy(), z["col"])
        assert np.array_equal(g.val.cpu().numpy(), z["val"])


@pytest.mark.parametrize("B,N,L,density", [(1, 1, 6, 1.0), (2, 7, 6, 0.3), (3, 33, 6, 0.1), (2, 64, 3, 0.05),
                                           (4, 200, 6, 0.04), (2, 500, 6, 0.016), (1, 37, 6, 0.0)])
def test_csr_pack_and_transpose_bit_exact(dev, B, N, L, density):
    from graph_representation_learning_b200 import PackedGraphBatch
    gen = torch.Generator().manual_seed(B * 1000 + N)
    A = (torch.rand(B, N, L, N, generator=gen) < density).float()
    A = A * torch.randint(1, 5, A.shape, generator=gen).float() / 4.0     # non-binary weights (fc_similarity style)
    rowptr, col, val = _scipy_like_csr(rp.preprocess_adj(A.permute(0, 1, 3, 2)))
    for dtype in (torch.float32, torch.float16):
        g = PackedGraphBatch.from_dense(A.to(dev, dtype))
        assert torch.equal(g.rowptr.cpu(), rowptr)
        assert torch.equal(g.col.cpu(), col)
        assert torch.equal(g.val.cpu(), val)
        assert g.nnz == val.numel()
    # CSC + perm against a stable sort of the CSR entries by global column
    R = B * N * (L + 1)
    rows = torch.repeat_interleave(torch.arange(R), (rowptr[1:] - rowptr[:-1]).long())
    gcol = (rows // (N * (L + 1))) * N + col.long()
    order = torch.sort(gcol, stable=True).indices
    colptr = torch.zeros(B * N + 1, dtype=torch.int64)
    colptr[1:] = torch.cumsum(torch.bincount(gcol, minlength=B * N), 0)
    g.build_transpose()                                               # the CSC is built on demand (legacy backward)
    assert torch.equal(g.colptr.cpu().long(), colptr)
    assert torch.equal(g.perm.cpu().long(), order)
    assert torch.equal(g.rowidx.cpu().long(), rows[order])
    # dense round trip
    assert torch.equal(g.to_dense().cpu(), rp.preprocess_adj(A.permute(0, 1, 3, 2)))


def test_edge_list_pack_equals_dense_pack(dev):
    from graph_representation_learning_b200 import PackedGraphBatch
    gen = torch.Generator().manual_seed(5)
    B, N, L = 3, 40, 6
    A = (torch.rand(B, N, L, N, generator=gen) < 0.1).float()
    gd = PackedGraphBatch.from_dense(A.to(dev))
    edges = A.nonzero()[:, [0, 1, 2, 3]].to(torch.int32)
    edges = edges[torch.randperm(edges.shape[0], generator=gen)]          # arrival order must not matter
    ge = PackedGraphBatch.from_edges(edges.to(dev), B, N, L)
    gd.build_transpose(), ge.build_transpose()
    for name in ("rowptr", "col", "val", "colptr", "rowidx", "perm"):
        assert torch.equal(getattr(gd, name), getattr(ge, name)), name
    empty = PackedGraphBatch.from_edges(torch.zeros(0, 4, dtype=torch.int32, device=dev), 2, 5, L)
    assert empty.nnz == 10 and torch.equal(empty.col.cpu(), torch.arange(5, dtype=torch.int32).repeat(2))


# ---------------------------------------------------------------------------------------- SpMM
@pytest.mark.parametrize("F", [256, 512, 128, 64, 24, 1024])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spmm_fwd_bwd_against_dense_matmul(dev, F, dtype):
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(F)
    B, N, L = 3, 50, 6
    A = (torch.rand(B, N, L, N, generator=gen) < 0.12).float()
    Ap = rp.preprocess_adj(A.permute(0, 1, 3, 2))
    keep = (torch.rand(Ap.shape, generator=gen) < 0.7)
    Ap_masked = Ap * keep / 0.7
    X = torch.randn(B, N, F, generator=gen)
    dY = torch.randn(B, N, (L + 1) * F, generator=gen)
    if dtype == torch.bfloat16:
        X, dY = X.bfloat16().float(), dY.bfloat16().float()
    ref_Y = torch.matmul(Ap_masked.double(), X.double()).reshape(B * N, (L + 1) * F)
    ref_dX = torch.matmul(Ap_masked.double().transpose(1, 2), dY.double().reshape(B, N * (L + 1), F)).reshape(B * N, F)

    g = PackedGraphBatch.from_dense(A.to(dev))
    w = g.edge_weights(0.3, mask=g.gather_dense_mask(keep.to(dev)))
    Xd = X.reshape(B * N, F).to(dev, dtype).requires_grad_(True)
    Y = ops.spmm_aggregate(g, w, [Xd])
    Y.backward(dY.reshape(B * N, -1).to(dev, dtype))
    tol = TOL_OUT if dtype == torch.float32 else TOL_BF16
    assert rel_err(Y.detach().cpu(), ref_Y) < tol
    assert rel_err(Xd.grad.cpu(), ref_dX) < tol


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,N,F,density", [(3, 50, 256, 0.12), (5, 37, 128, 0.6), (2, 300, 64, 0.03), (4, 129, 512, 0.0),
                                           (2, 513, 256, 0.02)])
def test_spmm_forward_kernels_agree_bit_for_bit(dev, dtype, B, N, F, density):
    """gather (node-per-warp), legacy (row-per-warp) and the TMA-staged shared-memory slab kernel sum every typed row
    in CSR order (dropped entries skipped), so they must agree exactly -- including nodes with more entries than one
    staged batch (density 0.6: the slab-v2 multi-batch path), rows longer than the register-staged 2*GL / 64 entries
    (density 0.6), empty typed rows (density 0), and N that is not a multiple of the TMA box."""
    from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops
    gen = torch.Generator().manual_seed(N * 7 + F)
    A = (torch.rand(B, N, 6, N, generator=gen) < density).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    X = torch.randn(B * N, F, generator=gen).to(dev, dtype)
    lib = _lib.load()
    for p_drop in (0.3, 0.9, 0.0):          # 0.9: typed rows whose entries are ALL dropped must still be written as zeros
        w = g.edge_weights(p_drop, seed=5, stream_id=3)
        assert getattr(w, "_grl_stream", None) is not None
        outs = {}
        try:
            for mode in (-2, -1, 0, 108, 116, 132, 208, 216, 100):
                stream_kernel = mode >= 100              # node-stream slab kernel (100 = its own choice of group width)
                ops._spmm_stream_enabled[0] = stream_kernel
                _lib.check(lib.grl_spmm_set_tile_mode(0 if mode == 100 else mode), "grl_spmm_set_tile_mode")
                Y = torch.full((g.num_rows, F), float("nan"), device=dev, dtype=dtype)
                ops.spmm_fwd_(g, w, X, Y)
                outs[mode] = Y
        finally:
            lib.grl_spmm_set_tile_mode(0)
            ops._spmm_stream_enabled[0] = True
        for mode, Y in outs.items():
            assert torch.equal(Y, outs[-2]), f"mode {mode} differs from the row kernel (p_drop={p_drop})"
    assert lib.grl_spmm_set_tile_mode(5) == -1


def test_spmm_virtual_concat_matches_cat(dev):
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(9)
    B, N, L = 2, 30, 6
    A = (torch.rand(B, N, L, N, generator=gen) < 0.15).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    x1 = torch.randn(B * N, 256, generator=gen).to(dev)
    x2 = torch.randn(B * N, 256, generator=gen).to(dev)
    both = ops.spmm_aggregate(g, g.val, [x1, x2])
    cat = ops.spmm_aggregate(g, g.val, [torch.cat([x1, x2], 1).contiguous()])
    assert torch.equal(both, cat)


def test_dropedge_philox_statistics_and_determinism(dev):
    from graph_representation_learning_b200 import PackedGraphBatch
    B, N, L = 8, 200, 6
    A = (torch.rand(B, N, L, N) < 0.04).half()
    g = PackedGraphBatch.from_dense(A.to(dev))
    w1 = g.edge_weights(0.3, seed=123, stream_id=1)
    w1b = g.edge_weights(0.3, seed=123, stream_id=1)
    w2 = g.edge_weights(0.3, seed=123, stream_id=2)
    assert torch.equal(w1, w1b) and not torch.equal(w1, w2)
    vals = torch.unique(w1)
    assert vals.numel() == 2 and vals[0] == 0 and abs(float(vals[1]) - 1 / 0.7) < 1e-6      # golden fact (iii)
    keep = float((w1 != 0).float().mean())
    assert abs(keep - 0.7) < 4 * (0.21 / g.nnz) ** 0.5 + 1e-3
    self_rows = g.rowptr[0:-1:(L + 1)].long()
    dropped_self = int((w1[self_rows] == 0).sum())
    assert 0 < dropped_self < B * N                                                          # self loops are dropped too
    assert torch.equal(g.edge_weights(0.0), g.val)


# ---------------------------------------------------------------------------------------- epilogue
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bias_relu_dropout_epilogue(dev, dtype):
    from graph_representation_learning_b200 import ops
    gen = torch.Generator().manual_seed(3)
    M, C = 301, 64
    x = torch.randn(M, C, generator=gen).to(dtype)
    b = torch.randn(C, generator=gen)
    keep = (torch.rand(M, C, generator=gen) < 0.5)
    ref = torch.relu(x.float() + b) * keep / 0.5
    xd = x.to(dev)
    y = torch.empty_like(xd)
    ops.epilogue_fwd_(xd, b.to(dev), y, True, 0.5, keep.to(dev, torch.uint8).contiguous(), 0, 0)
    assert rel_err(y.cpu(), ref) < (1e-6 if dtype == torch.float32 else TOL_BF16)
    dy = torch.randn(M, C, generator=gen).to(dtype)
    dx = torch.empty_like(xd)
    ops.epilogue_bwd_(dy.to(dev), y, dx, True, 0.5)
    ref_dx = dy.float() * ((x.float() + b) > 0) * keep / 0.5
    assert rel_err(dx.cpu(), ref_dx) < (1e-6 if dtype == torch.float32 else TOL_BF16)
    assert rel_err(ops.colsum(dy.to(dev)).cpu(), dy.float().sum(0)) < 1e-5
    # Philox path: keep rate and scale
    y2 = torch.empty(4096, 256, device=dev)
    ops.epilogue_fwd_(torch.ones(4096, 256, device=dev), None, y2, False, 0.5, None, 7, 3)
    assert set(torch.unique(y2).tolist()) == {0.0, 2.0}
    assert abs(float((y2 != 0).float().mean()) - 0.5) < 5e-3


# ---------------------------------------------------------------------------------------- whole model
def _model_from_case(c, dev):
    from graph_representation_learning_b200 import GraphCNNDropEdge
    m = c["meta"]
    model = GraphCNNDropEdge(m["in_dim"], m["out_dim"], m["L"], net_size=m["S"], use_attention=m["use_attention"])
    missing = model.load_state_dict(c["sd0"], strict=True)
    assert not missing.missing_keys and not missing.unexpected_keys
    return model.to(dev)


@pytest.mark.parametrize("name", MODEL_CASES)
def test_model_eval_forward_matches_reference_vectors(dev, name):
    c = load_model_case(name)
    model = _model_from_case(c, dev).eval()
    with torch.no_grad():
        out = model([c["V"].to(dev), c["A"].to(dev)])          # fp16 dense adjacency straight from the collate
        out32 = model([c["V"].to(dev), c["A"].float().to(dev)])
    assert out.dtype == torch.float32 and out.shape == c["logits_eval"].shape
    assert rel_err(out.cpu(), c["logits_eval"]) < TOL_OUT
    assert torch.equal(out, out32)


@pytest.mark.parametrize("name", MODEL_CASES)
def test_model_train_step_matches_reference_vectors(dev, name):
    from graph_representation_learning_b200 import PackedGraphBatch
    c = load_model_case(name)
    m = c["meta"]
    B, N = m["B"], m["N"]
    model = _model_from_case(c, dev).train()
    g = PackedGraphBatch.from_dense(c["A"].to(dev))
    masks = {"edge": [g.gather_dense_mask(x.to(dev)) for x in c["edge_masks"]],
             "feat": [x.to(dev).reshape(B * N, -1).contiguous() for x in c["feat_masks"]]}
    opt = torch.optim.Adam([p for p in model.parameters() if p.requires_grad])
    opt.zero_grad()
    logits = model([c["V"].to(dev), g], masks=masks)
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), c["y"].to(dev))
    loss.backward()
    assert rel_err(logits.detach().cpu(), c["logits_train"]) < TOL_OUT
    assert abs(float(loss) - float(c["loss"])) < 1e-5 * max(1.0, float(c["loss"]))
    for k, p in model.named_parameters():
        if p.requires_grad:
            assert rel_err(p.grad.cpu(), c["grad"][k]) < TOL_GRAD, k
        else:
            assert p.grad is None
    norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0)
    assert abs(float(norm) - float(c["grad_norm"])) < 1e-4 * float(c["grad_norm"])
    opt.step()
    for k, v in model.state_dict().items():
        assert rel_err(v.cpu(), c["sd1"][k]) < 1e-5, k


def test_model_bf16_within_stated_tolerance(dev):
    c = load_model_case("model_small")
    model = _model_from_case(c, dev).eval().set_compute_dtype(torch.bfloat16)
    with torch.no_grad():
        out = model([c["V"].to(dev), c["A"].to(dev)])
    assert out.dtype == torch.float32
    assert rel_err(out.cpu(), c["logits_eval"]) < TOL_BF16


def _production_errors(dev, name, dtype):
    """CUDA model (same seed -> the reference constructor's weights, checked) on the regenerated inputs and injected masks
    of a production-shape case; returns {tensor: (our norm-wise error vs the reference's fp64 run, the reference's own
    fp32 noise)} for eval logits, train logits, loss and every gradient (at the stored sample positions)."""
    from graph_representation_learning_b200 import PackedGraphBatch
    c = load_production_case(name)
    m = c["meta"]
    B, N = m["B"], m["N"]
    model = production_model(c).to(dev).set_compute_dtype(dtype)
    V, A, y = c["V"].to(dev), c["A"].to(dev), c["y"].to(dev)
    errs = {}
    model.eval()
    with torch.no_grad():
        errs["logits_eval"] = rel_err(model([V, A]).cpu(), c["logits_eval"])
    model.train()
    g = PackedGraphBatch.from_dense(A)
    masks = {"edge": [g.gather_dense_mask(x.to(dev)) for x in c["edge_masks"]],
             "feat": [x.to(dev).reshape(B * N, -1).contiguous() for x in c["feat_masks"]]}
    logits = model([V, g], masks=masks)
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), y)
    loss.backward()
    errs["logits_train"] = rel_err(logits.detach().cpu(), c["logits_train"])
    errs["loss"] = abs(float(loss) - float(c["loss"])) / float(c["loss"])
    for k, p in model.named_parameters():
        if p.requires_grad:
            idx = c["sample_index"](k, p.numel())
            errs["grad/" + k] = rel_err(p.grad.reshape(-1).cpu()[idx], c["grad"][k])
    table = {k: (v, c["err_ref32"][k]) for k, v in errs.items()}
    out_dir = os.path.join(os.path.dirname(GOLDEN), "..", "gpurun_out")
    try:                                              # the per-tensor table of DESIGN.md section 2 comes from here
        os.makedirs(out_dir, exist_ok=True)
        import json
        with open(os.path.join(out_dir, f"parity_{name}_{'f32' if dtype == torch.float32 else 'bf16'}.json"), "w") as f:
            json.dump({k: {"ours": a, "ref_fp32": b} for k, (a, b) in table.items()}, f, indent=1)
    except OSError:
        pass
    return c, table


@pytest.mark.parametrize("name", PRODUCTION_CASES)
def test_production_shape_fp32_against_reference_fp64_run(dev, name):
    """cfg1 EXACTLY (8 x 200, in_dim 4369, S=256: tiled attention (128,16), K=4369 emb1, long-K promoted accumulators,
    200-node slabs) and an S=512 case (attention (256,32), K=7168), fp32, against vectors of the unmodified reference.
    Tolerance per tensor: max(1e-5, 2 x the reference's own fp32-vs-fp64 noise on that tensor)."""
    c, table = _production_errors(dev, name, torch.float32)
    bad = {k: (ours, allowed(c, k, TOL_OUT)) for k, (ours, _) in table.items() if not ours < allowed(c, k, TOL_OUT)}
    assert not bad, f"(ours, allowed): {bad}"


# tensors whose bf16 gradient is well conditioned: everything downstream of the attention's score path
BF16_CONDITIONED = ("loss", "grad/classifier.bias")


@pytest.mark.parametrize("name", TAME_CASES)
def test_production_shape_bf16_train_step(dev, name):
    """bf16 activations (cfg3's dtype), fp32 master weights, on the well-conditioned (tame) cases, every tensor against the
    reference's fp64 run.  Bar per tensor: the stated 1e-2 -- or, where the reference's OWN bf16 run (torch.autocast on the
    CPU, same masks; stored as err_ref_bf16) is further away than that, twice that run's distance (the fp32 rule of SURVEY.md 8c).  The gradients upstream
    of the unscaled softmax attention are such tensors: dS = P (dP - D) cancels, and the reference's bf16 run is 1.1-2.4
    (norm-wise!) away from its fp64 run there; ours is 0.5-0.9."""
    c, table = _production_errors(dev, name, torch.bfloat16)
    bad = {}
    for k, (ours, _) in table.items():
        bar = max(TOL_BF16, 2.0 * c["err_ref_bf16"][k])
        if not ours < bar:
            bad[k] = (ours, bar)
    assert not bad, f"(ours, allowed): {bad}"
    for k in BF16_CONDITIONED:
        assert table[k][0] < TOL_BF16, (k, table[k])


@pytest.mark.parametrize("name", ["prod_cfg1", "prod_s512"])
def test_production_shape_bf16_at_initial_scale(dev, name):
    """At the constructor's scale of the frozen projection the logits are O(300) and the cross entropy is saturated: the
    gradient hinges on which class wins per node, and the reference's OWN fp32 run is 3-6e-3 away from its fp64 run at
    cfg1 (an amplification of ~5e4 over fp32 epsilon).  A 4e-3 (bf16) perturbation cannot hold 1e-2 on gradients there
    for ANY implementation; what is well defined -- eval logits and the loss -- is held to the stated 1e-2."""
    c, table = _production_errors(dev, name, torch.bfloat16)
    assert table["logits_eval"][0] < max(TOL_BF16, c["err_ref_bf16"]["logits_eval"]) and table["loss"][0] < TOL_BF16, table


def _medium_case_errors(dev, seed):
    """cfg-1-like shape (8 x 200, S=256; input_dim reduced 4369 -> 512 to keep the fp64 CPU oracle fast).
    Returns {tensor: (our error vs the fp64 oracle, allowed)} with
    allowed = max(tolerance, 2 x the error of the reference-style fp32 CPU run against the same fp64 truth)."""
    from graph_representation_learning_b200 import GraphCNNDropEdge
    torch.manual_seed(seed)
    B, N, L, in_dim, S, out_dim = 8, 200, 6, 512, 256, 53
    model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    gen = torch.Generator().manual_seed(seed + 100)
    V = torch.rand(B, N, in_dim, generator=gen)
    A = (torch.rand(B, N, L, N, generator=gen) < 8.0 / N).half()
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    ref_logits, _, ref_grads = rp.loss_and_grads({k: v.double() for k, v in sd.items()}, V.double(), A.double(), y,
                                                 training=False)
    f32_logits, _, f32_grads = rp.loss_and_grads(sd, V, A.float(), y, training=False)
    model = model.to(dev).eval()
    logits = model([V.to(dev), A.to(dev)])
    torch.nn.functional.cross_entropy(logits.transpose(1, 2), y.to(dev)).backward()
    errs = {"logits": (rel_err(logits.detach().cpu(), ref_logits), max(TOL_OUT, 2 * rel_err(f32_logits, ref_logits)))}
    for k, p in model.named_parameters():
        if p.requires_grad:
            errs[k] = (rel_err(p.grad.cpu(), ref_grads[k]), max(TOL_GRAD, 2 * rel_err(f32_grads[k], ref_grads[k])))
    return errs


def test_model_against_fp64_oracle_medium_shape(dev):
    """Realistic regime.  One pre-activation within fp32 noise of zero flips its ReLU derivative and alone moves a
    norm-wise gradient error by ~1e-3/sqrt(n) (measured: scripts/stage_errors.py -- 1 flip in 409 600 at z2 explains
    2e-4 on everything upstream of gcn2), for ANY fp32 implementation: the reference-style fp32 CPU run itself shows
    1.5e-4 .. 1e-3 on some seeds.  So this test takes the MEDIAN over three seeds and allows
    max(3e-4, 2 x reference-fp32 error); the tight 1e-4 gradient bar is held by the golden-vector tests above."""
    runs = [_medium_case_errors(dev, seed) for seed in (21, 22, 23)]
    bad = {}
    for k in runs[0]:
        ours = sorted(r[k][0] for r in runs)[1]
        allowed = max(sorted(r[k][1] for r in runs)[1], 3e-4 if k != "logits" else 0.0)
        if not ours < allowed:
            bad[k] = (ours, allowed)
    assert not bad, f"median (ours, allowed): {bad}; runs: {runs}"


def test_philox_training_is_reproducible_and_unbiased(dev):
    from graph_representation_learning_b200 import GraphCNNDropEdge
    torch.manual_seed(5)
    model = GraphCNNDropEdge(64, 8, 6, net_size=64).to(dev).train()
    V = torch.rand(4, 40, 64, device=dev)
    A = (torch.rand(4, 40, 6, 40, device=dev) < 0.1).half()
    model.rng_seed, model._rng_calls = 99, 0
    a = model([V, A])
    model._rng_calls = 0
    b = model([V, A])
    c = model([V, A])
    assert torch.equal(a, b) and not torch.equal(a, c)


def test_no_cpu_fallback(dev):
    from graph_representation_learning_b200 import GraphCNNDropEdge, GrlError
    model = GraphCNNDropEdge(16, 4, 6, net_size=32)
    with pytest.raises(GrlError):
        model([torch.rand(1, 5, 16), torch.zeros(1, 5, 6, 5)])


# ---------------------------------------------------------------------------------------- NodeSelfAtten core
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,N,H,D", [(3, 150, 128, 16), (2, 77, 256, 32), (2, 500, 128, 16), (2, 50, 32, 4), (3, 9, 16, 2),
                                     (1, 64, 128, 16), (1, 2100, 128, 16)])
def test_self_attention_matches_reference_math(dev, dtype, B, N, H, D):
    """gamma * (softmax(f g^T) h) + V and every gradient against an fp64 restatement of robust_gcn.py:94-96
    (tiled kernels for (128,16) / (256,32), generic kernels otherwise; N not a multiple of the 64-row tile)."""
    from graph_representation_learning_b200 import ops
    gen = torch.Generator().manual_seed(B * 1000 + N + H)
    f = torch.rand(B, N, D, generator=gen) * 1.5
    g = torch.rand(B, N, D, generator=gen) * 1.5
    h = torch.rand(B, N, H, generator=gen)
    V = torch.randn(B, N, H, generator=gen)
    gamma = torch.randn(H, generator=gen)
    dOut = torch.randn(B, N, H, generator=gen)
    if dtype == torch.bfloat16:
        f, g, h, V, dOut = (t.bfloat16().float() for t in (f, g, h, V, dOut))
    ref_in = [t.double().requires_grad_(True) for t in (f, g, h, gamma, V)]
    s = torch.softmax(torch.matmul(ref_in[0], ref_in[1].transpose(1, 2)), dim=-1)
    ref_out = ref_in[3] * torch.matmul(s, ref_in[2]) + ref_in[4]
    ref_out.backward(dOut.double())

    ins = [t.to(dev, dtype).requires_grad_(True) for t in (f, g, h)] + [gamma.to(dev).requires_grad_(True)] + \
          [V.to(dev, dtype).requires_grad_(True)]
    out = ops.self_attention(*ins)
    out.backward(dOut.to(dev, dtype))
    tol_o, tol_g = (TOL_OUT, TOL_GRAD) if dtype == torch.float32 else (TOL_BF16, 2 * TOL_BF16)
    assert rel_err(out.detach().float().cpu(), ref_out.detach()) < tol_o
    for name, a, r in zip("f g h gamma V".split(), ins, ref_in):
        assert rel_err(a.grad.float().cpu(), r.grad) < tol_g, name
    # eval: no N x N tensor is saved, same output
    with torch.no_grad():
        out2 = ops.self_attention(*[t.detach() for t in ins])
    assert torch.equal(out2, out.detach())


# ---------------------------------------------------------------------------------------- large graphs (cfg4 / cfg5 sizes)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_spmm_large_graph_against_sparse_library_and_adjoint(dev, dtype):
    """One 150 000-node graph (the slab does not fit: gather kernels), packed from an edge list with duplicates.
    Checker: torch.sparse on the same device in fp64 (test infrastructure only), plus the size-independent adjoint
    identity <dY, A X> == <A^T dY, X> that ties the transposed kernel to the forward one."""
    from graph_representation_learning_b200 import PackedGraphBatch, ops, synthetic
    wl = synthetic.Workload("large", 1, 150_000, input_dim=8, degree=12.0)
    _, edges, _ = synthetic.make_edge_batch(wl, 5, dev, features=False)
    g = PackedGraphBatch.from_edges(edges, 1, wl.N, wl.num_edges)
    assert g.nnz == edges.shape[0] + wl.N
    F = 128
    gen = torch.Generator(device=dev).manual_seed(3)
    X = torch.randn(wl.N, F, device=dev, generator=gen).to(dtype)
    dY = torch.randn(g.num_rows, F, device=dev, generator=gen).to(dtype)
    w = g.edge_weights(0.3, seed=9, stream_id=2)
    Y = torch.empty(g.num_rows, F, device=dev, dtype=dtype)
    ops.spmm_fwd_(g, w, X, Y)
    dX = torch.empty(wl.N, F, device=dev, dtype=dtype)
    ops.spmm_bwd_(g, w, dY, dX)
    A = torch.sparse_csr_tensor(g.rowptr.long(), g.col.long(), w.double(), size=(g.num_rows, wl.N))
    ref_Y = torch.sparse.mm(A, X.double())
    ref_dX = torch.sparse.mm(A.to_sparse_coo().t().coalesce(), dY.double())
    tol = TOL_OUT if dtype == torch.float32 else TOL_BF16
    assert rel_err(Y.float().cpu(), ref_Y.cpu()) < tol
    assert rel_err(dX.float().cpu(), ref_dX.cpu()) < tol
    lhs = float((dY.double() * Y.double()).sum())
    rhs = float((dX.double() * X.double()).sum())
    assert abs(lhs - rhs) <= (1e-5 if dtype == torch.float32 else 2e-2) * max(abs(lhs), abs(rhs), 1.0)


def test_model_eval_on_edge_list_equals_dense_path(dev):
    """from_edges (cfg4 / cfg5 entry) and from_dense (the reference collate's format) give the same packed graph and
    bit-identical eval logits."""
    from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch, synthetic
    wl = synthetic.Workload("eq", 3, 130, input_dim=40, output_dim=7, net_size=256)
    V, A, _ = synthetic.make_batch(wl, 11, dev)
    b, i, l, j = torch.nonzero(A, as_tuple=True)
    edges = torch.stack([b, i, l, j], 1).to(torch.int32)
    edges = edges[torch.randperm(edges.shape[0], device=dev)]                    # the pack must not depend on edge order
    g_d, g_e = PackedGraphBatch.from_dense(A), PackedGraphBatch.from_edges(edges, wl.B, wl.N, wl.num_edges)
    g_d.build_transpose(), g_e.build_transpose()
    for name in ("rowptr", "col", "val", "colptr", "rowidx", "perm"):
        assert torch.equal(getattr(g_d, name), getattr(g_e, name)), name
    torch.manual_seed(2)
    model = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size).to(dev).eval()
    with torch.no_grad():
        assert torch.equal(model([V, g_d]), model([V, g_e]))


# ---------------------------------------------------------------------------------------- sibling model
def test_robust_gcn_matches_reference_vectors(dev):
    """RobustGCN (gnn/models/networks/robust_gcn.py:173-231) on the same kernels: eval logits, and train-mode logits /
    loss / every gradient with the six injected feature-dropout masks, against vectors of the unmodified reference."""
    import json
    import os

    import numpy as np
    from _golden import GOLDEN
    from graph_representation_learning_b200 import RobustGCN
    z = np.load(os.path.join(GOLDEN, "robustgcn_small.npz"))
    m = json.loads(str(z["meta"]))
    B, N, S = m["B"], m["N"], m["S"]
    model = RobustGCN(m["in_dim"], m["out_dim"], m["L"], net_size=S)
    sd0 = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd0/")}
    res = model.load_state_dict(sd0, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    model = model.to(dev)
    V, A, y = (torch.from_numpy(z[k]).to(dev) for k in ("V", "A", "y"))
    model.eval()
    with torch.no_grad():
        out = model([V, A])
    assert rel_err(out.cpu(), torch.from_numpy(z["logits_eval"])) < TOL_OUT
    widths = [S, S, S, S, S // 2, S // 2]
    fm = [torch.from_numpy(np.unpackbits(z[f"feat_mask{i}"])[: B * N * w].reshape(B * N, w).copy()).to(dev).contiguous()
          for i, w in enumerate(widths)]
    model.train()
    logits = model([V, A], masks={"feat": fm})
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), y)
    loss.backward()
    assert rel_err(logits.detach().cpu(), torch.from_numpy(z["logits_train"])) < TOL_OUT
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * max(1.0, float(z["loss"]))
    for k, p in model.named_parameters():
        assert rel_err(p.grad.cpu(), torch.from_numpy(z["grad/" + k])) < TOL_GRAD, k


def test_rp_graphcnn_dropedge_matches_reference_vectors(dev):
    """RPGraphCNNDropEdge (gnn/models/networks/rp_robust_gcn.py:29-115): random projection in front of the attention,
    leaky-ReLU epilogues (GRL_EPI_LEAKY), attention over rp_size (generic-width kernels).  Eval logits and train-mode
    logits / loss / every gradient with injected DropEdge + dropout masks, against vectors of the unmodified reference."""
    import json
    import os

    import numpy as np
    from _golden import GOLDEN
    from graph_representation_learning_b200 import PackedGraphBatch, RPGraphCNNDropEdge
    z = np.load(os.path.join(GOLDEN, "rpdropedge_small.npz"))
    m = json.loads(str(z["meta"]))
    B, N, S, L, rp = m["B"], m["N"], m["S"], m["L"], m["rp_size"]
    model = RPGraphCNNDropEdge(m["in_dim"], m["out_dim"], L, net_size=S, rp_size=rp)
    sd0 = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd0/")}
    res = model.load_state_dict(sd0, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    assert not model.rp_emb.projection.weight.requires_grad and not model.rp_final.projection.weight.requires_grad
    model = model.to(dev)
    V, A, y = (torch.from_numpy(z[k]).to(dev) for k in ("V", "A", "y"))
    model.eval()
    with torch.no_grad():
        out = model([V, A])
    assert rel_err(out.cpu(), torch.from_numpy(z["logits_eval"])) < TOL_OUT
    g = PackedGraphBatch.from_dense(A)
    em = [g.gather_dense_mask(torch.from_numpy(np.unpackbits(z[f"edge_mask{i}"])[: B * (L + 1) * N * N]
                                               .reshape(B, (L + 1) * N, N).copy()).to(dev)) for i in range(3)]
    fm = [torch.from_numpy(np.unpackbits(z[f"feat_mask{i}"])[: B * N * w].reshape(B * N, w).copy()).to(dev).contiguous()
          for i, w in enumerate([S, S, S, S, rp])]
    model.train()
    logits = model([V, g], masks={"edge": em, "feat": fm})
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), y)
    loss.backward()
    assert rel_err(logits.detach().cpu(), torch.from_numpy(z["logits_train"])) < TOL_OUT
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * max(1.0, float(z["loss"]))
    n = 0
    for k, p in model.named_parameters():
        if p.requires_grad:
            assert rel_err(p.grad.cpu(), torch.from_numpy(z["grad/" + k])) < TOL_GRAD, k
            n += 1
    assert n == len([k for k in z.files if k.startswith("grad/")])


def test_deep_rp_robust_gcn_matches_reference_vectors(dev):
    """DeepRPRobustGCN (gnn/models/networks/deep_rp_robust_gcn.py:65-168): nine GraphConv + BatchNorm1d + leaky(0.2)
    blocks on the packed graph (csrc/batchnorm.cu).  Eval logits (running statistics), train-mode logits / loss / every
    gradient with injected DropEdge + dropout masks, and the BatchNorm running statistics after the train forward,
    against vectors of the unmodified reference."""
    import json
    import os

    import numpy as np
    from _golden import GOLDEN
    from graph_representation_learning_b200 import DeepRPRobustGCN, PackedGraphBatch
    z = np.load(os.path.join(GOLDEN, "deeprp_small.npz"))
    m = json.loads(str(z["meta"]))
    B, N, S, L = m["B"], m["N"], m["S"], m["L"]
    model = DeepRPRobustGCN(m["in_dim"], m["out_dim"], L, net_size=S, lambda_value=m["lambda_value"])
    sd0 = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd0/")}
    res = model.load_state_dict(sd0, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    model = model.to(dev)
    V, A, y = (torch.from_numpy(z[k]).to(dev) for k in ("V", "A", "y"))
    model.eval()
    with torch.no_grad():
        out = model([V, A])
    assert rel_err(out.cpu(), torch.from_numpy(z["logits_eval"])) < TOL_OUT
    g = PackedGraphBatch.from_dense(A)
    em = [g.gather_dense_mask(torch.from_numpy(np.unpackbits(z[f"edge_mask{i}"])[: B * (L + 1) * N * N]
                                               .reshape(B, (L + 1) * N, N).copy()).to(dev)) for i in range(4)]
    fm = [torch.from_numpy(np.unpackbits(z[f"feat_mask{i}"])[: B * N * S].reshape(B * N, S).copy()).to(dev).contiguous()
          for i in range(4)]
    model.train()
    logits = model([V, g], masks={"edge": em, "feat": fm})
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), y)
    loss.backward()
    assert rel_err(logits.detach().cpu(), torch.from_numpy(z["logits_train"])) < TOL_OUT
    assert abs(float(loss) - float(z["loss"])) < 1e-5 * max(1.0, float(z["loss"]))
    n = 0
    for k, p in model.named_parameters():
        if p.requires_grad:
            ref = torch.from_numpy(z["grad/" + k])
            if k.endswith(("emb.bias", "gcn.bias")):
                # a bias in front of BatchNorm has zero gradient (the batch mean removes it): both sides hold rounding noise
                assert float(ref.abs().max()) < 1e-6 and float(p.grad.abs().max()) < 1e-6, k
            else:
                assert rel_err(p.grad.cpu(), ref) < TOL_GRAD, k
            n += 1
    assert n == len([k for k in z.files if k.startswith("grad/")])
    sd = model.state_dict()
    for k in z.files:
        if k.startswith("sd1/"):
            name = k.split("/", 1)[1]
            if name.endswith("num_batches_tracked"):
                assert int(sd[name]) == int(z[k]), name
            else:
                assert rel_err(sd[name].cpu(), torch.from_numpy(z[k])) < 1e-6, name


def test_deep_rp_gcn_matches_reference_vectors(dev):
    """DeepRPGCN (gnn/models/networks/deep_rp_gcn.py:106-150): 29 GraphConv + BatchNorm1d + leaky(0.2) blocks, every third
    on the concat [feats, feats]; eval logits, train-mode logits / loss / gradients against vectors of the unmodified
    reference.  29 normalised layers amplify fp32 rounding: the bars are 10x the single-model ones (the reference's own
    fp32 run differs from its fp64 run by as much)."""
    import json
    import os

    import numpy as np
    from _golden import GOLDEN
    from graph_representation_learning_b200 import DeepRPGCN
    z = np.load(os.path.join(GOLDEN, "deeprpgcn_small.npz"))
    m = json.loads(str(z["meta"]))
    B, N, S, L = m["B"], m["N"], m["S"], m["L"]
    model = DeepRPGCN(m["in_dim"], m["out_dim"], L, net_size=S)
    sd0 = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith("sd0/")}
    res = model.load_state_dict(sd0, strict=True)
    assert not res.missing_keys and not res.unexpected_keys
    model = model.to(dev)
    V, A, y = (torch.from_numpy(z[k]).to(dev) for k in ("V", "A", "y"))
    model.eval()
    with torch.no_grad():
        out = model([V, A])
    assert rel_err(out.cpu(), torch.from_numpy(z["logits_eval"])) < 10 * TOL_OUT
    fm = [torch.from_numpy(np.unpackbits(z["feat_mask0"])[: B * N * S].reshape(B * N, S).copy()).to(dev).contiguous()]
    model.train()
    logits = model([V, A], masks={"feat": fm})
    loss = torch.nn.functional.cross_entropy(logits.transpose(1, 2), y)
    loss.backward()
    assert rel_err(logits.detach().cpu(), torch.from_numpy(z["logits_train"])) < 10 * TOL_OUT
    assert abs(float(loss.detach()) - float(z["loss"])) < 1e-4 * max(1.0, float(z["loss"]))
    worst = {}
    for k, p in model.named_parameters():
        ref = torch.from_numpy(z["grad/" + k])
        if k.endswith(("emb.bias", "gcn.bias")):
            assert float(p.grad.abs().max()) < 1e-5, k           # zero in exact arithmetic (BatchNorm removes the mean)
        else:
            worst[k] = rel_err(p.grad.cpu(), ref)
    assert max(worst.values()) < 10 * TOL_GRAD, sorted(worst.items(), key=lambda kv: -kv[1])[:3]


@pytest.mark.parametrize("B,N,L,dtype,density,weighted", [
    (3, 500, 6, torch.float16, 0.016, False), (2, 13, 6, torch.float32, 0.3, True), (1, 1030, 2, torch.float16, 0.05, True),
    (2, 64, 6, torch.float16, 0.2, False), (2, 96, 3, torch.float32, 0.0, False), (1, 37, 6, torch.float16, 1.0, False)])
def test_csr_pack_single_read_bitmap_equals_two_pass_pack(dev, B, N, L, dtype, density, weighted):
    """grl_csr_count_bitmap / grl_csr_fill_bitmap (one read of the dense adjacency) against grl_csr_count / grl_csr_fill:
    rowptr, col, val bit-exact -- 16-byte, 8-byte and scalar load paths, rows longer than 32 bitmap words, empty and full
    rows, fp16 similarity weights, -0.0 treated as zero."""
    from graph_representation_learning_b200 import _lib
    lib = _lib.load()
    gen = torch.Generator().manual_seed(B * 100 + N)
    A = (torch.rand(B, N, L, N, generator=gen) < density).float()
    if weighted:
        A = A * (0.25 + torch.rand(B, N, L, N, generator=gen))
    A[0, 0, 0, N - 1] = -0.0
    A = A.to(dtype).to(dev).contiguous()
    R = B * N * (L + 1)
    st, code = _lib.stream_ptr(dev), _lib.dtype_code(A)
    outs = []
    for use_bitmap in (False, True):
        rowptr = torch.empty(R + 1, dtype=torch.int32, device=dev)
        bitmap = torch.empty(lib.grl_csr_bitmap_bytes(B, N, L), dtype=torch.uint8, device=dev)
        if use_bitmap:
            _lib.check(lib.grl_csr_count_bitmap(A.data_ptr(), code, B, N, L, 1, rowptr.data_ptr(), bitmap.data_ptr(), st), "count")
        else:
            _lib.check(lib.grl_csr_count(A.data_ptr(), code, B, N, L, 1, rowptr.data_ptr(), st), "count")
        ws = torch.empty(lib.grl_scan_workspace_bytes(R), dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_scan_inclusive_i32(rowptr.data_ptr() + 4, R, ws.data_ptr(), ws.numel(), st), "scan")
        nnz = int(rowptr[-1])
        col = torch.full((nnz,), -1, dtype=torch.int32, device=dev)
        val = torch.full((nnz,), float("nan"), device=dev)
        if use_bitmap:
            _lib.check(lib.grl_csr_fill_bitmap(A.data_ptr(), code, bitmap.data_ptr(), B, N, L, 1, rowptr.data_ptr(),
                                               col.data_ptr(), val.data_ptr(), st), "fill")
        else:
            _lib.check(lib.grl_csr_fill(A.data_ptr(), code, B, N, L, 1, rowptr.data_ptr(), col.data_ptr(), val.data_ptr(), st), "fill")
        outs.append((rowptr, col, val))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
    assert int(outs[1][0][-1]) == int((A.float() != 0).sum()) + B * N


@pytest.mark.parametrize("dtype", [torch.float16, torch.float32])
@pytest.mark.parametrize("odd", [None, 2.0, -1.0, 0.5, 6.1e-5, 0.0625, 1.5])
def test_csr_pack_fetches_values_unless_every_one_is_1(dev, dtype, odd):
    """the count pass flags a batch that holds ANY stored value other than +1 (one element out of 1.2 M here: -1, a value
    whose bits are a subset of 1.0's, a near-subnormal half); only an all-ones batch lets the fill pass skip the dense tensor"""
    from graph_representation_learning_b200 import PackedGraphBatch
    gen = torch.Generator().manual_seed(11)
    B, N, L = 2, 100, 6
    A = (torch.rand(B, N, L, N, generator=gen) < 0.05).float()
    if odd is not None:
        A[1, 57, 3, 99] = odd
    A = A.to(dtype).to(dev)
    g = PackedGraphBatch.from_dense(A, with_transpose=False)
    dense = g.to_dense(g.val)                                       # [B, (L+1) N, N], identity block at t = 0
    eye = torch.eye(N, device=dev).view(1, N, 1, N).expand(B, N, 1, N)
    want = torch.cat([eye, A.float()], 2).reshape(B, N * (L + 1), N)
    assert torch.equal(dense, want)


# ---------------------------------------------------------------------------------------- advisor findings (round 1)
@pytest.mark.parametrize("L", [30, 31, 45])
def test_aggregation_with_many_edge_types(dev, L):
    """dropedge_stream_kernel keeps a node's L + 2 row pointers one per lane, so it serves L + 1 <= 31; wider models go
    through grl_dropedge_weights + the gather kernels.  Both sides of the boundary against the dense matmul."""
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    gen = torch.Generator().manual_seed(L)
    B, N, F = 2, 40, 64
    A = (torch.rand(B, N, L, N, generator=gen) < 0.1).float()
    Ap = rp.preprocess_adj(A.permute(0, 1, 3, 2))
    keep = torch.rand(Ap.shape, generator=gen) < 0.7
    X = torch.randn(B, N, F, generator=gen)
    dY = torch.randn(B * N, (L + 1) * F, generator=gen)
    ref_Y = torch.matmul((Ap * keep / 0.7).double(), X.double()).reshape(B * N, (L + 1) * F)
    ref_dX = torch.matmul((Ap * keep / 0.7).double().transpose(1, 2), dY.double().reshape(B, N * (L + 1), F)).reshape(B * N, F)
    g = PackedGraphBatch.from_dense(A.to(dev))
    w = g.edge_weights(0.3, mask=g.gather_dense_mask(keep.to(dev)))
    assert (getattr(w, "_grl_stream", None) is not None) == (L + 1 <= 31)
    Xd = X.reshape(B * N, F).to(dev).requires_grad_(True)
    Y = ops.spmm_aggregate(g, w, [Xd])
    Y.backward(dY.to(dev))
    assert rel_err(Y.detach().cpu(), ref_Y) < TOL_OUT and rel_err(Xd.grad.cpu(), ref_dX) < TOL_OUT
    # the Philox path writes every weight, too
    w2 = g.edge_weights(0.3, seed=9, stream_id=2)
    assert bool(torch.isfinite(w2).all()) and set(torch.unique(w2 * 0.7).round(decimals=4).tolist()) <= {0.0, 1.0}


def test_edge_list_duplicates_and_out_of_range_entries(dev):
    """Repeated (b, i, l, j) triples aggregate like the reference's idempotent dense build adj[start, label, end] = 1
    (graph_utils.py:821-829); triples outside the batch are neutralised on the device (no out-of-bounds write) and
    rejected by the host collate."""
    from graph_representation_learning_b200 import PackedGraphBatch, ops
    from graph_representation_learning_b200.collate import EdgeListCollate
    gen = torch.Generator().manual_seed(3)
    B, N, L, F = 2, 20, 6, 32
    A = (torch.rand(B, N, L, N, generator=gen) < 0.15).float()
    edges = A.nonzero().to(torch.int32)
    dup = torch.cat([edges, edges[::3], edges[::5]])[torch.randperm(edges.shape[0] + len(edges[::3]) + len(edges[::5]), generator=gen)]
    bad = torch.tensor([[0, N, 0, 1], [B, 0, 0, 0], [0, 1, L, 2], [1, 2, 3, -1]], dtype=torch.int32)
    X = torch.randn(B * N, F, generator=gen).to(dev)
    gd = PackedGraphBatch.from_dense(A.to(dev))
    ge = PackedGraphBatch.from_edges(torch.cat([dup, bad]).to(dev), B, N, L)
    assert ge.nnz == dup.shape[0] + 4 + B * N
    Yd = ops.spmm_aggregate(gd, gd.edge_weights(0.0), [X])
    Ye = ops.spmm_aggregate(ge, ge.edge_weights(0.0), [X])
    assert torch.equal(Yd, Ye)
    assert torch.equal(ge.to_dense().cpu(), rp.preprocess_adj(A.permute(0, 1, 3, 2)))
    sample = {"textline_encoding": np.zeros((5, 4), np.float32), "node_label": np.zeros(5, np.int64),
              "edges": np.array([[0, 0, 5]], np.int32)}
    with pytest.raises(ValueError):
        EdgeListCollate(num_edges=L, pin_memory=False)([sample])
