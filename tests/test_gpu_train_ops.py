"""GPU tests of the loss / clip / optimizer kernels (csrc/train_ops.cu) against the library calls the reference makes:
nn.CrossEntropyLoss (cross_entropy_loss.py:19-35), clip_grad_norm_ (kv_procedure.py:161), torch.optim.Adam
(builtin_optimizer.py:22-26)."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dev():
    from graph_representation_learning_b200 import _lib
    assert _lib.load().grl_device_ok() == 1
    return torch.device("cuda:0")


@pytest.mark.parametrize("weighted", [False, True])
@pytest.mark.parametrize("B,N,C", [(3, 50, 53), (2, 7, 5), (1, 1, 2), (4, 333, 70), (2, 40, 64), (2, 40, 33), (40, 500, 32)])
def test_cross_entropy_matches_torch(dev, B, N, C, weighted):
    from graph_representation_learning_b200.trainer import CrossEntropyLoss
    gen = torch.Generator().manual_seed(B * 100 + N)
    pred = (torch.randn(B, N, C, generator=gen) * 4).to(dev).requires_grad_(True)
    y = torch.randint(0, C, (B, N), generator=gen)
    y[torch.rand(B, N, generator=gen) < 0.3] = -100                      # the collate's label padding
    y[0, 0] = 1                                                          # at least one valid row
    w = (torch.rand(C, generator=gen) + 0.5) if weighted else None
    crit = CrossEntropyLoss(None if w is None else w.tolist()).to(dev)
    loss = crit(pred, y.to(dev))
    (loss * 1.7).backward()
    ref_in = pred.detach().double().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_in.transpose(1, 2), y.to(dev), weight=None if w is None else w.double().to(dev))
    (ref * 1.7).backward()
    assert abs(float(loss) - float(ref)) <= 2e-6 * max(1.0, abs(float(ref)))
    assert float((pred.grad.double() - ref_in.grad).norm() / ref_in.grad.norm()) < 1e-6
    assert torch.equal(pred.grad[(y == -100).to(dev)], torch.zeros_like(pred.grad[(y == -100).to(dev)]))


def test_grad_norm_and_clip(dev):
    from graph_representation_learning_b200.trainer import FlatGradBucket
    torch.manual_seed(3)
    model = torch.nn.Sequential(torch.nn.Linear(301, 77), torch.nn.Linear(77, 5)).to(dev)
    b = FlatGradBucket(model)
    b.flat.normal_()
    ref = torch.linalg.vector_norm(b.flat.double())
    n = b.norm()
    assert abs(float(n) - float(ref)) < 1e-6 * float(ref)
    assert torch.equal(b.norm(), n)                                      # deterministic
    before = b.flat.clone()
    total = b.clip_(1.5)
    assert abs(float(total) - float(ref)) < 1e-6 * float(ref)
    assert float((b.flat - before * (1.5 / (float(ref) + 1e-6))).abs().max()) < 1e-6


@pytest.mark.parametrize("wd", [0.0, 0.01])
def test_flat_adam_matches_torch_adam(dev, wd):
    from graph_representation_learning_b200.trainer import FlatAdam
    torch.manual_seed(5)
    m1 = torch.nn.Sequential(torch.nn.Linear(40, 33), torch.nn.ReLU(), torch.nn.Linear(33, 7)).to(dev)
    m2 = copy.deepcopy(m1)
    o1 = FlatAdam(m1.parameters(), weight_decay=wd)
    o2 = torch.optim.Adam(m2.parameters(), weight_decay=wd)
    x = torch.randn(64, 40, device=dev)
    for it in range(4):
        if it == 2:
            for o in (o1, o2):
                o.param_groups[0]["lr"] = 3e-4                           # the per-epoch scheduler pokes lr like this
        for m, o in ((m1, o1), (m2, o2)):
            o.zero_grad()
            m(x).square().mean().backward()
            o.step()
    for p1, p2 in zip(m1.parameters(), m2.parameters()):
        assert float((p1 - p2).abs().max()) < 2e-6
    # state_dict has torch.optim.Adam's layout and round-trips (also INTO a stock torch Adam)
    sd = o1.state_dict()
    assert set(sd["state"][0]) == {"step", "exp_avg", "exp_avg_sq"} and int(sd["state"][0]["step"]) == 4
    o3 = torch.optim.Adam(m2.parameters(), weight_decay=wd)
    o3.load_state_dict(sd)
    o4 = FlatAdam(m1.parameters(), weight_decay=wd)
    o4.load_state_dict(o2.state_dict())
    assert o4._step == 4
    for a, b in zip(o4.state_dict()["state"].values(), o2.state_dict()["state"].values()):
        assert torch.allclose(a["exp_avg"], b["exp_avg"]) and torch.allclose(a["exp_avg_sq"], b["exp_avg_sq"])


def test_train_step_with_own_loss_clip_adam_equals_library_sequence(dev):
    """KVTrainStep (CE kernel -> flat bucket -> norm kernel -> FlatAdam with the clip folded in) against the reference's
    sequence spelled with library calls (F.cross_entropy, clip_grad_norm_, torch.optim.Adam) on the same model, data and
    Philox DropEdge / dropout streams."""
    from graph_representation_learning_b200 import GraphCNNDropEdge, synthetic
    from graph_representation_learning_b200.trainer import FlatAdam, KVTrainStep
    wl = synthetic.Workload("t", 3, 60, input_dim=48, output_dim=9, net_size=64)
    V, A, y = synthetic.make_batch(wl, 4, dev)
    y[0, :7] = -100
    torch.manual_seed(1)
    ma = GraphCNNDropEdge(wl.input_dim, wl.output_dim, wl.num_edges, wl.net_size).to(dev)
    mb = copy.deepcopy(ma)
    ma.rng_seed = mb.rng_seed = 77
    step = KVTrainStep(ma, max_grad_norm=0.5)
    assert isinstance(step.optimizer, FlatAdam)
    opt = torch.optim.Adam([p for p in mb.parameters() if p.requires_grad])
    before = {k: p.detach().clone() for k, p in mb.named_parameters()}
    la = step.run_train_step({"textline_encoding": V, "adjacency_matrix": A, "node_label": y})
    mb.train()
    opt.zero_grad()
    lb = torch.nn.functional.cross_entropy(mb([V, A]).transpose(1, 2), y)
    lb.backward()
    norm_b = torch.nn.utils.clip_grad_norm_(mb.parameters(), 0.5)
    opt.step()
    assert abs(float(la) - float(lb)) <= 2e-6 * max(1.0, abs(float(lb)))
    # the bucket holds the CLIPPED gradients after the step, like clip_grad_norm_ leaves them in p.grad
    assert abs(float(step.bucket.flat.norm()) - min(float(norm_b), 0.5)) <= 1e-4
    lr = 1e-3
    n_all = n_far = 0
    for (k, pa), (_, pb) in zip(ma.named_parameters(), mb.named_parameters()):
        if not pb.requires_grad:
            assert torch.equal(pa, pb), k
            continue
        d = (pa - pb).abs()
        # Adam's first step is lr * g / (|g| + eps): an entry whose gradient is at rounding level may step the other
        # way (CE kernel vs library differ by ~1e-7 relative), never further than 2 * lr; everything else agrees
        assert float(d.max()) <= 2 * lr * 1.001, k
        assert float((pb - before[k]).abs().max()) <= lr * 1.001, k
        n_all += d.numel()
        n_far += int((d > 1e-6).sum())
    assert n_far <= 0.002 * n_all, (n_far, n_all)


@pytest.mark.parametrize("M,C,p", [(1000, 256, 0.5), (130, 16, 0.0), (4097, 128, 0.5), (33, 512, 0.5)])
def test_dz_split_matches_separate_kernels(M, C, p):
    """grl_dz_split (one pass: dZ, tf32 / bf16 split of dZ, bias gradient) against grl_bias_act_dropout_bwd +
    grl_split_act_mn (bit-exact) and an fp64 column sum (fixed-order fp32 reduction: 1e-6)."""
    from graph_representation_learning_b200 import _lib, ops
    lib = _lib.load()
    dev = torch.device("cuda:0")
    gen = torch.Generator().manual_seed(M + C)
    dY = torch.randn(M, C, generator=gen).to(dev)
    Y = (torch.randn(M, C, generator=gen).clamp_min(0) * (torch.rand(M, C, generator=gen) < 0.5)).to(dev)
    st = _lib.stream_ptr(dev)
    flags = _lib.EPI_RELU | (_lib.EPI_DROPOUT if p > 0 else 0)
    dZ_ref = torch.empty_like(dY)
    ops.epilogue_bwd_(dY, Y, dZ_ref, True, p)
    rows_p = (M + 31) // 32 * 64
    hi_ref = torch.empty(M, C, device=dev)
    pk_ref = torch.empty(rows_p, C, dtype=torch.bfloat16, device=dev)
    _lib.check(lib.grl_split_act_mn(dZ_ref.data_ptr(), C, hi_ref.data_ptr(), C, pk_ref.data_ptr(), C, M, C, st), "split")
    dZ = torch.full_like(dY, float("nan"))
    hi = torch.full((M, C), float("nan"), device=dev)
    pk = torch.full((rows_p, C), float("nan"), dtype=torch.bfloat16, device=dev)
    db = torch.full((C,), float("nan"), device=dev)
    ws = torch.empty(lib.grl_dz_split_workspace_bytes(M, C), dtype=torch.uint8, device=dev)
    _lib.check(lib.grl_dz_split(dY.data_ptr(), C, Y.data_ptr(), C, dZ.data_ptr(), C, hi.data_ptr(), pk.data_ptr(), C, M, C,
                                flags, p, db.data_ptr(), ws.data_ptr(), ws.numel(), st), "grl_dz_split")
    assert torch.equal(dZ, dZ_ref) and torch.equal(hi, hi_ref)
    assert torch.equal(pk.view(torch.int16), pk_ref.view(torch.int16))
    ref = dZ_ref.double().sum(0)
    assert float((db.double() - ref).norm() / ref.norm().clamp_min(1e-30)) < 1e-6
    db2 = torch.empty_like(db)
    _lib.check(lib.grl_dz_split(dY.data_ptr(), C, Y.data_ptr(), C, dZ.data_ptr(), C, 0, 0, C, M, C, flags, p,
                                db2.data_ptr(), ws.data_ptr(), ws.numel(), st), "grl_dz_split")
    assert torch.equal(db, db2)                      # fixed summation order
    assert lib.grl_dz_split(dY.data_ptr(), 53, Y.data_ptr(), 53, dZ.data_ptr(), 53, 0, 0, 56, M, 53, flags, p, 0, 0, 0, st) != 0


@pytest.mark.parametrize("M,K,C,p,dtype", [(300, 40, 64, 0.0, torch.float32), (257, 33, 53, 0.0, torch.float32),
                                           (500, 64, 128, 0.5, torch.float32), (300, 64, 64, 0.5, torch.bfloat16)])
def test_leaky_relu_epilogue_matches_torch(dev, M, K, C, p, dtype):
    """GRL_EPI_LEAKY (F.leaky_relu's slope 0.01; rp_robust_gcn.py:100,109) fused in the GEMM epilogue, with and
    without an injected dropout mask, forward and backward (derivative read off the sign of the output) vs torch."""
    from graph_representation_learning_b200 import ops
    gen = torch.Generator().manual_seed(M + C)
    x = torch.randn(M, K, generator=gen).to(dev).to(dtype).requires_grad_(True)
    w = (torch.randn(C, K, generator=gen) / K ** 0.5).to(dev).requires_grad_(True)
    mask = (torch.rand(M, C, generator=gen) < 0.5).to(torch.uint8).to(dev) if p > 0 else None
    y = ops.linear_act([x], w, None, relu="leaky", p=p, mask=mask)
    dy = torch.randn(M, C, generator=gen).to(dev).to(dtype)
    y.backward(dy)
    xr = x.detach().double().requires_grad_(True)
    # the bf16 path multiplies by bf16(w): give the reference the same weights, or sign flips of near-zero
    # pre-activations (1 vs 0.01 derivative) dominate the comparison
    wr = w.detach().to(dtype).double().requires_grad_(True)
    yr = torch.nn.functional.leaky_relu(xr @ wr.t())
    if p > 0:
        yr = yr * mask.double() / (1.0 - p)
    yr.backward(dy.double())
    tol = 2e-6 if dtype == torch.float32 else 1e-2
    rel = lambda a, b: float((a.double() - b).norm() / b.norm())
    assert rel(y.detach(), yr.detach()) < tol
    assert rel(x.grad, xr.grad) < (5e-6 if dtype == torch.float32 else 2e-2)
    assert rel(w.grad, wr.grad) < (5e-6 if dtype == torch.float32 else 2e-2)
    assert bool((y.detach() < 0).any())                     # the negative branch is exercised


@pytest.mark.parametrize("M,C,training", [(30, 32, True), (1000, 53, True), (4097, 256, True), (300, 64, False)])
def test_batchnorm_leaky_matches_torch(dev, M, C, training):
    """grl_bn_act_fwd / grl_bn_act_bwd (csrc/batchnorm.cu) against nn.BatchNorm1d + F.leaky_relu(0.2) in fp64:
    output, dX, dgamma, dbeta, running statistics and num_batches_tracked; rows with a large common offset
    (variance taken around the mean, no cancellation)."""
    from graph_representation_learning_b200 import ops
    gen = torch.Generator().manual_seed(M + C)
    x0 = torch.randn(M, C, generator=gen) * (0.5 + torch.rand(C, generator=gen)) + 30.0 * torch.randn(C, generator=gen)
    bn = torch.nn.BatchNorm1d(C)
    with torch.no_grad():
        bn.weight.copy_(0.5 + torch.rand(C, generator=gen)); bn.bias.copy_(torch.randn(C, generator=gen))
        bn.running_mean.copy_(x0.mean(0) + 0.1 * torch.randn(C, generator=gen)); bn.running_var.copy_(0.5 + torch.rand(C, generator=gen))
    ref = copy.deepcopy(bn).double()
    bn = bn.to(dev)
    bn.train(training); ref.train(training)
    x = x0.to(dev).requires_grad_(True)
    y = ops.batchnorm_act(x, bn, 0.2, training)
    dy = torch.randn(M, C, generator=gen)
    y.backward(dy.to(dev))
    xr = x0.double().requires_grad_(True)
    yr = torch.nn.functional.leaky_relu(ref(xr), 0.2)
    yr.backward(dy.double())
    rel = lambda a, b: float((a.detach().cpu().double() - b.detach()).norm() / b.detach().norm())
    errs = {"y": rel(y, yr), "dx": rel(x.grad, xr.grad), "dgamma": rel(bn.weight.grad, ref.weight.grad),
            "dbeta": rel(bn.bias.grad, ref.bias.grad), "running_mean": rel(bn.running_mean, ref.running_mean),
            "running_var": rel(bn.running_var, ref.running_var)}
    bars = {"y": 5e-6, "dx": 2e-5, "dgamma": 2e-5, "dbeta": 2e-5, "running_mean": 1e-6, "running_var": 1e-5}
    assert all(errs[k] < bars[k] for k in bars), errs
    assert int(bn.num_batches_tracked) == int(ref.num_batches_tracked)
