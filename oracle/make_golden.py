"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Generates ``tests/golden/*`` from the UNMODIFIED reference.

Run in the authoring container only (``/root/reference`` does not exist on the GPU box):

    python oracle/make_golden.py

It imports the reference's own ``GraphCNNDropEdge`` / ``NumpyPadding`` (with two in-memory shims for the missing
``colorlog`` and ``decouple`` packages, nothing else is touched), replaces ``model.dropout`` / ``model.edge_dropout``
by mask-injecting modules, and stores inputs, masks, weights, logits, loss, gradients and the parameters after
one ``_run_train_step`` (gnn/trainer/training_procedures/kv_procedure.py:143-164: clip 5.0, Adam defaults).
"""
from __future__ import annotations

import json
import os
import sys
import types

import numpy as np
import torch

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def import_reference():
    if not os.path.isdir(REF):
        raise SystemExit(f"{REF} not present: golden vectors can only be regenerated in the authoring container")
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle import ref_loader
    ref_loader._shims()
    sys.path.insert(0, REF)
    from gnn.models import GraphCNNDropEdge  # noqa: E402
    from gnn.data_generator.data_collate.numpy_padding import NumpyPadding  # noqa: E402
    return GraphCNNDropEdge, NumpyPadding


class InjectedDropout(torch.nn.Module):
    """Stands in for nn.Dropout on the reference instance: pops pre-drawn keep masks in call order."""

    def __init__(self, p, masks):
        super().__init__()
        self.p, self.masks, self.i = p, masks, 0

    def forward(self, x):
        if not self.training:
            return x
        m = self.masks[self.i]
        self.i += 1
        assert m.shape == x.shape, (m.shape, x.shape)
        return x * m.to(x.dtype) / (1.0 - self.p)


def synth_graph(gen, B, N, L, deg, weighted=False, empty_types=()):
    """A[b, i, l, j] with ~deg neighbours per (node, type), j != i; fp16 like graph_utils.py:782-784."""
    A = torch.zeros(B, N, L, N)
    for b in range(B):
        for l in range(L):
            if l in empty_types:
                continue
            prob = torch.full((N, N), deg / max(N - 1, 1))
            prob.fill_diagonal_(0.0)
            hit = torch.bernoulli(prob, generator=gen)
            if weighted:
                hit = hit * (torch.randint(1, 9, (N, N), generator=gen).float() / 8.0)
            A[b, :, l, :] = hit
    return A.half()


def small_model_case(GraphCNNDropEdge, name, seed, B, N, in_dim, S, out_dim, L=6, deg=2.0, weighted=False,
                     empty_types=(), use_attention=True, pad=0):
    gen = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S, use_attention=use_attention)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V = torch.rand(B, N, in_dim, generator=gen)
    A = synth_graph(gen, B, N, L, deg, weighted, empty_types)
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    if pad:  # emulate collate padding: last graph has `pad` padded nodes on each side
        V[-1, :pad] = 0; V[-1, -pad:] = 0
        A[-1, :pad] = 0; A[-1, -pad:] = 0; A[-1, :, :, :pad] = 0; A[-1, :, :, -pad:] = 0
        y[-1, :pad] = -100; y[-1, -pad:] = -100
    H = S // 2
    edge_masks = [torch.bernoulli(torch.full((B, (L + 1) * N, N), 0.7), generator=gen).to(torch.uint8) for _ in range(3)]
    feat_masks = [torch.bernoulli(torch.full((B, N, S), 0.5), generator=gen).to(torch.uint8) for _ in range(4)]
    feat_masks.append(torch.bernoulli(torch.full((B, N, 10 * H), 0.5), generator=gen).to(torch.uint8))

    # eval forward
    model.eval()
    with torch.no_grad():
        logits_eval = model([V, A.float()])

    # train step with injected masks; dropout call order in drop_robust_gcn.py: emb(0) g1(1) g2(2) g3(3) rp(4)
    model.train()
    model.dropout = InjectedDropout(0.5, feat_masks)
    model.edge_dropout = InjectedDropout(0.3, edge_masks)
    opt = torch.optim.Adam([p for p in model.parameters() if p.requires_grad])
    opt.zero_grad()
    logits_train = model.forward([V, A.float()])
    loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), y)
    loss.backward()
    grads = {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}
    total_norm = torch.nn.utils.clip_grad_norm_(model.parameters(), 5.0)
    opt.step()
    sd1 = {k: v.detach().clone() for k, v in model.state_dict().items()}

    out = {"V": V.numpy(), "A": A.numpy(), "y": y.numpy(), "logits_eval": logits_eval.numpy(),
           "logits_train": logits_train.detach().numpy(), "loss": loss.detach().numpy(),
           "grad_norm": total_norm.detach().numpy(),
           "meta": np.array(json.dumps(dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed,
                                            use_attention=use_attention)))}
    for i, m in enumerate(edge_masks):
        out[f"edge_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for i, m in enumerate(feat_masks):
        out[f"feat_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for k, v in sd0.items():
        out["sd0/" + k] = v.numpy()
    for k, v in grads.items():
        out["grad/" + k] = v.numpy()
    for k, v in sd1.items():
        out["sd1/" + k] = v.numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(loss):.6f} grad-norm {float(total_norm):.4f} nnz(A) {int((A != 0).sum())}")


def preprocess_case(GraphCNNDropEdge):
    """Golden fact (i) of SURVEY 8c: the exact A' of preprocess_adj, stored as CSR."""
    gen = torch.Generator().manual_seed(7)
    B, N, L = 2, 9, 6
    A = synth_graph(gen, B, N, L, 2.0, weighted=True, empty_types=(4,))
    torch.manual_seed(0)
    model = GraphCNNDropEdge(8, 3, L, net_size=8)
    Ap = model.gcn1.preprocess_adj(A.float().permute(0, 1, 3, 2))     # B x 7N x N
    csr = Ap.reshape(B * N * (L + 1), N).to_sparse_csr()
    np.savez_compressed(os.path.join(OUT, "preprocess_adj.npz"), A=A.numpy(),
                        rowptr=csr.crow_indices().numpy().astype(np.int32),
                        col=csr.col_indices().numpy().astype(np.int32), val=csr.values().numpy())
    print("preprocess_adj: nnz", csr.values().numel())


def init_case(GraphCNNDropEdge):
    """Same seed -> same initial weights as the reference constructor (full-size model): per-tensor checksums."""
    torch.manual_seed(1234)
    model = GraphCNNDropEdge(4369, 53, 6, net_size=256)
    rec = {}
    for k, v in model.state_dict().items():
        v64 = v.double().reshape(-1)
        rec[k] = {"shape": list(v.shape), "sum": float(v64.sum()), "abs_sum": float(v64.abs().sum()),
                  "first": [float(x) for x in v64[:4]], "last": [float(x) for x in v64[-4:]]}
    with open(os.path.join(OUT, "init_seed1234.json"), "w") as f:
        json.dump(rec, f, indent=1)
    print("init: ", len(rec), "tensors")


def collate_case(NumpyPadding):
    """NumpyPadding symmetric padding (numpy_padding.py:63-103) on ragged samples."""
    rng = np.random.default_rng(5)
    sizes = [5, 9, 6, 9, 2]
    samples = []
    for n in sizes:
        samples.append({"textline_encoding": rng.random((n, 11), dtype=np.float32),
                        "adjacency_matrix": (rng.random((n, 6, n)) < 0.3).astype(np.float16),
                        "node_label": rng.integers(0, 7, size=(n,)).astype(np.int64),
                        "other": "dropped"})
    raw = {f"in{i}/{k}": v for i, s in enumerate(samples) for k, v in s.items() if k != "other"}
    pad = NumpyPadding({"textline_encoding": 0.0, "adjacency_matrix": 0.0, "node_label": -100.0}, True)
    out = pad([dict(s) for s in samples])
    from torch.utils.data._utils.collate import default_collate
    batch = default_collate(out)
    res = {f"out/{k}": v.numpy() for k, v in batch.items()}
    np.savez_compressed(os.path.join(OUT, "collate.npz"), sizes=np.array(sizes), **raw, **res)
    print("collate:", {k: tuple(v.shape) for k, v in batch.items()}, {k: v.dtype for k, v in batch.items()})


def robust_gcn_case(name, seed, B, N, in_dim, S, out_dim, L=6, deg=2.0):
    """Sibling model that reuses GraphConv / NodeSelfAtten: gnn/models/networks/robust_gcn.py:173-231.
    No DropEdge; six feature dropouts in call order emb(0) g1(1) g2(2) g3(3) attention-out(4) g4(5)."""
    from gnn.models.networks.robust_gcn import RobustGCN
    gen = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    model = RobustGCN(in_dim, out_dim, L, net_size=S)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V = torch.rand(B, N, in_dim, generator=gen)
    A = synth_graph(gen, B, N, L, deg)
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    H = S // 2
    feat_masks = [torch.bernoulli(torch.full((B, N, w), 0.5), generator=gen).to(torch.uint8) for w in (S, S, S, S, H, H)]
    model.eval()
    with torch.no_grad():
        logits_eval = model([V, A.float()])
    model.train()
    model.dropout = InjectedDropout(0.5, feat_masks)
    logits_train = model.forward([V, A.float()])
    loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), y)
    loss.backward()
    out = {"V": V.numpy(), "A": A.numpy(), "y": y.numpy(), "logits_eval": logits_eval.numpy(),
           "logits_train": logits_train.detach().numpy(), "loss": loss.detach().numpy(),
           "meta": np.array(json.dumps(dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed)))}
    for i, m in enumerate(feat_masks):
        out[f"feat_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for k, v in sd0.items():
        out["sd0/" + k] = v.numpy()
    for k, p in model.named_parameters():
        if p.grad is not None:
            out["grad/" + k] = p.grad.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(loss):.6f}")


def rp_dropedge_case(name, seed, B, N, in_dim, S, out_dim, rp_size, L=6, deg=2.0):
    """Sibling model: gnn/models/networks/rp_robust_gcn.py:29-115 -- random projection in front of the
    attention, leaky ReLU.  Dropout call order: emb(0) g1(1) g2(2) g3(3) rp_final(4); three DropEdge masks."""
    from gnn.models.networks.rp_robust_gcn import RPGraphCNNDropEdge
    gen = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    model = RPGraphCNNDropEdge(in_dim, out_dim, L, net_size=S, rp_size=rp_size)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V = torch.rand(B, N, in_dim, generator=gen)
    A = synth_graph(gen, B, N, L, deg)
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    edge_masks = [torch.bernoulli(torch.full((B, (L + 1) * N, N), 0.7), generator=gen).to(torch.uint8) for _ in range(3)]
    feat_masks = [torch.bernoulli(torch.full((B, N, w), 0.5), generator=gen).to(torch.uint8) for w in (S, S, S, S, rp_size)]
    model.eval()
    with torch.no_grad():
        logits_eval = model([V, A.float()])
    model.train()
    model.dropout = InjectedDropout(0.5, feat_masks)
    model.edge_dropout = InjectedDropout(0.3, edge_masks)
    logits_train = model.forward([V, A.float()])
    loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), y)
    loss.backward()
    out = {"V": V.numpy(), "A": A.numpy(), "y": y.numpy(), "logits_eval": logits_eval.numpy(),
           "logits_train": logits_train.detach().numpy(), "loss": loss.detach().numpy(),
           "meta": np.array(json.dumps(dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed,
                                            rp_size=rp_size)))}
    for i, m in enumerate(edge_masks):
        out[f"edge_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for i, m in enumerate(feat_masks):
        out[f"feat_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for k, v in sd0.items():
        out["sd0/" + k] = v.numpy()
    for k, p in model.named_parameters():
        if p.grad is not None:
            out["grad/" + k] = p.grad.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(loss):.6f}")


def deep_rp_case(name, seed, B, N, in_dim, S, out_dim, L=6, deg=2.0):
    """Sibling model: gnn/models/networks/deep_rp_robust_gcn.py:65-168 -- nine GraphConv + BatchNorm1d +
    leaky-ReLU(0.2) blocks.  DropEdge(0.2) masks in call order: blocks 3, 6, 8, 9; dropout(0.3): behind blocks 3, 6, 9
    and the attention.  BatchNorm parameters / running statistics are randomised so they matter; the state after the
    train-mode forward (running statistics, num_batches_tracked) is stored too."""
    from gnn.models.networks.deep_rp_robust_gcn import DeepRPRobustGCN
    gen = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    model = DeepRPRobustGCN(in_dim, out_dim, L, net_size=S, lambda_value=0.05)
    with torch.no_grad():
        for k, v in model.state_dict().items():
            if ".bn." in k and k.endswith("weight"):
                v.copy_(0.5 + torch.rand(v.shape, generator=gen))
            elif ".bn." in k and k.endswith(("bias", "running_mean")):
                v.copy_(0.2 * torch.randn(v.shape, generator=gen))
            elif ".bn." in k and k.endswith("running_var"):
                v.copy_(0.5 + torch.rand(v.shape, generator=gen))
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V = torch.rand(B, N, in_dim, generator=gen)
    A = synth_graph(gen, B, N, L, deg)
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    edge_masks = [torch.bernoulli(torch.full((B, (L + 1) * N, N), 0.8), generator=gen).to(torch.uint8) for _ in range(4)]
    feat_masks = [torch.bernoulli(torch.full((B, N, S), 0.7), generator=gen).to(torch.uint8) for _ in range(4)]
    model.eval()
    with torch.no_grad():
        logits_eval = model([V, A.float()])
    model.train()
    model.dropout = InjectedDropout(0.3, feat_masks)
    model.edge_dropout = InjectedDropout(0.2, edge_masks)
    logits_train = model.forward([V, A.float()])
    loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), y)
    loss.backward()
    out = {"V": V.numpy(), "A": A.numpy(), "y": y.numpy(), "logits_eval": logits_eval.numpy(),
           "logits_train": logits_train.detach().numpy(), "loss": loss.detach().numpy(),
           "meta": np.array(json.dumps(dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed,
                                            lambda_value=0.05)))}
    for i, m in enumerate(edge_masks):
        out[f"edge_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for i, m in enumerate(feat_masks):
        out[f"feat_mask{i}"] = np.packbits(m.numpy().reshape(-1))
    for k, v in sd0.items():
        out["sd0/" + k] = v.numpy()
    for k, v in model.state_dict().items():
        if ".bn.running" in k or k.endswith("num_batches_tracked"):
            out["sd1/" + k] = v.detach().numpy()
    for k, p in model.named_parameters():
        if p.grad is not None:
            out["grad/" + k] = p.grad.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(loss):.6f}")


def deep_rp_gcn_case(name, seed, B, N, in_dim, S, out_dim, L=6, deg=2.0):
    """Sibling model: gnn/models/networks/deep_rp_gcn.py:106-150 -- 29 GraphConv + BatchNorm1d +
    leaky(0.2) blocks, no DropEdge, one dropout(0.3) in front of the classifier."""
    from gnn.models.networks.deep_rp_gcn import DeepRPGCN
    gen = torch.Generator().manual_seed(seed)
    torch.manual_seed(seed)
    model = DeepRPGCN(in_dim, out_dim, L, net_size=S)
    sd0 = {k: v.detach().clone() for k, v in model.state_dict().items()}
    V = torch.rand(B, N, in_dim, generator=gen)
    A = synth_graph(gen, B, N, L, deg)
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    feat_masks = [torch.bernoulli(torch.full((B, N, S), 0.7), generator=gen).to(torch.uint8)]
    model.eval()
    with torch.no_grad():
        logits_eval = model([V, A.float()])
    model.train()
    model.dropout = InjectedDropout(0.3, feat_masks)
    logits_train = model.forward([V, A.float()])
    loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), y)
    loss.backward()
    out = {"V": V.numpy(), "A": A.numpy(), "y": y.numpy(), "logits_eval": logits_eval.numpy(),
           "logits_train": logits_train.detach().numpy(), "loss": loss.detach().numpy(),
           "feat_mask0": np.packbits(feat_masks[0].numpy().reshape(-1)),
           "meta": np.array(json.dumps(dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed)))}
    for k, v in sd0.items():
        out["sd0/" + k] = v.numpy()
    for k, p in model.named_parameters():
        if p.grad is not None:
            out["grad/" + k] = p.grad.detach().numpy()
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(loss):.6f}")


REF32_THREADS = (16, 4, 1)      # OpenMP thread counts of the reference fp32 runs (summation orders)


def production_case(GraphCNNDropEdge, name, seed, B, N, in_dim, S, out_dim, L=6, deg=8.0, w_rand_scale=None):
    """Production-shape vectors: the UNMODIFIED reference run in fp64 (the truth every tensor is judged
    against) and in fp32 (its own rounding noise, stored as one norm-wise number per tensor: SURVEY.md 8c's
    ``max(1e-5, 2 * err_ref32)`` rule).  Inputs / masks are NOT stored: ``oracle/golden_inputs.production_inputs``
    regenerates them from the seed.  Gradient tensors above 65 536 elements keep 32 768 seeded sample positions.

    ``w_rand_scale``: the frozen N(0,1) projection (drop_robust_gcn.py:20-25) is multiplied by this factor in the state
    dict of BOTH sides.  At its initial scale the logits are O(300) and the cross entropy is saturated (loss ~ 4e2): the
    gradient then hinges on which class wins at each node and is chaotic in the last bits -- the reference's own fp32 run
    is 3-6e-3 away from its fp64 run at cfg1.  The "tame" cases (scale 0.005, logits O(1)) are the well-conditioned
    check of the same kernels at the same shapes (fp32 at ~1e-5, bf16 at 1e-2)."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    from oracle.golden_inputs import production_inputs, sample_index
    meta = dict(B=B, N=N, in_dim=in_dim, S=S, out_dim=out_dim, L=L, seed=seed, deg=deg, use_attention=True,
                w_rand_scale=w_rand_scale)
    inp = production_inputs(meta)
    torch.manual_seed(seed)
    model32 = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S)
    sd0 = {k: v.detach().clone() for k, v in model32.state_dict().items()}
    if w_rand_scale is not None:
        sd0["w_rand.projection.weight"] = sd0["w_rand.projection.weight"] * w_rand_scale

    def run(dtype):
        torch.manual_seed(seed)
        model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S).to(dtype)
        model.load_state_dict({k: v.to(dtype) for k, v in sd0.items()})
        V, A = inp["V"].to(dtype), inp["A"].to(dtype)
        model.eval()
        with torch.no_grad():
            logits_eval = model([V, A])
        model.train()
        model.dropout = InjectedDropout(0.5, inp["feat_masks"])
        model.edge_dropout = InjectedDropout(0.3, inp["edge_masks"])
        logits_train = model.forward([V, A])
        loss = torch.nn.CrossEntropyLoss()(logits_train.transpose(1, 2), inp["y"])
        loss.backward()
        grads = {k: p.grad.detach().clone() for k, p in model.named_parameters() if p.grad is not None}
        return logits_eval, logits_train.detach(), loss.detach(), grads

    e64, t64, l64, g64 = run(torch.float64)

    def rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm())

    out = {"meta": np.array(json.dumps(meta)), "logits_eval": e64.numpy(), "logits_train": t64.numpy(),
           "loss": l64.numpy()}
    for k, g in g64.items():
        out["grad/" + k] = g.reshape(-1)[sample_index(k, g.numel(), seed)].numpy()
    # The reference's fp32 noise is ONE draw per summation order: the same unmodified fp32 run lands 1.1e-4 from fp64 with
    # 16 OpenMP threads and 2.2-2.7e-4 with 1 / 4 (prod_s512_tame, gradients upstream of the attention).  The yardstick
    # is therefore the per-tensor MAX over the thread counts below; every run is kept in err_ref32_runs.
    runs32 = {}
    threads0 = torch.get_num_threads()
    for nt in REF32_THREADS:
        torch.set_num_threads(nt)
        e32, t32, l32, g32 = run(torch.float32)
        r = {"logits_eval": rel(e32, e64), "logits_train": rel(t32, t64), "loss": abs(float(l32) - float(l64)) / float(l64)}
        for k, g in g64.items():
            idx = sample_index(k, g.numel(), seed)
            r["grad/" + k] = rel(g32[k].reshape(-1)[idx], g.reshape(-1)[idx])
        runs32[str(nt)] = r
    torch.set_num_threads(threads0)
    err32 = {k: max(r[k] for r in runs32.values()) for k in next(iter(runs32.values()))}
    out["err_ref32_runs"] = np.array(json.dumps(runs32))
    # the reference's OWN bf16 run (torch.autocast on the CPU: nn.Linear / matmul / bmm in bf16), same masks: the
    # yardstick for our bf16 mode where 1e-2 is out of reach for any bf16 implementation (gradients upstream of the
    # unscaled softmax attention: dS = P (dP - D) cancels, a 4e-3 perturbation of its inputs moves them by O(1))
    torch.manual_seed(seed)
    model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S)
    model.load_state_dict(sd0)
    model.train()
    model.dropout = InjectedDropout(0.5, inp["feat_masks"])
    model.edge_dropout = InjectedDropout(0.3, inp["edge_masks"])
    with torch.autocast("cpu", dtype=torch.bfloat16):
        lb = model.forward([inp["V"], inp["A"].float()])
    loss_b = torch.nn.CrossEntropyLoss()(lb.float().transpose(1, 2), inp["y"])
    loss_b.backward()
    errb = {"logits_train": rel(lb.detach().float(), t64), "loss": abs(float(loss_b) - float(l64)) / float(l64)}
    model.eval()
    with torch.no_grad(), torch.autocast("cpu", dtype=torch.bfloat16):
        errb["logits_eval"] = rel(model([inp["V"], inp["A"].float()]).float(), e64)
    for k, pp in model.named_parameters():
        if pp.grad is not None:
            idx = sample_index(k, pp.numel(), seed)
            errb["grad/" + k] = rel(pp.grad.reshape(-1)[idx], g64[k].reshape(-1)[idx])
    out["err_ref_bf16"] = np.array(json.dumps(errb))
    out["err_ref32"] = np.array(json.dumps(err32))
    out["weights_check"] = np.array(json.dumps({k: [float(v.double().sum()), float(v.double().abs().sum())]
                                                for k, v in sd0.items()}))
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **out)
    print(f"{name}: loss {float(l64):.6f}; reference fp32-vs-fp64 noise: " +
          ", ".join(f"{k.split('/')[-1]} {v:.1e}" for k, v in err32.items()))


def output_schema_case():
    """The reference's JSON schema of KVInference's output (tests/assets/schemas/output_schema.json, a data asset of its
    test-suite) as a fixture: tests/test_gpu_procedures.py validates what our KVInference writes back against it."""
    src = os.path.join(REF, "tests", "assets", "schemas", "output_schema.json")
    with open(src) as f:
        schema = json.load(f)
    with open(os.path.join(OUT, "output_schema.json"), "w") as f:
        json.dump(schema, f, indent=1, sort_keys=True)
    print("output_schema.json: required =", schema["items"]["required"])


def main():
    os.makedirs(OUT, exist_ok=True)
    GraphCNNDropEdge, NumpyPadding = import_reference()
    if len(sys.argv) > 1 and sys.argv[1] == "production":       # cfg1 exactly + one S=512 case (no stored inputs)
        production_case(GraphCNNDropEdge, "prod_cfg1", seed=41, B=8, N=200, in_dim=4369, S=256, out_dim=53)
        production_case(GraphCNNDropEdge, "prod_s512", seed=42, B=2, N=300, in_dim=600, S=512, out_dim=53)
        production_case(GraphCNNDropEdge, "prod_cfg1_tame", seed=43, B=8, N=200, in_dim=4369, S=256, out_dim=53,
                        w_rand_scale=0.005)
        production_case(GraphCNNDropEdge, "prod_s512_tame", seed=44, B=2, N=300, in_dim=600, S=512, out_dim=53,
                        w_rand_scale=0.005)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "output_schema":
        output_schema_case()
        return
    if len(sys.argv) > 1 and sys.argv[1] == "robust_gcn":       # add the sibling-model vectors without touching the rest
        robust_gcn_case("robustgcn_small", seed=21, B=3, N=10, in_dim=24, S=32, out_dim=7)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "deep_rp":
        deep_rp_case("deeprp_small", seed=23, B=3, N=10, in_dim=24, S=32, out_dim=7)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "deep_rp_gcn":
        deep_rp_gcn_case("deeprpgcn_small", seed=24, B=2, N=9, in_dim=16, S=16, out_dim=5)
        return
    if len(sys.argv) > 1 and sys.argv[1] == "rp_dropedge":
        rp_dropedge_case("rpdropedge_small", seed=22, B=3, N=10, in_dim=24, S=32, out_dim=7, rp_size=48)
        return
    small_model_case(GraphCNNDropEdge, "model_small", seed=11, B=3, N=10, in_dim=24, S=32, out_dim=7)
    small_model_case(GraphCNNDropEdge, "model_weighted_pad", seed=12, B=2, N=12, in_dim=20, S=32, out_dim=5,
                     weighted=True, empty_types=(4, 5), pad=2)
    small_model_case(GraphCNNDropEdge, "model_noatt", seed=13, B=2, N=8, in_dim=16, S=32, out_dim=4,
                     use_attention=False)
    preprocess_case(GraphCNNDropEdge)
    init_case(GraphCNNDropEdge)
    collate_case(NumpyPadding)
    robust_gcn_case("robustgcn_small", seed=21, B=3, N=10, in_dim=24, S=32, out_dim=7)
    rp_dropedge_case("rpdropedge_small", seed=22, B=3, N=10, in_dim=24, S=32, out_dim=7, rp_size=48)
    deep_rp_case("deeprp_small", seed=23, B=3, N=10, in_dim=24, S=32, out_dim=7)
    deep_rp_gcn_case("deeprpgcn_small", seed=24, B=2, N=9, in_dim=16, S=16, out_dim=5)
    output_schema_case()


if __name__ == "__main__":
    main()
