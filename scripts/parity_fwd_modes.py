#!/usr/bin/env python
"""Worst ratio (our error) / (allowed = max(1e-5, 2 x reference fp32 noise)) over all tensors of the four production-shape
golden cases, for different forward GEMM arithmetic: three tf32 products (default) against the tf32 + bf16-correction form
at several TMEM promotion intervals.  GPU box only."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _golden import allowed, load_production_case, production_model, rel_err  # noqa: E402
from graph_representation_learning_b200 import PackedGraphBatch, _lib, ops  # noqa: E402


def run(name):
    dev = torch.device("cuda:0")
    c = load_production_case(name)
    m = c["meta"]
    B, N = m["B"], m["N"]
    model = production_model(c).to(dev)
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
    for k, p in model.named_parameters():
        if p.requires_grad:
            idx = c["sample_index"](k, p.numel())
            errs["grad/" + k] = rel_err(p.grad.reshape(-1).cpu()[idx], c["grad"][k])
    worst = max(errs, key=lambda k: errs[k] / allowed(c, k, 1e-5))
    return errs[worst] / allowed(c, worst, 1e-5), worst, errs["logits_train"], errs["grad/emb1.0.weight"]


if __name__ == "__main__":
    lib = _lib.load()
    for fwd, bc_promo in (("tf32x3", 4), ("tf32bc", 4), ("tf32bc", 2), ("tf32bc", 1)):
        ops.GemmConfig.fp32_mode = fwd
        lib.grl_gemm_bc_set_promo(bc_promo)
        row = {"fwd": fwd, "bc_promo": bc_promo}
        for name in ("prod_cfg1", "prod_cfg1_tame", "prod_s512", "prod_s512_tame"):
            r, k, lt, ge = run(name)
            row[name] = {"worst_ratio": round(r, 3), "tensor": k, "logits_train": float(f"{lt:.2e}"), "grad_emb1": float(f"{ge:.2e}")}
        print(json.dumps(row), flush=True)
    lib.grl_gemm_bc_set_promo(4)
