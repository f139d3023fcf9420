#!/usr/bin/env python
"""Per-tensor gradient error of a production-shape golden case under different arithmetic choices (forward / backward GEMM
mode, tensor-core vs FFMA attention): which part of the error is ReLU-flip noise and which is ours.  GPU box only."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _golden import load_production_case, production_model, rel_err  # noqa: E402
from graph_representation_learning_b200 import PackedGraphBatch, ops  # noqa: E402


def run(name, fwd, bwd, tc, promo=2):
    from graph_representation_learning_b200 import _lib
    _lib.load().grl_gemm_x3_set_promo(promo)
    ops.GemmConfig.fp32_mode, ops.GemmConfig.fp32_mode_bwd, ops.AttentionConfig.tensor_cores = fwd, bwd, tc
    dev = torch.device("cuda:0")
    c = load_production_case(name)
    m = c["meta"]
    B, N = m["B"], m["N"]
    model = production_model(c).to(dev).train()
    g = PackedGraphBatch.from_dense(c["A"].to(dev))
    masks = {"edge": [g.gather_dense_mask(x.to(dev)) for x in c["edge_masks"]],
             "feat": [x.to(dev).reshape(B * N, -1).contiguous() for x in c["feat_masks"]]}
    logits = model([c["V"].to(dev), g], masks=masks)
    torch.nn.functional.cross_entropy(logits.transpose(1, 2), c["y"].to(dev)).backward()
    out = {"logits_train": rel_err(logits.detach().cpu(), c["logits_train"])}
    for k, p in model.named_parameters():
        if p.requires_grad:
            idx = c["sample_index"](k, p.numel())
            out[k] = rel_err(p.grad.reshape(-1).cpu()[idx], c["grad"][k])
    return out, c["err_ref32"]


if __name__ == "__main__":
    names = sys.argv[1:] or ["prod_s512_tame", "prod_cfg1_tame", "prod_s512", "prod_cfg1"]
    for name in names:
        rows = {}
        for fwd, bwd, tc, promo in (("tf32x3", "tf32bc", True, 2), ("tf32x3", "tf32bc", True, 1), ("tf32x3", "tf32bc", True, 4),
                                    ("tf32x3", "tf32bc", False, 1)):
            rows[f"{fwd}/{bwd}/{'tc' if tc else 'ffma'}/p{promo}"], ref = run(name, fwd, bwd, tc, promo)
        print(name)
        keys = list(next(iter(rows.values())))
        print(f"{'tensor':28s} {'ref32':>9s} " + " ".join(f"{k:>22s}" for k in rows))
        for k in keys:
            r = ref.get("grad/" + k, ref.get(k, float("nan")))
            print(f"{k:28s} {r:9.2e} " + " ".join(f"{rows[c][k]:22.2e}" for c in rows))
