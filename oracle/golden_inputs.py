"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Seeded inputs of the production-shape golden cases.

The production-shape cases (``tests/golden/prod_*.npz``, written by ``oracle/make_golden.py production``) do not store
their inputs: features, adjacency, labels, the injected DropEdge / dropout masks and the sample positions of the large
gradient tensors are regenerated here from the stored seed, by the SAME function on both sides (the generating script
next to the unmodified reference, and the tests next to the CUDA path).  CPU ``torch.Generator`` streams are
platform-independent, and the GPU box runs the same image.
"""
from __future__ import annotations

from typing import Any, Dict

import torch

SAMPLE = 32768          # entries kept of a gradient tensor with more than 2 * SAMPLE elements


def production_inputs(meta: Dict[str, Any]) -> Dict[str, Any]:
    """meta: {B, N, in_dim, S, out_dim, L, seed, deg}.  Returns V [B,N,in_dim] f32, A [B,N,L,N] f16 (j != i, ~deg
    neighbours per (node, type): SURVEY.md 8d), y [B,N] i64, edge_masks 3 x uint8 [B,(L+1)N,N] (keep = 0.7),
    feat_masks 5 x uint8 [B,N,w] (keep = 0.5; w = S, S, S, S, 10 * S/2) in the reference's dropout call order
    emb(0) g1(1) g2(2) g3(3) rp(4) (drop_robust_gcn.py:64,77,81,86,100)."""
    B, N, L, S = meta["B"], meta["N"], meta["L"], meta["S"]
    gen = torch.Generator().manual_seed(int(meta["seed"]) + 7919)
    V = torch.rand(B, N, meta["in_dim"], generator=gen)
    A = torch.rand(B, N, L, N, generator=gen) < (meta["deg"] / max(N - 1, 1))
    A &= ~torch.eye(N, dtype=torch.bool).view(1, N, 1, N)
    y = torch.randint(0, meta["out_dim"], (B, N), generator=gen)
    edge_masks = [(torch.rand(B, (L + 1) * N, N, generator=gen) < 0.7).to(torch.uint8) for _ in range(3)]
    H = S // 2
    feat_masks = [(torch.rand(B, N, w, generator=gen) < 0.5).to(torch.uint8) for w in (S, S, S, S, 10 * H)]
    return {"V": V, "A": A.half(), "y": y, "edge_masks": edge_masks, "feat_masks": feat_masks}


def sample_index(name: str, numel: int, seed: int) -> torch.Tensor:
    """positions of tensor ``name`` that the golden file keeps (all of them for small tensors), sorted"""
    if numel <= 2 * SAMPLE:
        return torch.arange(numel)
    h = sum((i + 1) * ord(c) for i, c in enumerate(name)) % 100003
    gen = torch.Generator().manual_seed(int(seed) * 131 + h)
    return torch.randperm(numel, generator=gen)[:SAMPLE].sort().values
