"""Seeded synthetic document-graph batches of the shapes BASELINE.json names.

Graph: for every graph, node i and edge type l, each j != i is a neighbour with probability d/(N-1)
(so the out-degree is Binomial(N-1, d/(N-1)) with distinct uniform targets) -> ``A[b, i, l, j] = 1`` in fp16,
the dtype the reference's graph builder emits (gnn/data_generator/data_process/utils/graph_utils.py:782-784).
Features: ``dense`` = U[0,1) on all columns (headline variant); ``bow`` = 7 ones in the first input_dim-4
binary columns + 4 spatial floats in U[0.09,1) (what TextlineEncoding produces,
gnn/data_generator/data_process/textline_encoding.py:33-55).  Labels uniform over the classes.
Everything is generated with a torch.Generator on the requested device, so CPU (oracle) and GPU runs that
need identical data generate on CPU and copy.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch


@dataclass
class Workload:
    name: str
    B: int
    N: int
    input_dim: int = 4369
    output_dim: int = 53
    num_edges: int = 6
    net_size: int = 256
    degree: float = 8.0
    dtype: torch.dtype = torch.float32
    use_attention: bool = True


CONFIGS = {
    "cfg1": Workload("cfg1: 8 graphs x 200 nodes, S=256, fp32", 8, 200),
    "cfg2": Workload("cfg2: 256 graphs x 500 nodes, ~8 nbrs/type, S=256, fp32", 256, 500),
    "cfg3": Workload("cfg3: 128 graphs/GPU x 1000 nodes, S=512, bf16", 128, 1000, net_size=512, dtype=torch.bfloat16),
}


def make_batch(wl: Workload, seed: int, device, features: str = "dense", B: Optional[int] = None,
               feature_dtype: torch.dtype = torch.float32):
    """returns (V [B,N,input_dim], A [B,N,L,N] fp16, y [B,N] int64) on ``device``"""
    B = wl.B if B is None else B
    N, L = wl.N, wl.num_edges
    dev = torch.device(device)
    gen = torch.Generator(device=dev).manual_seed(seed)
    A = torch.empty(B, N, L, N, dtype=torch.float16, device=dev)
    prob = wl.degree / max(N - 1, 1)
    eye = torch.eye(N, dtype=torch.bool, device=dev).view(N, 1, N)
    for b in range(B):  # per graph to bound the fp32 temporary
        hit = torch.rand(N, L, N, generator=gen, device=dev) < prob
        hit &= ~eye
        A[b] = hit.to(torch.float16)
    # The collate emits V with a 16-byte-aligned row pitch (leading dimension rounded up to 8 elements) so that
    # TMA can read it in place; callers see the logical [B, N, input_dim] view.
    ld = (wl.input_dim + 7) // 8 * 8
    Vp = torch.zeros(B, N, ld, device=dev, dtype=feature_dtype)
    V = Vp[:, :, :wl.input_dim]
    if features == "dense":
        for b in range(B):
            V[b] = torch.rand(N, wl.input_dim, generator=gen, device=dev, dtype=torch.float32).to(feature_dtype)
    elif features == "bow":
        idx = torch.randint(0, wl.input_dim - 4, (B, N, 7), generator=gen, device=dev)
        V.scatter_(2, idx, 1.0)
        V[:, :, -4:] = (0.09 + 0.91 * torch.rand(B, N, 4, generator=gen, device=dev)).to(feature_dtype)
    else:
        raise ValueError(features)
    y = torch.randint(0, wl.output_dim, (B, N), generator=gen, device=dev)
    return V, A, y


CONFIGS["cfg4"] = Workload("cfg4: 1 graph x 1M nodes, degree 16/type, S=256, fp32, no attention", 1, 1_000_000,
                           degree=16.0, use_attention=False)


def make_edge_batch(wl: Workload, seed: int, device, B: Optional[int] = None, N: Optional[int] = None,
                    feature_dtype: torch.dtype = torch.float32, features: bool = True):
    """Same graph distribution as ``make_batch`` without the O(N^2) dense adjacency (cfg4: one 1M-node graph, cfg5:
    10 000-node documents): for every (graph, node, type) a Poisson(degree) number of neighbours j != i, uniform.
    Returns (V [B,N,input_dim] or None, edges int32 [E,4] = (b, i, l, j), y [B,N]) on ``device``; duplicate
    (b,i,l,j) are possible with probability ~degree^2 / 2N per row and stay separate entries, as the pack defines."""
    B = wl.B if B is None else B
    N = wl.N if N is None else N
    L = wl.num_edges
    dev = torch.device(device)
    gen = torch.Generator(device=dev).manual_seed(seed)
    parts = []
    for b in range(B):
        k = torch.poisson(torch.full((N * L,), float(wl.degree), device=dev), generator=gen).to(torch.int64)
        k.clamp_(max=N - 1)
        src = torch.repeat_interleave(torch.arange(N * L, device=dev), k)        # (i, l) flattened
        i, l = src // L, src % L
        j = (i + 1 + torch.randint(0, max(N - 1, 1), (src.numel(),), generator=gen, device=dev)) % N
        parts.append(torch.stack([torch.full_like(i, b), i, l, j], dim=1).to(torch.int32))
    edges = torch.cat(parts, dim=0)
    V = None
    if features:
        ld = (wl.input_dim + 7) // 8 * 8
        Vp = torch.empty(B, N, ld, device=dev, dtype=feature_dtype)
        for b in range(B):
            for r0 in range(0, N, 65536):                                        # bounded fp32 temporary
                r1 = min(N, r0 + 65536)
                Vp[b, r0:r1] = torch.rand(r1 - r0, ld, generator=gen, device=dev, dtype=torch.float32).to(feature_dtype)
        V = Vp[:, :, :wl.input_dim]
    y = torch.randint(0, wl.output_dim, (B, N), generator=gen, device=dev)
    return V, edges, y
