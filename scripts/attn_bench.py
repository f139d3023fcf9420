#!/usr/bin/env python
"""NodeSelfAtten core (ops.self_attention) forward and forward + backward time at the cfg2 / cfg3 shapes."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import ops  # noqa: E402


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


dev = torch.device("cuda:0")
for B, N, H, D, dt in ((256, 500, 128, 16, torch.float32), (128, 1000, 256, 32, torch.bfloat16), (64, 100, 128, 16, torch.float32)):
    g = torch.Generator().manual_seed(1)
    f = torch.rand(B, N, D, generator=g).to(dev, dt).requires_grad_(True)
    gg = torch.rand(B, N, D, generator=g).to(dev, dt).requires_grad_(True)
    h = torch.randn(B, N, H, generator=g).to(dev, dt).requires_grad_(True)
    V = torch.randn(B, N, H, generator=g).to(dev, dt).requires_grad_(True)
    gamma = torch.full((H,), 0.5, device=dev, requires_grad=True)
    dO = torch.randn(B, N, H, generator=g).to(dev, dt)

    def fwd():
        with torch.no_grad():
            return ops.self_attention(f, gg, h, gamma, V)

    def fwd_bwd():
        out = ops.self_attention(f, gg, h, gamma, V)
        out.backward(dO)

    t_f, t_fb = timed(fwd), timed(fwd_bwd)
    print(json.dumps({"B": B, "N": N, "H": H, "D": D, "dtype": str(dt), "fwd_ms": round(t_f, 4), "fwd_bwd_ms": round(t_fb, 4)}), flush=True)
