#!/usr/bin/env python
"""A/B of the fp32-grade NT GEMMs: single-CTA kernel (gemm_bc_kernel) against the CTA-pair kernel (gemm_bc_2sm_kernel,
tcgen05 cta_group::2; grl_gemm_bc_set_pairs).  Checks that both give the same bits and prints ms / useful TF/s."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import _lib, ops  # noqa: E402


def timed(fn, iters=10):
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


def main():
    dev = torch.device("cuda:0")
    lib = _lib.load()
    shapes = [(128000, 256, 1792, "tf32x3"), (128000, 256, 3584, "tf32x3"), (128000, 256, 4369, "tf32x3"), (128000, 1280, 128, "tf32x3"),
              (128000, 256, 1792, "tf32bc"), (128000, 512, 1792, "tf32bc"), (128000, 1792, 256, "tf32bc"), (19000, 256, 1792, "tf32x3")]
    if len(sys.argv) > 1 and sys.argv[1] == "small":
        shapes = [(9500 + 77, 256, 160, "tf32x3"), (9500 + 77, 200, 96, "tf32bc")]
    for M, N, K, mode in shapes:
        g = torch.Generator().manual_seed(M + N + K)
        A = torch.randn(M, K, generator=g).to(dev)
        W = (torch.randn(N, K, generator=g) / K ** 0.5).to(dev)
        bias = torch.randn(N, generator=g).to(dev)
        res = {"M": M, "N": N, "K": K, "mode": mode}
        outs = {}
        for pairs in (0, 1):
            _lib.check(lib.grl_gemm_bc_set_pairs(pairs), "grl_gemm_bc_set_pairs")
            outs[pairs] = ops.gemm_nt([A], [W], bias=bias, relu=True, fp32_mode=mode)
            torch.cuda.synchronize()
            ms = timed(lambda: ops.gemm_nt([A], [W], bias=bias, relu=True, fp32_mode=mode))
            res["pairs_ms" if pairs else "single_ms"] = round(ms, 4)
            res["pairs_tflops" if pairs else "single_tflops"] = round(2.0 * M * N * K / ms / 1e9, 1)
        lib.grl_gemm_bc_set_pairs(0)
        ref = torch.relu(A.double() @ W.double().t() + bias.double())
        res["identical"] = bool(torch.equal(outs[0], outs[1]))
        res["err_pairs_vs_fp64"] = float((outs[1].double() - ref).norm() / ref.norm())
        res["err_single_vs_fp64"] = float((outs[0].double() - ref).norm() / ref.norm())
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
