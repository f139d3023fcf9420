#!/usr/bin/env python
"""bf16 NT GEMM (N > 128): CTA-pair kernel (gemm_nt_bf16_2sm_kernel, tcgen05 cta_group::2) against the single-CTA
persistent kernel on the cfg3 per-GPU shapes: time, useful TF/s, bit-equality.  python scripts/gemm_bf16_2sm_ab.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
lib = _lib.load()


def timeit(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for (M, N, K, relu, p) in ((128000, 512, 4376, True, 0.5), (128000, 512, 3584, True, 0.5), (128000, 512, 7168, True, 0.5),
                           (128000, 512, 3584, False, 0.0), (128000, 2560, 256, True, 0.5), (128000, 256, 1024, True, 0.0),
                           (19001, 300, 100, True, 0.0), (19001, 515, 1000, False, 0.0)):
    a = torch.randn(M, K, device=dev).bfloat16()
    b = torch.randn(N, K, device=dev) / K ** 0.5
    bias = torch.randn(N, device=dev) if relu else None
    out, ms = {}, {}
    for on in (0, 1):
        lib.grl_gemm_bc_set_pairs(on)
        f = lambda: ops.gemm_nt([a], [b], bias=bias, relu=relu, p=p, seed=7, stream_id=3)
        out[on] = f()
        ms[on] = timeit(f)
    lib.grl_gemm_bc_set_pairs(1)
    print(json.dumps({"M": M, "N": N, "K": K, "single_ms": round(ms[0], 4), "pairs_ms": round(ms[1], 4),
                      "single_tflops": round(2.0 * M * N * K / ms[0] / 1e9, 1), "pairs_tflops": round(2.0 * M * N * K / ms[1] / 1e9, 1),
                      "identical": bool(torch.equal(out[0], out[1]))}), flush=True)
