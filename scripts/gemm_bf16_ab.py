"""bf16 NT GEMM: persistent kernel (double-buffered TMEM accumulators) vs the one-tile-per-CTA kernel on the cfg3
per-GPU shapes (128 graphs x 1000 nodes, S = 512): time, useful TF/s, and bit-equality of the outputs.
Run on the GPU box: python scripts/gemm_bf16_ab.py"""
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


M = 128000
for (N, K, relu, p) in ((512, 4376, True, 0.5), (512, 3584, True, 0.5), (512, 7168, True, 0.5), (7168, 512, False, 0.0),
                        (3584, 512, False, 0.0), (2560, 256, True, 0.5), (2560, 64, False, 0.0), (256, 2560, False, 0.0),
                        (300, 100, True, 0.0)):
    a = torch.randn(M, K, device=dev).bfloat16()
    b = torch.randn(N, K, device=dev) / K ** 0.5
    bias = torch.randn(N, device=dev) if relu else None
    out, ms = {}, {}
    for on in (0, 1):
        lib.grl_gemm_nt_bf16_set_persistent(on)
        f = lambda: ops.gemm_nt([a], [b], bias=bias, relu=relu, p=p, seed=7, stream_id=3)
        out[on] = f()
        ms[on] = timeit(f)
    lib.grl_gemm_nt_bf16_set_persistent(1)
    ref = a[:2048].double() @ b.bfloat16().double().t()
    if bias is not None:
        ref = torch.relu(ref + bias.double())
    chk = None
    if p == 0.0:
        chk = float((out[1][:2048].double() - ref).norm() / ref.norm())
    fl = 2.0 * M * N * K
    print(json.dumps({"nt_bf16": [M, N, K], "epilogue": {"bias_relu": relu, "p": p}, "one_tile_ms": round(ms[0], 4),
                      "persistent_ms": round(ms[1], 4), "useful_TFs": [round(fl / ms[0] / 1e9, 1), round(fl / ms[1] / 1e9, 1)],
                      "bit_identical": bool(torch.equal(out[0], out[1])), "rel_err_vs_fp64": chk}), flush=True)
    del a, b, out
