"""Accuracy (vs fp64) and speed of the fp32 GEMM modes: tf32x3 (three tf32 products) vs tf32bc (tf32 + one bf16
correction product).  Run on the GPU box: python scripts/gemm_mix_check.py"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)


def rel(x, ref):
    return float((x.double() - ref).norm() / ref.norm())


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


def both(fn):
    out = {}
    for mode in ("tf32x3", "tf32bc"):
        ops.GemmConfig.fp32_mode = ops.GemmConfig.fp32_mode_bwd = mode
        out[mode] = fn()
    ops.GemmConfig.fp32_mode, ops.GemmConfig.fp32_mode_bwd = "tf32x3", "tf32bc"
    return out


print("== accuracy NT")
for (M, N, K, pos) in ((300, 53, 37, False), (512, 256, 32, False), (512, 256, 4369, False), (1000, 256, 4369, True),
                       (777, 1280, 128, False), (512, 256, 1792, True), (640, 128, 512, False)):
    a = torch.rand(M, K, device=dev) if pos else torch.randn(M, K, device=dev)
    b = torch.randn(N, K, device=dev) / K ** 0.5
    ref = a.double() @ b.double().t()
    r = both(lambda: rel(ops.gemm_nt([a], [b]), ref))
    print(json.dumps({"nt": [M, N, K], "positive_a": pos, **r, "torch_fp32": rel(a @ b.t(), ref)}))
# two K segments + bias + relu
a0, a1 = torch.randn(500, 256, device=dev), torch.randn(500, 100, device=dev)
b0, b1 = torch.randn(128, 256, device=dev) / 16, torch.randn(128, 100, device=dev) / 10
bias = torch.randn(128, device=dev)
ref = torch.relu(a0.double() @ b0.double().t() + a1.double() @ b1.double().t() + bias.double())
print(json.dumps({"nt_2seg_bias_relu": both(lambda: rel(ops.gemm_nt([a0, a1], [b0, b1], bias=bias, relu=True), ref))}))

print("== accuracy TN")
for (R, Ma, Nb, scale) in ((1000, 256, 53, 1.0), (4096, 1792, 256, 1.0), (128000, 256, 128, 1e-5), (5000, 100, 36, 1.0)):
    a = torch.randn(R, Ma, device=dev)
    b = torch.randn(R, Nb, device=dev) * scale
    ref = a.double().t() @ b.double()
    r = both(lambda: rel(ops.gemm_tn(a, b), ref))
    print(json.dumps({"tn": [R, Ma, Nb], **r, "torch_fp32": rel(a.t() @ b, ref)}))

print("== speed (cfg2 shapes, M = 128000)")
M = 128000
for (N, K) in ((256, 4369), (256, 1792), (256, 3584), (1792, 256), (3584, 256), (1280, 128), (128, 512)):
    a = torch.randn(M, K, device=dev)
    b = torch.randn(N, K, device=dev) / K ** 0.5
    fl = 2.0 * M * N * K
    r = both(lambda: timeit(lambda: ops.gemm_nt([a], [b])))
    print(json.dumps({"nt": [M, N, K], **{k: round(v, 4) for k, v in r.items()},
                      "useful_TFs": {k: round(fl / v / 1e9, 1) for k, v in r.items()}}))
    del a, b
for (Ma, Nb) in ((4369, 256), (1792, 256), (3584, 256), (256, 128), (1280, 53)):
    a = torch.randn(M, Ma, device=dev)
    b = torch.randn(M, Nb, device=dev)
    fl = 2.0 * M * Ma * Nb
    r = both(lambda: timeit(lambda: ops.gemm_tn(a, b)))
    print(json.dumps({"tn": [M, Ma, Nb], **{k: round(v, 4) for k, v in r.items()},
                      "useful_TFs": {k: round(fl / v / 1e9, 1) for k, v in r.items()}}))
    del a, b
