"""Diagnostic: norm-wise error of the tcgen05 GEMM modes vs fp64 as a function of K, and TF32X3 speed vs the
promotion interval (run on the GPU box:  python scripts/gemm_accuracy.py)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import _lib, ops  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
lib = _lib.load()


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


promos = (1, 2, 4, 1 << 20)
print(f"{'K':>6} " + " ".join(f"{'x3 promo=' + (str(p) if p < 1000 else 'inf'):>13}" for p in promos) +
      f" {'tf32':>10} {'bf16 exact':>11} {'cublas fp32':>12} {'x3 p=2 pos':>11}")
for K in (32, 128, 512, 1024, 2048, 4368, 8192):
    M, N = 512, 256
    a = torch.randn(M, K, device=dev)
    b = torch.randn(N, K, device=dev) / K ** 0.5
    ref = a.double() @ b.double().t()
    row = []
    for promo in promos:
        lib.grl_gemm_x3_set_promo(promo)
        row.append(rel(ops.gemm_nt([a], [b]), ref))
    lib.grl_gemm_x3_set_promo(2)
    ops.GemmConfig.fp32_mode = "tf32"
    row.append(rel(ops.gemm_nt([a], [b]), ref))
    ops.GemmConfig.fp32_mode = "tf32x3"
    ab, bb = a.bfloat16(), b.bfloat16()
    row.append(rel(ops.gemm_nt([ab], [bb], out_dtype=torch.float32), ab.double() @ bb.double().t()))
    row.append(rel(a @ b.t(), ref))
    ap = torch.rand(M, K, device=dev)           # post-ReLU-like non-negative activations
    row.append(rel(ops.gemm_nt([ap], [b]), ap.double() @ b.double().t()))
    print(f"{K:6d} " + " ".join(f"{v:13.2e}" for v in row[:len(promos)]) + " " + " ".join(f"{v:11.2e}" for v in row[len(promos):]))

print("\nspeed, M=128000 (cfg2 node count): ms and TFLOP/s (useful 2MNK flops)")
for (N, K) in ((256, 4368), (256, 1792), (256, 3584), (128, 512), (1280, 128)):
    M = 128000
    a = torch.randn(M, K, device=dev)
    b = torch.randn(N, K, device=dev) / K ** 0.5
    fl = 2.0 * M * N * K
    cols = []
    for promo in (1, 2, 4, 1 << 20):
        lib.grl_gemm_x3_set_promo(promo)
        ms = timeit(lambda: ops.gemm_nt([a], [b]))
        cols.append(f"x3 p={promo if promo < 1000 else 'inf'}: {ms:7.3f} ms {fl / ms / 1e9:6.1f} TF")
    lib.grl_gemm_x3_set_promo(2)
    ops.GemmConfig.fp32_mode = "tf32"
    ms = timeit(lambda: ops.gemm_nt([a], [b]))
    cols.append(f"tf32: {ms:7.3f} ms {fl / ms / 1e9:6.1f} TF")
    ops.GemmConfig.fp32_mode = "tf32x3"
    ab, bb = a.bfloat16(), b.bfloat16()
    ms = timeit(lambda: ops.gemm_nt([ab], [bb]))
    cols.append(f"bf16: {ms:7.3f} ms {fl / ms / 1e9:6.1f} TF")
    ms = timeit(lambda: a @ b.t())
    cols.append(f"cublas fp32: {ms:7.3f} ms {fl / ms / 1e9:6.1f} TF")
    ms = timeit(lambda: ab @ bb.t())
    cols.append(f"cublas bf16: {ms:7.3f} ms {fl / ms / 1e9:6.1f} TF")
    print(f"N={N} K={K}: " + " | ".join(cols))
