"""Speed of both fp32 GEMM modes vs the promotion interval (diagnostic)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from graph_representation_learning_b200 import _lib, ops
dev = torch.device("cuda:0")
lib = _lib.load()
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
M = 128000
for (N, K) in ((256, 1792), (1792, 256)):
    a = torch.randn(M, K, device=dev); b = torch.randn(N, K, device=dev) / K ** 0.5
    x = torch.randn(M, K, device=dev); dz = torch.randn(M, N, device=dev)
    ref = (a[:2048].double() @ b.double().t())
    for mode in ("tf32x3", "tf32bc"):
        ops.GemmConfig.fp32_mode = ops.GemmConfig.fp32_mode_bwd = mode
        for promo in (1, 2, 4, 8, 1 << 20):
            lib.grl_gemm_x3_set_promo(promo)
            t_nt = timeit(lambda: ops.gemm_nt([a], [b]))
            t_tn = timeit(lambda: ops.gemm_tn(x, dz))
            err = float((ops.gemm_nt([a[:2048]], [b]).double() - ref).norm() / ref.norm())
            print(f"N={N} K={K} {mode} promo={promo if promo < 1000 else 'inf'}: nt {t_nt:.4f} ms  tn {t_tn:.4f} ms  nt err {err:.2e}")
    lib.grl_gemm_x3_set_promo(2)
