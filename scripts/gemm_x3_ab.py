"""Three-product fp32 GEMM (NT): activation operand through shared memory (SS) vs through TMEM (TS), promo 2 / 4."""
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
ops.GemmConfig.fp32_mode = "tf32x3"
for (N, K) in ((256, 4368), (256, 1792), (256, 3584), (1280, 128), (128, 512), (53, 1280)):
    a = torch.rand(M, K, device=dev); b = torch.randn(N, K, device=dev) / K ** 0.5
    ref = (a[:2048].double() @ b.double().t())
    for sw, name in ((-4, "SS"), (-3, "TS")):
        lib.grl_gemm_x3_set_promo(sw)
        for promo in (2, 4):
            lib.grl_gemm_x3_set_promo(promo)
            t = timeit(lambda: ops.gemm_nt([a], [b]))
            err = float((ops.gemm_nt([a[:2048]], [b]).double() - ref).norm() / ref.norm())
            print(f"N={N} K={K} x3-{name} promo={promo}: {t:.4f} ms {2.0*M*N*K/t/1e9:.0f} TF/s useful  err {err:.2e}")
    lib.grl_gemm_x3_set_promo(2); lib.grl_gemm_x3_set_promo(-3)
    del a, b
