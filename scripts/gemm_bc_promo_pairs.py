import json, os, sys, torch
sys.path.insert(0, "/root/repo")
from graph_representation_learning_b200 import _lib, ops
dev = torch.device("cuda:0"); lib = _lib.load()
def timed(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
g = torch.Generator().manual_seed(1)
A = torch.randn(128000, 1792, generator=g).to(dev); W = (torch.randn(256, 1792, generator=g) / 42).to(dev)
dZ = torch.randn(128000, 256, generator=g).to(dev)
ref = (A.double() @ W.double().t()); refw = (A.double().t() @ dZ.double())
for promo in (2, 4, 8, 16, 56):
    lib.grl_gemm_bc_set_promo(promo)
    y = ops.gemm_nt([A], [W], fp32_mode="tf32bc"); dw = ops.gemm_tn(A, dZ, "tf32bc")
    t = timed(lambda: ops.gemm_nt([A], [W], fp32_mode="tf32bc")); t2 = timed(lambda: ops.gemm_tn(A, dZ, "tf32bc"))
    print(json.dumps({"promo": promo, "nt_ms": round(t, 4), "nt_err": float((y.double() - ref).norm() / ref.norm()),
                      "tn_ms": round(t2, 4), "tn_err": float((dw.double() - refw).norm() / refw.norm())}), flush=True)
lib.grl_gemm_bc_set_promo(4)
