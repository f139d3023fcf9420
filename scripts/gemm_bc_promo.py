"""Second-generation fp32 GEMMs (tf32 + bf16 correction): accumulator promotion interval (K-blocks of 32 kept in TMEM
before the partial sum goes to fp32 registers) vs time and error, on the cfg2 backward shapes.  Run on the GPU box."""
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
for kind, (P, Q) in (("nt", (1792, 256)), ("nt", (3584, 256)), ("nt", (256, 1792)), ("tn", (1792, 256)), ("tn", (3584, 256)),
                     ("tn", (4369, 256))):
    if kind == "nt":
        a = torch.randn(M, Q, device=dev)
        b = torch.randn(P, Q, device=dev) / Q ** 0.5
        ref = a[:2048].double() @ b.double().t()
        run = lambda: ops.gemm_nt([a], [b], fp32_mode="tf32bc")
        err = lambda: float((ops.gemm_nt([a[:2048]], [b], fp32_mode="tf32bc").double() - ref).norm() / ref.norm())
    else:
        a = torch.randn(M, P, device=dev)
        b = torch.randn(M, Q, device=dev)
        ref = a.double().t() @ b.double()
        run = lambda: ops.gemm_tn(a, b, "tf32bc")
        err = lambda: float((ops.gemm_tn(a, b, "tf32bc").double() - ref).norm() / ref.norm())
    row = {"kind": kind, "shape": [M, P, Q]}
    for promo in (2, 4, 8, 16):
        lib.grl_gemm_bc_set_promo(promo)
        row[f"promo{promo}"] = {"ms": round(timeit(run), 4), "err": float(f"{err():.2e}")}
    lib.grl_gemm_bc_set_promo(4)
    print(json.dumps(row), flush=True)
    del a, b
