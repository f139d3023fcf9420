import json, os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import _lib, ops
lib = _lib.load(); dev = torch.device("cuda:0")
def timed(fn, iters=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters
for R, Ma, Nb in [(128000, 1792, 256), (128000, 3584, 256), (128000, 4372, 256), (128000, 256, 128), (128000, 1280, 53), (5000, 1792, 256), (777, 53, 1280), (20000, 300, 256), (1000, 129, 7)]:
    g = torch.Generator().manual_seed(R + Ma)
    a = torch.randn(R, Ma, generator=g).to(dev); b = torch.randn(R, Nb, generator=g).to(dev)
    res = {"R": R, "Ma": Ma, "Nb": Nb}; outs = {}
    for pairs in (0, 1):
        lib.grl_gemm_bc_set_pairs(pairs)
        outs[pairs] = ops.gemm_tn(a, b, "tf32bc"); torch.cuda.synchronize()
        ms = timed(lambda: ops.gemm_tn(a, b, "tf32bc"))
        res["pairs_ms" if pairs else "single_ms"] = round(ms, 4)
    lib.grl_gemm_bc_set_pairs(1)
    ref = a.double().t() @ b.double()
    res["identical"] = bool(torch.equal(outs[0], outs[1]))
    res["err_pairs"] = float((outs[1].double() - ref).norm() / ref.norm()); res["err_single"] = float((outs[0].double() - ref).norm() / ref.norm())
    res["tflops_pairs"] = round(2.0 * R * Ma * Nb / res["pairs_ms"] / 1e9, 1); res["tflops_single"] = round(2.0 * R * Ma * Nb / res["single_ms"] / 1e9, 1)
    print(json.dumps(res), flush=True)
