"""Second-generation fp32 NT GEMM (tf32 + bf16 correction): clusters of 2 CTAs sharing weight tiles by TMA multicast
vs single-CTA loads, on the cfg2 backward (dX) shapes: time, useful TF/s, bit-equality.  Run on the GPU box."""
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


for (M, N, K) in ((128000, 1792, 256), (128000, 3584, 256), (128000, 128, 1280), (128000, 256, 128), (128000, 256, 3584),
                  (127900, 53, 1280), (19000, 300, 100), (128000 + 128, 512, 64)):
    a = torch.randn(M, K, device=dev)
    b = torch.randn(N, K, device=dev) / K ** 0.5
    bias = torch.randn(N, device=dev)
    out, ms = {}, {}
    for on in (0, 1):
        lib.grl_gemm_bc_set_multicast(on)
        f = lambda: ops.gemm_nt([a], [b], bias=bias, relu=True, fp32_mode="tf32bc")
        out[on] = f()
        torch.cuda.synchronize()
        ms[on] = timeit(f)
    lib.grl_gemm_bc_set_multicast(1)
    ref = torch.relu(a[:4096].double() @ b.double().t() + bias.double())
    fl = 2.0 * M * N * K
    print(json.dumps({"nt_tf32bc": [M, N, K], "single_ms": round(ms[0], 4), "multicast_ms": round(ms[1], 4),
                      "useful_TFs": [round(fl / ms[0] / 1e9, 1), round(fl / ms[1] / 1e9, 1)],
                      "bit_identical": bool(torch.equal(out[0], out[1])),
                      "rel_err_vs_fp64": float((out[1][:4096].double() - ref).norm() / ref.norm())}), flush=True)
    del a, b, out
