"""One GEMM shape through both fp32 modes, for `ncu -k regex:gemm_`:  python scripts/gemm_one.py [N K]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import ops  # noqa: E402

N, K = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (256, 1792)
M = 128000
dev = torch.device("cuda:0")
a = torch.randn(M, K, device=dev)
b = torch.randn(N, K, device=dev) / K ** 0.5
x = torch.randn(M, K, device=dev)
dz = torch.randn(M, N, device=dev)
for mode in ("tf32x3", "tf32bc"):
    ops.GemmConfig.fp32_mode = ops.GemmConfig.fp32_mode_bwd = mode
    for _ in range(2):
        ops.gemm_nt([a], [b])
        ops.gemm_tn(x, dz)
torch.cuda.synchronize()
