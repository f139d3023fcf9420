import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from graph_representation_learning_b200 import ops
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(1)
A = torch.randn(128000, 3584, generator=g).to(dev); W = (torch.randn(256, 3584, generator=g) / 60).to(dev); b = torch.randn(256, generator=g).to(dev)
dZ = torch.randn(128000, 256, generator=g).to(dev)
for _ in range(3):
    y = ops.gemm_nt([A], [W], bias=b, relu=True, fp32_mode="tf32x3")
    u = ops.gemm_nt([A[:, :1792]], [W[:, :1792]], fp32_mode="tf32bc")
    dw = ops.gemm_tn(A, dZ, "tf32bc")
torch.cuda.synchronize()
print("ok")
