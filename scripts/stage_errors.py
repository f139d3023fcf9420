"""Diagnostic (GPU box): per-stage forward error and ReLU sign flips of the CUDA path vs the fp64 oracle,
next to the reference-style fp32 CPU run, on a cfg-1-like shape."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch, ops  # noqa: E402
from oracle import reference_port as rp  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(21)
B, N, L, in_dim, S, out_dim = 8, 200, 6, 512, 256, 53
model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S)
sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
gen = torch.Generator().manual_seed(2)
V = torch.rand(B, N, in_dim, generator=gen)
A = (torch.rand(B, N, L, N, generator=gen) < 8.0 / N).half()


def stages(sd, V, A):
    lin = torch.nn.functional.linear
    Ap = rp.preprocess_adj(A.permute(0, 1, 3, 2))
    z0 = lin(V, sd["emb1.0.weight"], sd["emb1.0.bias"])
    emb = torch.relu(z0)
    z1 = rp.graph_conv(emb, Ap, sd["gcn1.h_weights"], sd["gcn1.bias"])
    g1 = torch.relu(z1)
    z2 = rp.graph_conv(g1, Ap, sd["gcn2.h_weights"], sd["gcn2.bias"])
    g2 = torch.relu(z2)
    z3 = rp.graph_conv(torch.cat([g1, g2], -1), Ap, sd["gcn3.h_weights"], sd["gcn3.bias"])
    return {"z0": z0, "z1": z1, "z2": z2, "z3": z3}


ref = stages({k: v.double() for k, v in sd.items()}, V.double(), A.double())
cpu32 = stages(sd, V, A.float())

g = PackedGraphBatch.from_dense(A.to(dev))
m = model.to(dev)
x = V.reshape(B * N, in_dim).to(dev)
ours = {}
ours["z0"] = ops.linear_act([x], m.emb1[0].weight, m.emb1[0].bias)
emb = torch.relu(ours["z0"])
ours["z1"] = ops.graph_conv([emb], g, g.val, m.gcn1.h_weights, m.gcn1.bias)
g1 = torch.relu(ours["z1"])
ours["z2"] = ops.graph_conv([g1], g, g.val, m.gcn2.h_weights, m.gcn2.bias)
g2 = torch.relu(ours["z2"])
ours["z3"] = ops.graph_conv([g1, g2], g, g.val, m.gcn3.h_weights, m.gcn3.bias)


def rel(a, b):
    return float((a.double().reshape(-1) - b.reshape(-1)).norm() / b.norm())


print(f"{'stage':6} {'ours rel':>10} {'cpu32 rel':>10} {'ours flips':>11} {'cpu32 flips':>12} {'n':>9}  |z|<1e-5 frac")
for k in ref:
    r = ref[k].reshape(-1)
    o = ours[k].detach().cpu().reshape(-1)
    c = cpu32[k].reshape(-1)
    print(f"{k:6} {rel(o, r):10.2e} {rel(c, r):10.2e} {int(((o > 0) != (r > 0)).sum()):11d} "
          f"{int(((c > 0) != (r > 0)).sum()):12d} {r.numel():9d}  {float((r.abs() < 1e-5).float().mean()):.2e} "
          f"std {float(r.std()):.3f}")
