"""Medium-shape model errors vs the fp64 oracle over several seeds, for both fp32 GEMM modes and for the
reference-style fp32 CPU run: are outliers (ReLU-derivative flips) mode-specific or seed noise?"""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from _golden import rel_err
from oracle import reference_port as rp
from graph_representation_learning_b200 import GraphCNNDropEdge, ops
dev = torch.device("cuda:0")
keys = ["logits", "emb1.0.weight", "self_atten.f.0.weight", "self_atten.h.0.weight"]
for seed in range(21, 21 + int(sys.argv[1]) if len(sys.argv) > 1 else 29):
    torch.manual_seed(seed)
    B, N, L, in_dim, S, out_dim = 8, 200, 6, 512, 256, 53
    model = GraphCNNDropEdge(in_dim, out_dim, L, net_size=S)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    gen = torch.Generator().manual_seed(seed + 100)
    V = torch.rand(B, N, in_dim, generator=gen)
    A = (torch.rand(B, N, L, N, generator=gen) < 8.0 / N).half()
    y = torch.randint(0, out_dim, (B, N), generator=gen)
    ref_logits, _, ref_grads = rp.loss_and_grads({k: v.double() for k, v in sd.items()}, V.double(), A.double(), y, training=False)
    f32_logits, _, f32_grads = rp.loss_and_grads(sd, V, A.float(), y, training=False)
    row = {"seed": seed, "cpu_fp32": {k: rel_err(f32_logits if k == "logits" else f32_grads[k], ref_logits if k == "logits" else ref_grads[k]) for k in keys}}
    model = model.to(dev).eval()
    from graph_representation_learning_b200 import _lib
    grid = (("tf32x3", "tf32x3", 2), ("tf32bc", "tf32bc", 2), ("tf32x3", "tf32bc", 2), ("tf32bc", "tf32x3", 2))
    if os.environ.get("NOISE_X3_PROMO"):      # forward three-product accumulator promotion interval: 2 (default) vs 4 vs 8
        grid = (("tf32x3", "tf32bc", 2), ("tf32x3", "tf32bc", 4), ("tf32x3", "tf32bc", 8))
    for mode, mode_b, x3p in grid:
        ops.GemmConfig.fp32_mode, ops.GemmConfig.fp32_mode_bwd = mode, mode_b
        _lib.load().grl_gemm_x3_set_promo(x3p)
        promo = mode_b[4:] + (f"_x3promo{x3p}" if x3p != 2 else "")
        model.zero_grad(set_to_none=True)
        logits = model([V.to(dev), A.to(dev)])
        torch.nn.functional.cross_entropy(logits.transpose(1, 2), y.to(dev)).backward()
        g = dict(model.named_parameters())
        row[f"{mode}_p{promo}"] = {k: rel_err(logits.detach().cpu() if k == "logits" else g[k].grad.cpu(), ref_logits if k == "logits" else ref_grads[k]) for k in keys}
    print(json.dumps({k: ({kk: float(f"{vv:.2e}") for kk, vv in v.items()} if isinstance(v, dict) else v) for k, v in row.items()}))
