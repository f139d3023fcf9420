"""Loader for the committed reference vectors (tests/golden, written by oracle/make_golden.py)."""
import json
import os

import numpy as np
import torch

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
MODEL_CASES = ["model_small", "model_weighted_pad", "model_noatt"]


def load_model_case(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    B, N, S, L = meta["B"], meta["N"], meta["S"], meta["L"]
    H = S // 2
    case = {"meta": meta}
    for k in ("V", "A", "y", "logits_eval", "logits_train", "loss", "grad_norm"):
        case[k] = torch.from_numpy(z[k])
    case["edge_masks"] = [torch.from_numpy(np.unpackbits(z[f"edge_mask{i}"])[: B * (L + 1) * N * N]
                                           .reshape(B, (L + 1) * N, N).copy()) for i in range(3)]
    widths = [S, S, S, S, 10 * H]
    case["feat_masks"] = [torch.from_numpy(np.unpackbits(z[f"feat_mask{i}"])[: B * N * w].reshape(B, N, w).copy())
                          for i, w in enumerate(widths)]
    for group in ("sd0", "grad", "sd1"):
        case[group] = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith(group + "/")}
    return case


def rel_err(x, ref):
    """norm-wise relative error against a reference tensor (SURVEY.md 8c tolerance definition)."""
    x, ref = x.double().reshape(-1), ref.double().reshape(-1)
    return float((x - ref).norm() / ref.norm().clamp_min(1e-30))


PRODUCTION_CASES = ["prod_cfg1", "prod_s512", "prod_cfg1_tame", "prod_s512_tame"]
TAME_CASES = ["prod_cfg1_tame", "prod_s512_tame"]      # w_rand scaled to 0.005: logits O(1), CE not saturated


def load_production_case(name):
    """Production-shape vectors of the unmodified reference (oracle/make_golden.py production): fp64 outputs, the
    reference's own fp32-vs-fp64 noise per tensor, inputs regenerated from the seed (oracle/golden_inputs.py)."""
    from oracle.golden_inputs import production_inputs, sample_index
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    meta = json.loads(str(z["meta"]))
    case = {"meta": meta, "err_ref32": json.loads(str(z["err_ref32"])),
            "err_ref_bf16": json.loads(str(z["err_ref_bf16"])),
            "weights_check": json.loads(str(z["weights_check"]))}
    for k in ("logits_eval", "logits_train", "loss"):
        case[k] = torch.from_numpy(z[k])
    case["grad"] = {k.split("/", 1)[1]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad/")}
    case.update(production_inputs(meta))
    case["sample_index"] = lambda key, numel: sample_index(key, numel, meta["seed"])
    return case


def production_model(case):
    """``GraphCNNDropEdge`` built under the case's seed (= the reference constructor's weights, verified against the
    stored checksums), with the frozen projection rescaled when the case says so."""
    from graph_representation_learning_b200 import GraphCNNDropEdge
    m = case["meta"]
    torch.manual_seed(m["seed"])
    model = GraphCNNDropEdge(m["in_dim"], m["out_dim"], m["L"], net_size=m["S"])
    if m.get("w_rand_scale") is not None:
        with torch.no_grad():
            model.w_rand.projection.weight.mul_(m["w_rand_scale"])
    check_weights(model, case)
    return model


def check_weights(model, case):
    """same seed -> same initial weights as the reference constructor (sum / abs-sum per tensor of the golden file)"""
    for k, v in model.state_dict().items():
        s, a = case["weights_check"][k]
        v = v.double()
        assert abs(float(v.sum()) - s) <= 1e-9 * max(1.0, a) and abs(float(v.abs().sum()) - a) <= 1e-9 * max(1.0, a), k


def allowed(case, key, floor):
    """our norm-wise error against the fp64 reference may be max(floor, 2 x the reference's own fp32 noise).
    That noise is one draw per summation order (the same unmodified fp32 run is 1.1e-4 from fp64 with 16 OpenMP threads and
    2.2-2.7e-4 with 1 or 4): err_ref32 is the per-tensor max over the runs of oracle/make_golden.REF32_THREADS, all of
    them stored as err_ref32_runs."""
    return max(floor, 2.0 * case["err_ref32"][key])
