"""The oracle (oracle/reference_port.py) against the vectors produced by the UNMODIFIED reference."""
import os
import sys

import pytest
import torch

from _golden import MODEL_CASES, load_model_case, rel_err
from oracle import reference_port as rp

TOL = 2e-6  # same ATen kernels on the same CPU class; only summation-order noise is expected


@pytest.mark.parametrize("name", MODEL_CASES)
def test_eval_forward_matches_reference(name):
    c = load_model_case(name)
    out = rp.forward(c["sd0"], c["V"], c["A"].float(), training=False, use_attention=c["meta"]["use_attention"])
    assert rel_err(out, c["logits_eval"]) < TOL


@pytest.mark.parametrize("name", MODEL_CASES)
def test_train_step_matches_reference(name):
    c = load_model_case(name)
    masks = {"edge": c["edge_masks"], "feat": c["feat_masks"]}
    sd = {k: v.clone() for k, v in c["sd0"].items()}
    logits, loss, grads = rp.loss_and_grads(sd, c["V"], c["A"].float(), c["y"], True, masks,
                                            c["meta"]["use_attention"])
    assert rel_err(logits, c["logits_train"]) < TOL
    assert abs(float(loss) - float(c["loss"])) < 1e-5 * max(1.0, abs(float(c["loss"])))
    assert set(grads) == set(c["grad"])
    for k, g in grads.items():
        assert rel_err(g, c["grad"][k]) < 2e-5, k
    norm = rp.clip_grad_norm_(grads, 5.0)
    assert abs(float(norm) - float(c["grad_norm"])) < 1e-4 * float(c["grad_norm"])
    rp.adam_step(sd, grads, {}, step=1)
    for k, v in c["sd1"].items():
        assert rel_err(sd[k], v) < 1e-5, k


def test_preprocess_adj_structure():
    """Golden fact (i): A' == cat([I, A]) reshaped; rows i*(L+1)+t."""
    import numpy as np
    from _golden import GOLDEN
    z = np.load(os.path.join(GOLDEN, "preprocess_adj.npz"))
    A = torch.from_numpy(z["A"]).float()
    Ap = rp.preprocess_adj(A.permute(0, 1, 3, 2))
    B, N, L, _ = A.shape
    csr = Ap.reshape(B * N * (L + 1), N).to_sparse_csr()
    assert np.array_equal(csr.crow_indices().numpy().astype(np.int32), z["rowptr"])
    assert np.array_equal(csr.col_indices().numpy().astype(np.int32), z["col"])
    assert np.array_equal(csr.values().numpy(), z["val"])


def test_dropedge_value_set():
    """Golden fact (iii): after DropEdge, entries are in {0, a/0.7}; self loops are dropped too."""
    torch.manual_seed(0)
    Ap = rp.preprocess_adj(torch.zeros(1, 50, 50, 6))
    dropped = rp._drop(Ap, rp.P_EDGE, True, None)
    vals = torch.unique(dropped)
    assert torch.allclose(vals, torch.tensor([0.0, 1 / 0.7]))
    diag = dropped.reshape(50, 7, 50)[:, 0, :].diagonal()
    assert 0 < int((diag == 0).sum()) < 50


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="live reference only in the authoring container")
def test_against_live_reference():
    sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "oracle"))
    from make_golden import import_reference
    GraphCNNDropEdge, _ = import_reference()
    torch.manual_seed(3)
    m = GraphCNNDropEdge(30, 6, 6, net_size=64).eval()
    V = torch.rand(2, 15, 30)
    A = (torch.rand(2, 15, 6, 15) < 0.2).float()
    with torch.no_grad():
        ref = m([V, A])
    out = rp.forward(m.state_dict(), V, A)
    assert rel_err(out, ref) < TOL


@pytest.mark.parametrize("name", ["prod_cfg1", "prod_s512", "prod_cfg1_tame", "prod_s512_tame"])
def test_port_matches_reference_fp64_run_at_production_shapes(name):
    """cfg1 exactly (8 x 200, in_dim 4369, S=256) and an S=512 case: the port in fp64, fed the regenerated inputs and
    masks, against the fp64 run of the unmodified reference (logits, loss, every gradient at the stored positions)."""
    from _golden import load_production_case, production_model
    c = load_production_case(name)
    model = production_model(c)                              # same seed -> the reference constructor's weights (checked)
    sd = {k: v.detach().double() for k, v in model.state_dict().items()}
    V, A = c["V"].double(), c["A"].double()
    assert rel_err(rp.forward(sd, V, A, training=False), c["logits_eval"]) < 1e-12
    logits, loss, grads = rp.loss_and_grads(sd, V, A, c["y"], True, {"edge": c["edge_masks"], "feat": c["feat_masks"]})
    assert rel_err(logits, c["logits_train"]) < 1e-12
    assert abs(float(loss) - float(c["loss"])) < 1e-12 * float(c["loss"])
    assert set(grads) == set(c["grad"])
    for k, g in grads.items():
        idx = c["sample_index"](k, g.numel())
        assert rel_err(g.reshape(-1)[idx], c["grad"][k]) < 1e-10, k
