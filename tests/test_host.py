"""CPU tests of the host-side mirror of the reference interface: collate, factories, metrics, config, checkpoint,
procedures and the GNNLearningWarper control flow (run with a tiny CPU nn.Module standing in for the CUDA model --
the product model itself refuses to run without a GPU, which is tested too)."""
import os

import numpy as np
import pytest
import torch
from torch import nn

from _golden import GOLDEN

from graph_representation_learning_b200 import GNNLearningWarper, GraphCNNDropEdge, GrlError
from graph_representation_learning_b200 import collate as C
from graph_representation_learning_b200 import procedures as P
from graph_representation_learning_b200 import trainer as T
from graph_representation_learning_b200.config import AttrDict, load_config, plain


def _golden_samples():
    z = np.load(os.path.join(GOLDEN, "collate.npz"))
    n = len({k.split("/")[0] for k in z.files if k.startswith("in")})
    samples = [{k.split("/", 1)[1]: z[k] for k in z.files if k.startswith(f"in{i}/")} for i in range(n)]
    return samples, {k.split("/", 1)[1]: z[k] for k in z.files if k.startswith("out/")}


def test_numpy_padding_matches_reference_vectors():
    samples, want = _golden_samples()
    for s in samples:
        s["other"] = "dropped"
    chain = C.load_collate_processors({"NumpyPadding": {
        "name_value_pairs": {"textline_encoding": 0.0, "adjacency_matrix": 0.0, "node_label": -100.0},
        "only_selected_items": True}})
    batch = chain([dict(s) for s in samples])
    assert set(batch) == set(want)
    for k, v in want.items():
        assert batch[k].dtype == torch.from_numpy(v).dtype, k           # golden fact (v): f32 / f16 / i64
        assert np.array_equal(batch[k].numpy(), v), k


def test_edge_list_collate_places_documents_like_numpy_padding():
    samples, want = _golden_samples()
    col = C.EdgeListCollate(num_edges=6, pin_memory=False)
    out = col([dict(s) for s in samples])
    assert np.array_equal(out["textline_encoding"].numpy(), want["textline_encoding"])
    assert np.array_equal(out["node_label"].numpy(), want["node_label"])
    assert out["features_padded"].stride(1) % 4 == 0                     # 16-byte row pitch for TMA
    B, N, L = out["graph_shape"]
    dense = np.zeros((B, N, L, N), dtype=np.float16)
    e = out["edges"].numpy()
    dense[e[:, 0], e[:, 1], e[:, 2], e[:, 3]] = 1
    assert np.array_equal(dense, want["adjacency_matrix"])
    # explicit triples give the same batch as the dense per-document matrices
    tri = [dict(s, edges=C.dense_to_edges(s["adjacency_matrix"])[0]) for s in samples]
    for s in tri:
        del s["adjacency_matrix"]
    assert torch.equal(col(tri)["edges"], out["edges"])
    assert C.pad_offset(5, 9) == 2 and C.pad_offset(2, 9) == 3           # golden fact (iv)


def test_macro_scores_match_sklearn():
    from sklearn.metrics import classification_report
    g = torch.Generator().manual_seed(0)
    for n_cls_true, n_cls_pred in ((7, 9), (3, 3), (5, 2)):
        y = torch.randint(0, n_cls_true, (400,), generator=g)
        y[::9] = -100
        p = torch.randint(0, n_cls_pred, (400,), generator=g)
        got = T.macro_scores(p, y, (-100, None), "x")
        sel = y != -100
        ref = classification_report([str(int(a)) for a in y[sel]], [str(int(a)) for a in p[sel]], output_dict=True,
                                    zero_division=0)["macro avg"]
        for k, v in ref.items():
            assert abs(got[f"x_{k}"] - v) < 1e-12, k
    assert T.macro_scores(torch.zeros(3, dtype=torch.long), torch.full((3,), -100), (-100,), "x")["x_f1-score"] == 0.0


def test_factories_follow_the_reference_quirks():
    opt = T.from_registry(T.OPTIMIZERS, "optimizer", "BuitlinOptimizer", {"type_optimizer": "Adam", "lr": 0.5})
    o = opt.get_optimizer([nn.Parameter(torch.zeros(2))])
    assert o.param_groups[0]["lr"] == 1e-3                               # configured lr is swallowed (builtin_optimizer.py)
    lr = T.from_registry(T.LR_SCHEDULERS, "lr", "DecayLearningRate", {"lr": 0.001, "factor": 0.9995, "num_epochs": 100})
    assert lr._step_lr(0) == 0.001
    assert abs(lr._step_lr(10) - 0.001 * (1 - 10 / 101.0) ** 0.9995) < 1e-15
    assert T.from_registry(T.LOSSES, "loss", None, None) is None
    with pytest.raises(ValueError):
        T.from_registry(T.LOSSES, "loss", "Nope", {})
    d = T.Dictlist()
    d._update({"a": 1.0})
    d._update({"a": 2.0, "b": 0.12345})
    assert d._result() == {"a": 1.5, "b": 0.123}
    ce = T.CrossEntropyLoss()
    pred, y = torch.randn(2, 5, 4), torch.tensor([[0, 1, -100, 3, 2], [1, 1, 0, -100, -100]])
    assert torch.allclose(ce(pred, y), nn.functional.cross_entropy(pred.reshape(-1, 4), y.reshape(-1), ignore_index=-100))


def test_config_and_checkpoint_round_trip(tmp_path):
    path = tmp_path / "c.yaml"
    path.write_text("seed: 7\nprocedure:\n  type: KVProcedure\n  args: {metrics_interval: 2}\nnum_epochs: 3\n")
    cfg = load_config(str(path))
    assert cfg.seed == 7 and cfg.procedure.args.metrics_interval == 2 and cfg.optimizer.args.type_optimizer == "Adam"
    assert isinstance(cfg.loss, AttrDict) and plain(cfg)["loss"] == {"type": "CrossEntropyLoss", "args": {}}
    with pytest.raises(AttributeError):
        cfg.nope
    h = P.CheckpointHandler()
    assert h.make_checkpoint_name("model") == "model_latest.pt"
    assert h.make_checkpoint_name("model", 2, 5) == "model_epoch_2_minibatch_5.pt"
    m = nn.Linear(3, 2)
    f = h.save_checkpoint({"epoch": 1, "config": plain(cfg), "meta_data": {}, "state_dict": m.state_dict()}, str(tmp_path))
    back = h.restore_checkpoint(f)
    assert set(back) == {"epoch", "config", "meta_data", "state_dict"} and torch.equal(back["state_dict"]["weight"], m.weight)


class TinyDenseModel(nn.Module):
    """CPU stand-in with the model interface the procedures use: forward([V, A]) -> B x N x classes."""

    def __init__(self, d, c):
        super().__init__()
        self.lin = nn.Linear(d, c)
        self.num_edges = 6

    def forward(self, inputs):
        V, A = inputs
        agg = torch.einsum("biln,bnd->bid", A.float(), V)
        return self.lin(V + 0.1 * agg)


def _cfg(tmp_path, **over):
    data = {"num_documents": 12, "n_min": 5, "n_max": 9, "input_dim": 8, "num_classes": 4, "batch_size": 4,
            "shuffle": False, "drop_last": True,
            "data_collate": {"NumpyPadding": {"name_value_pairs": {"textline_encoding": 0.0, "adjacency_matrix": 0.0,
                                                                   "node_label": -100.0}, "only_selected_items": True}}}
    cfg = {"output_dir": str(tmp_path), "num_gpus": 0, "num_epochs": 2,
           "data_config": {"dataset": {"type": "SyntheticDocumentDataset", "args": {"node_label_padding_value": -100,
                                                                                    "other_class_index": None}},
                           "training": dict(data, seed=1), "validation": dict(data, seed=2, batch_size=2)}}
    cfg.update(over)
    return cfg


def test_warper_trains_on_cpu_stand_in_and_checkpoints(tmp_path):
    model = TinyDenseModel(8, 4)
    w = GNNLearningWarper(model, config=_cfg(tmp_path))
    before = model.lin.weight.detach().clone()
    f1 = w.train()
    assert 0.0 <= f1 <= 1.0 and not torch.equal(before, model.lin.weight)
    assert w.trainer.global_step == 2 * 3 and hasattr(model, "lambda_value")
    assert abs(w.trainer.optimizer.param_groups[0]["lr"] - 0.001 * (1 - 1 / 101.0) ** 0.9995) < 1e-12
    ckpt = os.path.join(str(tmp_path), "GNN-B200", "models", "model_latest.pt")
    assert os.path.exists(ckpt)
    # restore through the config, like base_procedure.py:150-170
    m2 = TinyDenseModel(8, 4)
    GNNLearningWarper(m2, config=_cfg(tmp_path, checkpoint_path=ckpt, num_epochs=0))
    saved = torch.load(ckpt, weights_only=False)["state_dict"]
    assert torch.equal(m2.lin.weight, saved["lin.weight"])
    with pytest.raises(AttributeError):
        GNNLearningWarper(model, config=_cfg(tmp_path, procedure={"type": "Nope", "args": {}}))


def test_kv_inference_batches_equal_sized_documents(tmp_path):
    rng = np.random.default_rng(0)
    docs = []
    for n in (4, 6, 4, 5, 6):
        docs.append([{"text": f"t{i}", "location": [[0, 0]] * 4} for i in range(n)])

    def featurizer(doc):
        n = len(doc)
        r = np.random.default_rng(n * 100 + len(doc[0]["text"]))
        return {"textline_encoding": r.random((n, 8), dtype=np.float32),
                "adjacency_matrix": (r.random((n, 6, n)) < 0.3).astype(np.float16)}

    model = TinyDenseModel(8, 3)
    cfg = _cfg(tmp_path, is_train=False, procedure={"type": "KVInference", "args": {}})
    ids = {1: ("total", "key"), 2: ("total", "value")}
    w = GNNLearningWarper(model, config=cfg, featurizer=featurizer, id_to_class=ids)
    out = w.predict(docs)
    assert [len(o) for o in out] == [4, 6, 4, 5, 6]
    one = GNNLearningWarper(model, config=cfg, featurizer=featurizer, id_to_class=ids, max_batch_documents=1)
    single = one.predict([[dict(b) for b in d] for d in docs])
    for a, b in zip(out, single):
        for x, y in zip(a, b):
            assert x["key_type"] == y["key_type"] and x["formal_key"] == y["formal_key"]
            assert abs(x["confidence"] - y["confidence"]) < 1e-6 and 0 < x["confidence"] <= 1
    assert set(out[0][0]) >= {"text", "location", "key_type", "formal_key", "confidence"}   # output_schema.json keys
    assert len(w.predict(docs[0])) == 1                                                      # single document input


def test_product_model_refuses_cpu():
    m = GraphCNNDropEdge(8, 3, 6, net_size=16)
    with pytest.raises(GrlError):
        m([torch.zeros(1, 4, 8), torch.zeros(1, 4, 6, 4)])


def test_fused_graphconv_policy_is_one_wave_of_clusters():
    """ops.FusedGraphConv.wants: "auto" runs a no-grad GraphConv layer fused when its CTA clusters fit one wave of the
    148 SMs (B * ceil(N / 128) tiles x ceil(C / 128) CTAs); True / False force it (host logic only)."""
    from types import SimpleNamespace
    from graph_representation_learning_b200 import ops
    prev = ops.FusedGraphConv.enabled
    try:
        ops.FusedGraphConv.enabled = "auto"
        assert ops.FusedGraphConv.wants(SimpleNamespace(B=64, N=100), 256)            # 64 tiles x 2 CTAs
        assert not ops.FusedGraphConv.wants(SimpleNamespace(B=64, N=200), 256)        # 128 tiles x 2
        assert ops.FusedGraphConv.wants(SimpleNamespace(B=8, N=200), 512)             # 16 tiles x 4
        assert not ops.FusedGraphConv.wants(SimpleNamespace(B=256, N=500), 128)
        ops.FusedGraphConv.enabled = True
        assert ops.FusedGraphConv.wants(SimpleNamespace(B=256, N=500), 256)
        ops.FusedGraphConv.enabled = False
        assert not ops.FusedGraphConv.wants(SimpleNamespace(B=1, N=10), 64)
    finally:
        ops.FusedGraphConv.enabled = prev
