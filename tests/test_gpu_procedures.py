"""GPU tests of the callers around the kernels: collate -> device CSR, KVProcedure / GNNLearningWarper train,
KVInference predict -- all through the CUDA model (C-ABI kernels)."""
import os

import numpy as np
import pytest
import torch

from _golden import GOLDEN

pytestmark = pytest.mark.gpu

REFERENCE_KEYS = {"emb1.0.weight", "emb1.0.bias", "gcn1.h_weights", "gcn1.bias", "gcn2.h_weights", "gcn2.bias",
                  "gcn3.h_weights", "gcn3.bias", "emb2.0.weight", "emb2.0.bias", "self_atten.gamma",
                  "self_atten.f.0.weight", "self_atten.f.0.bias", "self_atten.g.0.weight", "self_atten.g.0.bias",
                  "self_atten.h.0.weight", "self_atten.h.0.bias", "w_rand.projection.weight", "classifier.weight",
                  "classifier.bias"}


def _samples(seed=3, sizes=(17, 30, 22, 30, 9, 26), D=24, C=5):
    rng = np.random.default_rng(seed)
    out = []
    for n in sizes:
        A = (rng.random((n, 6, n)) < 0.15).astype(np.float16)
        out.append({"textline_encoding": rng.random((n, D), dtype=np.float32), "adjacency_matrix": A,
                    "node_label": rng.integers(0, C, n).astype(np.int64)})
    return out


def test_edge_list_collate_packs_the_same_graph_as_the_dense_collate():
    from graph_representation_learning_b200 import GraphCNNDropEdge, PackedGraphBatch
    from graph_representation_learning_b200 import collate as C
    dev = torch.device("cuda:0")
    samples = _samples()
    dense = C.load_collate_processors({"NumpyPadding": {"name_value_pairs": {
        "textline_encoding": 0.0, "adjacency_matrix": 0.0, "node_label": -100.0}, "only_selected_items": True}})(
        [dict(s) for s in samples])
    host = C.EdgeListCollate(num_edges=6)([dict(s) for s in samples])
    packed = C.EdgeListCollate.to_device(host, dev)
    g_edge, g_dense = packed["adjacency_matrix"], PackedGraphBatch.from_dense(dense["adjacency_matrix"].to(dev))
    g_edge.build_transpose(), g_dense.build_transpose()
    for name in ("rowptr", "col", "val", "colptr", "rowidx", "perm"):
        assert torch.equal(getattr(g_edge, name), getattr(g_dense, name)), name            # integer work: bit-exact
    assert torch.equal(packed["node_label"].cpu(), dense["node_label"])
    assert torch.equal(packed["textline_encoding"].cpu(), dense["textline_encoding"])
    assert packed["textline_encoding"].stride(1) % 4 == 0                                  # pitch kept through H2D
    torch.manual_seed(0)
    model = GraphCNNDropEdge(24, 5, 6, net_size=64).to(dev).eval()
    with torch.no_grad():
        a = model([packed["textline_encoding"], g_edge])
        b = model([dense["textline_encoding"].to(dev), dense["adjacency_matrix"].to(dev)])
    assert torch.equal(a, b)


def _cfg(tmp_path, **over):
    data = {"num_documents": 24, "n_min": 20, "n_max": 40, "input_dim": 48, "num_classes": 6, "batch_size": 8,
            "shuffle": False, "drop_last": True, "pin_memory": True,
            "data_collate": {"NumpyPadding": {"name_value_pairs": {"textline_encoding": 0.0, "adjacency_matrix": 0.0,
                                                                   "node_label": -100.0}, "only_selected_items": True}}}
    cfg = {"output_dir": str(tmp_path), "num_gpus": 1, "num_epochs": 3, "seed": 5,
           "data_config": {"dataset": {"type": "SyntheticDocumentDataset", "args": {"node_label_padding_value": -100,
                                                                                    "other_class_index": None}},
                           "training": dict(data, seed=1), "validation": dict(data, seed=2, batch_size=4)}}
    cfg.update(over)
    return cfg


def test_warper_trains_the_cuda_model_and_writes_a_reference_compatible_checkpoint(tmp_path):
    from graph_representation_learning_b200 import GNNLearningWarper, GraphCNNDropEdge, _lib
    torch.manual_seed(1)
    model = GraphCNNDropEdge(48, 6, 6, net_size=64)
    w = GNNLearningWarper(model, config=_cfg(tmp_path))
    n0 = _lib.launches()
    seen = []
    inner = w.trainer._optimize_per_epoch

    def recording(epoch):
        out = inner(epoch)
        seen.append((epoch, out["loss"]))
        return out
    w.trainer._optimize_per_epoch = recording
    f1 = w.train()
    assert _lib.launches() - n0 > 100                     # our kernels ran, not a fallback
    assert 0.0 <= f1 <= 1.0
    ckpt = torch.load(os.path.join(str(tmp_path), "GNN-B200", "models", "model_latest.pt"), weights_only=False)
    assert set(ckpt) == {"epoch", "config", "meta_data", "state_dict"}
    assert set(ckpt["state_dict"]) == REFERENCE_KEYS      # loadable by the reference's load_state_dict(strict=False)
    best_epoch, best_loss = min(seen, key=lambda t: (t[1], t[0]))
    assert len(seen) == 3 and ckpt["meta_data"]["loss"] == best_loss and ckpt["epoch"] == best_epoch   # kv_procedure.py:358-368
    assert w.trainer.global_step == 3 * 3 and next(model.parameters()).is_cuda


def test_kv_inference_batched_equals_per_document_on_the_cuda_model(tmp_path):
    from graph_representation_learning_b200 import GNNLearningWarper, GraphCNNDropEdge
    feats = {}

    def featurizer(doc):
        key = id(doc)
        n = len(doc)
        r = np.random.default_rng(n * 7 + int(doc[0]["text"][1:]))
        f = {"textline_encoding": r.random((n, 24), dtype=np.float32)}
        A = (r.random((n, 6, n)) < 0.2).astype(np.float16)
        if n % 2:
            f["adjacency_matrix"] = A
        else:
            s, l, e = np.nonzero(A)
            f["edges"] = np.stack([s, l, e], 1).astype(np.int32)
        feats[key] = f
        return f

    docs = [[{"text": f"t{k}", "location": [[0, 0]] * 4} for _ in range(n)] for k, n in enumerate((12, 15, 12, 15, 15, 8))]
    torch.manual_seed(2)
    model = GraphCNNDropEdge(24, 5, 6, net_size=64)
    cfg = _cfg(tmp_path, is_train=False, procedure={"type": "KVInference", "args": {}})
    ids = {i: (f"field{i}", "key" if i % 2 else "value") for i in range(1, 5)}
    batched = GNNLearningWarper(model, config=cfg, featurizer=featurizer, id_to_class=ids).predict(docs)
    single = GNNLearningWarper(model, config=cfg, featurizer=featurizer, id_to_class=ids, max_batch_documents=1).predict(
        [[dict(b) for b in d] for d in docs])
    for a, b in zip(batched, single):
        assert len(a) == len(b)
        for x, y in zip(a, b):
            assert x["formal_key"] == y["formal_key"] and x["key_type"] == y["key_type"]
            assert abs(x["confidence"] - y["confidence"]) < 1e-5


def test_kv_inference_output_matches_the_oracle_and_the_reference_schema(tmp_path):
    """predict() (kv_inference.py:57-75): the annotated boxes validate against the reference's own output schema
    (tests/golden/output_schema.json, a copy of its tests/assets/schemas/output_schema.json made by oracle/make_golden.py)
    and carry the classes / confidences of the ORACLE's forward (oracle/reference_port.py) on the same documents."""
    import json
    import jsonschema
    from _golden import GOLDEN
    from graph_representation_learning_b200 import GNNLearningWarper, GraphCNNDropEdge
    from oracle import reference_port as rp     # checker only
    feats = {}

    def featurizer(doc):
        n = len(doc)
        r = np.random.default_rng(n * 11 + int(doc[0]["text"][1:]))
        f = {"textline_encoding": r.random((n, 24), dtype=np.float32)}
        A = (r.random((n, 6, n)) < 0.2).astype(np.float16)
        if n % 2:
            f["adjacency_matrix"] = A
        else:
            s_, l_, e_ = np.nonzero(A)
            f["edges"] = np.stack([s_, l_, e_], 1).astype(np.int32)
        feats[doc[0]["text"]] = (f["textline_encoding"], A)
        return f

    docs = [[{"text": f"t{k}", "location": [[0, 0], [4, 0], [4, 2], [0, 2]], "type": "textline"} for _ in range(n)]
            for k, n in enumerate((12, 15, 12, 9, 15, 40))]
    torch.manual_seed(3)
    model = GraphCNNDropEdge(24, 5, 6, net_size=64)
    sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
    cfg = _cfg(tmp_path, is_train=False, procedure={"type": "KVInference", "args": {}})
    ids = {i: (f"field{i}", "key" if i % 2 else "value") for i in range(1, 5)}
    out = GNNLearningWarper(model, config=cfg, featurizer=featurizer, id_to_class=ids).predict(docs)
    schema = json.load(open(os.path.join(GOLDEN, "output_schema.json")))
    ids[0] = ("other", "other")
    assert len(out) == len(docs)
    for doc_out, doc in zip(out, docs):
        jsonschema.validate(doc_out, schema)                         # the reference's output contract
        V, A = feats[doc[0]["text"]]
        logits = rp.forward(sd, torch.from_numpy(V)[None], torch.from_numpy(A.astype(np.float32))[None], training=False)
        prob = torch.softmax(logits, dim=2)[0]
        top2 = prob.topk(2, dim=1).values
        score, cls = prob.max(dim=1)
        assert len(doc_out) == len(doc)
        for i, box in enumerate(doc_out):
            assert abs(box["confidence"] - float(score[i])) < 1e-5
            if float(top2[i, 0] - top2[i, 1]) > 1e-4:                # away from ties the class is the oracle's
                assert (box["formal_key"], box["key_type"]) == ids[int(cls[i])]


def test_graphed_eval_forward_matches_eager():
    """CUDA-graph replay of pack + eval forward == the eager path, bit for bit, on fresh batches after capture"""
    from graph_representation_learning_b200 import GraphCNNDropEdge
    from graph_representation_learning_b200.procedures import GraphedEvalForward
    dev = torch.device("cuda:0")
    torch.manual_seed(2)
    B, N, L, D, S, C = 4, 50, 6, 96, 64, 7
    model = GraphCNNDropEdge(D, C, L, net_size=S).to(dev).eval()
    gf = GraphedEvalForward(model, B, N, D, L)
    gen = torch.Generator().manual_seed(1)
    for density in (0.1, 0.3, 0.0):
        V = torch.rand(B, N, D, generator=gen)
        A = (torch.rand(B, N, L, N, generator=gen) < density).half()
        out = gf(V, A).clone()
        with torch.no_grad():
            ref = model([V.to(dev), A.to(dev)])
        assert torch.equal(out, ref)
    full = (torch.ones(B, N, L, N)).half()                 # the dense worst case fits the buffers
    with torch.no_grad():                                  # same path as the capture (fused GraphConv layers need no_grad)
        ref = model([V.to(dev), full.to(dev)])
    assert torch.equal(gf(V, full), ref)


def test_graphed_train_step_matches_eager_step_bit_for_bit():
    """trainer.GraphedTrainStep: forward + CE + backward captured in ONE CUDA graph (dense pack included), optimizer eager
    behind it.  Replays draw fresh DropEdge / dropout masks through the device-side Philox offset
    (grl_rng_set_offset_ptr) -- the masks of the eager step at the same step index -- so losses and weights after three
    steps on changing batches equal the eager KVTrainStep's bit for bit."""
    from graph_representation_learning_b200 import GraphCNNDropEdge
    from graph_representation_learning_b200.trainer import GraphedTrainStep, KVTrainStep
    dev = torch.device("cuda:0")
    B, N, L, D, S, C = 4, 60, 6, 96, 64, 7

    def make():
        torch.manual_seed(5)
        m = GraphCNNDropEdge(D, C, L, net_size=S).to(dev)
        m.rng_seed = 1234
        return m

    gen = torch.Generator().manual_seed(3)
    batches = []
    for k in range(3):
        A = (torch.rand(B, N, L, N, generator=gen) < (0.10 - 0.01 * k)).half()      # (the eager sync-free pack sizes its buffers by the first batch)
        batches.append({"textline_encoding": torch.rand(B, N, D, generator=gen), "adjacency_matrix": A,
                        "node_label": torch.randint(0, C, (B, N), generator=gen)})
    m_g, m_e = make(), make()
    step_g, step_e = KVTrainStep(m_g), KVTrainStep(m_e)
    graphed = GraphedTrainStep(step_g, B, N, D, L)
    for k, v in m_g.state_dict().items():                          # warm-up and capture did not touch the weights
        assert torch.equal(v, m_e.state_dict()[k]), k
    captured_calls = m_g._rng_calls - 1                            # the forward call the graph captured
    masks_seen = []
    for r, batch in enumerate(batches, start=1):
        loss_g = graphed.run_train_step(batch).clone()
        m_e._rng_calls = captured_calls + r                        # same Philox streams as replay r
        dense = {k: v.to(dev) for k, v in batch.items()}
        loss_e = step_e.run_train_step(dense)
        assert torch.equal(loss_g, loss_e), (r, float(loss_g), float(loss_e))
        masks_seen.append(float(loss_g))
    for k, v in m_g.state_dict().items():
        assert torch.equal(v, m_e.state_dict()[k]), k
    # replaying the SAME batch twice gives different losses: the masks are redrawn per replay
    a = float(graphed.run_train_step(batches[0]))
    b = float(graphed.run_train_step(batches[0]))
    assert a != b
