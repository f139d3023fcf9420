"""World-size-2 gloo test of the data-parallel step (host logic of trainer.KVTrainStep / FlatGradBucket):
N-rank averaged gradients == 1-rank gradients on the concatenated batch, replicas stay bit-identical after the step.
Runs on CPU with a stand-in model (the CUDA model needs a GPU); the collective code path is the one NCCL runs."""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp
from torch import nn

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class TinyDenseModel(nn.Module):
    def __init__(self, d, c):
        super().__init__()
        self.a = nn.Linear(d, 16)
        self.b = nn.Linear(16, c)
        self.frozen = nn.Linear(4, 4)
        for p in self.frozen.parameters():
            p.requires_grad = False

    def forward(self, inputs):
        V, A = inputs
        agg = torch.einsum("biln,bnd->bid", A.float(), V)
        return self.b(torch.relu(self.a(V + 0.1 * agg)))


def _batch(seed, B, N=6, D=5, C=3):
    """RAGGED labels: graph b has b % 5 padded (-100) nodes, so the two ranks own different numbers of valid nodes and
    a mean of per-rank means would NOT equal the mean over the concatenated batch (cross_entropy_loss.py:33-35)."""
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, C, (B, N), generator=g)
    for b in range(B):
        y[b, : (b * b) % 5] = -100
    return {"textline_encoding": torch.rand(B, N, D, generator=g),
            "adjacency_matrix": (torch.rand(B, N, 6, N, generator=g) < 0.3).half(),
            "node_label": y}


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    from graph_representation_learning_b200.trainer import KVTrainStep
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)
    model = TinyDenseModel(5, 3)
    step = KVTrainStep(model, max_grad_norm=1e9, device=torch.device("cpu"))
    full = _batch(11, 4 * world)
    shard = {k: v[rank * 4:(rank + 1) * 4] for k, v in full.items()}        # rank r owns graphs [r*B/G, (r+1)*B/G)
    assert step.bucket.n_head_params == 2                                    # module `a` = the overlapped head bucket
    loss = step.run_train_step(shard)
    assert step.bucket._hooks and step.bucket._tail_work is None             # the overlapped path ran and was joined
    torch.save({"grad": step.bucket.flat.clone(), "sd": model.state_dict(), "loss": loss},
               os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.timeout(180)
def test_two_rank_gradient_average_equals_single_rank_on_concatenated_batch(tmp_path):
    sys.path.insert(0, ROOT)
    from graph_representation_learning_b200.trainer import KVTrainStep
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0 = torch.load(tmp_path / "r0.pt")
    r1 = torch.load(tmp_path / "r1.pt")
    assert torch.equal(r0["grad"], r1["grad"])                              # identical after the all-reduce
    for k in r0["sd"]:
        assert torch.equal(r0["sd"][k], r1["sd"][k]), k                     # replicas stay in lock step

    torch.manual_seed(0)
    model = TinyDenseModel(5, 3)
    single = KVTrainStep(model, max_grad_norm=1e9, device=torch.device("cpu"))
    full = _batch(11, 4 * world)
    counts = [(full["node_label"][r * 4:(r + 1) * 4] != -100).sum().item() for r in range(world)]
    assert counts[0] != counts[1]                                           # really ragged
    loss = single.run_train_step(full)
    assert torch.allclose(single.bucket.flat, r0["grad"], rtol=1e-5, atol=1e-7)
    assert torch.allclose(r0["loss"] + r1["loss"], loss, rtol=1e-5)         # per-rank shares add up to the global mean
    for k, v in model.state_dict().items():
        assert torch.allclose(v, r0["sd"][k], rtol=1e-5, atol=1e-6), k


def test_flat_bucket_views_and_clip():
    sys.path.insert(0, ROOT)
    from graph_representation_learning_b200.trainer import FlatGradBucket
    m = TinyDenseModel(5, 3)
    b = FlatGradBucket(m)
    n = sum(p.numel() for p in m.parameters() if p.requires_grad)
    assert b.flat.numel() == n and all(p.grad.data_ptr() >= b.flat.data_ptr() for p in b.params)
    assert m.frozen.weight.grad is None
    b.flat.fill_(2.0)
    ref = [p.grad.clone() for p in b.params]
    total_ref = torch.nn.utils.clip_grad_norm_([p for p in b.params], 5.0)   # on the same views
    b.flat.fill_(2.0)
    total = b.clip_(5.0)
    assert torch.allclose(total, torch.linalg.vector_norm(torch.cat([r.reshape(-1) for r in ref])))
    assert torch.allclose(total, total_ref)
    assert abs(float(torch.linalg.vector_norm(b.flat)) - 5.0) < 1e-4
    b.zero()
    assert float(b.flat.abs().sum()) == 0.0


def test_bucket_survives_model_zero_grad():
    """model.zero_grad() sets p.grad = None on torch 2.x; the bucket must fold the fresh gradient
    tensors back in, so that the norm, the clip and the optimizer see them and the weights move."""
    sys.path.insert(0, ROOT)
    from graph_representation_learning_b200.trainer import KVTrainStep
    torch.manual_seed(0)
    model = TinyDenseModel(5, 3)
    step = KVTrainStep(model, max_grad_norm=5.0, device=torch.device("cpu"))
    before = {k: v.clone() for k, v in model.state_dict().items()}
    model.zero_grad()                                                        # what kv_procedure.py:349 does
    assert all(p.grad is None for p in step.bucket.params)
    batch = _batch(3, 4)
    loss, _ = step.step_process(batch)
    loss.backward()
    total = step.clip_and_step()
    assert float(total) > 0 and float(step.bucket.flat.abs().sum()) > 0
    assert all(p.grad.data_ptr() >= step.bucket.flat.data_ptr() for p in step.bucket.params)
    moved = [k for k, v in model.state_dict().items() if not torch.equal(v, before[k])]
    assert "a.weight" in moved and "b.bias" in moved and "frozen.weight" not in moved
