"""2-GPU NCCL test of the data-parallel train step with the CUDA model (run with `gpurun --gpus 2`; skipped on one device):
rank-identical parameters after the step, and summed gradients == the single-GPU gradients on the concatenated RAGGED
batch (different numbers of valid, non -100, nodes per rank: cross_entropy_loss.py:33-35 is a mean over the valid nodes of
the whole batch).  Dropout / DropEdge probabilities are set to 0 so that both runs draw nothing."""
import os
import socket
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _batch(B, N, in_dim, out_dim, L=6):
    g = torch.Generator().manual_seed(77)
    V = torch.rand(B, N, in_dim, generator=g)
    A = (torch.rand(B, N, L, N, generator=g) < 6.0 / N).half()
    y = torch.randint(0, out_dim, (B, N), generator=g)
    for b in range(B):                               # ragged: graph b has (7 b) % 23 padded nodes on each side
        pad = (7 * b) % 23
        if pad:
            y[b, :pad] = -100; y[b, -pad:] = -100
            V[b, :pad] = 0; V[b, -pad:] = 0
            A[b, :pad] = 0; A[b, -pad:] = 0; A[b, :, :, :pad] = 0; A[b, :, :, -pad:] = 0
    return V, A, y


def _model(dev, in_dim, out_dim, S):
    from graph_representation_learning_b200 import GraphCNNDropEdge
    torch.manual_seed(3)
    m = GraphCNNDropEdge(in_dim, out_dim, 6, net_size=S).to(dev)
    m.dropout.p = 0.0
    m.edge_dropout.p = 0.0
    return m


CFG = dict(B=8, N=96, in_dim=200, out_dim=11, S=256)


def _worker(rank, world, port, out_dir):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from graph_representation_learning_b200.trainer import KVTrainStep
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dev = torch.device("cuda", rank)
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    model = _model(dev, CFG["in_dim"], CFG["out_dim"], CFG["S"])
    step = KVTrainStep(model, max_grad_norm=5.0)
    V, A, y = _batch(CFG["B"], CFG["N"], CFG["in_dim"], CFG["out_dim"])
    per = CFG["B"] // world
    sl = slice(rank * per, (rank + 1) * per)
    loss = step.run_train_step({"textline_encoding": V[sl], "adjacency_matrix": A[sl], "node_label": y[sl]})
    assert step.bucket._hooks and step.bucket._tail_work is None      # overlapped tail bucket was launched and joined
    torch.cuda.synchronize()
    torch.save({"grad": step.bucket.flat.cpu(), "loss": loss.cpu(),
                "sd": {k: v.cpu() for k, v in model.state_dict().items()}}, os.path.join(out_dir, f"r{rank}.pt"))
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


@pytest.mark.timeout(600)
def test_two_gpu_nccl_step_equals_single_gpu_on_concatenated_ragged_batch(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs (gpurun --gpus 2)")
    import torch.multiprocessing as mp
    from graph_representation_learning_b200.trainer import KVTrainStep
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    r0, r1 = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(r0["grad"], r1["grad"])
    for k in r0["sd"]:
        assert torch.equal(r0["sd"][k], r1["sd"][k]), k               # replicas stay bit-identical

    dev = torch.device("cuda:0")
    model = _model(dev, CFG["in_dim"], CFG["out_dim"], CFG["S"])
    single = KVTrainStep(model, max_grad_norm=5.0)
    V, A, y = _batch(CFG["B"], CFG["N"], CFG["in_dim"], CFG["out_dim"])
    n_valid = [(y[r * 4:(r + 1) * 4] != -100).sum().item() for r in range(world)]
    assert n_valid[0] != n_valid[1]
    loss = single.run_train_step({"textline_encoding": V, "adjacency_matrix": A, "node_label": y})
    torch.cuda.synchronize()

    def rel(a, b):
        return float((a.double() - b.double()).norm() / b.double().norm())
    assert abs(float(r0["loss"] + r1["loss"]) - float(loss)) < 1e-5 * abs(float(loss))
    assert rel(r0["grad"], single.bucket.flat.cpu()) < 1e-5
    for k, v in model.state_dict().items():
        assert rel(r0["sd"][k], v.cpu()) < 1e-5, k
