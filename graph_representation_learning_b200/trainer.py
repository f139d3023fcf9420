"""Train / validation step of the KV node-classification procedure on the B200 path.

Mirrors ``KVProcedure._step_process`` / ``_run_train_step`` / ``_run_val_step``
(gnn/trainer/training_procedures/kv_procedure.py:106-178): H2D of the batch, ``model.forward([V, A])``,
``CrossEntropyLoss`` over ``pred.transpose(1, 2)`` (gnn/trainer/losses/cross_entropy_loss.py:33-35, ignore_index
-100 = the collate's label padding), backward, ``clip_grad_norm_(max_grad_norm)``, ``optimizer.step()``.
The optimizer is built like ``BuitlinOptimizer`` does (gnn/trainer/optimizers/builtin_optimizer.py:18-26):
``getattr(torch.optim, type)(params, **kwargs)`` with the configured ``lr`` swallowed, i.e. Adam starts at 1e-3.

Data parallelism (new; the reference's DDP wrapper is never applied -- SURVEY.md 0.6): one process per GPU,
gradients of the replicated 12-40 MB model are averaged with ONE NCCL all-reduce over a flat bucket per step.
Per-document graphs shard with no halo, so this is the only collective on the path.
"""
from __future__ import annotations

from typing import Any, Dict, Optional, Tuple

import torch
import torch.distributed as dist
from torch import nn


class CrossEntropyLoss(nn.Module):
    """gnn/trainer/losses/cross_entropy_loss.py:9-35"""

    def __init__(self, weight=None):
        super().__init__()
        w = None if weight is None else torch.as_tensor(weight, dtype=torch.float32)
        self.criterion = nn.CrossEntropyLoss(w)

    def forward(self, pred: torch.Tensor, target: torch.Tensor, **kwargs) -> torch.Tensor:
        return self.criterion(pred.transpose(1, 2), target)


class FlatGradBucket:
    """All trainable gradients live in one flat fp32 buffer (``p.grad`` are views), so the data-parallel
    reduction is a single in-place NCCL all-reduce and ``clip_grad_norm_`` is one norm over one tensor."""

    def __init__(self, model: nn.Module):
        self.params = [p for p in model.parameters() if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(n, dtype=torch.float32, device=self.params[0].device)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def zero(self) -> None:
        self.flat.zero_()

    def all_reduce_mean(self) -> None:
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
            self.flat.div_(dist.get_world_size())

    def clip_(self, max_norm: float) -> torch.Tensor:
        total = torch.linalg.vector_norm(self.flat)
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        self.flat.mul_(coef)
        return total


class KVTrainStep:
    def __init__(self, model: nn.Module, max_grad_norm: float = 5.0, optimizer_type: str = "Adam",
                 optimizer_kwargs: Optional[Dict[str, Any]] = None, device: Optional[torch.device] = None):
        self.model = model
        self.device = device or next(model.parameters()).device
        self.criterion = CrossEntropyLoss()
        self.max_grad_norm = max_grad_norm
        self.bucket = FlatGradBucket(model)
        kwargs = dict(optimizer_kwargs or {})
        kwargs.pop("lr", None)  # the reference never forwards lr (builtin_optimizer.py:10,18-23)
        if optimizer_type == "Adam" and self.device.type == "cuda":
            kwargs.setdefault("fused", True)
        self.optimizer = getattr(torch.optim, optimizer_type)(self.bucket.params, **kwargs)

    def _to_device(self, batch_sample: Dict[str, torch.Tensor]) -> Tuple[list, torch.Tensor]:
        V = batch_sample["textline_encoding"].to(self.device, non_blocking=True)
        A = batch_sample["adjacency_matrix"]
        if isinstance(A, torch.Tensor):
            A = A.to(self.device, non_blocking=True)   # fp16 stays fp16: the CSR pack reads it directly
        y = batch_sample["node_label"].to(self.device, non_blocking=True)
        return [V, A], y

    def step_process(self, batch_sample: Dict[str, torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
        inputs, y = self._to_device(batch_sample)
        pred = self.model.forward(inputs)
        return self.criterion(pred, y), pred

    def run_train_step(self, batch_sample: Dict[str, torch.Tensor]) -> torch.Tensor:
        """returns the (device) loss; no host sync here -- callers read it when they need it"""
        self.model.train()
        self.bucket.zero()                       # optimizer.zero_grad() with grads kept as bucket views
        with torch.set_grad_enabled(True):
            loss, _ = self.step_process(batch_sample)
            loss.backward()
            self.bucket.all_reduce_mean()
            self.bucket.clip_(self.max_grad_norm)
            self.optimizer.step()
        return loss.detach()

    @torch.no_grad()
    def run_val_step(self, batch_sample: Dict[str, torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
        self.model.eval()
        loss, pred = self.step_process(batch_sample)
        return loss, pred.argmax(dim=-1)
