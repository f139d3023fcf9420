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

from .features import SparseNodeFeatures


_ws_cache: Dict[Any, torch.Tensor] = {}


def bind_process_to_gpu_numa(device_index: int) -> Optional[list]:
    """Pin this process to the CPU cores NVML reports as local to GPU ``device_index`` (its NUMA node), so that the
    pinned host buffers allocated afterwards (first touch) and the H2D copies of the data loader stay on the socket
    the GPU hangs off.  One process per GPU on an 8-GPU box otherwise shares one node's memory controllers and PCIe
    root: measured 98 ms per 2.34 GB batch at 8 ranks against 40 ms at 1.  Call before allocating pinned memory.
    Returns the CPU list, or None when NVML / the cpuset does not allow it (nothing is changed then)."""
    import os
    try:
        import pynvml
        pynvml.nvmlInit()
        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        handle = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid) if not uuid.startswith("GPU-") else uuid)
        n_cpu = os.cpu_count() or 1
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, (n_cpu + 63) // 64)
        local = {c for c in range(n_cpu) if (int(mask[c // 64]) >> (c % 64)) & 1}
        cpus = sorted(local & os.sched_getaffinity(0))
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return cpus
    except Exception:
        return None


def _reduce_workspace(device: torch.device) -> torch.Tensor:
    """scratch for the block partial sums of the loss / norm reductions (one per device and stream)"""
    from . import _lib
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    ws = _ws_cache.get(key)
    if ws is None:
        ws = torch.empty(_lib.load().grl_reduce_workspace_bytes(), dtype=torch.uint8, device=device)
        _ws_cache[key] = ws
    return ws


def _dp_world() -> int:
    return dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1


class _CrossEntropyFn(torch.autograd.Function):
    """mean (weighted) cross entropy over the rows whose label is not ignore_index -- csrc/train_ops.cu.

    ``global_mean`` (data parallel): the denominator is the all-reduced weight sum, so the returned value is this
    rank's SHARE of the mean over the global batch's valid nodes (cross_entropy_loss.py:33-35 applied to the
    concatenated batch): shares add up to the global loss and gradients add up to the single-process gradient, whatever
    the number of padded (-100) nodes per rank.  One 4-byte all-reduce between forward and backward."""

    @staticmethod
    def forward(ctx, logits2d, target1d, weight, ignore_index, global_mean, stats):
        from . import _lib
        M, C = logits2d.shape
        dev = logits2d.device
        lse = torch.empty(M, dtype=torch.float32, device=dev)
        out = torch.empty(4, dtype=torch.float32, device=dev)
        _lib.check(_lib.load().grl_ce_fwd(logits2d.data_ptr(), logits2d.stride(0), target1d.data_ptr(), _lib.ptr(weight),
                                          int(ignore_index), M, C, lse.data_ptr(), out.data_ptr(),
                                          _reduce_workspace(dev).data_ptr(), _lib.stream_ptr(dev)), "grl_ce_fwd")
        _lib.count_launch(2)
        stats["bad_labels"] = out[2]
        loss = out[0]
        denom = out
        if global_mean and _dp_world() > 1:
            denom = out.clone()
            dist.all_reduce(denom[1:2], op=dist.ReduceOp.SUM)            # weight sum of the global batch
            loss = torch.where(out[2] > 0, out[0], out[3] / denom[1].clamp_min(1e-30))
        ctx.save_for_backward(logits2d, target1d, weight, lse, denom)
        ctx.ignore_index = int(ignore_index)
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        from . import _lib
        logits2d, target1d, weight, lse, out = ctx.saved_tensors
        M, C = logits2d.shape
        d = torch.empty(M, C, dtype=torch.float32, device=logits2d.device)
        g = grad_out.detach().reshape(1).float().contiguous()
        _lib.check(_lib.load().grl_ce_bwd(logits2d.data_ptr(), logits2d.stride(0), target1d.data_ptr(), _lib.ptr(weight),
                                          ctx.ignore_index, M, C, lse.data_ptr(), out.data_ptr(), g.data_ptr(),
                                          d.data_ptr(), C, _lib.stream_ptr(d.device)), "grl_ce_bwd")
        _lib.count_launch(1)
        return d, None, None, None, None, None


class CrossEntropyLoss(nn.Module):
    """gnn/trainer/losses/cross_entropy_loss.py:9-35.  CUDA fp32 logits run the kernels of csrc/train_ops.cu; anything
    else (the CPU tensors of the host-logic tests) goes through nn.CrossEntropyLoss like the reference.

    ``global_mean`` is switched on by ``KVTrainStep`` under torch.distributed: the mean then runs over the valid nodes
    of the GLOBAL batch and the gradient all-reduce is a plain sum (see ``_CrossEntropyFn``).
    A label outside [0, C) that is not the ignore index makes the loss NaN and is reported by ``check_labels()`` (the
    reference's nn.CrossEntropyLoss raises on the spot; raising here would cost a host sync per step)."""

    def __init__(self, weight=None):
        super().__init__()
        w = None if weight is None else torch.as_tensor(weight, dtype=torch.float32)
        self.criterion = nn.CrossEntropyLoss(w)
        self.global_mean = False
        self._stats: Dict[str, Any] = {}

    def check_labels(self) -> None:
        """host read of the last forward's out-of-range label count; raises like nn.CrossEntropyLoss would have"""
        bad = self._stats.get("bad_labels")
        if bad is not None and float(bad) > 0:
            raise ValueError(f"{int(float(bad))} target labels are outside [0, n_classes) and not ignore_index="
                             f"{self.criterion.ignore_index}")

    def forward(self, pred: torch.Tensor, target: torch.Tensor, **kwargs) -> torch.Tensor:
        if pred.is_cuda and pred.dtype == torch.float32 and pred.dim() >= 2 and target.dtype == torch.int64:
            C = pred.shape[-1]
            logits2d = pred.reshape(-1, C)
            if logits2d.stride(1) != 1:
                logits2d = logits2d.contiguous()
            w = self.criterion.weight
            if w is not None:
                w = w.to(device=pred.device, dtype=torch.float32).contiguous()
            return _CrossEntropyFn.apply(logits2d, target.reshape(-1).contiguous(), w, self.criterion.ignore_index,
                                         self.global_mean, self._stats)
        if self.global_mean and _dp_world() > 1:
            # host-logic tests (gloo, CPU tensors): the same global mean, spelled with torch ops
            w = self.criterion.weight
            ls = nn.functional.cross_entropy(pred.transpose(1, 2), target, weight=w, reduction="sum",
                                             ignore_index=self.criterion.ignore_index)
            valid = target != self.criterion.ignore_index
            wsum = (w.to(pred.device)[target[valid]].sum() if w is not None else valid.sum()).to(ls.dtype).reshape(1)
            dist.all_reduce(wsum, op=dist.ReduceOp.SUM)
            return ls / wsum[0]
        return self.criterion(pred.transpose(1, 2), target)


class FlatGradBucket:
    """All trainable gradients live in one flat fp32 buffer (``p.grad`` are views), so ``clip_grad_norm_`` is one norm
    over one tensor and the data-parallel reduction is NCCL all-reduces over contiguous slices of it.

    Data parallel ("bucketed in reverse layer order, overlapped with backward"): the buffer is cut in two
    at the end of the FIRST parameterised sub-module (``emb1``: its weight gradient [S, 4369] is the last and largest
    GEMM of the backward pass).  Post-accumulate-grad hooks count the tail's gradients in; when the last one has landed
    the tail slice is all-reduced asynchronously (NCCL's own stream) while the compute stream still runs emb1's
    weight-gradient GEMM; ``all_reduce()`` then reduces the head slice and joins both."""

    def __init__(self, model: nn.Module):
        self.params = [p for p in model.parameters() if p.requires_grad]
        n = sum(p.numel() for p in self.params)
        self.flat = torch.zeros(n, dtype=torch.float32, device=self.params[0].device)
        off = 0
        for p in self.params:
            p.grad = self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()
        # head = trainable parameters of the first child module that has any (their gradients arrive last)
        self.n_head_params, self.head_numel = 0, 0
        for child in model.children():
            own = [p for p in child.parameters() if p.requires_grad]
            if own:
                if len(self.params) > len(own) and all(a is b for a, b in zip(own, self.params)):
                    self.n_head_params, self.head_numel = len(own), sum(p.numel() for p in own)
                break
        self._hooks: list = []
        self._pending = 0
        self._tail_work = None
        self._armed = False

    def _views(self, lo: int = 0, hi: Optional[int] = None):
        off = 0
        for k, p in enumerate(self.params):
            if k >= lo and (hi is None or k < hi):
                yield p, self.flat[off:off + p.numel()].view_as(p)
            off += p.numel()

    def reattach(self, lo: int = 0, hi: Optional[int] = None) -> None:
        """Fold gradients that are no longer views of the bucket back in: ``model.zero_grad()`` /
        ``optimizer.zero_grad()`` (set_to_none=True on torch 2.x) leave ``p.grad = None`` and autograd then allocates
        fresh tensors.  None counts as zero.  Cheap when nothing strayed (pointer compares only)."""
        for p, view in self._views(lo, hi):
            g = p.grad
            if g is None:
                view.zero_()
                p.grad = view
            elif g.data_ptr() != view.data_ptr():
                view.copy_(g)
                p.grad = view

    def zero(self) -> None:
        self.flat.zero_()
        for p, view in self._views():
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view

    # ------------------------------------------------------------------ data-parallel reduction
    def arm_overlap(self) -> None:
        """call before ``backward()``: the tail slice is reduced as soon as its last gradient has been accumulated"""
        if _dp_world() <= 1 or self.n_head_params == 0:
            return
        if not self._hooks:
            for p in self.params[self.n_head_params:]:
                self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        self._pending = len(self.params) - self.n_head_params
        self._tail_work = None
        self._armed = True

    def _on_grad(self, _p) -> None:
        if not self._armed:
            return
        self._pending -= 1
        if self._pending == 0:
            self._launch_tail()

    def _launch_tail(self) -> None:
        self.reattach(self.n_head_params, None)
        self._tail_work = dist.all_reduce(self.flat[self.head_numel:], op=dist.ReduceOp.SUM, async_op=True)

    def all_reduce(self, average: bool = True) -> None:
        """sum (``average=False``: the loss was already normalised by the GLOBAL valid-node count) or mean of the
        gradients over the ranks; joins the overlapped tail reduction when ``arm_overlap()`` started one"""
        self.reattach()
        world = _dp_world()
        if world <= 1:
            return
        if self._armed:
            self._armed = False
            if self._tail_work is None:          # some tail parameter got no gradient this step: nothing was launched
                self._launch_tail()
            head = dist.all_reduce(self.flat[:self.head_numel], op=dist.ReduceOp.SUM, async_op=True)
            self._tail_work.wait()
            head.wait()
            self._tail_work = None
        else:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM)
        if average:
            self.flat.div_(world)

    def all_reduce_mean(self) -> None:
        self.all_reduce(average=True)

    def norm(self) -> torch.Tensor:
        """total 2-norm of the gradients as a device scalar (deterministic two-stage reduction on CUDA)"""
        self.reattach()
        if self.flat.is_cuda:
            from . import _lib
            out = torch.empty(1, dtype=torch.float32, device=self.flat.device)
            _lib.check(_lib.load().grl_l2_norm(self.flat.data_ptr(), self.flat.numel(), out.data_ptr(),
                                               _reduce_workspace(self.flat.device).data_ptr(),
                                               _lib.stream_ptr(self.flat.device)), "grl_l2_norm")
            _lib.count_launch(2)
            return out[0]
        return torch.linalg.vector_norm(self.flat)

    def clip_(self, max_norm: float) -> torch.Tensor:
        """clip_grad_norm_ (kv_procedure.py:161): scale by min(1, max_norm / (norm + 1e-6)); returns the norm"""
        total = self.norm()
        coef = torch.clamp(max_norm / (total + 1e-6), max=1.0)
        self.flat.mul_(coef)
        return total


class FlatAdam(torch.optim.Optimizer):
    """torch.optim.Adam (amsgrad=False) over ONE flat parameter / moment buffer: a single kernel per step
    (csrc/train_ops.cu grl_adam_step), optionally with the clip_grad_norm_ scaling folded in from a device-resident
    norm.  ``param_groups`` / ``state_dict()`` keep torch.optim.Adam's layout (per-parameter ``step`` / ``exp_avg`` /
    ``exp_avg_sq``, here views of the flat buffers), so schedulers that poke ``param_groups[*]["lr"]`` and the
    checkpoint handler work unchanged.  Parameters are re-pointed at views of the flat buffer."""

    def __init__(self, params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 amsgrad: bool = False, flat_grad: Optional[torch.Tensor] = None, **ignored: Any):
        if amsgrad:
            raise ValueError("FlatAdam implements amsgrad=False only")
        # frozen parameters (w_rand / rp_emb / rp_final: requires_grad=False) never get a gradient; torch.optim.Adam skips
        # them at every step, so they are left out of the flat buffers (weight decay must not touch them)
        params = list(params)
        if params and isinstance(params[0], dict):
            params = [dict(g, params=[p for p in g["params"] if p.requires_grad]) for g in params]
            params = [g for g in params if g["params"]]
        else:
            params = [p for p in params if p.requires_grad]
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self._params = [p for g in self.param_groups for p in g["params"]]
        if not self._params or not all(p.is_cuda and p.dtype == torch.float32 for p in self._params):
            raise ValueError("FlatAdam needs fp32 CUDA parameters")
        dev = self._params[0].device
        self._sizes = [p.numel() for p in self._params]
        n = sum(self._sizes)
        self.flat_p = torch.empty(n, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(n, dtype=torch.float32, device=dev)
        self.flat_g = flat_grad if flat_grad is not None else torch.zeros(n, dtype=torch.float32, device=dev)
        self._own_grad = flat_grad is None
        self._step = 0
        self._adopt_params()

    def _views(self, flat):
        off = 0
        for p, k in zip(self._params, self._sizes):
            yield p, flat[off:off + k].view_as(p)
            off += k

    def _adopt_params(self) -> None:
        with torch.no_grad():
            for p, view in self._views(self.flat_p):
                if p.data.data_ptr() != view.data_ptr():
                    view.copy_(p.data)
                    p.data = view

    def _publish_state(self) -> None:
        step = torch.tensor(float(self._step))
        for (p, m), (_, v) in zip(self._views(self.flat_m), self._views(self.flat_v)):
            self.state[p] = {"step": step.clone(), "exp_avg": m, "exp_avg_sq": v}

    @torch.no_grad()
    def step(self, closure=None, grad_norm: Optional[torch.Tensor] = None, max_norm: float = 0.0):
        from . import _lib
        loss = closure() if closure is not None else None
        self._adopt_params()                     # a .to() / load that re-allocated a parameter is folded back in
        skipped = set()
        if self._own_grad:                       # gradients are not views of one bucket: gather them
            for p, view in self._views(self.flat_g):
                if p.grad is None:               # torch.optim.Adam skips these (no moment update, no decay)
                    skipped.add(id(p))
                elif p.grad.data_ptr() != view.data_ptr():
                    view.copy_(p.grad)
        self._step += 1
        lib, off = _lib.load(), 0
        for g in self.param_groups:
            b1, b2 = g["betas"]
            # maximal runs of parameters that have a gradient: one launch per run (one per group when none is skipped)
            runs, start, pos = [], None, off
            for p in g["params"]:
                if id(p) in skipped:
                    if start is not None:
                        runs.append((start, pos))
                        start = None
                elif start is None:
                    start = pos
                pos += p.numel()
            if start is not None:
                runs.append((start, pos))
            for lo, hi in runs:
                sl = slice(lo, hi)
                _lib.check(lib.grl_adam_step(self.flat_p[sl].data_ptr(), self.flat_g[sl].data_ptr(),
                                             self.flat_m[sl].data_ptr(), self.flat_v[sl].data_ptr(), hi - lo,
                                             float(g["lr"]), float(b1), float(b2), float(g["eps"]),
                                             float(g["weight_decay"]), self._step, _lib.ptr(grad_norm), float(max_norm),
                                             _lib.stream_ptr(self.flat_p.device)), "grl_adam_step")
                _lib.count_launch(1)
            off = pos
        if not self.state:
            self._publish_state()
        else:
            for st in self.state.values():
                st["step"] += 1
        return loss

    def load_state_dict(self, state_dict) -> None:
        super().load_state_dict(state_dict)
        steps = [int(st["step"]) for st in self.state.values() if "step" in st]
        self._step = max(steps) if steps else 0
        with torch.no_grad():
            for (p, m), (_, v) in zip(self._views(self.flat_m), self._views(self.flat_v)):
                st = self.state.get(p)
                if st:
                    m.copy_(st["exp_avg"])
                    v.copy_(st["exp_avg_sq"])
        if self.state:
            self._publish_state()


class KVTrainStep:
    def __init__(self, model: nn.Module, max_grad_norm: float = 5.0, optimizer_type: str = "Adam",
                 optimizer_kwargs: Optional[Dict[str, Any]] = None, device: Optional[torch.device] = None,
                 criterion: Optional[nn.Module] = None, optimizer_factory=None):
        """``optimizer_factory(params) -> torch.optim.Optimizer`` (e.g. ``BuitlinOptimizer.get_optimizer``) overrides
        ``optimizer_type`` / ``optimizer_kwargs``; ``criterion(pred, target)`` defaults to the CE of the KV task."""
        self.model = model
        self.device = device or next(model.parameters()).device
        self.criterion = criterion if criterion is not None else CrossEntropyLoss()
        if isinstance(self.criterion, nn.Module):
            self.criterion = self.criterion.to(self.device)
        self.max_grad_norm = max_grad_norm
        self.bucket = FlatGradBucket(model)
        if optimizer_factory is not None:
            self.optimizer = optimizer_factory(self.bucket.params)
        else:
            kwargs = dict(optimizer_kwargs or {})
            kwargs.pop("lr", None)  # the reference never forwards lr (builtin_optimizer.py:10,18-23)
            if optimizer_type == "Adam" and self.device.type == "cuda" and not kwargs.get("amsgrad"):
                kwargs.pop("fused", None)
                self.optimizer = FlatAdam(self.bucket.params, flat_grad=self.bucket.flat, **kwargs)
            else:
                self.optimizer = getattr(torch.optim, optimizer_type)(self.bucket.params, **kwargs)
        if isinstance(self.optimizer, FlatAdam) and self.optimizer._own_grad:
            # a factory-built FlatAdam: let it read the bucket directly instead of gathering p.grad every step
            self.optimizer.flat_g, self.optimizer._own_grad = self.bucket.flat, False
        # data parallel: CE normalised by the valid-node count of the GLOBAL batch -> gradients are summed, not averaged
        # (ragged batches: N-GPU gradients == 1-GPU gradients on the concatenated batch, SURVEY.md 8e "Check")
        self.sum_gradients = isinstance(self.criterion, CrossEntropyLoss)
        if self.sum_gradients:
            self.criterion.global_mean = True

    def _to_device(self, batch_sample: Dict[str, torch.Tensor]) -> Tuple[list, torch.Tensor]:
        if "adjacency_matrix" not in batch_sample and "edges" in batch_sample:      # EdgeListCollate host batch
            from .collate import EdgeListCollate
            batch_sample = EdgeListCollate.to_device(batch_sample, self.device)
        V = batch_sample["textline_encoding"]
        if not isinstance(V, SparseNodeFeatures) and not V.is_floating_point():
            V = V.float()                               # kv_procedure.py:117 (.float() of whatever the dataset gave)
        V = V.to(self.device, non_blocking=True)
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
            self.bucket.arm_overlap()
            loss.backward()
            self.clip_and_step()
        return loss.detach()

    def clip_and_step(self) -> torch.Tensor:
        """gradient all-reduce (N > 1) -> clip_grad_norm_ -> optimizer.step (kv_procedure.py:161-162); returns the
        gradient norm as a device scalar.  With FlatAdam the clip scaling is folded into the step kernel."""
        if self.device.type == "cuda":
            from .graph import check_dense_overflow
            check_dense_overflow()               # a truncated sync-free dense pack must not reach the weights
        self.bucket.all_reduce(average=not self.sum_gradients)
        if isinstance(self.optimizer, FlatAdam):
            total = self.bucket.norm()
            self.optimizer.step(grad_norm=total.reshape(1), max_norm=float(self.max_grad_norm))
        else:
            total = self.bucket.clip_(self.max_grad_norm)
            self.optimizer.step()
        return total

    @torch.no_grad()
    def run_val_step(self, batch_sample: Dict[str, torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
        self.model.eval()
        loss, pred = self.step_process(batch_sample)
        return loss, pred.argmax(dim=-1)


class GraphedTrainStep:
    """CUDA-graph replay of the forward + loss + backward of ``KVProcedure._run_train_step``
    (gnn/trainer/training_procedures/kv_procedure.py:143-164) for one fixed batch signature ``(B, N)`` (SURVEY.md 7.1
    step 10).

    Small batches make the train step launch-bound (cfg1, 8 x 200 nodes: ~170 kernels of a few microseconds each).  The
    path -- CSR / CSC / typed-transpose pack of the dense adjacency, model forward, cross entropy, the whole backward into
    the flat gradient bucket -- is stream-ordered, allocation-stable and free of host reads once the CSR capacity is
    fixed, so it is captured ONCE; ``run_train_step`` copies the batch into static buffers, bumps ONE device counter and
    replays.  That counter is the Philox stream offset every DropEdge / dropout kernel adds to its stream id
    (``grl_rng_set_offset_ptr``: the pointer is baked into the captured kernel nodes), so each replay draws fresh masks --
    the same masks the eager step draws at the same step index (``tests/test_gpu_procedures.py``).  ``clip_grad_norm_``,
    the optimizer step and (N > 1) the gradient all-reduce stay eager behind the graph: their scalars (Adam bias
    corrections, learning rate) change per step.  Single process only: the global valid-node count of the
    data-parallel cross entropy is a collective inside the forward."""

    def __init__(self, step: KVTrainStep, B: int, N: int, input_dim: int, num_edges: int,
                 adj_dtype: torch.dtype = torch.float16, nnz_capacity: Optional[int] = None):
        from . import _lib
        from .graph import PackedGraphBatch
        if _dp_world() > 1:
            raise RuntimeError("GraphedTrainStep captures a single-process step; run KVTrainStep.run_train_step under torch.distributed")
        dev = step.device
        if dev.type != "cuda":
            raise RuntimeError("GraphedTrainStep needs the model on a CUDA device (no CPU fallback)")
        self.step, self.B, self.N, self.L = step, B, N, num_edges
        self.capacity = int(nnz_capacity) if nnz_capacity is not None else B * N * (num_edges * N + 1)   # dense worst case
        ld = (input_dim + 3) // 4 * 4                              # 16-byte row pitch: TMA reads V in place
        self._Vp = torch.zeros(B, N, ld, device=dev)
        self.V = self._Vp[:, :, :input_dim]
        self.A = torch.zeros(B, N, num_edges, N, dtype=adj_dtype, device=dev)
        self.y = torch.zeros(B, N, dtype=torch.int64, device=dev)
        self.rng_offset = torch.zeros(1, dtype=torch.int64, device=dev)    # read as uint64 by the kernels
        model = step.model.train()

        def fwd_bwd():
            step.bucket.zero()
            g = PackedGraphBatch.from_dense(self.A, nnz_capacity=self.capacity, with_transpose=True)
            loss = step.criterion(model.forward([self.V, g]), self.y)
            loss.backward()
            return loss.detach()

        lib = _lib.load()
        with torch.enable_grad():
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):                          # warm-up: library load, function attributes, pools
                for _ in range(2):
                    fwd_bwd()
            torch.cuda.current_stream(dev).wait_stream(side)
            self.graph = torch.cuda.CUDAGraph()
            _lib.check(lib.grl_rng_set_offset_ptr(self.rng_offset.data_ptr()), "grl_rng_set_offset_ptr")
            try:
                with torch.cuda.graph(self.graph):
                    self.loss = fwd_bwd()
            finally:
                lib.grl_rng_set_offset_ptr(0)
        step.bucket.zero()
        self.replays = 0

    def run_train_step(self, batch_sample: Dict[str, torch.Tensor]) -> torch.Tensor:
        """batch_sample: {"textline_encoding" [B,N,D], "adjacency_matrix" dense [B,N,L,N], "node_label" [B,N]} (host or
        device).  Returns the static device loss of this step (no host sync here)."""
        self.V.copy_(batch_sample["textline_encoding"], non_blocking=True)
        self.A.copy_(batch_sample["adjacency_matrix"], non_blocking=True)
        self.y.copy_(batch_sample["node_label"], non_blocking=True)
        self.rng_offset.add_(16)                                   # one forward's worth of Philox stream ids
        self.graph.replay()
        self.replays += 1
        self.step.clip_and_step()
        return self.loss


# ----------------------------------------------------------------------------------------------------------
# getattr-by-name factories of the reference trainer (losses / optimizers / lr_schedulers), restated
# ----------------------------------------------------------------------------------------------------------
class BuitlinOptimizer:
    """gnn/trainer/optimizers/builtin_optimizer.py:8-26 (the class name is the reference's spelling).  ``lr`` is kept
    on the object and NOT forwarded to torch.optim, exactly like the reference: Adam starts at its default 1e-3 and
    the scheduler overwrites ``param_groups[*]["lr"]`` at the end of each epoch."""

    def __init__(self, type_optimizer: str, lr: float, **kwargs: Any):
        self.learning_rate = lr
        self.type_optimizer = type_optimizer
        self.optimizer_args = kwargs

    @classmethod
    def _from_config(cls, opt: Dict[str, Any]) -> "BuitlinOptimizer":
        return cls(**opt)

    def get_optimizer(self, parameters) -> torch.optim.Optimizer:
        parameters = list(parameters)
        args = dict(self.optimizer_args)
        if (self.type_optimizer == "Adam" and parameters and isinstance(parameters[0], torch.Tensor)
                and all(p.is_cuda and p.dtype == torch.float32 for p in parameters) and not args.get("amsgrad")):
            for k in ("fused", "foreach", "capturable", "differentiable", "maximize"):
                args.pop(k, None)
            optimizer = FlatAdam(parameters, **args)         # same update rule, one kernel over a flat buffer
        else:
            optimizer = getattr(torch.optim, self.type_optimizer)(parameters, **args)
        if hasattr(self.learning_rate, "prepare"):
            self.learning_rate.prepare(optimizer)
        return optimizer


class DecayLearningRate:
    """gnn/trainer/lr_schedulers/decay_lr.py:6-25: lr * (1 - epoch / (num_epochs + 1)) ** factor, set per epoch."""

    def __init__(self, lr: float = 0.002, factor: float = 0.9, num_epochs: int = 100):
        self.lr = lr
        self.initial_lr = lr
        self.factor = factor
        self.epochs = num_epochs

    @classmethod
    def _from_config(cls, config: Dict[str, Any]) -> "DecayLearningRate":
        return cls(**config)

    def _step_lr(self, epoch: int, step: Optional[int] = None) -> float:
        rate = (1.0 - epoch / float(self.epochs + 1)) ** self.factor
        self.lr = self.initial_lr * rate
        return self.lr


LOSSES = {"CrossEntropyLoss": CrossEntropyLoss}
OPTIMIZERS = {"BuitlinOptimizer": BuitlinOptimizer}
LR_SCHEDULERS = {"DecayLearningRate": DecayLearningRate}


def from_registry(registry: Dict[str, Any], kind: str, type_name: Optional[str], args: Optional[Dict[str, Any]]):
    """``getattr(<package>, cfg.type)._from_config(cfg.args)`` of base_procedure.py:95-140; None type -> None"""
    if not type_name:
        return None
    if type_name not in registry:
        raise ValueError(f"unknown {kind} {type_name!r}; available: {sorted(registry)}")
    cls = registry[type_name]
    args = dict(args or {})
    return cls._from_config(args) if hasattr(cls, "_from_config") else cls(**args)


class Dictlist(dict):
    """gnn/utils/metric_tracker.py:6-29: running lists per key, averages rounded to 3 decimals."""

    def _update(self, dict_item: Dict[str, Any]) -> None:
        for key, value in dict_item.items():
            self.setdefault(key, []).append(value)

    def _avg(self, key: str) -> float:
        vals = self[key]
        return round(sum(vals) / len(vals), 3)

    def _result(self) -> Dict[str, Any]:
        return {key: self._avg(key) for key in self.keys()}


def macro_scores(pred: torch.Tensor, target: torch.Tensor, ignore: Tuple[Any, ...] = (-100,),
                 item_name: str = "Node classification") -> Dict[str, float]:
    """Macro precision / recall / F1 / support of ``classification_report(..., zero_division=0)["macro avg"]`` as
    KVProcedure._get_metric_scores computes it (kv_procedure.py:61-104): entries whose label is in ``ignore`` are
    dropped, the average runs over the classes present in labels OR predictions.  One confusion matrix on the device
    and a single small D2H copy instead of sklearn over host lists."""
    pred, target = pred.reshape(-1), target.reshape(-1)
    keep = torch.ones_like(target, dtype=torch.bool)
    for v in ignore:
        if v is not None:
            keep &= target != v
    pred, target = pred[keep], target[keep]
    names = ("precision", "recall", "f1-score", "support")
    if target.numel() == 0:
        return {f"{item_name}_{k}": 0.0 for k in names}
    C = int(max(int(pred.max()), int(target.max()))) + 1
    cm = torch.bincount(target * C + pred, minlength=C * C).view(C, C).double()
    tp, n_true, n_pred = cm.diagonal(), cm.sum(1), cm.sum(0)
    present = (n_true + n_pred) > 0
    prec = torch.where(n_pred > 0, tp / n_pred.clamp_min(1), torch.zeros_like(tp))
    rec = torch.where(n_true > 0, tp / n_true.clamp_min(1), torch.zeros_like(tp))
    f1 = torch.where((prec + rec) > 0, 2 * prec * rec / (prec + rec).clamp_min(1e-300), torch.zeros_like(tp))
    k = present.sum().clamp_min(1)
    vals = torch.stack([prec[present].sum() / k, rec[present].sum() / k, f1[present].sum() / k, n_true.sum()]).tolist()
    return {f"{item_name}_{n}": float(v) for n, v in zip(names, vals)}


class DevicePrefetcher:
    """Iterates a host loader one batch ahead: batch i+1 is copied host -> device (from pinned memory, on a side
    stream) and, for ``EdgeListCollate`` batches, packed into CSR/CSC there while the main stream runs step i.
    Yields device batches in the dict format ``KVTrainStep`` takes.  The reference moves every batch synchronously
    inside the step (kv_procedure.py:116-123); this is the same copy, overlapped."""

    def __init__(self, loader, device: torch.device):
        self.loader = loader
        self.device = torch.device(device)
        self.stream = torch.cuda.Stream(self.device)

    def _load(self, batch):
        with torch.cuda.stream(self.stream):
            if "adjacency_matrix" not in batch and "edges" in batch:
                from .collate import EdgeListCollate
                out = EdgeListCollate.to_device(batch, self.device)
            else:
                out = {}
                for k, v in batch.items():
                    if isinstance(v, SparseNodeFeatures):
                        out[k] = v.to(self.device, non_blocking=True)
                        out[k].build_transpose()           # CSC for the emb1 weight gradient, on the copy stream
                        out[k].identity_perm()
                    elif isinstance(v, torch.Tensor):
                        if k == "textline_encoding":
                            if not v.is_floating_point():
                                v = v.float()
                            # land the features with a 16-byte row pitch (TMA reads them in place); the re-lay of an
                            # unaligned width (4369) happens here, on the copy stream, not inside the step
                            D, m = v.shape[-1], 16 // v.element_size()
                            if D % m:
                                buf = torch.empty(*v.shape[:-1], (D + m - 1) // m * m, dtype=v.dtype, device=self.device)
                                buf[..., :D].copy_(v, non_blocking=True)
                                out[k] = buf[..., :D]
                                continue
                        out[k] = v.to(self.device, non_blocking=True)
                    else:
                        out[k] = v
        return out

    def __iter__(self):
        main = torch.cuda.current_stream(self.device)
        it = iter(self.loader)
        try:
            nxt = self._load(next(it))
        except StopIteration:
            return
        while nxt is not None:
            main.wait_stream(self.stream)                 # batch i is on the device
            cur = nxt
            for v in cur.values():                        # allocated on the side stream, consumed on the main one
                if isinstance(v, torch.Tensor):
                    v.record_stream(main)
                elif isinstance(v, SparseNodeFeatures):
                    for t in v.tensors():
                        t.record_stream(main)
                elif hasattr(v, "rowptr"):
                    gT = getattr(v, "_typed_T", None)      # built on the copy stream by prepare_backward()
                    for t in (v.rowptr, v.col, v.val, v.colptr, v.rowidx, v.perm,
                              *((gT.rowptr, gT.col, gT.t_perm) if gT is not None else ())):
                        if t is not None:
                            t.record_stream(main)
            try:
                nxt = self._load(next(it))                # enqueue the copy of batch i+1 before step i is launched
            except StopIteration:
                nxt = None
            yield cur
