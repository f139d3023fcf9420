"""Training / inference procedures of the KV node-classification task on the B200 path.

Mirrors (same class names, constructor arguments, method names and control flow):

* ``CheckpointHandler``          gnn/utils/checkpoint_handler.py:9-64     ({"epoch","config","meta_data","state_dict"})
* ``BaseProcedure``              gnn/trainer/training_procedures/base_procedure.py:14-197
* ``KVProcedure``                gnn/trainer/training_procedures/kv_procedure.py:18-377
* ``KVInference``                gnn/inferencer/inference_procedures/kv_inference.py:13-118

What differs: the step itself runs through ``KVTrainStep`` (flat gradient bucket, one NCCL all-reduce when
torch.distributed is initialised, fused Adam, no per-step host synchronisation), metric scores come from one
confusion matrix on the device, batches can be prefetched to the device on a copy stream, and experiment trackers
(tensorboardX / neptune) are optional instead of being imported -- and contacted -- at module import.
The reference's datasets (JSON -> BoW features + heuristic graph, CPU string / geometry code) are out of scope
(SURVEY.md section 2.1); any ``torch.utils.data.Dataset`` yielding the three arrays plugs in through ``DATASETS``.
"""
from __future__ import annotations

import logging
import math
import os
from typing import Any, Dict, Iterable, List, Optional, Tuple

import numpy as np
import torch
from torch import nn
from torch.utils.data import DataLoader, Dataset

from . import collate as collate_mod
from .config import plain
from .graph import PackedGraphBatch
from .trainer import (LOSSES, LR_SCHEDULERS, OPTIMIZERS, Dictlist, KVTrainStep, from_registry, macro_scores)

logger = logging.getLogger(__name__)


class CheckpointHandler:
    def __init__(self):
        self.device = torch.device("cuda" if torch.cuda.is_available() else "cpu")

    @staticmethod
    def make_checkpoint_name(name: str, epoch: Optional[int] = None, step: Optional[int] = None) -> str:
        return name + "_latest.pt" if epoch is None or step is None else f"{name}_epoch_{epoch}_minibatch_{step}.pt"

    def save_checkpoint(self, checkpoint: Dict[str, Any], output_path: str, epoch: Optional[int] = None,
                        step: Optional[int] = None) -> str:
        os.makedirs(output_path, exist_ok=True)
        path = os.path.join(output_path, self.make_checkpoint_name("model", epoch, step))
        torch.save(checkpoint, path)
        return path

    def restore_checkpoint(self, checkpoint_path: str) -> Dict[str, Any]:
        return torch.load(checkpoint_path, map_location=self.device, weights_only=False)


class _NullWriter:
    """stands in for tensorboardX.SummaryWriter when it is not installed / not wanted"""

    def add_scalar(self, *a, **k):
        pass

    def add_histogram(self, *a, **k):
        pass

    def close(self):
        pass


def _make_writer(config) -> Any:
    if not config.get("logging", {}).get("use_tensorboard", False):
        return _NullWriter()
    try:
        from tensorboardX import SummaryWriter
    except ImportError:
        logger.warning("tensorboardX is not installed: scalars are not written")
        return _NullWriter()
    return SummaryWriter(os.path.join(config.output_dir, config.logging.summary_dir_name))


# ------------------------------------------------------------------------------------------------ datasets
class TensorDictDataset(Dataset):
    """A list of per-document samples ``{"textline_encoding": [n, D] f32, "adjacency_matrix": [n, L, n] f16 (or
    "edges" [e, 3]), "node_label": [n] i64}`` -- the dict DatapileDataset.__getitem__ returns
    (gnn/data_generator/datasets/datapile_dataset.py:260-273) after ``only_selected_items``."""

    def __init__(self, samples: List[Dict[str, np.ndarray]], id_to_class: Optional[Dict[int, Tuple[str, ...]]] = None,
                 data_config: Optional[Dict[str, Any]] = None, **_: Any):
        self.samples = samples
        self.id_to_class = id_to_class or {}
        self.data_config = data_config

    @classmethod
    def _from_config(cls, args: Dict[str, Any], **kwargs: Any) -> "TensorDictDataset":
        return cls(args["samples"], args.get("id_to_class"), data_config=args, **kwargs)

    def __len__(self) -> int:
        return len(self.samples)

    def __getitem__(self, idx: int) -> Dict[str, np.ndarray]:
        return dict(self.samples[idx])


class SyntheticDocumentDataset(TensorDictDataset):
    """Seeded random documents of ragged size (n in [n_min, n_max]) with the reference's dtypes; for tests, smoke runs
    and benchmarks of the whole procedure."""

    def __init__(self, num_documents: int = 32, n_min: int = 20, n_max: int = 40, input_dim: int = 64,
                 num_classes: int = 7, num_edges: int = 6, degree: float = 3.0, seed: int = 0,
                 data_config: Optional[Dict[str, Any]] = None, **_: Any):
        rng = np.random.default_rng(seed)
        samples = []
        for _i in range(num_documents):
            n = int(rng.integers(n_min, n_max + 1))
            A = (rng.random((n, num_edges, n)) < degree / max(n - 1, 1)).astype(np.float16)
            A[np.arange(n), :, np.arange(n)] = 0
            samples.append({"textline_encoding": rng.random((n, input_dim), dtype=np.float32),
                            "adjacency_matrix": A, "node_label": rng.integers(0, num_classes, n).astype(np.int64)})
        id_to_class = {i: (f"class{i}", "value") for i in range(1, num_classes)}
        super().__init__(samples, id_to_class, data_config)

    @classmethod
    def _from_config(cls, args: Dict[str, Any], **kwargs: Any) -> "SyntheticDocumentDataset":
        known = ("num_documents", "n_min", "n_max", "input_dim", "num_classes", "num_edges", "degree", "seed")
        return cls(**{k: args[k] for k in known if k in args}, data_config=args)


DATASETS = {"TensorDictDataset": TensorDictDataset, "SyntheticDocumentDataset": SyntheticDocumentDataset}


class BaseDataLoader:
    """gnn/data_generator/base_dataloader.py:15-112 (dataset by name, collate chain, per-rank batch split)."""

    def __init__(self, config):
        self.config = config

    def _load_dataset(self, dataset_type: str, args: Dict[str, Any], **kwargs: Any) -> Dataset:
        if dataset_type not in DATASETS:
            raise KeyError(f"dataset {dataset_type!r} is not registered (the reference's JSON datasets are out of scope: "
                           f"register a torch Dataset in procedures.DATASETS); available: {sorted(DATASETS)}")
        return DATASETS[dataset_type]._from_config(args, **kwargs)

    def _get_dataloader(self, dataset: Dataset, data_config, custom_collate=None) -> DataLoader:
        collates = collate_mod.load_collate_processors(data_config.get("data_collate", {}) or {})
        batch_size = int(data_config.get("batch_size", 1))
        sampler = None
        if self.config.get("distributed", False):
            # the reference combines DistributedSampler with shuffle=True, which torch rejects:
            # here the sampler owns the shuffling
            from torch.utils.data.distributed import DistributedSampler
            world = torch.distributed.get_world_size() if torch.distributed.is_initialized() else int(self.config.num_gpus)
            rank = torch.distributed.get_rank() if torch.distributed.is_initialized() else int(self.config.local_rank)
            batch_size = max(1, batch_size // world)
            sampler = DistributedSampler(dataset, num_replicas=world, rank=rank,
                                         shuffle=bool(data_config.get("shuffle", False)))
        return DataLoader(dataset, batch_size=batch_size, num_workers=int(data_config.get("num_workers", 0)),
                          drop_last=bool(data_config.get("drop_last", False)),
                          pin_memory=bool(data_config.get("pin_memory", False)) and torch.cuda.is_available(),
                          collate_fn=custom_collate if custom_collate is not None else collates,
                          shuffle=bool(data_config.get("shuffle", False)) if sampler is None else False, sampler=sampler)


# ------------------------------------------------------------------------------------------------ training
class BaseProcedure:
    def __init__(self, model: nn.Module, config, ems_exp: Any = None, **kwargs: Any):
        self.logger = logger
        self.config = config
        self.ems_exp = ems_exp
        self.model_dir = self._make_output_dir()
        self.checkpointer = CheckpointHandler()
        self.device, self.device_ids = self._prepare_device(int(config.num_gpus))
        self.model = self._load_prev_checkpoint(model).to(self.device)
        self.criterion = self._init_criterion()
        self.lr_scheduler = self._init_lr_scheduler()
        # optimizer + flat gradient bucket + (when torch.distributed is up) the gradient all-reduce live in the step
        opt = from_registry(OPTIMIZERS, "optimizer", config.optimizer.type, config.optimizer.args)
        self.step = KVTrainStep(self.model, max_grad_norm=float(config.max_grad_norm), device=self.device,
                                criterion=self.criterion,
                                optimizer_factory=(lambda params: self._fused(opt).get_optimizer(params)) if opt else None)
        self.optimizer = self.step.optimizer
        self.tb_writer = _make_writer(config)

    def _fused(self, opt):
        if self.device.type == "cuda" and opt.type_optimizer in ("Adam", "AdamW"):
            opt.optimizer_args.setdefault("fused", True)
        return opt

    @classmethod
    def _from_config(cls, model: nn.Module, config, ems_exp: Any = None, **kwargs: Any) -> "BaseProcedure":
        return cls(model, config, ems_exp, **kwargs)

    def _prepare_device(self, n_gpu_use: int) -> Tuple[torch.device, List[int]]:
        """base_procedure.py:54-77, except that under torchrun each rank takes cuda:LOCAL_RANK instead of cuda:0."""
        n_gpu = torch.cuda.device_count()
        if n_gpu_use > 0 and n_gpu == 0:
            raise RuntimeError("config.num_gpus > 0 but no CUDA device is visible: the B200 path has no CPU fallback "
                               "(run the reference itself for CPU)")
        n_gpu_use = min(n_gpu_use, n_gpu)
        local = int(os.environ.get("LOCAL_RANK", self.config.get("local_rank", 0))) if self.config.get("distributed") else 0
        return torch.device(f"cuda:{local}" if n_gpu_use > 0 else "cpu"), list(range(n_gpu_use))

    def _init_criterion(self):
        return from_registry(LOSSES, "loss", self.config.loss.type, self.config.loss.args)

    def _init_lr_scheduler(self):
        return from_registry(LR_SCHEDULERS, "lr_scheduler", self.config.lr_scheduler.type, self.config.lr_scheduler.args)

    def _make_output_dir(self) -> str:
        out = os.path.join(self.config.output_dir, self.config.model_dir_name)
        os.makedirs(out, exist_ok=True)
        return out

    def _load_prev_checkpoint(self, model: nn.Module) -> nn.Module:
        path = self.config.get("checkpoint_path")
        if path:
            ckpt = CheckpointHandler().restore_checkpoint(path)
            state, prev = ckpt.get("state_dict"), ckpt.get("config")
            if state and prev:
                model.load_state_dict(state, strict=False)          # base_procedure.py:163-167
        return model

    def _update_learning_rate(self, epoch: int, step: int) -> float:
        lr = self.lr_scheduler._step_lr(epoch, step)
        for group in self.optimizer.param_groups:
            group["lr"] = lr
        return lr


class KVProcedure(BaseProcedure):
    """Epoch / step loops of kv_procedure.py:180-244,346-377.  ``metrics_interval`` (procedure arg, default 1 = the
    reference's behaviour) spaces out the per-step metric read-back, the only host synchronisation of a step."""

    def __init__(self, model: nn.Module, config, ems_exp: Any = None, metrics_interval: int = 1, **kwargs: Any):
        super().__init__(model, config, ems_exp, **kwargs)
        self.global_step = 0
        self.metrics_interval = max(1, int(metrics_interval))
        self.train_loader, self.val_loader, self.class_names = self._init_dataloaders()
        ds_args = config.data_config.dataset.get("args", {}) or {}
        self._ignore = (ds_args.get("node_label_padding_value", -100), ds_args.get("other_class_index"))

    def _init_dataloaders(self):
        dl = BaseDataLoader(self.config)
        dc = self.config.data_config
        train_ds = dl._load_dataset(dc.dataset.type, dc.training, data_type="training")
        val_ds = dl._load_dataset(dc.dataset.type, dc.validation, data_type="validation")
        names = [(idx, "_".join(n)) for idx, n in getattr(train_ds, "id_to_class", {}).items()]
        class_names = tuple(["other"] + [n for _, n in sorted(names)])
        return dl._get_dataloader(train_ds, dc.training), dl._get_dataloader(val_ds, dc.validation), class_names

    def _get_metric_scores(self, preds: torch.Tensor, gts: torch.Tensor, item_name: str = "item"):
        return macro_scores(preds, gts, self._ignore, item_name)

    def _step_process(self, batch_sample, want_metrics: bool = True):
        loss, pred = self.step.step_process(batch_sample)
        scores: Dict[str, Any] = {}
        if want_metrics:
            y = batch_sample["node_label"].to(self.device, non_blocking=True)
            scores = self._get_metric_scores(pred.argmax(dim=-1), y, item_name="Node classification")
            scores["loss"] = float(loss.detach())
            if hasattr(self.step.criterion, "check_labels"):
                self.step.criterion.check_labels()        # out-of-range targets raise here (the step's one host read)
        return loss, scores

    def _run_train_step(self, batch_sample) -> Dict[str, Any]:
        want = self.global_step % self.metrics_interval == 0
        self.model.train()
        self.step.bucket.zero()
        with torch.set_grad_enabled(True):
            loss, scores = self._step_process(batch_sample, want)
            self.step.bucket.arm_overlap()
            loss.backward()
            self.step.clip_and_step()
        return scores

    @torch.no_grad()
    def _run_val_step(self, batch_sample) -> Tuple[Dict[str, Any], torch.Tensor, torch.Tensor]:
        self.model.eval()
        loss, pred = self.step.step_process(batch_sample)
        y = batch_sample["node_label"].to(self.device)
        scores = self._get_metric_scores(pred.argmax(dim=-1), y, item_name="Node classification")
        scores["loss"] = float(loss.detach())
        return scores, pred.argmax(dim=-1).reshape(-1), y.reshape(-1)

    def cosine_schedule_lambda(self, step: int, epoch: int, total_steps: int, base_value: float, max_value: float,
                               warmup_steps: int = 0) -> float:
        """kv_procedure.py:252-281 (the value is poked into ``model.lambda_value`` every step)."""
        step = max(0, min(step, total_steps))
        warmup_steps = min(warmup_steps, total_steps)
        if step < warmup_steps:
            val = base_value + (max_value - base_value) * (step / warmup_steps)
        else:
            progress = float(step - warmup_steps) / float(max(1, total_steps - warmup_steps))
            val = base_value + 0.5 * (max_value - base_value) * (1 + math.cos(math.pi * progress))
        self.tb_writer.add_scalar("RP/Lambda", val, epoch)
        return val

    def _optimize_per_epoch(self, epoch: int) -> Dict[str, Any]:
        train_metrics = Dictlist()
        steps_per_epoch = max(1, len(self.train_loader))
        for train_batch in self.train_loader:
            scores = self._run_train_step(train_batch)
            if scores:
                train_metrics._update(scores)
                self.tb_writer.add_scalar("Train_step_loss", scores["loss"], self.global_step)
                if self.ems_exp:
                    self.ems_exp["Train/step_loss"].append(scores["loss"])
            self.global_step += 1
            self.model.lambda_value = self.cosine_schedule_lambda(
                self.global_step, epoch, int(self.config.num_epochs) * steps_per_epoch, base_value=1e-4, max_value=1.0,
                warmup_steps=5 * steps_per_epoch)
        train_result = train_metrics._result()
        self.logger.info(f"Training epoch: {epoch} step: {self.global_step} metrics: {train_result}")
        for name, score in train_result.items():
            self.tb_writer.add_scalar(f"Train {name}", score, epoch)

        val_metrics = Dictlist()
        preds, labels = [], []
        for val_batch in self.val_loader:
            scores, p, y = self._run_val_step(val_batch)
            val_metrics._update(scores)
            preds.append(p)
            labels.append(y)
        val_result = val_metrics._result() if len(val_metrics) else {"loss": float("inf")}
        for name, score in val_result.items():
            self.tb_writer.add_scalar(f"Val {name}", score, epoch)
        macro = {"precision": 0.0, "recall": 0.0, "f1-score": 0.0, "support": 0.0}
        if preds:
            m = macro_scores(torch.cat(preds), torch.cat(labels), self._ignore, item_name="m")
            macro = {k[2:]: v for k, v in m.items()}
        for name, score in macro.items():
            self.tb_writer.add_scalar(f"Macro Val {name}", score, epoch)
        macro["loss"] = val_result["loss"]
        return macro

    def __call__(self) -> float:
        best_loss = float("inf")
        metrics: Dict[str, Any] = {"f1-score": 0.0}
        # kv_procedure.py:349 calls model.zero_grad(); on torch 2.x that sets p.grad = None and would detach the
        # parameters from the flat bucket, so the bucket itself is cleared (and re-attaches stray gradients anyway)
        self.step.bucket.zero()
        for epoch in range(int(self.config.num_epochs)):
            metrics = self._optimize_per_epoch(epoch)
            self._update_learning_rate(epoch, self.global_step)
            is_rank0 = not (torch.distributed.is_available() and torch.distributed.is_initialized()) \
                or torch.distributed.get_rank() == 0
            if metrics["loss"] < best_loss and is_rank0:
                best_loss = metrics["loss"]
                self.checkpointer.save_checkpoint({"epoch": epoch, "config": plain(self.config), "meta_data": metrics,
                                                   "state_dict": self.model.state_dict()}, self.model_dir)
        self.tb_writer.close()
        return metrics["f1-score"]


TRAINING_PROCEDURES = {"KVProcedure": KVProcedure}


# ------------------------------------------------------------------------------------------------ inference
class KVInference:
    """kv_inference.py:13-118: documents in, the same documents out with ``key_type`` / ``formal_key`` /
    ``confidence`` written into every text-line box.

    The reference runs one forward per document (batch = 1, no ``no_grad``).  Here documents with the SAME node count
    are stacked into one batch -- bit-compatible with batch = 1 because no padding is introduced (padded nodes would
    be live in the unmasked attention softmax) -- and the forward runs under ``no_grad``.
    ``featurizer(document) -> {"textline_encoding": [n, D], "adjacency_matrix": [n, L, n] | "edges": [e, 3]}`` stands
    in for the reference's dataset processors (out of scope); ``id_to_class[idx] = (formal_key, key_type)``.
    """

    def __init__(self, model: nn.Module, config, featurizer=None, id_to_class: Optional[Dict[int, Tuple[str, str]]] = None,
                 max_batch_documents: int = 64, **kwargs: Any):
        self.config = config
        infer = config.get("inference_settings", {}) or {}
        n_gpu = int(infer.get("num_gpus", config.get("num_gpus", 1)))
        if n_gpu > 0 and torch.cuda.device_count() == 0:
            raise RuntimeError("inference_settings.num_gpus > 0 but no CUDA device is visible (no CPU fallback)")
        self.device = torch.device("cuda:0" if n_gpu > 0 else "cpu")
        path = config.get("checkpoint_path")
        if path:
            ckpt = CheckpointHandler().restore_checkpoint(path)
            if ckpt.get("state_dict") and ckpt.get("config"):
                model.load_state_dict(ckpt["state_dict"], strict=False)
        self.model = model.to(self.device)
        act = infer.get("activation", {"type": "Softmax", "args": {"dim": 2}})
        self.activator = getattr(torch.nn, act["type"])(**dict(act.get("args", {})))
        self.featurizer = featurizer
        self.id_to_class = dict(id_to_class or {})
        self.id_to_class[0] = ("other", "other")
        self.max_batch = max(1, int(max_batch_documents))

    @classmethod
    def _from_config(cls, model: nn.Module, config, **kwargs: Any) -> "KVInference":
        return cls(model, config, **kwargs)

    @torch.no_grad()
    def step_process(self, items: List[Dict[str, Any]]) -> List[List[Dict[str, Any]]]:
        """items: featurized documents of EQUAL node count; returns their annotated boxes."""
        V = torch.as_tensor(np.stack([np.asarray(it["textline_encoding"], dtype=np.float32) for it in items]))
        V = V.to(self.device)
        n = V.shape[1]
        if "edges" in items[0]:
            parts = []
            for b, it in enumerate(items):
                tri = np.asarray(it["edges"], dtype=np.int32).reshape(-1, 3)
                parts.append(np.concatenate([np.full((tri.shape[0], 1), b, np.int32), tri], axis=1))
            L = int(self.model.num_edges)
            A = PackedGraphBatch.from_edges(torch.from_numpy(np.concatenate(parts)).to(self.device), len(items), n, L)
        else:
            A = torch.as_tensor(np.stack([np.asarray(it["adjacency_matrix"]) for it in items])).to(self.device)
            if not A.is_floating_point():
                A = A.float()
        scores, classes = self.activator(self.model([V, A])).max(dim=-1)
        classes, scores = classes.cpu().tolist(), scores.cpu().tolist()
        outputs = []
        for b, it in enumerate(items):
            boxes = it["label"]
            assert len(boxes) == n
            for idx, box in enumerate(boxes):
                formal_key, key_type = self.id_to_class[classes[b][idx]]
                box["key_type"], box["formal_key"], box["confidence"] = key_type, formal_key, scores[b][idx]
            outputs.append(boxes)
        return outputs

    def __call__(self, samples) -> List[List[Dict[str, Any]]]:
        """samples: one document (list of boxes) or a list of documents (handle_single_input of the reference)."""
        if samples and isinstance(samples[0], dict):
            samples = [samples]
        self.model.eval()
        items = []
        for doc in samples:
            it = dict(self.featurizer(doc)) if self.featurizer is not None else dict(doc)
            it.setdefault("label", doc)
            items.append(it)
        outputs: List[Any] = [None] * len(items)
        by_n: Dict[int, List[int]] = {}
        for k, it in enumerate(items):
            by_n.setdefault(int(np.asarray(it["textline_encoding"]).shape[0]), []).append(k)
        for _, idxs in sorted(by_n.items()):
            for s in range(0, len(idxs), self.max_batch):
                group = idxs[s:s + self.max_batch]
                for k, out in zip(group, self.step_process([items[k] for k in group])):
                    outputs[k] = out
        return outputs


INFERENCE_PROCEDURES = {"KVInference": KVInference}


class GraphedEvalForward:
    """CUDA-graph replay of ``model.eval()([V, A])`` for one fixed batch signature ``(B, N)``.

    Small documents make the eval forward launch-bound (64 x 100 nodes: ~60 kernels of a few microseconds each, 1.3 ms
    of host launch latency against 0.3 ms of device work).  The whole path -- CSR pack of the dense adjacency, the model
    -- is stream-ordered, allocation-stable and free of host reads once the CSR capacity is fixed, so it is captured once
    and replayed: ``__call__`` copies the new batch into the static input buffers and launches ONE graph.
    The CSR buffers are sized for the dense worst case ``B*N*(L*N+1)`` entries, so no batch can overflow them and no host
    read of the entry count is needed (64 x 200 nodes: 0.3 GB); meant for the small-document regime where launches dominate.
    The KV inference of the reference (gnn/inference/kv_inference.py:43-49) runs one eager forward per document."""

    def __init__(self, model: nn.Module, B: int, N: int, input_dim: int, num_edges: int,
                 adj_dtype: torch.dtype = torch.float16):
        from .graph import PackedGraphBatch
        self.model = model.eval()
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise RuntimeError("GraphedEvalForward needs the model on a CUDA device (no CPU fallback)")
        self.B, self.N, self.L = B, N, num_edges
        self.capacity = B * N * (num_edges * N + 1)
        ld = (input_dim + 3) // 4 * 4                              # 16-byte row pitch: TMA reads V in place
        self._Vp = torch.zeros(B, N, ld, device=dev)
        self.V = self._Vp[:, :, :input_dim]
        self.A = torch.zeros(B, N, num_edges, N, dtype=adj_dtype, device=dev)

        def run():
            g = PackedGraphBatch.from_dense(self.A, nnz_capacity=self.capacity, with_transpose=False)
            return self.model([self.V, g])

        with torch.no_grad():
            side = torch.cuda.Stream(dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):                          # warm-up: library load, function attributes, pools
                for _ in range(2):
                    run()
            torch.cuda.current_stream(dev).wait_stream(side)
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self.logits = run()

    @torch.no_grad()
    def __call__(self, V: torch.Tensor, A: torch.Tensor) -> torch.Tensor:
        """V [B, N, input_dim] float, A [B, N, L, N] (host or device); returns the static logits buffer [B, N, out]"""
        self.V.copy_(V, non_blocking=True)
        self.A.copy_(A, non_blocking=True)
        self.graph.replay()
        return self.logits
