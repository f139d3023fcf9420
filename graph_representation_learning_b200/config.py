"""YAML config -> attribute dict, the shape ``GNNLearningWarper._from_config`` produces with anyconfig + munch
(gnn/cl_warper.py:61-79).  Neither package is needed: PyYAML plus a 20-line attribute dict give the same access
patterns (``cfg.procedure.type``, ``cfg["loss"]``, ``dict(cfg)``, ``cfg.get(...)``)."""
from __future__ import annotations

from typing import Any, Dict, Mapping

import yaml


class AttrDict(dict):
    """dict with attribute access, recursive like ``munch.munchify``."""

    def __getattr__(self, key: str) -> Any:
        try:
            return self[key]
        except KeyError:
            raise AttributeError(key) from None

    def __setattr__(self, key: str, value: Any) -> None:
        self[key] = value

    def __delattr__(self, key: str) -> None:
        try:
            del self[key]
        except KeyError:
            raise AttributeError(key) from None


def attrify(obj: Any) -> Any:
    if isinstance(obj, Mapping):
        return AttrDict({k: attrify(v) for k, v in obj.items()})
    if isinstance(obj, (list, tuple)):
        return type(obj)(attrify(v) for v in obj)
    return obj


def plain(obj: Any) -> Any:
    """back to plain dicts / lists (what goes into a checkpoint's "config" entry)"""
    if isinstance(obj, Mapping):
        return {k: plain(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return [plain(v) for v in obj]
    return obj


# defaults for the keys the procedures read, so a minimal config (or a plain dict) is enough
DEFAULTS: Dict[str, Any] = {
    "experiment_name": "GNN-B200", "seed": 1111, "is_train": True, "output_dir": "./assets/outputs",
    "checkpoint_path": None, "num_epochs": 1, "max_grad_norm": 5.0, "benchmark": False, "deterministic": True,
    "local_rank": 0, "num_gpus": 1, "distributed": False, "model_dir_name": "models",
    "procedure": {"type": "KVProcedure", "args": {}},
    "loss": {"type": "CrossEntropyLoss", "args": {}},
    "lr_scheduler": {"type": "DecayLearningRate", "args": {"lr": 0.001, "factor": 0.9995, "num_epochs": 100}},
    "optimizer": {"type": "BuitlinOptimizer", "args": {"type_optimizer": "Adam", "lr": 0.001}},
    "logging": {"summary_dir_name": "summary", "use_tensorboard": False},
}


def _merge(base: Dict[str, Any], over: Mapping) -> Dict[str, Any]:
    out = dict(base)
    for k, v in over.items():
        if isinstance(v, Mapping) and isinstance(out.get(k), Mapping):
            out[k] = _merge(dict(out[k]), v)
        else:
            out[k] = v
    return out


def load_config(path_or_dict: Any) -> AttrDict:
    """path to a YAML file (the reference's configs/sumi_node_classification.yaml schema) or a mapping"""
    if isinstance(path_or_dict, Mapping):
        raw = path_or_dict
    else:
        with open(path_or_dict) as fh:
            raw = yaml.safe_load(fh) or {}
    return attrify(_merge(DEFAULTS, raw))
