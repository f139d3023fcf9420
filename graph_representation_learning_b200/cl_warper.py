"""``GNNLearningWarper`` -- the reference's train / infer entry point (gnn/cl_warper.py:19-115), same constructor and
methods: ``GNNLearningWarper(model, config_path=None, config=None)``, ``.train()``, ``.predict(samples)``.

Differences, all forced by the environment or by bugs of the reference:
* the config is YAML -> attribute dict without anyconfig / munch; no experiment tracker is contacted at import;
* ``distributed: true`` really runs data parallel: the process group comes from torchrun's environment
  (``init_method="env://"``, backend nccl on GPUs / gloo on CPU), every rank takes ``cuda:LOCAL_RANK``, and the
  gradient all-reduce is the flat-bucket one of ``trainer.FlatGradBucket`` -- the reference defines DDP wrappers but
  never applies them.
"""
from __future__ import annotations

import os
import random
from typing import Any, Dict, List, Optional

import numpy as np
import torch
from torch import nn

from . import procedures
from .config import AttrDict, load_config


class GNNLearningWarper:
    def __init__(self, model: nn.Module, config_path: Optional[str] = None, config: Optional[Dict[str, Any]] = None,
                 **procedure_kwargs: Any):
        assert config_path or config is not None
        self.model = model
        self.device = torch.device("cuda" if torch.cuda.is_available() else "cpu")
        self.config = self._from_config(config_path if config_path else config)

        random.seed(self.config.seed)
        np.random.seed(self.config.seed)
        torch.manual_seed(self.config.seed)
        if torch.cuda.is_available():
            torch.cuda.manual_seed_all(self.config.seed)

        self.config.output_dir = self._make_output_dir()
        os.environ["OUTPUT_DIR"] = self.config.output_dir
        self.checkpointer = procedures.CheckpointHandler()

        ptype = self.config.procedure.type
        pargs = dict(self.config.procedure.get("args", {}) or {})
        pargs.update(procedure_kwargs)
        if self.config.is_train:
            if ptype not in procedures.TRAINING_PROCEDURES:
                raise AttributeError(f"training procedure {ptype!r} not found; available: "
                                     f"{sorted(procedures.TRAINING_PROCEDURES)}")
            self.trainer = procedures.TRAINING_PROCEDURES[ptype]._from_config(self.model, self.config, ems_exp=None, **pargs)
        else:
            if ptype not in procedures.INFERENCE_PROCEDURES:
                raise AttributeError(f"inference procedure {ptype!r} not found; available: "
                                     f"{sorted(procedures.INFERENCE_PROCEDURES)}")
            self.inferencer = procedures.INFERENCE_PROCEDURES[ptype]._from_config(self.model, self.config, **pargs)

    @staticmethod
    def _from_config(config_path) -> AttrDict:
        cfg = load_config(config_path)
        if cfg.distributed and not torch.distributed.is_initialized():
            use_cuda = torch.cuda.is_available() and int(cfg.num_gpus) > 0
            if use_cuda:
                torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", cfg.local_rank)))
            torch.distributed.init_process_group(backend="nccl" if use_cuda else "gloo", init_method="env://")
        torch.backends.cudnn.benchmark = bool(cfg.benchmark)
        torch.backends.cudnn.deterministic = bool(cfg.deterministic)
        return cfg

    def _make_output_dir(self) -> str:
        out = os.path.join(self.config.output_dir, self.config.experiment_name)
        os.makedirs(out, exist_ok=True)
        return out

    def train(self) -> Any:
        return self.trainer()

    def predict(self, samples: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        return self.inferencer(samples)
