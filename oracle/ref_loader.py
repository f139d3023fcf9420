"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Locates and imports the UNMODIFIED reference model for the CPU baseline.

The reference is pure Python (no native code), so "building" it is staging: ``stage_reference()`` -- called by
``__graft_entry__.build()`` where ``/root/reference`` exists -- copies the ``gnn/models`` package and the one logger module
it imports, byte for byte, into the git-ignored ``oracle/_ref/`` so that they travel to the GPU box like the built
``.so``.  ``load_reference()`` imports ``GraphCNNDropEdge`` from there (or straight from ``/root/reference`` in the
authoring container) behind two in-memory shims for packages the image lacks (``colorlog``, ``decouple``: SURVEY.md 8c);
nothing of the reference is edited.  When neither location exists it returns ``None`` and callers fall back to the port
(``oracle/reference_port.py``), saying so (``kind: "port"``).

Only ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs, ``oracle/make_golden.py`` and tests import this module.
"""
from __future__ import annotations

import os
import shutil
import sys
import types
from typing import Optional, Tuple

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref")
SOURCE = "/root/reference"
# what gnn.models needs at import time (gnn/models/base_network.py:5-6 -> gnn.utils.logger.color_logger)
_FILES = ["gnn/__init__.py", "gnn/utils/__init__.py", "gnn/utils/logger/__init__.py", "gnn/utils/logger/color_logger.py"]
_TREES = ["gnn/models"]


def stage_reference(source: str = SOURCE, dest: str = STAGED) -> bool:
    """copy the reference's model package (unmodified) into oracle/_ref; False when the reference is not present"""
    if not os.path.isdir(os.path.join(source, "gnn", "models")):
        return False
    for rel in _FILES:
        os.makedirs(os.path.dirname(os.path.join(dest, rel)), exist_ok=True)
        shutil.copyfile(os.path.join(source, rel), os.path.join(dest, rel))
    for rel in _TREES:
        shutil.copytree(os.path.join(source, rel), os.path.join(dest, rel), dirs_exist_ok=True,
                        ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    return True


def _shims() -> None:
    colorlog = types.ModuleType("colorlog")
    colorlog.basicConfig = lambda **kw: None
    sys.modules.setdefault("colorlog", colorlog)
    # color_logger.py:36-40 opens <OUTPUT_DIR>/output.log for every model instance: keep that out of the working directory
    log_dir = os.path.join(STAGED, "_logs")
    os.makedirs(log_dir, exist_ok=True)
    decouple = types.ModuleType("decouple")
    decouple.config = lambda key, default=None, **kw: log_dir if key == "OUTPUT_DIR" else default
    sys.modules.setdefault("decouple", decouple)
    import numpy as np
    if not hasattr(np, "float"):
        np.float = float  # removed numpy alias used by textline_encoding.py:71


def reference_root() -> Optional[str]:
    for root in (STAGED, SOURCE):
        if os.path.isfile(os.path.join(root, "gnn", "models", "networks", "drop_robust_gcn.py")):
            return root
    return None


def load_reference() -> Optional[Tuple[type, str]]:
    """(the reference's GraphCNNDropEdge class, where it was imported from) or None"""
    root = reference_root()
    if root is None:
        return None
    _shims()
    if root not in sys.path:
        sys.path.insert(0, root)
    from gnn.models import GraphCNNDropEdge  # noqa: E402
    return GraphCNNDropEdge, root


def reference_train_step(model, optimizer, V, A, y, max_grad_norm: float = 5.0):
    """KVProcedure._run_train_step (gnn/trainer/training_procedures/kv_procedure.py:143-164) around the imported model,
    without the sklearn report and the neptune / tensorboard plumbing (gnn.trainer cannot be imported here: SURVEY.md 8c):
    zero_grad -> forward([V, A]) -> CrossEntropyLoss(pred.transpose(1, 2), y) -> backward -> clip_grad_norm_ -> step."""
    import torch
    model.train()
    optimizer.zero_grad()
    pred = model.forward([V, A])
    loss = torch.nn.CrossEntropyLoss()(pred.transpose(1, 2), y)
    loss.backward()
    torch.nn.utils.clip_grad_norm_(model.parameters(), max_grad_norm)
    optimizer.step()
    return loss.detach()
