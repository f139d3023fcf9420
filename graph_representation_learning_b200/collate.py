"""Batch construction (collate) for the GraphCNNDropEdge path.

Two collates share one contract -- a dict ``{"textline_encoding": V, "adjacency_matrix": A, "node_label": y}``
that ``KVProcedure._step_process`` hands to ``model.forward([V, A])``:

* ``NumpyPadding`` + ``default_collate``: the reference's host collate, restated
  (gnn/data_generator/data_collate/numpy_padding.py:9-103, gnn/data_generator/base_dataloader.py:41-58).  Every
  array is zero-padded **symmetrically** to the shape of the sample with the largest element count, so an n-node
  document lands at ``[(Nmax-n)//2, (Nmax-n)//2 + n)`` on *both* node axes; labels pad with -100.  Output: dense
  ``V [B,N,4369] f32``, ``A [B,N,6,N] f16``, ``y [B,N] i64``; the model packs ``A`` into CSR on the device.
* ``EdgeListCollate``: the same placement, but the adjacency never exists densely on the host --
  each document contributes its ``(start, type, end)`` triples (what ``Graph`` writes into ``adj[start, label, end]``,
  gnn/data_generator/data_process/utils/graph_utils.py:821-829) and the batch carries one ``int32 [E,4]`` edge list
  ``(b, i+off_b, l, j+off_b)``; ``adjacency_matrix`` is then a ``PackedGraphBatch`` built by the CUDA pack kernels.
  Host-to-device traffic for A drops from ``B*N*6*N*2`` bytes to ``16*E``.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data._utils.collate import default_collate

from .features import SparseNodeFeatures
from .graph import PackedGraphBatch


class BaseCollate:
    """gnn/data_generator/data_collate/base_collate.py"""

    @classmethod
    def _from_config(cls, config: Dict[str, Any]) -> "BaseCollate":
        return cls(**config)

    def __call__(self, inputs):
        raise NotImplementedError


class NumpyPadding(BaseCollate):
    """gnn/data_generator/data_collate/numpy_padding.py:9-103 (same arguments, same placement rule)."""

    def __init__(self, name_value_pairs: dict, only_selected_items: bool = False):
        self.name_value_pairs = name_value_pairs
        self.only_selected_items = only_selected_items

    def __call__(self, inputs: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        inputs = self._normalize_shape(inputs)
        if self.only_selected_items:
            inputs = self._return_selected_items(inputs)
        return inputs

    @staticmethod
    def _set_padding(padding_size: Sequence[int], pad_type: str = "symmetric") -> Tuple[Tuple[int, int], ...]:
        if pad_type == "symmetric":
            return tuple((int(s) // 2, int(s) - int(s) // 2) for s in padding_size)
        if pad_type == "edge":
            return tuple((int(s), 0) for s in padding_size)
        raise ValueError(f"{pad_type} not found")

    def _return_selected_items(self, inputs: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        keys = self.name_value_pairs.keys()
        return [{name: item for name, item in sample.items() if name in keys} for sample in inputs]

    def _normalize_shape(self, inputs: List[Dict[str, Any]]) -> List[Dict[str, Any]]:
        for in_name, value in self.name_value_pairs.items():
            present = [s.get(in_name, None) for s in inputs]
            arrays = [a for a in present if a is not None]
            if not arrays or not all(isinstance(a, np.ndarray) for a in arrays):
                continue
            shapes = [list(a.shape) for a in arrays]
            # the reference takes the shape with the largest PRODUCT, not the per-axis maximum (numpy_padding.py:88)
            max_size = max(shapes, key=lambda sh: np.prod(np.array(sh)))
            padded = []
            for a, sh in zip(arrays, shapes):
                pads = self._set_padding(np.subtract(max_size, sh), pad_type="symmetric")
                padded.append(np.pad(a, pads, constant_values=value))
            # the reference writes back positionally (numpy_padding.py:100-101): samples lacking the key break it
            # there too; keep the same pairing rule for samples that have it
            it = iter(padded)
            for idx in range(len(inputs)):
                if inputs[idx].get(in_name, None) is not None:
                    inputs[idx][in_name] = next(it)
        return inputs


class Compose:
    """torchvision.transforms.Compose as used at base_dataloader.py:89 (torchvision is not needed for that)."""

    def __init__(self, fns: Sequence[Callable]):
        self.fns = list(fns)

    def __call__(self, x):
        for fn in self.fns:
            x = fn(x)
        return x


COLLATES = {"NumpyPadding": NumpyPadding}


def load_collate_processors(collate_config: Dict[str, Any]) -> Compose:
    """BaseDataLoader._load_collate_processors (base_dataloader.py:41-58): configured collates, then default_collate."""
    procs: List[Callable] = []
    for name in collate_config:
        if name not in COLLATES:
            raise KeyError(f"unknown data collate {name!r}; registered: {sorted(COLLATES)}")
        procs.append(COLLATES[name]._from_config(dict(collate_config[name])))
    procs.append(default_collate)
    return Compose(procs)


def pad_offset(n: int, n_max: int) -> int:
    """first index of an n-node document inside an n_max-node padded batch (symmetric padding, numpy_padding.py:39-40)"""
    return (n_max - n) // 2


def dense_to_edges(adjacency: np.ndarray) -> Tuple[np.ndarray, Optional[np.ndarray]]:
    """One document's dense ``[n, L, n]`` adjacency -> (int32 [e, 3] triples (start, type, end) in row-major order,
    float32 weights or None when all ones).  For datasets that still build the dense matrix per document."""
    s, l, e = np.nonzero(adjacency)
    vals = adjacency[s, l, e].astype(np.float32)
    tri = np.stack([s, l, e], axis=1).astype(np.int32)
    return tri, (None if np.all(vals == 1.0) else vals)


class EdgeListCollate(BaseCollate):
    """Pads features / labels like ``NumpyPadding`` and ships the adjacency as one edge list (module docstring).

    Each sample: ``textline_encoding [n, D]``, ``node_label [n]`` and either ``edges [e, 3]`` (+ optional
    ``edge_weights [e]``) or a dense ``adjacency_matrix [n, L, n]`` (converted with ``dense_to_edges``).
    Returns tensors on the host (pinned when ``pin_memory``); ``to_device`` moves them and packs the CSR/CSC batch.
    """

    def __init__(self, num_edges: int = 6, feature_padding_value: float = 0.0, label_padding_value: int = -100,
                 pin_memory: bool = True, feature_ld_multiple: int = 4):
        self.L = num_edges
        self.fpad, self.lpad = feature_padding_value, label_padding_value
        self.pin = pin_memory and torch.cuda.is_available()
        self.ld_mult = feature_ld_multiple      # V rows keep a 16-byte pitch so TMA reads them in place

    def __call__(self, samples: List[Dict[str, Any]]) -> Dict[str, Any]:
        B = len(samples)
        sizes = [int(s["textline_encoding"].shape[0]) for s in samples]
        # NumpyPadding pads to the sample with the largest element count; for [n, D] features that is the largest n
        n_max = max(sizes)
        D = int(samples[0]["textline_encoding"].shape[1])
        ld = (D + self.ld_mult - 1) // self.ld_mult * self.ld_mult
        Vp = torch.full((B, n_max, ld), self.fpad, dtype=torch.float32)
        y = torch.full((B, n_max), self.lpad, dtype=torch.int64)
        edge_parts, weight_parts, any_w = [], [], False
        for b, s in enumerate(samples):
            n, off = sizes[b], pad_offset(sizes[b], n_max)
            Vp[b, off:off + n, :D] = torch.as_tensor(np.asarray(s["textline_encoding"], dtype=np.float32))
            if "node_label" in s:
                y[b, off:off + n] = torch.as_tensor(np.asarray(s["node_label"]).astype(np.int64))
            if "edges" in s:
                tri, wts = np.asarray(s["edges"], dtype=np.int32).reshape(-1, 3), s.get("edge_weights")
            else:
                tri, wts = dense_to_edges(np.asarray(s["adjacency_matrix"]))
            if tri.shape[0] and (tri.min() < 0 or tri[:, 0].max() >= n or tri[:, 2].max() >= n or tri[:, 1].max() >= self.L):
                raise ValueError(f"sample {b}: edge triple (start, label, end) outside [0, {n}) x [0, {self.L}) x [0, {n})")
            e4 = np.empty((tri.shape[0], 4), dtype=np.int32)
            e4[:, 0] = b
            e4[:, 1] = tri[:, 0] + off
            e4[:, 2] = tri[:, 1]
            e4[:, 3] = tri[:, 2] + off
            edge_parts.append(e4)
            any_w = any_w or wts is not None
            weight_parts.append(wts)
        edges = torch.from_numpy(np.concatenate(edge_parts, axis=0) if edge_parts else np.zeros((0, 4), np.int32))
        weights = None
        if any_w:
            weights = torch.from_numpy(np.concatenate([
                (w if w is not None else np.ones(p.shape[0], np.float32)) for w, p in zip(weight_parts, edge_parts)
            ]).astype(np.float32))
        if self.pin:
            Vp = Vp.pin_memory()
        # "features_padded" is the storage (16-byte row pitch); "textline_encoding" the logical [B, N, D] view of it
        out = {"textline_encoding": Vp[:, :, :D], "features_padded": Vp, "node_label": y, "edges": edges,
               "edge_weights": weights, "graph_shape": (B, n_max, self.L)}
        if self.pin:
            out["node_label"] = y.pin_memory()
            out["edges"] = edges.pin_memory()
            if weights is not None:
                out["edge_weights"] = weights.pin_memory()
        return out

    @staticmethod
    def to_device(batch: Dict[str, Any], device: torch.device) -> Dict[str, Any]:
        """H2D (async from pinned memory) + CSR/CSC pack on the device; the result feeds ``model.forward([V, A])``."""
        B, N, L = batch["graph_shape"]
        D = batch["textline_encoding"].shape[-1]
        if isinstance(batch["textline_encoding"], SparseNodeFeatures):            # bag-of-words features (features.py)
            V = batch["textline_encoding"].to(device, non_blocking=True)
            V.build_transpose()
            V.identity_perm()
        else:
            V = batch["features_padded"].to(device, non_blocking=True)[:, :, :D]     # the pitch survives the copy
        y = batch["node_label"].to(device, non_blocking=True)
        edges = batch["edges"].to(device, non_blocking=True)
        wts = batch["edge_weights"]
        if wts is not None:
            wts = wts.to(device, non_blocking=True)
        g = PackedGraphBatch.from_edges(edges, B, N, L, weights=wts)
        return {"textline_encoding": V, "adjacency_matrix": g, "node_label": y}


COLLATES["EdgeListCollate"] = EdgeListCollate
