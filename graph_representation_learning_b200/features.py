"""Sparse node features for ``emb1``.

The reference's text-line encoding is a bag of characters: ``TextlineEncoding.get_bow_matrix``
(gnn/data_generator/data_process/textline_encoding.py:23-42) sets a handful of ones in ``input_dim - 4`` binary
columns and appends 4 spatial floats (:53-55), so a ``[B, N, 4369]`` fp32 batch carries ~11 non-zeros per row.
``SparseNodeFeatures`` ships exactly those: a CSR over the *light* columns (few rows use them) plus a small dense
block over the *heavy* columns (the spatial floats, very frequent characters).  ``GraphCNNDropEdge.forward`` accepts
it in place of the dense ``V``; ``emb1`` then runs as

    Z = V_heavy @ W[:, heavy]^T  +  sum_{e in row} val[e] * W^T[col[e], :]        (ops.sparse_linear_act)

i.e. a K = n_heavy tensor-core GEMM plus a gather-sum over the rows of ``W^T`` (csrc/spmm.cu, fp32 FMA in CSR
order), and its weight gradient as the same two pieces transposed (CSC built by csrc/csr_pack.cu).  The result
equals the dense path within fp32 rounding (tests/test_gpu_features.py).  Host-to-device traffic for cfg2 drops from
2.24 GB to 14 MB per batch.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib


class SparseNodeFeatures:
    """``[B, N, D]`` node features as CSR (light columns) + dense block (heavy columns).

    rowptr int32 [B*N+1], col int32 [nnz] (original column ids, ascending per row), val f32 [nnz];
    heavy_cols int64 [Dh] (ascending), heavy f32 [B*N, ld_h] (``ld_h`` = Dh rounded up to 4, zero padded);
    colptr / rowidx / perm: CSC of the light part (device only, built on demand).
    """

    def __init__(self, shape: Tuple[int, int, int], rowptr: torch.Tensor, col: torch.Tensor, val: torch.Tensor,
                 heavy_cols: torch.Tensor, heavy: torch.Tensor):
        self.shape = tuple(int(x) for x in shape)
        self.rowptr, self.col, self.val = rowptr, col, val
        self.heavy_cols, self.heavy = heavy_cols, heavy
        self.colptr: Optional[torch.Tensor] = None
        self.rowidx: Optional[torch.Tensor] = None
        self.perm: Optional[torch.Tensor] = None
        self._identity: Optional[torch.Tensor] = None

    # ------------------------------------------------------------------ construction
    @staticmethod
    def from_dense(V: torch.Tensor, heavy_fraction: float = 1.0 / 128, max_light_per_column: int = 1024
                   ) -> "SparseNodeFeatures":
        """Convert a dense ``[B, N, D]`` batch (host or device; a data-format step of the collate, plain tensor ops).
        A column is *heavy* when more than ``min(max_light_per_column, heavy_fraction * B*N)`` rows use it: heavy
        columns go to the dense block, so no CSC column is longer than that bound (the weight gradient walks CSC
        columns one warp each)."""
        if V.dim() != 3:
            raise _lib.GrlError(f"from_dense expects [B, N, D], got {tuple(V.shape)}")
        B, N, D = V.shape
        x = V.reshape(B * N, D)
        nz = x != 0
        per_col = nz.sum(dim=0)
        bound = min(max_light_per_column, max(1, int(heavy_fraction * B * N)))
        heavy_mask = per_col > bound
        heavy_cols = torch.nonzero(heavy_mask).flatten()
        Dh = int(heavy_cols.numel())
        ld_h = max(4, (Dh + 3) // 4 * 4)
        heavy = torch.zeros(B * N, ld_h, dtype=torch.float32, device=V.device)
        if Dh:
            heavy[:, :Dh] = x[:, heavy_cols].float()
        light = nz & ~heavy_mask
        rows, cols = torch.nonzero(light, as_tuple=True)               # row-major: ascending rows, then columns
        counts = torch.bincount(rows, minlength=B * N)
        rowptr = torch.zeros(B * N + 1, dtype=torch.int32, device=V.device)
        rowptr[1:] = torch.cumsum(counts, 0).to(torch.int32)
        val = x[rows, cols].float().contiguous()
        return SparseNodeFeatures((B, N, D), rowptr, cols.to(torch.int32).contiguous(), val, heavy_cols, heavy)

    def to_dense(self) -> torch.Tensor:
        B, N, D = self.shape
        out = torch.zeros(B * N, D, dtype=torch.float32, device=self.val.device)
        rows = torch.repeat_interleave(torch.arange(B * N, device=self.val.device),
                                       (self.rowptr[1:] - self.rowptr[:-1]).long())
        out[rows, self.col.long()] = self.val
        if self.heavy_cols.numel():
            out[:, self.heavy_cols] = self.heavy[:, :self.heavy_cols.numel()]
        return out.view(B, N, D)

    # ------------------------------------------------------------------ plumbing
    @property
    def device(self) -> torch.device:
        return self.val.device

    @property
    def nnz(self) -> int:
        return int(self.col.numel())

    @property
    def num_nodes(self) -> int:
        return self.shape[0] * self.shape[1]

    def nbytes(self) -> int:
        return sum(t.numel() * t.element_size() for t in (self.rowptr, self.col, self.val, self.heavy_cols, self.heavy))

    def tensors(self):
        return [t for t in (self.rowptr, self.col, self.val, self.heavy_cols, self.heavy, self.colptr, self.rowidx,
                            self.perm, self._identity) if t is not None]

    def pin_memory(self) -> "SparseNodeFeatures":
        return SparseNodeFeatures(self.shape, *(t.pin_memory() for t in (self.rowptr, self.col, self.val, self.heavy_cols,
                                                                         self.heavy)))

    def to(self, device, non_blocking: bool = False) -> "SparseNodeFeatures":
        return SparseNodeFeatures(self.shape, *(t.to(device, non_blocking=non_blocking)
                                                for t in (self.rowptr, self.col, self.val, self.heavy_cols, self.heavy)))

    def identity_perm(self) -> torch.Tensor:
        """perm of the forward gather (entries are read in CSR order)"""
        if self._identity is None:
            self._identity = torch.arange(self.nnz, dtype=torch.int32, device=self.device)
        return self._identity

    def build_transpose(self) -> None:
        """CSC of the light part (csrc/csr_pack.cu): colptr over the D feature columns, rowidx = node, perm = CSR slot"""
        if self.colptr is not None:
            return
        if self.device.type != "cuda":
            raise _lib.GrlError("SparseNodeFeatures.build_transpose runs on the GPU; there is no CPU path")
        lib = _lib.load()
        dev, D = self.device, self.shape[2]
        self.colptr = torch.empty(D + 1, dtype=torch.int32, device=dev)
        self.rowidx = torch.empty(max(self.nnz, 1), dtype=torch.int32, device=dev)
        self.perm = torch.empty(max(self.nnz, 1), dtype=torch.int32, device=dev)
        ws_bytes = lib.grl_csr_transpose_workspace_bytes(D, self.nnz)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_csr_transpose_general(self.rowptr.data_ptr(), self.col.data_ptr(), self.num_nodes, D, self.nnz,
                                                 self.colptr.data_ptr(), self.rowidx.data_ptr(), self.perm.data_ptr(),
                                                 ws.data_ptr(), ws_bytes, _lib.stream_ptr(dev)),
                   "grl_csr_transpose_general")
        _lib.count_launch(4)
