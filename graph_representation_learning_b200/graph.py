"""Device-resident packed batch of typed document graphs (CSR + CSC of the reference's ``A'``).

The reference keeps the batch adjacency dense -- ``B x N x L x N`` from the collate
(gnn/data_generator/base_dataloader.py:57) turned into ``A' = B x (L+1)N x N`` by
``GraphConv.preprocess_adj`` (gnn/models/networks/robust_gcn.py:53-72).  Here the same matrix is
held as CSR (forward gather) and CSC+perm (transposed gather for the gradient):

    row r = (b*N + i)*(L+1) + t     t = 0: self loop, t = l+1: edge type l
    col   = local node j of graph b (ascending inside a row)
    val   = adjacency value as fp32 (1.0 for binary graphs and for self loops)

All packing runs in the CUDA library (csrc/csr_pack.cu); nothing here touches the CPU.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import torch

from . import _lib


# Dense-adjacency packing without a per-batch host read (`rowptr[R].item()` serialised host and device in
# every training forward of the dense entry).  Training forwards only (eval packs exactly: its result goes to the host
# anyway).  The first batch of a (B, N, L, device) signature is counted exactly (one read) and fixes a slot capacity with
# 25 % head-room; later batches are packed into that capacity with clamped row pointers and their true count is copied
# to pinned memory behind an event.  ``check_dense_overflow()`` -- called by ``KVTrainStep`` between backward and the
# optimizer step, when that early event has long fired, and by the next pack of the signature -- raises if a batch did
# not fit (and grows the capacity), so a truncated batch never reaches the weights.
_dense_capacity: Dict[Tuple, int] = {}
_dense_pending: Dict[Tuple, Tuple] = {}


def check_dense_overflow() -> None:
    """raise if any sync-free dense pack since the last call was truncated (see above); cheap when nothing is pending"""
    for key in list(_dense_pending):
        _check_pending_overflow(key)


def _check_pending_overflow(key) -> None:
    pend = _dense_pending.pop(key, None)
    if pend is None:
        return
    ev, host, cap = pend
    ev.synchronize()
    true_nnz = int(host[0])
    if true_nnz > cap:
        _dense_capacity[key] = int(true_nnz * 1.25) + 1024
        raise _lib.GrlError(f"the previous dense adjacency batch held {true_nnz} entries but was packed into {cap} slots "
                            "(truncated); the capacity has been raised -- re-run the step, or pass nnz_capacity / use "
                            "EdgeListCollate batches, whose entry count is known on the host")


@dataclass
class PackedGraphBatch:
    B: int
    N: int
    L: int
    nnz: int                      # number of stored entries (or capacity when packed without a host sync)
    rowptr: torch.Tensor          # int32 [R+1]
    col: torch.Tensor             # int32 [nnz]
    val: torch.Tensor             # float32 [nnz]
    colptr: Optional[torch.Tensor] = None   # int32 [B*N+1]
    rowidx: Optional[torch.Tensor] = None   # int32 [nnz]   global CSR row of each CSC entry
    perm: Optional[torch.Tensor] = None     # int32 [nnz]   CSR position of each CSC entry
    _eval_weights: Optional[torch.Tensor] = None   # cached edge_weights(p=0) (val + its node streams)
    _typed_T: Optional["PackedGraphBatch"] = None  # typed_transpose()
    t_perm: Optional[torch.Tensor] = None          # (on a typed transpose) int32 [nnz] CSR position in the forward packing

    @property
    def num_nodes(self) -> int:
        return self.B * self.N

    @property
    def num_rows(self) -> int:
        return self.B * self.N * (self.L + 1)

    @property
    def device(self) -> torch.device:
        return self.rowptr.device

    # ------------------------------------------------------------------ builders
    @staticmethod
    def from_dense(adj: torch.Tensor, nnz_capacity: Optional[int] = None, with_transpose: bool = True,
                   self_loop: bool = True, sync_free: bool = False) -> "PackedGraphBatch":
        """Pack the collate's dense ``[B, N, L, N]`` adjacency (fp16 or fp32, layout [b, start, type, end]).

        ``nnz_capacity``: give an upper bound to skip the one host read of ``rowptr[R]``.
        ``sync_free`` (what ``GraphCNNDropEdge.forward`` uses for dense inputs): only the first batch of a
        ``(B, N, L)`` signature is counted on the host; later ones are packed into a remembered capacity (+25 %)
        with clamped row pointers, and an overflow is reported by the next pack (see ``_check_pending_overflow``).
        ``self_loop=False``: ``adj`` is an already preprocessed ``A'`` viewed as ``[B, N, L+1, N]`` (its
        identity block, possibly DropEdge-masked, is taken as data).
        """
        _lib.require_cuda(adj, "adjacency")
        if adj.dim() != 4 or adj.shape[1] != adj.shape[3]:
            raise _lib.GrlError(f"adjacency must be [B, N, L, N], got {tuple(adj.shape)}")
        if adj.dtype not in (torch.float16, torch.float32):
            adj = adj.float()
        adj = adj.contiguous()
        B, N, T, _ = adj.shape
        sl = 1 if self_loop else 0
        L = T if self_loop else T - 1
        lib = _lib.load()
        dev = adj.device
        st = _lib.stream_ptr(dev)
        R = B * N * (L + 1)
        rowptr = torch.empty(R + 1, dtype=torch.int32, device=dev)
        code = _lib.dtype_code(adj)
        # one read of the dense tensor: the count pass leaves a bitmap (1/16 of an fp16 A) that the fill pass walks
        bitmap = torch.empty(lib.grl_csr_bitmap_bytes(B, N, T), dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_csr_count_bitmap(adj.data_ptr(), code, B, N, T, sl, rowptr.data_ptr(), bitmap.data_ptr(), st),
                   "grl_csr_count_bitmap")
        ws_bytes = lib.grl_scan_workspace_bytes(R)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_scan_inclusive_i32(rowptr.data_ptr() + 4, R, ws.data_ptr(), ws_bytes, st),
                   "grl_scan_inclusive_i32")
        _lib.count_launch(2)
        exact = False
        key = (B, N, T, sl, dev.index)
        if nnz_capacity is None and sync_free and not torch.cuda.is_current_stream_capturing():
            _check_pending_overflow(key)
        cap = _dense_capacity.get(key) if (nnz_capacity is None and sync_free) else nnz_capacity
        if cap is None:
            nnz = int(rowptr[-1].item())                      # counted exactly: one host read
            exact = True
            if sync_free:                                     # first batch of this signature fixes the capacity
                _dense_capacity[key] = int(nnz * 1.25) + 1024
        elif nnz_capacity is not None:
            nnz = int(nnz_capacity)
        else:
            nnz = int(cap)
            nnz_dev = torch.empty(1, dtype=torch.int32, device=dev)
            _lib.check(lib.grl_csr_clamp_rowptr(rowptr.data_ptr(), R, nnz, nnz_dev.data_ptr(), st), "grl_csr_clamp_rowptr")
            _lib.count_launch(1)
            if not torch.cuda.is_current_stream_capturing():
                host = torch.empty(1, dtype=torch.int32, pin_memory=True)
                host.copy_(nnz_dev, non_blocking=True)
                ev = torch.cuda.Event()
                ev.record()
                _dense_pending[key] = (ev, host, nnz)
        col = torch.empty(nnz, dtype=torch.int32, device=dev)
        val = torch.empty(nnz, dtype=torch.float32, device=dev) if exact else \
            torch.zeros(nnz, dtype=torch.float32, device=dev)
        _lib.check(lib.grl_csr_fill_bitmap(adj.data_ptr(), code, bitmap.data_ptr(), B, N, T, sl, rowptr.data_ptr(),
                                           col.data_ptr(), val.data_ptr(), st), "grl_csr_fill_bitmap")
        _lib.count_launch(1)
        g = PackedGraphBatch(B, N, L, nnz, rowptr, col, val)
        if with_transpose:
            g.prepare_backward()
        return g

    @staticmethod
    def from_edges(edges: torch.Tensor, B: int, N: int, L: int, weights: Optional[torch.Tensor] = None,
                   with_transpose: bool = True) -> "PackedGraphBatch":
        """Pack an edge list ``int32 [E, 4] = (b, i, l, j)`` (node indices already include the collate's pad
        offset).  Self loops are added here, exactly once per node, as type-0 rows.

        Repeated ``(b, i, l, j)`` triples keep one slot each but only the first (largest weight) carries a value, the
        others 0: the aggregation equals the reference's idempotent dense build ``adj[start, label, end] = w``
        (graph_utils.py:821-829).  Triples outside the batch are neutralised on the device (value 0, no out-of-bounds
        write); ``collate.EdgeListCollate`` range-checks on the host and raises."""
        _lib.require_cuda(edges, "edges")
        if edges.dim() != 2 or edges.shape[1] != 4:
            raise _lib.GrlError(f"edges must be [E, 4], got {tuple(edges.shape)}")
        edges = edges.to(torch.int32).contiguous()
        E = edges.shape[0]
        if weights is not None:
            weights = weights.to(device=edges.device, dtype=torch.float32).contiguous()
        lib = _lib.load()
        dev = edges.device
        st = _lib.stream_ptr(dev)
        R = B * N * (L + 1)
        rowptr = torch.zeros(R + 1, dtype=torch.int32, device=dev)
        _lib.check(lib.grl_coo_count(_lib.ptr(edges) if E else 0, E, B, N, L, rowptr.data_ptr(), st), "grl_coo_count")
        ws_bytes = lib.grl_scan_workspace_bytes(R)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_scan_inclusive_i32(rowptr.data_ptr() + 4, R, ws.data_ptr(), ws_bytes, st),
                   "grl_scan_inclusive_i32")
        nnz = E + B * N
        col = torch.empty(nnz, dtype=torch.int32, device=dev)
        val = torch.empty(nnz, dtype=torch.float32, device=dev)
        cursor = torch.empty(R, dtype=torch.int32, device=dev)
        _lib.check(lib.grl_coo_fill(_lib.ptr(edges) if E else 0, _lib.ptr(weights), E, B, N, L, rowptr.data_ptr(),
                                    cursor.data_ptr(), col.data_ptr(), val.data_ptr(), st), "grl_coo_fill")
        _lib.count_launch(5)
        g = PackedGraphBatch(B, N, L, nnz, rowptr, col, val)
        if with_transpose:
            g.prepare_backward()
        return g

    def prepare_backward(self) -> None:
        """Build now what the backward aggregation will read: the typed transpose (aggregate-then-project form,
        ops._GraphConv) when the model's edge types fit its kernels, else the CSC of the legacy transposed SpMM.  Both are
        also built on first use."""
        if self.L + 1 <= 31:
            self.typed_transpose()
        else:
            self.build_transpose()

    def build_transpose(self) -> None:
        """CSC + perm for the transposed SpMM of the backward pass (csrc/csr_pack.cu)."""
        if self.colptr is not None:
            return
        lib = _lib.load()
        dev = self.device
        n_nodes = self.num_nodes
        self.colptr = torch.empty(n_nodes + 1, dtype=torch.int32, device=dev)
        self.rowidx = torch.empty(self.nnz, dtype=torch.int32, device=dev)
        self.perm = torch.empty(self.nnz, dtype=torch.int32, device=dev)
        ws_bytes = lib.grl_csr_transpose_workspace_bytes(n_nodes, self.nnz)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        _lib.check(lib.grl_csr_transpose(self.rowptr.data_ptr(), self.col.data_ptr(), self.B, self.N, self.L,
                                         self.nnz, self.colptr.data_ptr(), self.rowidx.data_ptr(),
                                         self.perm.data_ptr(), ws.data_ptr(), ws_bytes, _lib.stream_ptr(dev)),
                   "grl_csr_transpose")
        _lib.count_launch(4)

    def typed_transpose(self) -> "PackedGraphBatch":
        """The batch with every typed adjacency transposed, A_t^T: rows (j, t) list the source nodes i of the entries
        (i, t) -> j.  Built once per batch from the CSC (csrc/csr_pack.cu csc_typed_kernel); ``t_perm`` maps its
        entries to CSR positions of this packing.  The backward aggregation U = A_t^T dZ is a FORWARD aggregation over
        it (ops._GraphConv: dX = sum_t (A_t^T dZ) W_t^T, autograd of robust_gcn.py:45-50)."""
        if self._typed_T is not None:
            return self._typed_T
        dev = self.device
        lib = _lib.load()
        t_rowptr = torch.empty(self.num_rows + 1, dtype=torch.int32, device=dev)
        t_col = torch.empty(self.nnz, dtype=torch.int32, device=dev)
        t_perm = torch.empty(self.nnz, dtype=torch.int32, device=dev)
        rc = _lib.GRL_E_UNSUP
        if self.direct_typed_transpose and self.colptr is None:
            # one CTA per graph, counters in shared memory (csrc/typed_transpose.cu); graphs too large for that take the CSC
            rc = lib.grl_typed_transpose(self.rowptr.data_ptr(), self.col.data_ptr(), self.B, self.N, self.L,
                                         t_rowptr.data_ptr(), t_col.data_ptr(), t_perm.data_ptr(), _lib.stream_ptr(dev))
            if rc != _lib.GRL_E_UNSUP:
                _lib.check(rc, "grl_typed_transpose")
                _lib.count_launch(1)
        if rc == _lib.GRL_E_UNSUP:
            self._typed_from_csc(t_rowptr, t_col, t_perm)
        gT = PackedGraphBatch(self.B, self.N, self.L, self.nnz, t_rowptr, t_col, self.val)
        gT.t_perm = t_perm
        self._typed_T = gT
        return gT

    direct_typed_transpose = True      # tests switch it off to compare the two builds

    def _typed_from_csc(self, t_rowptr, t_col, t_perm) -> None:
        self.build_transpose()
        dev = self.device
        _lib.check(_lib.load().grl_csc_typed_regroup(self.colptr.data_ptr(), self.rowidx.data_ptr(), self.perm.data_ptr(),
                                                     self.B, self.N, self.L, t_rowptr.data_ptr(), t_col.data_ptr(),
                                                     t_perm.data_ptr(), _lib.stream_ptr(dev)), "grl_csc_typed_regroup")
        _lib.count_launch(1)

    def transposed_weights(self, w: torch.Tensor) -> torch.Tensor:
        """per-entry weights ``w`` of this packing (one GraphConv layer's DropEdge weights) in the entry order of
        ``typed_transpose()``, with the node streams of the forward slab kernel attached (grl_stream_from_weights)"""
        gT = self.typed_transpose()
        wT = torch.empty_like(w)
        if self.L + 1 <= 31 and self.N < (1 << 24) - 2:
            ew = torch.empty(max(self.nnz, 1), 2, dtype=torch.int32, device=self.device)
            nodeinfo = torch.empty(self.num_nodes, 2, dtype=torch.int32, device=self.device)
            _lib.check(_lib.load().grl_stream_from_weights(gT.rowptr.data_ptr(), gT.col.data_ptr(), gT.t_perm.data_ptr(),
                                                           w.data_ptr(), self.num_nodes, self.L, wT.data_ptr(),
                                                           ew.data_ptr(), nodeinfo.data_ptr(), _lib.stream_ptr(self.device)),
                       "grl_stream_from_weights")
            _lib.count_launch(1)
            wT._grl_stream = (nodeinfo, ew)
        else:
            torch.index_select(w, 0, gT.t_perm.long(), out=wT)
        return wT

    # ------------------------------------------------------------------ DropEdge
    def edge_weights(self, p: float, seed: int = 0, stream_id: int = 0,
                     mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Per-entry weights of one layer: ``val * keep / (1-p)`` (reference: nn.Dropout(p=0.3) on A',
        gnn/models/networks/drop_robust_gcn.py:38,76,80,85).  ``mask`` (uint8 [nnz], CSR order) injects
        the keep bits for parity tests; otherwise Philox(seed, stream_id, entry)."""
        if p == 0.0 and mask is None and self._eval_weights is not None:
            return self._eval_weights
        if mask is not None:
            _lib.require_cuda(mask, "edge mask")
            if mask.dtype != torch.uint8 or mask.numel() != self.nnz:
                raise _lib.GrlError("edge mask must be uint8 with one entry per stored non-zero")
            mask = mask.contiguous()
        w = torch.empty_like(self.val)
        if self.L + 1 <= 31 and self.N < (1 << 24) - 2:      # one lane per row pointer of a node (csrc/spmm_slab2.cu)
            # forward fast path: the same w plus the compacted, row-tagged node streams of csrc/spmm_slab2.cu
            ew = torch.empty(max(self.nnz, 1), 2, dtype=torch.int32, device=self.device)
            nodeinfo = torch.empty(self.num_nodes, 2, dtype=torch.int32, device=self.device)
            _lib.check(_lib.load().grl_dropedge_stream(self.rowptr.data_ptr(), self.col.data_ptr(), self.val.data_ptr(),
                                                       _lib.ptr(mask), self.num_nodes, self.L, float(p),
                                                       int(seed) & (2**64 - 1), int(stream_id), w.data_ptr(),
                                                       ew.data_ptr(), nodeinfo.data_ptr(), _lib.stream_ptr(self.device)),
                       "grl_dropedge_stream")
            w._grl_stream = (nodeinfo, ew)       # rides along with the weight tensor (ops.spmm_fwd_ looks for it)
        else:
            _lib.check(_lib.load().grl_dropedge_weights(self.val.data_ptr(), _lib.ptr(mask), self.nnz, float(p),
                                                        int(seed) & (2**64 - 1), int(stream_id), w.data_ptr(),
                                                        _lib.stream_ptr(self.device)), "grl_dropedge_weights")
        _lib.count_launch(1)
        if p == 0.0 and mask is None:
            self._eval_weights = w               # eval: one stream per batch, shared by the three layers
        return w

    def to_dense(self, w: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Dense ``A'`` ``[B, (L+1)N, N]`` fp32 (what the reference's preprocess_adj returns); tests / debugging."""
        R = self.num_rows
        counts = (self.rowptr[1:] - self.rowptr[:-1]).long()
        rows = torch.repeat_interleave(torch.arange(R, device=self.device), counts)
        n = rows.numel()
        out = torch.zeros(R, self.N, dtype=torch.float32, device=self.device)
        out.index_put_((rows, self.col.long()[:n]), (self.val if w is None else w)[:n], accumulate=True)
        return out.view(self.B, self.N * (self.L + 1), self.N)

    def gather_dense_mask(self, dense_mask: torch.Tensor) -> torch.Tensor:
        """Test helper: pick the keep bits of a reference-style dense mask ``[B, (L+1)N, N]`` at the stored
        entries (CSR order) -> uint8 [nnz]."""
        R = self.num_rows
        counts = (self.rowptr[1:] - self.rowptr[:-1]).long()
        rows = torch.repeat_interleave(torch.arange(R, device=self.device), counts)
        rows_per_graph = self.N * (self.L + 1)
        b = rows // rows_per_graph
        rr = rows % rows_per_graph
        return (dense_mask[b, rr, self.col.long()[: rows.numel()]] != 0).to(torch.uint8)
