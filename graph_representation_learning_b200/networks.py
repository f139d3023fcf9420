"""Drop-in ``nn.Module`` mirror of the reference's GraphCNNDropEdge stack, running on the B200 kernels.

Reference interface mirrored here (same constructor arguments, sub-module names, parameter shapes, init
order -- so ``torch.manual_seed(s)`` gives identical initial weights -- and ``state_dict`` keys):

* ``make_linear_relu``            gnn/models/networks/robust_gcn.py:10-11
* ``GraphConv``                   gnn/models/networks/robust_gcn.py:14-75
* ``NodeSelfAtten``               gnn/models/networks/robust_gcn.py:78-99
* ``RanPACLayer``                 gnn/models/networks/drop_robust_gcn.py:13-28
* ``GraphCNNDropEdge``            gnn/models/networks/drop_robust_gcn.py:31-103
* ``BaseNetwork``                 gnn/models/base_network.py:9-56

What differs is *how* forward/backward run: the dense ``B x (L+1)N x N`` adjacency and ``nn.Dropout`` on it
are replaced by a packed CSR/CSC batch (graph.py) and the kernels behind ``ops``.  All compute requires a
CUDA device; there is no CPU path (the reference itself is the CPU implementation).
"""
from __future__ import annotations

import logging
from typing import Any, Dict, Optional, Sequence, Union

import torch
from torch import nn

from . import _lib, ops
from .features import SparseNodeFeatures
from .graph import PackedGraphBatch

RP_FACTOR = 10  # drop_robust_gcn.py:10

logger = logging.getLogger(__name__)

_REGISTRY: Dict[str, type] = {}


def register_model(cls):
    _REGISTRY[cls.__name__] = cls
    return cls


class BaseNetwork(nn.Module):
    """gnn/models/base_network.py:9-56 (factory by class name + parameter counter)."""

    def __init__(self):
        super().__init__()
        self.logger = logger

    @classmethod
    def _from_config(cls, config: Dict[str, Any]) -> "BaseNetwork":
        model_class = _REGISTRY.get(config["type"])
        if model_class is None:
            raise KeyError(f"Cannot find {config['type']} class. Registered models: {sorted(_REGISTRY)}.")
        if cls is BaseNetwork:
            return model_class._from_config(config)
        args = config.get("args", {})
        try:
            model = model_class(**args)
        except TypeError as er:
            raise TypeError(f"{er}. Check `args` fields defined in config with the actual keyword args "
                            f"required in {model_class.__name__} `__init__` method.")
        model.logger.info(f"Num parameters of {model.__class__.__name__}: {model._count_parameters()}")
        return model

    def _count_parameters(self) -> str:
        return f"{sum(pp.numel() for pp in self.parameters() if pp.requires_grad):,}"

    def forward(self):
        raise NotImplementedError


def make_linear_relu(input_dim: int, output_dim: int) -> nn.Sequential:
    return nn.Sequential(nn.Linear(input_dim, output_dim), nn.ReLU())


def _flat(V: torch.Tensor) -> torch.Tensor:
    return V.reshape(-1, V.shape[-1])


class GraphConv(nn.Module):
    """Multi-relational graph convolution: aggregate over (L+1) typed adjacencies, then one projection.

    ``h_weights`` is ``[(L+1)*F, C]`` (type-major blocks of F rows, block 0 = self loop) and ``bias`` ``[C]``,
    initialised exactly like the reference (xavier_normal_ / normal_(1e-4, 5e-5), robust_gcn.py:22-30).
    """

    def __init__(self, input_dim: int, output_dim: int, num_edges: int, with_bias: bool = True):
        super().__init__()
        self.C, self.L, self.F = output_dim, num_edges, input_dim
        self.gpu = torch.cuda.is_available()
        self.h_weights = nn.Parameter(torch.empty(self.F * (self.L + 1), self.C))
        self.bias = nn.Parameter(torch.empty(self.C)) if with_bias else None
        nn.init.xavier_normal_(self.h_weights)
        if self.bias is not None:
            nn.init.normal_(self.bias, mean=0.0001, std=0.00005)

    def forward(self, V: torch.Tensor, A: Union[torch.Tensor, PackedGraphBatch], preprocess_A: bool = True,
                edge_weights: Optional[torch.Tensor] = None) -> torch.Tensor:
        """V: B x N x F.  A: ``PackedGraphBatch`` (native) -- or, for call-compatibility with the reference,
        the dense ``B x N x N x L`` tensor (``preprocess_A=True``) or the dense preprocessed
        ``B x (L+1)N x N`` tensor, possibly already DropEdge-masked (``preprocess_A=False``)."""
        B, N = V.shape[0], V.shape[1]
        if isinstance(A, PackedGraphBatch):
            g = A
        elif preprocess_A:
            g = self.preprocess_adj(A)
        else:
            g = PackedGraphBatch.from_dense(A.reshape(B, N, self.L + 1, N), self_loop=False)
        w = g.edge_weights(0.0) if edge_weights is None else edge_weights
        out = ops.graph_conv([_flat(V)], g, w, self.h_weights, self.bias)
        return out.view(B, N, self.C)

    def preprocess_adj(self, adj: torch.Tensor) -> PackedGraphBatch:
        """adj: dense ``B x N x N x L`` ([b, i, j, l], i.e. after the model's ``A.permute(0, 1, 3, 2)``).
        Returns the packed ``A'`` (``.to_dense()`` gives the reference's ``B x (L+1)N x N`` tensor)."""
        return PackedGraphBatch.from_dense(adj.permute(0, 1, 3, 2))

    def __repr__(self):
        return f"GraphConv(in_dim={self.F}, out_dim={self.C}, num_edges={self.L}, bias={self.bias is not None})"


class NodeSelfAtten(nn.Module):
    """Dense per-graph self attention, unscaled and unmasked: softmax(relu(Vf) relu(Vg)^T) relu(Vh) * gamma + V."""

    def __init__(self, input_dim: int):
        super().__init__()
        self.F = input_dim
        self.f = make_linear_relu(input_dim, int(self.F // 8))
        self.g = make_linear_relu(input_dim, int(self.F // 8))
        self.h = make_linear_relu(input_dim, self.F)
        self.softmax = torch.nn.Softmax(-1)
        self.gamma = nn.Parameter(torch.empty(input_dim))
        torch.nn.init.normal_(self.gamma)

    def forward(self, V: torch.Tensor) -> torch.Tensor:
        B, N, H = V.shape
        x = _flat(V)
        f_out = ops.linear_act([x], self.f[0].weight, self.f[0].bias, relu=True).view(B, N, -1)
        g_out = ops.linear_act([x], self.g[0].weight, self.g[0].bias, relu=True).view(B, N, -1)
        h_out = ops.linear_act([x], self.h[0].weight, self.h[0].bias, relu=True).view(B, N, H)
        return ops.self_attention(f_out, g_out, h_out, self.gamma, V)

    def __repr__(self):
        return f"NodeSelfAttention(input_dim={self.F})"


class RanPACLayer(nn.Module):
    """Frozen random projection (weights ~ N(0,1), requires_grad=False, still part of the state dict)."""

    def __init__(self, input_dim: int, output_dim: int, lambda_value: Optional[float] = None):
        super().__init__()
        self.projection = nn.Linear(input_dim, output_dim, bias=False)
        for param in self.projection.parameters():
            param.requires_grad = False
        nn.init.normal_(self.projection.weight, mean=0, std=1.0)
        if lambda_value is not None:        # rp_robust_gcn.py:13-24: N(0,1) scaled by sqrt(output_dim) * lambda
            with torch.no_grad():
                self.projection.weight *= float(output_dim) ** 0.5 * lambda_value

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.linear_act([_flat(x)], self.projection.weight, None).view(*x.shape[:-1], -1)


def _pack_dense(A: torch.Tensor, training: bool) -> PackedGraphBatch:
    """CSR pack of a dense [B, N, L, N] batch adjacency; the transposes of the backward only when a backward can follow"""
    grad = torch.is_grad_enabled()
    return PackedGraphBatch.from_dense(A, sync_free=training and grad, with_transpose=grad)


@register_model
class GraphCNNDropEdge(BaseNetwork):
    def __init__(self, input_dim: int, output_dim: int, num_edges: int, net_size: int = 256,
                 use_attention: bool = True):
        super().__init__()
        self.output_dim = output_dim
        self.net_size = net_size
        self.num_edges = num_edges
        self.emb1 = make_linear_relu(input_dim, self.net_size)
        self.dropout = nn.Dropout(p=0.5)
        self.edge_dropout = nn.Dropout(p=0.3)
        self.gcn1 = GraphConv(self.net_size, self.net_size, num_edges)
        self.gcn2 = GraphConv(self.net_size, self.net_size, num_edges)
        self.gcn3 = GraphConv(self.net_size * 2, self.net_size, num_edges)
        half_net_size = self.net_size // 2
        self.emb2 = make_linear_relu(self.net_size * 2, half_net_size)
        self.use_attention = use_attention
        if use_attention:
            self.self_atten = NodeSelfAtten(half_net_size)
        rp_size = half_net_size * RP_FACTOR
        self.w_rand = RanPACLayer(half_net_size, rp_size)
        self.classifier = nn.Linear(rp_size, output_dim)
        # B200-path state (not parameters, not in the state dict)
        self.compute_dtype = torch.float32
        self.rng_seed: Optional[int] = None   # None -> torch.initial_seed() at call time
        self._rng_calls = 0

    # ------------------------------------------------------------------ B200-path knobs
    def set_compute_dtype(self, dtype: torch.dtype) -> "GraphCNNDropEdge":
        """fp32 (reference numerics) or bf16 activations with fp32 master weights / fp32 accumulation."""
        if dtype not in (torch.float32, torch.bfloat16):
            raise _lib.GrlError(f"compute dtype must be float32 or bfloat16, got {dtype}")
        self.compute_dtype = dtype
        return self

    def _next_streams(self):
        seed = torch.initial_seed() if self.rng_seed is None else self.rng_seed
        base = self._rng_calls * 16
        self._rng_calls += 1
        return seed, base

    # ------------------------------------------------------------------ forward
    def forward(self, inputs: Sequence[Any], efficient_mode: bool = True, *,
                masks: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        """inputs = (V, A): V ``B x N x input_dim`` (or the same batch as ``SparseNodeFeatures``); A the collate's dense ``B x N x L x N`` adjacency
        ([b, start, type, end], fp16/fp32) or a ``PackedGraphBatch``.  Returns fp32 logits ``B x N x output_dim``.

        ``efficient_mode`` is accepted for signature compatibility (the adjacency is always packed once).
        ``masks`` (tests only) injects keep-bits: {"edge": [3 x uint8[nnz]], "feat": [5 x uint8[B*N, C]]}.
        """
        V, A = inputs
        sparse_v = isinstance(V, SparseNodeFeatures)
        _lib.require_cuda(V.val if sparse_v else V, "V")
        B, N = V.shape[0], V.shape[1]
        g = A if isinstance(A, PackedGraphBatch) else _pack_dense(A, self.training)
        if g.B != B or g.N != N or g.L != self.num_edges:
            raise _lib.GrlError(f"graph batch (B={g.B}, N={g.N}, L={g.L}) does not match V {tuple(V.shape)} / "
                                f"num_edges={self.num_edges}")
        train = self.training
        p_feat = self.dropout.p if train else 0.0
        p_edge = self.edge_dropout.p if train else 0.0
        seed, sid = self._next_streams()
        em = masks["edge"] if masks else [None] * 3
        fm = masks["feat"] if masks else [None] * 5

        lin = self.emb1[0]
        if sparse_v:     # bag-of-words features as CSR + heavy-column block (features.py): gather-sum instead of the GEMM
            emb = ops.sparse_linear_act(V, lin.weight, lin.bias, relu=True, p=p_feat, mask=fm[0], seed=seed,
                                        stream_id=sid + 0, compute_dtype=self.compute_dtype)
        else:
            x = _flat(V)
            if x.dtype != self.compute_dtype:
                x = x.to(self.compute_dtype)
            emb = ops.linear_act([x], lin.weight, lin.bias, relu=True, p=p_feat, mask=fm[0], seed=seed, stream_id=sid + 0)

        w1 = g.edge_weights(p_edge, seed, sid + 1, em[0])
        g1 = ops.graph_conv([emb], g, w1, self.gcn1.h_weights, self.gcn1.bias, relu=True, p=p_feat, mask=fm[1],
                            seed=seed, stream_id=sid + 2)
        w2 = g.edge_weights(p_edge, seed, sid + 3, em[1])
        g2 = ops.graph_conv([g1], g, w2, self.gcn2.h_weights, self.gcn2.bias, relu=True, p=p_feat, mask=fm[2],
                            seed=seed, stream_id=sid + 4)
        w3 = g.edge_weights(p_edge, seed, sid + 5, em[2])
        g3 = ops.graph_conv([g1, g2], g, w3, self.gcn3.h_weights, self.gcn3.bias, relu=True, p=p_feat, mask=fm[3],
                            seed=seed, stream_id=sid + 6)

        lin = self.emb2[0]
        new_v = ops.linear_act([g1, g3], lin.weight, lin.bias, relu=True)
        if self.use_attention:
            new_v = _flat(self.self_atten(new_v.view(B, N, -1)))
        wr = self.w_rand.projection.weight
        new_v = ops.linear_act([new_v], wr, None, relu=True, p=p_feat, mask=fm[4], seed=seed, stream_id=sid + 7)
        logits = ops.linear_act([new_v], self.classifier.weight, self.classifier.bias)
        return logits.float().view(B, N, self.output_dim)


class _FeatureDropout(torch.autograd.Function):
    """nn.Dropout(p) without an activation in front (RobustGCN drops the attention output, robust_gcn.py:222): the keep
    factors come from the same Philox / injected-mask epilogue kernel as everywhere else, applied to ones."""

    @staticmethod
    def forward(ctx, x, p, mask, seed, stream_id):
        k = torch.empty_like(x)
        ops.epilogue_fwd_(torch.ones_like(x), None, k, False, p, mask, seed, stream_id)
        ctx.save_for_backward(k)
        return x * k

    @staticmethod
    def backward(ctx, dy):
        (k,) = ctx.saved_tensors
        return dy * k, None, None, None, None


@register_model
class RobustGCN(BaseNetwork):
    """gnn/models/networks/robust_gcn.py:173-231 -- the sibling of GraphCNNDropEdge without DropEdge and random
    projection, with two more GraphConv layers behind the attention.  Same kernels: one packed graph per forward, virtual
    concats ([g2, g1], [g3, g1] in the reference's order), fused bias / ReLU / dropout epilogues."""

    def __init__(self, input_dim: int, output_dim: int, num_edges: int, net_size: int = 256, use_attention: bool = True):
        super().__init__()
        self.output_dim, self.net_size, self.num_edges = output_dim, net_size, num_edges
        self.emb1 = make_linear_relu(input_dim, net_size)
        self.dropout = nn.Dropout(p=0.5)
        self.gcn1 = GraphConv(net_size, net_size, num_edges)
        self.gcn2 = GraphConv(net_size, net_size, num_edges)
        self.gcn3 = GraphConv(net_size * 2, net_size, num_edges)
        half = net_size // 2
        self.emb2 = make_linear_relu(net_size * 2, half)
        self.use_attention = use_attention
        if use_attention:
            self.self_atten = NodeSelfAtten(half)
        self.gcn4 = GraphConv(half, half, num_edges)
        self.gcn5 = GraphConv(half, half, num_edges)
        self.classifier = nn.Linear(half, output_dim)
        self.compute_dtype = torch.float32
        self.rng_seed: Optional[int] = None
        self._rng_calls = 0

    def forward(self, inputs: Sequence[Any], efficient_mode: bool = True, *,
                masks: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        V, A = inputs
        _lib.require_cuda(V, "V")
        B, N = V.shape[0], V.shape[1]
        g = A if isinstance(A, PackedGraphBatch) else _pack_dense(A, self.training)
        if g.B != B or g.N != N or g.L != self.num_edges:
            raise _lib.GrlError(f"graph batch (B={g.B}, N={g.N}, L={g.L}) does not match V {tuple(V.shape)} / "
                                f"num_edges={self.num_edges}")
        p = self.dropout.p if self.training else 0.0
        seed = torch.initial_seed() if self.rng_seed is None else self.rng_seed
        sid = self._rng_calls * 16
        self._rng_calls += 1
        fm = masks["feat"] if masks else [None] * 6
        w = g.edge_weights(0.0)                                  # no DropEdge in this model
        lin = self.emb1[0]
        emb = ops.linear_act([_flat(V).float()], lin.weight, lin.bias, relu=True, p=p, mask=fm[0], seed=seed, stream_id=sid)
        g1 = ops.graph_conv([emb], g, w, self.gcn1.h_weights, self.gcn1.bias, relu=True, p=p, mask=fm[1], seed=seed,
                            stream_id=sid + 1)
        g2 = ops.graph_conv([g1], g, w, self.gcn2.h_weights, self.gcn2.bias, relu=True, p=p, mask=fm[2], seed=seed,
                            stream_id=sid + 2)
        g3 = ops.graph_conv([g2, g1], g, w, self.gcn3.h_weights, self.gcn3.bias, relu=True, p=p, mask=fm[3], seed=seed,
                            stream_id=sid + 3)
        lin = self.emb2[0]
        new_v = ops.linear_act([g3, g1], lin.weight, lin.bias, relu=True)
        if self.use_attention:
            new_v = _flat(self.self_atten(new_v.view(B, N, -1)))
        if p > 0.0:
            new_v = _FeatureDropout.apply(new_v.contiguous(), p, fm[4], seed, sid + 4)
        g4 = ops.graph_conv([new_v], g, w, self.gcn4.h_weights, self.gcn4.bias, relu=True, p=p, mask=fm[5], seed=seed,
                            stream_id=sid + 5)
        g5 = ops.graph_conv([g4], g, w, self.gcn5.h_weights, self.gcn5.bias, relu=True)
        logits = ops.linear_act([g5], self.classifier.weight, self.classifier.bias)
        return logits.float().view(B, N, self.output_dim)


@register_model
class RPGraphCNNDropEdge(BaseNetwork):
    """gnn/models/networks/rp_robust_gcn.py:29-115 -- GraphCNNDropEdge with the random projection in FRONT of the
    attention: emb2 -> frozen ``rp_emb`` (H -> rp_size, leaky ReLU) -> attention over rp_size -> frozen ``rp_final``
    (rp_size -> rp_size, leaky ReLU, dropout) -> classifier.  Same kernels as GraphCNNDropEdge; leaky ReLU (slope 0.01)
    is the ``GRL_EPI_LEAKY`` epilogue of the GEMMs."""

    def __init__(self, input_dim: int, output_dim: int, num_edges: int, net_size: int = 256, use_attention: bool = True,
                 rp_size: int = 10000, lambda_value: float = 0.05):
        super().__init__()
        self.output_dim, self.net_size, self.num_edges = output_dim, net_size, num_edges
        self.rp_size, self.lambda_value, self.use_attention = rp_size, lambda_value, use_attention
        self.emb1 = make_linear_relu(input_dim, net_size)
        self.dropout = nn.Dropout(p=0.5)
        self.edge_dropout = nn.Dropout(p=0.3)
        self.gcn1 = GraphConv(net_size, net_size, num_edges)
        self.gcn2 = GraphConv(net_size, net_size, num_edges)
        self.gcn3 = GraphConv(net_size * 2, net_size, num_edges)
        self.half_net_size = net_size // 2
        self.emb2 = make_linear_relu(net_size * 2, self.half_net_size)
        self.rp_emb = RanPACLayer(self.half_net_size, rp_size, lambda_value)
        if use_attention:
            self.self_atten = NodeSelfAtten(rp_size)
        self.rp_final = RanPACLayer(rp_size, rp_size, lambda_value)
        self.classifier = nn.Linear(rp_size, output_dim)
        self.rng_seed: Optional[int] = None
        self._rng_calls = 0

    def forward(self, inputs: Sequence[Any], efficient_mode: bool = True, *,
                masks: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        """Same inputs as GraphCNNDropEdge.forward; ``masks`` = {"edge": 3 x uint8[nnz], "feat": 5 x uint8[B*N, C]}
        in call order (emb1, g1, g2, g3, rp_final)."""
        V, A = inputs
        sparse_v = isinstance(V, SparseNodeFeatures)
        _lib.require_cuda(V.val if sparse_v else V, "V")
        B, N = V.shape[0], V.shape[1]
        g = A if isinstance(A, PackedGraphBatch) else _pack_dense(A, self.training)
        if g.B != B or g.N != N or g.L != self.num_edges:
            raise _lib.GrlError(f"graph batch (B={g.B}, N={g.N}, L={g.L}) does not match V {tuple(V.shape)} / "
                                f"num_edges={self.num_edges}")
        p_feat = self.dropout.p if self.training else 0.0
        p_edge = self.edge_dropout.p if self.training else 0.0
        seed = torch.initial_seed() if self.rng_seed is None else self.rng_seed
        sid = self._rng_calls * 16
        self._rng_calls += 1
        em = masks["edge"] if masks else [None] * 3
        fm = masks["feat"] if masks else [None] * 5
        lin = self.emb1[0]
        if sparse_v:
            emb = ops.sparse_linear_act(V, lin.weight, lin.bias, relu=True, p=p_feat, mask=fm[0], seed=seed, stream_id=sid)
        else:
            emb = ops.linear_act([_flat(V).float()], lin.weight, lin.bias, relu=True, p=p_feat, mask=fm[0], seed=seed,
                                 stream_id=sid)
        w1 = g.edge_weights(p_edge, seed, sid + 1, em[0])
        g1 = ops.graph_conv([emb], g, w1, self.gcn1.h_weights, self.gcn1.bias, relu=True, p=p_feat, mask=fm[1],
                            seed=seed, stream_id=sid + 2)
        w2 = g.edge_weights(p_edge, seed, sid + 3, em[1])
        g2 = ops.graph_conv([g1], g, w2, self.gcn2.h_weights, self.gcn2.bias, relu=True, p=p_feat, mask=fm[2],
                            seed=seed, stream_id=sid + 4)
        w3 = g.edge_weights(p_edge, seed, sid + 5, em[2])
        g3 = ops.graph_conv([g1, g2], g, w3, self.gcn3.h_weights, self.gcn3.bias, relu=True, p=p_feat, mask=fm[3],
                            seed=seed, stream_id=sid + 6)
        lin = self.emb2[0]
        new_v = ops.linear_act([g1, g3], lin.weight, lin.bias, relu=True)
        new_v = ops.linear_act([new_v], self.rp_emb.projection.weight, None, relu="leaky")
        if self.use_attention:
            new_v = _flat(self.self_atten(new_v.view(B, N, -1)))
        new_v = ops.linear_act([new_v], self.rp_final.projection.weight, None, relu="leaky", p=p_feat, mask=fm[4],
                               seed=seed, stream_id=sid + 7)
        logits = ops.linear_act([new_v], self.classifier.weight, self.classifier.bias)
        return logits.float().view(B, N, self.output_dim)


class GCNBlock(nn.Module):
    """GraphConv + BatchNorm1d + leaky ReLU(0.2) (deep_rp_robust_gcn.py:30-45); parameter container -- the arithmetic
    runs in ``DeepRPRobustGCN.forward`` on the packed graph."""

    def __init__(self, in_channels: int, out_channels: int, num_edges: int):
        super().__init__()
        self.gcn = GraphConv(in_channels, out_channels, num_edges)
        self.bn = nn.BatchNorm1d(out_channels)


class EmbeddingBlock(nn.Module):
    """Linear + BatchNorm1d + leaky ReLU(0.2) (deep_rp_robust_gcn.py:48-62)."""

    def __init__(self, in_channels: int, out_channels: int):
        super().__init__()
        self.emb = nn.Linear(in_channels, out_channels)
        self.bn = nn.BatchNorm1d(out_channels)


@register_model
class DeepRPRobustGCN(BaseNetwork):
    """gnn/models/networks/deep_rp_robust_gcn.py:65-168 -- nine GraphConv + BatchNorm + leaky-ReLU(0.2) blocks with two
    skip concats, DropEdge(0.2) in front of blocks 3, 6, 8, 9, plain dropout(0.3) behind blocks 3, 6, 9 and the
    attention, a frozen random projection (scaled by ``lambda_value``, leaky 0.2) in front of the attention.  Same
    SpMM / GEMM / attention kernels as GraphCNNDropEdge; BatchNorm + leaky ReLU is csrc/batchnorm.cu.  fp32."""

    LEAKY = 0.2

    def __init__(self, input_dim: int, output_dim: int, num_edges: int, net_size: int = 256, use_attention: bool = True,
                 rp_size: int = 10000, lambda_value: float = 0.01):
        super().__init__()
        self.output_dim, self.net_size, self.num_edges = output_dim, net_size, num_edges
        self.rp_size, self.lambda_value, self.use_attention = rp_size, lambda_value, use_attention
        self.dropout = nn.Dropout(p=0.3)
        self.edge_dropout = nn.Dropout(p=0.2)
        self.emb1 = EmbeddingBlock(input_dim, net_size)
        self.gcn1 = GCNBlock(net_size, net_size, num_edges)
        self.gcn2 = GCNBlock(net_size, net_size, num_edges)
        self.gcn3 = GCNBlock(net_size * 2, net_size, num_edges)
        self.gcn4 = GCNBlock(net_size, net_size, num_edges)
        self.gcn5 = GCNBlock(net_size, net_size, num_edges)
        self.gcn6 = GCNBlock(net_size * 2, net_size, num_edges)
        self.gcn7 = GCNBlock(net_size, net_size, num_edges)
        self.gcn8 = GCNBlock(net_size, net_size, num_edges)
        self.gcn9 = GCNBlock(net_size, net_size, num_edges)
        self.emb2 = EmbeddingBlock(net_size * 2, net_size)
        self.rp_embed2 = RanPACLayer(net_size, net_size, 1)
        self.self_atten = NodeSelfAtten(net_size)
        self.classifier = nn.Linear(net_size, output_dim)
        self.rng_seed: Optional[int] = None
        self._rng_calls = 0

    def _block(self, blk: GCNBlock, xs, g, w):
        z = ops.graph_conv(xs, g, w, blk.gcn.h_weights, blk.gcn.bias)
        return ops.batchnorm_act(z, blk.bn, self.LEAKY, self.training)

    def forward(self, inputs: Sequence[Any], efficient_mode: bool = True, *,
                masks: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        """Same inputs as GraphCNNDropEdge.forward (dense fp32 V).  ``masks`` = {"edge": 4 x uint8[nnz] (blocks 3, 6,
        8, 9), "feat": 4 x uint8[B*N, net_size] (behind blocks 3, 6, 9 and the attention)}."""
        V, A = inputs
        _lib.require_cuda(V, "V")
        B, N = V.shape[0], V.shape[1]
        g = A if isinstance(A, PackedGraphBatch) else _pack_dense(A, self.training)
        if g.B != B or g.N != N or g.L != self.num_edges:
            raise _lib.GrlError(f"graph batch (B={g.B}, N={g.N}, L={g.L}) does not match V {tuple(V.shape)} / "
                                f"num_edges={self.num_edges}")
        p_feat = self.dropout.p if self.training else 0.0
        p_edge = self.edge_dropout.p if self.training else 0.0
        seed = torch.initial_seed() if self.rng_seed is None else self.rng_seed
        sid = self._rng_calls * 16
        self._rng_calls += 1
        em = masks["edge"] if masks else [None] * 4
        fm = masks["feat"] if masks else [None] * 4

        def drop(x, k):
            return _FeatureDropout.apply(x.contiguous(), p_feat, fm[k], seed, sid + 8 + k) if p_feat > 0.0 else x

        w0 = g.edge_weights(0.0)
        z = ops.linear_act([_flat(V).float()], self.emb1.emb.weight, self.emb1.emb.bias)
        emb = ops.batchnorm_act(z, self.emb1.bn, self.LEAKY, self.training)
        g1 = self._block(self.gcn1, [emb], g, w0)
        g2 = self._block(self.gcn2, [g1], g, w0)
        g3 = drop(self._block(self.gcn3, [g1, g2], g, g.edge_weights(p_edge, seed, sid + 0, em[0])), 0)
        g4 = self._block(self.gcn4, [g3], g, w0)
        g5 = self._block(self.gcn5, [g4], g, w0)
        g6 = drop(self._block(self.gcn6, [g4, g5], g, g.edge_weights(p_edge, seed, sid + 1, em[1])), 1)
        g7 = self._block(self.gcn7, [g6], g, w0)
        g8 = self._block(self.gcn8, [g7], g, g.edge_weights(p_edge, seed, sid + 2, em[2]))
        g9 = drop(self._block(self.gcn9, [g8], g, g.edge_weights(p_edge, seed, sid + 3, em[3])), 2)
        z = ops.linear_act([g8, g9], self.emb2.emb.weight, self.emb2.emb.bias)
        feats = ops.batchnorm_act(z, self.emb2.bn, self.LEAKY, self.training)
        # RanPACLayer.forward: leaky(projection(x) * lambda, 0.2) = lambda * leaky(projection(x), 0.2) for lambda > 0
        if self.lambda_value <= 0:
            raise _lib.GrlError("DeepRPRobustGCN needs lambda_value > 0")
        feats = ops.linear_act([feats], self.rp_embed2.projection.weight, None, relu="leaky0.2") * self.lambda_value
        feats = drop(_flat(self.self_atten(feats.view(B, N, -1))), 3)
        logits = ops.linear_act([feats], self.classifier.weight, self.classifier.bias)
        return logits.float().view(B, N, self.output_dim)


class CustomGCNSequential(nn.Module):
    """deep_rp_gcn.py:70-103 with the file's constants (NUM_GCN_LAYERS = 29, SKIP_CONNECTION_POS = 3, no RP layers):
    every third block takes the concat [feats, feats] (the reference re-binds ``prev_feats`` right before using it)."""

    NUM_GCN_LAYERS = 29
    SKIP_CONNECTION_POS = 3

    def __init__(self, net_size: int, num_edges: int, lambda_value: float):
        super().__init__()
        self.layers = nn.ModuleList()
        for ii in range(self.NUM_GCN_LAYERS):
            wide = ii % self.SKIP_CONNECTION_POS == 0
            self.layers.append(GCNBlock(net_size * 2 if wide else net_size, net_size, num_edges))


@register_model
class DeepRPGCN(BaseNetwork):
    """gnn/models/networks/deep_rp_gcn.py:106-150 -- embedding block, 29 GraphConv + BatchNorm + leaky(0.2) blocks on the
    un-dropped adjacency, embedding block, dropout(0.3), classifier.  fp32."""

    LEAKY = 0.2

    def __init__(self, input_dim: int, output_dim: int, num_edges: int, net_size: int = 256, rp_size: int = 10000,
                 lambda_value: float = 0.01):
        super().__init__()
        self.output_dim, self.net_size, self.num_edges = output_dim, net_size, num_edges
        self.rp_size, self.lambda_value = rp_size, lambda_value
        self.dropout = nn.Dropout(p=0.3)
        self.emb1 = EmbeddingBlock(input_dim, net_size)
        self.gcn_layers = CustomGCNSequential(net_size, num_edges, lambda_value)
        self.emb2 = EmbeddingBlock(net_size, net_size)
        self.classifier = nn.Linear(net_size, output_dim)
        self.rng_seed: Optional[int] = None
        self._rng_calls = 0

    def forward(self, inputs: Sequence[Any], efficient_mode: bool = True, *,
                masks: Optional[Dict[str, Any]] = None) -> torch.Tensor:
        """``masks`` = {"feat": [uint8[B*N, net_size]]} (the one dropout in front of the classifier)."""
        V, A = inputs
        _lib.require_cuda(V, "V")
        B, N = V.shape[0], V.shape[1]
        g = A if isinstance(A, PackedGraphBatch) else _pack_dense(A, self.training)
        if g.B != B or g.N != N or g.L != self.num_edges:
            raise _lib.GrlError(f"graph batch (B={g.B}, N={g.N}, L={g.L}) does not match V {tuple(V.shape)} / "
                                f"num_edges={self.num_edges}")
        p_feat = self.dropout.p if self.training else 0.0
        seed = torch.initial_seed() if self.rng_seed is None else self.rng_seed
        sid = self._rng_calls * 16
        self._rng_calls += 1
        w0 = g.edge_weights(0.0)
        z = ops.linear_act([_flat(V).float()], self.emb1.emb.weight, self.emb1.emb.bias)
        feats = ops.batchnorm_act(z, self.emb1.bn, self.LEAKY, self.training)
        for idx, blk in enumerate(self.gcn_layers.layers):
            xs = [feats, feats] if idx % CustomGCNSequential.SKIP_CONNECTION_POS == 0 else [feats]
            z = ops.graph_conv(xs, g, w0, blk.gcn.h_weights, blk.gcn.bias)
            feats = ops.batchnorm_act(z, blk.bn, self.LEAKY, self.training)
        z = ops.linear_act([feats], self.emb2.emb.weight, self.emb2.emb.bias)
        feats = ops.batchnorm_act(z, self.emb2.bn, self.LEAKY, self.training)
        if p_feat > 0.0:
            feats = _FeatureDropout.apply(feats.contiguous(), p_feat, masks["feat"][0] if masks else None, seed, sid)
        logits = ops.linear_act([feats], self.classifier.weight, self.classifier.bias)
        return logits.float().view(B, N, self.output_dim)
