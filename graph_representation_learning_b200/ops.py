"""Autograd operators of the hot path, each a thin launcher of C-ABI kernels (include/grl_b200.h).

* ``spmm_aggregate``  -- K3+K4 of SURVEY.md: DropEdge-weighted multi-relational aggregation
  (reference: torch.matmul(A', V), gnn/models/networks/robust_gcn.py:45-47) and its transposed gradient.
* ``linear_act``      -- Linear (+bias) (+ReLU) (+dropout) over a *virtual concat* of inputs
  (reference: make_linear_relu / nn.Linear / F.relu / nn.Dropout / torch.cat,
  gnn/models/networks/robust_gcn.py:10-11, drop_robust_gcn.py:64,84-89,99-103).
* ``graph_conv``      -- aggregate-then-project GraphConv with the ReLU/dropout epilogue fused
  (reference: GraphConv.forward, robust_gcn.py:32-51 + drop_robust_gcn.py:76-77).

Dense contractions go through ``gemm`` below.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from .graph import PackedGraphBatch


# ----------------------------------------------------------------------------------------------
# kernel launchers (no autograd)
# ----------------------------------------------------------------------------------------------
def _row_major_2d(t: torch.Tensor, name: str) -> int:
    """returns the leading dimension (elements) of a 2-D row-major view"""
    if t.dim() != 2 or t.stride(1) != 1:
        raise _lib.GrlError(f"{name} must be a 2-D row-major view, got shape {tuple(t.shape)} stride {t.stride()}")
    return t.stride(0) if t.shape[0] > 1 else max(t.stride(0), t.shape[1])


_spmm_stream_enabled = [True]     # tests / A-B scripts switch the node-stream kernel off to reach the gather kernels


def spmm_fwd_(g: PackedGraphBatch, w: torch.Tensor, X: torch.Tensor, Y: torch.Tensor) -> None:
    """Y[r, :F] = sum_e w[e] X[base(r)+col[e], :F];  Y is a [R, F] view (possibly a column slice)."""
    ldx, ldy = _row_major_2d(X, "X"), _row_major_2d(Y, "Y")
    F = X.shape[1]
    if X.shape[0] != g.num_nodes or Y.shape[0] != g.num_rows or Y.shape[1] != F or X.dtype != Y.dtype:
        raise _lib.GrlError(f"spmm_fwd shape mismatch: X {tuple(X.shape)} Y {tuple(Y.shape)} for {g.num_nodes} nodes")
    stream = getattr(w, "_grl_stream", None)
    if stream is not None and _spmm_stream_enabled[0]:
        # node-stream slab kernel (csrc/spmm_slab2.cu); GRL_E_UNSUP = the graph's slab does not fit -> gather kernels
        nodeinfo, ew = stream
        rc = _lib.load().grl_spmm_fwd_stream(nodeinfo.data_ptr(), ew.data_ptr(), X.data_ptr(), ldx, Y.data_ptr(), ldy,
                                             _lib.dtype_code(X), g.B, g.N, g.L, F, _lib.stream_ptr(X.device))
        if rc == 0:
            _lib.count_launch(1)
            return
        if rc != _lib.GRL_E_UNSUP:
            _lib.check(rc, "grl_spmm_fwd_stream")
    _lib.check(_lib.load().grl_spmm_fwd(g.rowptr.data_ptr(), g.col.data_ptr(), w.data_ptr(), X.data_ptr(), ldx,
                                        Y.data_ptr(), ldy, _lib.dtype_code(X), g.B, g.N, g.L, F,
                                        _lib.stream_ptr(X.device)), "grl_spmm_fwd")
    _lib.count_launch(1)


def spmm_bwd_(g: PackedGraphBatch, w: torch.Tensor, dY: torch.Tensor, dX: torch.Tensor, accumulate: bool = False) -> None:
    """dX[c, :F] (+)= sum_k w[perm[k]] dY[rowidx[k], :F]"""
    if g.colptr is None:
        g.build_transpose()
    lddy, lddx = _row_major_2d(dY, "dY"), _row_major_2d(dX, "dX")
    F = dX.shape[1]
    if dX.shape[0] != g.num_nodes or dY.shape[0] != g.num_rows or dY.shape[1] != F or dX.dtype != dY.dtype:
        raise _lib.GrlError(f"spmm_bwd shape mismatch: dY {tuple(dY.shape)} dX {tuple(dX.shape)}")
    _lib.check(_lib.load().grl_spmm_bwd(g.colptr.data_ptr(), g.rowidx.data_ptr(), g.perm.data_ptr(), w.data_ptr(),
                                        dY.data_ptr(), lddy, dX.data_ptr(), lddx, _lib.dtype_code(dX), g.B, g.N, g.L, F,
                                        1 if accumulate else 0, _lib.stream_ptr(dX.device)), "grl_spmm_bwd")
    _lib.count_launch(1)


def _act_flag(relu) -> int:
    """``relu`` arguments take True (ReLU), False (none), "leaky" (leaky ReLU, slope 0.01) or "leaky0.2" """
    if relu == "leaky":
        return _lib.EPI_LEAKY
    if relu == "leaky0.2":
        return _lib.EPI_LEAKY02
    return _lib.EPI_RELU if relu else 0


def epilogue_fwd_(X: torch.Tensor, bias: Optional[torch.Tensor], Y: torch.Tensor, relu: bool, p: float,
                  mask: Optional[torch.Tensor], seed: int, stream_id: int) -> None:
    flags = (_lib.EPI_BIAS if bias is not None else 0) | _act_flag(relu) | \
            (_lib.EPI_DROPOUT if p > 0.0 else 0)
    M, C = X.shape
    if mask is not None:
        if mask.dtype != torch.uint8 or mask.numel() != M * C or not mask.is_contiguous():
            raise _lib.GrlError("dropout mask must be contiguous uint8 [M, C]")
    if bias is not None and (bias.dtype != torch.float32 or not bias.is_contiguous()):
        raise _lib.GrlError("bias must be contiguous fp32")
    _lib.check(_lib.load().grl_bias_act_dropout_fwd(X.data_ptr(), _row_major_2d(X, "X"), _lib.ptr(bias), Y.data_ptr(),
                                                    _row_major_2d(Y, "Y"), _lib.dtype_code(X), M, C, flags, float(p),
                                                    _lib.ptr(mask), int(seed) & (2**64 - 1), int(stream_id),
                                                    _lib.stream_ptr(X.device)), "grl_bias_act_dropout_fwd")
    _lib.count_launch(1)


def epilogue_bwd_(dY: torch.Tensor, Y: Optional[torch.Tensor], dX: torch.Tensor, relu: bool, p: float) -> None:
    flags = _act_flag(relu) | (_lib.EPI_DROPOUT if p > 0.0 else 0)
    M, C = dY.shape
    _lib.check(_lib.load().grl_bias_act_dropout_bwd(dY.data_ptr(), _row_major_2d(dY, "dY"), _lib.ptr(Y),
                                                    _row_major_2d(Y, "Y") if Y is not None else C, dX.data_ptr(),
                                                    _row_major_2d(dX, "dX"), _lib.dtype_code(dY), M, C, flags, float(p),
                                                    _lib.stream_ptr(dY.device)), "grl_bias_act_dropout_bwd")
    _lib.count_launch(1)


def colsum(X: torch.Tensor) -> torch.Tensor:
    out = torch.zeros(X.shape[1], dtype=torch.float32, device=X.device)
    _lib.check(_lib.load().grl_colsum(X.data_ptr(), _row_major_2d(X, "X"), _lib.dtype_code(X), X.shape[0], X.shape[1],
                                      out.data_ptr(), _lib.stream_ptr(X.device)), "grl_colsum")
    _lib.count_launch(1)
    return out


# ----------------------------------------------------------------------------------------------
# dense contractions
# ----------------------------------------------------------------------------------------------
class GemmConfig:
    """How fp32 operands go through the tensor cores: "tf32x3" (error-compensated, three tf32 products, 4-5e-7
    norm-wise), "tf32bc" (tf32 main product + the two correction products as one bf16 product, activation operand
    through TMEM: 6-7e-7, 1.2-1.6x faster) or "tf32" (single pass, ~1e-3).  bf16 operands always use kind::f16.

    ``fp32_mode`` serves the FORWARD products, whose outputs meet ReLU thresholds: a pre-activation within the GEMM's
    error of zero flips its derivative and moves whole gradient tensors by 1e-4 .. 1e-3 (tests/test_gpu_parity.py,
    scripts/mode_noise.py), so the forward keeps the most accurate mode.  ``fp32_mode_bwd`` serves dX = dZ W and
    dW = X^T dZ: the backward pass is linear in dZ, its GEMM error stays a 7e-7 relative perturbation of the gradient,
    so it takes the faster mode."""
    fp32_mode = "tf32x3"
    fp32_mode_bwd = "tf32bc"
    _modes = {"tf32x3": 2, "tf32bc": 3, "tf32": 1}


def set_x3_promotion(k_blocks: int) -> None:
    """K-blocks (of 32) the three-product forward kernels accumulate in TMEM before the partial sum is promoted to fp32
    registers (grl_gemm_x3_set_promo; default 2).  TMEM accumulation truncates: fewer blocks = less rounding noise in
    the pre-activations = fewer ReLU-derivative flips against an fp64 run, at the price of more accumulator drains."""
    _lib.check(_lib.load().grl_gemm_x3_set_promo(int(k_blocks)), "grl_gemm_x3_set_promo")


def _mode_for(dtype: torch.dtype, fp32_mode: Optional[str] = None) -> int:
    if dtype == torch.bfloat16:
        return _lib.GEMM_BF16
    if dtype == torch.float32:
        return GemmConfig._modes[fp32_mode or GemmConfig.fp32_mode]
    raise _lib.GrlError(f"GEMM operands must be fp32 or bf16, got {dtype}")


def _ld_multiple(dtype: torch.dtype) -> int:
    return 16 // torch.empty(0, dtype=dtype).element_size()


def _tma_ready(x: torch.Tensor) -> torch.Tensor:
    """TMA needs a 16-byte aligned base and row pitch.  Aligned row-major views pass through untouched; anything
    else (e.g. a dense [*, 4369] fp32 batch handed over without a padded leading dimension) is re-laid once."""
    m = _ld_multiple(x.dtype)
    if x.dim() == 2 and x.stride(1) == 1 and x.stride(0) % m == 0 and x.data_ptr() % 16 == 0 and x.stride(0) >= x.shape[1]:
        return x
    K = x.shape[1]
    buf = torch.empty(x.shape[0], (K + m - 1) // m * m, dtype=x.dtype, device=x.device)
    buf[:, :K].copy_(x)
    return buf[:, :K]


def _prep_weight(b_view: torch.Tensor, dtype: torch.dtype, mode: int):
    """[N, K] view of a (small) weight -> contiguous, 16-byte-pitched operand(s): (hi, lo) for TF32X3 else (b, None)."""
    N, K = b_view.shape
    m = _ld_multiple(dtype)
    Kp = (K + m - 1) // m * m
    # fp32 weight views are split where they lie (any strides: h_weights and nn.Linear weights are wanted in both
    # orientations); the contiguous copy that used to come first was one more launch per operand
    if mode in (_lib.GEMM_TF32_BF16C, _lib.GEMM_TF32X3):
        src = b_view if b_view.dtype == torch.float32 else b_view.float()
    if mode == _lib.GEMM_TF32_BF16C:
        hi = torch.empty(N, Kp, dtype=torch.float32, device=b_view.device)
        packed = torch.empty(N, 2 * Kp, dtype=torch.bfloat16, device=b_view.device)
        _lib.check(_lib.load().grl_split_tf32_mix_strided(src.data_ptr(), src.stride(0), src.stride(1), hi.data_ptr(), Kp,
                                                          packed.data_ptr(), 2 * Kp, N, K, _lib.stream_ptr(src.device)),
                   "grl_split_tf32_mix_strided")
        _lib.count_launch(1)
        return hi[:, :K], packed
    if mode != _lib.GEMM_TF32X3:
        if Kp == K:
            return b_view.to(dtype).contiguous(), None
        buf = torch.empty(N, Kp, dtype=dtype, device=b_view.device)
        buf[:, :K].copy_(b_view)
        return buf[:, :K], None
    hi = torch.empty(N, Kp, dtype=torch.float32, device=b_view.device)
    lo = torch.empty(N, Kp, dtype=torch.float32, device=b_view.device)
    _lib.check(_lib.load().grl_split_tf32_strided(src.data_ptr(), src.stride(0), src.stride(1), hi.data_ptr(), lo.data_ptr(), Kp,
                                                  N, K, _lib.stream_ptr(src.device)), "grl_split_tf32_strided")
    _lib.count_launch(1)
    return hi[:, :K], lo[:, :K]


def gemm_nt(a_list: Sequence[torch.Tensor], b_list: Sequence[torch.Tensor], out_dtype: Optional[torch.dtype] = None,
            bias: Optional[torch.Tensor] = None, relu: bool = False, p: float = 0.0, mask: Optional[torch.Tensor] = None,
            seed: int = 0, stream_id: int = 0, fp32_mode: Optional[str] = None) -> torch.Tensor:
    """C[M, N] = epilogue(sum_k a_k[M, K_k] @ b_k[N, K_k]^T) on the tcgen05 kernels (csrc/gemm_tcgen05.cu,
    gemm_x3_tcgen05.cu, gemm_bc_tcgen05.cu).  ``a_k`` are activations (row-major views), ``b_k`` are [N, K_k] views
    of weights in any layout.  ``fp32_mode`` overrides ``GemmConfig.fp32_mode`` for this product."""
    if not 1 <= len(a_list) <= 2 or len(a_list) != len(b_list):
        raise _lib.GrlError("gemm_nt takes one or two K segments")
    cdt = a_list[0].dtype
    mode = _mode_for(cdt, fp32_mode)
    M, N = a_list[0].shape[0], b_list[0].shape[0]
    out_dtype = out_dtype or cdt
    C = torch.empty(M, N, dtype=out_dtype, device=a_list[0].device)
    if M == 0:
        return C
    a_ops = [_tma_ready(a) for a in a_list]
    b_ops = [_prep_weight(b, cdt, mode) for b in b_list]
    if bias is not None and (bias.dtype != torch.float32 or not bias.is_contiguous()):
        bias = bias.float().contiguous()
    if mask is not None and (mask.dtype != torch.uint8 or mask.numel() != M * N or not mask.is_contiguous()):
        raise _lib.GrlError("dropout mask must be contiguous uint8 [M, N]")
    flags = (_lib.EPI_BIAS if bias is not None else 0) | _act_flag(relu) | \
            (_lib.EPI_DROPOUT if p > 0.0 else 0)
    two = len(a_list) == 2
    a1, (b1, b1lo) = (a_ops[1], b_ops[1]) if two else (None, (None, None))
    _lib.check(_lib.load().grl_gemm_nt(
        a_ops[0].data_ptr(), a_ops[0].stride(0), b_ops[0][0].data_ptr(), _lib.ptr(b_ops[0][1]), b_ops[0][0].stride(0),
        a_ops[0].shape[1],
        _lib.ptr(a1), a1.stride(0) if two else 0, _lib.ptr(b1), _lib.ptr(b1lo), b1.stride(0) if two else 0,
        a1.shape[1] if two else 0,
        C.data_ptr(), C.stride(0), _lib.dtype_code(C), M, N, mode, _lib.ptr(bias), flags, float(p), _lib.ptr(mask),
        int(seed) & (2**64 - 1), int(stream_id), _lib.stream_ptr(C.device)), "grl_gemm_nt")
    _lib.count_launch(1)
    return C


_tn_workspace = {}


def gemm_tn(a: torch.Tensor, b: torch.Tensor, fp32_mode: Optional[str] = None,
            split_cache: Optional[dict] = None) -> torch.Tensor:
    """dW[Ma, Nb] = a[R, Ma]^T @ b[R, Nb], fp32 output: the weight-gradient contraction over the node dimension.

    fp32 operands, split-K over R with a deterministic reduce: "tf32bc" (default, ``GemmConfig.fp32_mode_bwd``) = the
    TMEM-operand kernel (csrc/gemm_bc_tcgen05.cu, grl_split_act_mn + grl_gemm_tn_bc); "tf32x3" = the MN-major
    three-product kernel (csrc/gemm_x3_tcgen05.cu, grl_gemm_tn).
    bf16 operands: MN-major kind::f16 kernel, fp32 accumulation and output (csrc/gemm_tcgen05.cu, grl_gemm_tn_bf16).
    ``split_cache``: a dict shared by the calls of one backward step so that the same narrow operand (dZ against the
    segments of a virtual concat) is split once."""
    if a.dtype not in (torch.float32, torch.bfloat16) or b.dtype != a.dtype:
        raise _lib.GrlError(f"gemm_tn operands must both be fp32 or both bf16, got {a.dtype} / {b.dtype}")
    a, b = _tma_ready(a), _tma_ready(b)
    R, Ma = a.shape
    Nb = b.shape[1]
    lib = _lib.load()
    ws_bytes = lib.grl_gemm_tn_workspace_bytes(Ma, Nb)
    key = (a.device, torch.cuda.current_stream(a.device).cuda_stream)
    ws = _tn_workspace.get(key)
    if ws is None or ws.numel() < ws_bytes:          # one reusable split-K workspace per (device, stream)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=a.device)
        _tn_workspace[key] = ws
    C = torch.empty(Ma, Nb, dtype=torch.float32, device=a.device)
    if a.dtype == torch.bfloat16:
        _lib.check(lib.grl_gemm_tn_bf16(a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), C.data_ptr(), Nb, Ma, Nb, R,
                                        ws.data_ptr(), ws.numel(), _lib.stream_ptr(a.device)), "grl_gemm_tn_bf16")
        _lib.count_launch(2)
        return C
    if (fp32_mode or GemmConfig.fp32_mode_bwd) == "tf32bc":
        # the WIDE operand goes through TMEM, the narrow one is split once in HBM; the product is stored transposed
        # when the caller's first operand is the narrow one
        swap = Nb > Ma
        wide, narrow = (b, a) if swap else (a, b)
        Wn, Nn = wide.shape[1], narrow.shape[1]
        ldp = (Nn + 7) // 8 * 8
        key = (narrow.data_ptr(), narrow.stride(0), R, Nn)
        cached = split_cache.get(key) if split_cache is not None else None
        if cached is None:
            hi = torch.empty(R, ldp, dtype=torch.float32, device=a.device)
            packed = torch.empty((R + 31) // 32 * 64, ldp, dtype=torch.bfloat16, device=a.device)
            _lib.check(lib.grl_split_act_mn(narrow.data_ptr(), narrow.stride(0), hi.data_ptr(), ldp, packed.data_ptr(),
                                            ldp, R, Nn, _lib.stream_ptr(a.device)), "grl_split_act_mn")
            _lib.count_launch(1)
            if split_cache is not None:
                split_cache[key] = (hi, packed, narrow)        # keeps `narrow` alive: the key is its address
        else:
            hi, packed, _ = cached
        _lib.check(lib.grl_gemm_tn_bc(wide.data_ptr(), wide.stride(0), Wn, hi.data_ptr(), ldp, packed.data_ptr(), ldp, Nn,
                                      R, C.data_ptr(), 1 if swap else 0, ws.data_ptr(), ws.numel(),
                                      _lib.stream_ptr(a.device)), "grl_gemm_tn_bc")
        _lib.count_launch(2)
        return C
    _lib.check(lib.grl_gemm_tn(a.data_ptr(), a.stride(0), b.data_ptr(), b.stride(0), C.data_ptr(), Nb, Ma, Nb, R,
                               ws.data_ptr(), ws.numel(), _lib.stream_ptr(a.device)), "grl_gemm_tn")
    _lib.count_launch(2)
    return C


# ----------------------------------------------------------------------------------------------
# autograd functions
# ----------------------------------------------------------------------------------------------
class _SpmmAggregate(torch.autograd.Function):
    @staticmethod
    def forward(ctx, g: PackedGraphBatch, w: torch.Tensor, *xs: torch.Tensor):
        F_total = sum(x.shape[1] for x in xs)
        Y = torch.empty(g.num_nodes, (g.L + 1) * F_total, dtype=xs[0].dtype, device=xs[0].device)
        Yr = Y.view(g.num_rows, F_total)
        off = 0
        for x in xs:
            spmm_fwd_(g, w, x, Yr[:, off:off + x.shape[1]])
            off += x.shape[1]
        ctx.g, ctx.w = g, w
        ctx.widths = [x.shape[1] for x in xs]
        return Y

    @staticmethod
    def backward(ctx, dY: torch.Tensor):
        g, w = ctx.g, ctx.w
        dY = dY.contiguous()
        F_total = sum(ctx.widths)
        dYr = dY.view(g.num_rows, F_total)
        grads: List[Optional[torch.Tensor]] = []
        off = 0
        for k, Fk in enumerate(ctx.widths):
            if ctx.needs_input_grad[2 + k]:
                dX = torch.empty(g.num_nodes, Fk, dtype=dY.dtype, device=dY.device)
                spmm_bwd_(g, w, dYr[:, off:off + Fk], dX)
                grads.append(dX)
            else:
                grads.append(None)
            off += Fk
        return (None, None, *grads)


def spmm_aggregate(g: PackedGraphBatch, w: torch.Tensor, xs: Sequence[torch.Tensor]) -> torch.Tensor:
    """new_V [B*N, (L+1)*F] of the virtual concat ``xs`` (each [B*N, F_k] row-major)."""
    return _SpmmAggregate.apply(g, w, *xs)


class _LinearAct(torch.autograd.Function):
    """y = dropout(relu(sum_k x_k @ W[:, slice_k]^T + b)); weight is [out, in] like nn.Linear."""

    @staticmethod
    def forward(ctx, weight, bias, relu, p, mask, seed, stream_id, weight_is_kn, *xs):
        # weight_is_kn: weight stored [in, out] (GraphConv.h_weights) instead of nn.Linear's [out, in]
        b_views, off = [], 0
        for x in xs:
            K = x.shape[1]
            b_views.append(weight[off:off + K, :].t() if weight_is_kn else weight[:, off:off + K])   # [N, K_k]
            off += K
        z = gemm_nt(xs, b_views, None, bias, relu, p, mask, seed, stream_id)     # epilogue fused in the kernel
        ctx.save_for_backward(weight, z if relu else None, *xs)
        ctx.cfg = (relu, p, weight_is_kn, bias is not None)
        return z

    @staticmethod
    def backward(ctx, dY):
        weight, y, *xs = ctx.saved_tensors
        relu, p, weight_is_kn, has_bias = ctx.cfg
        dZ, d_weight, d_bias = _linear_grads(dY, y, xs, relu, p, weight_is_kn, ctx.needs_input_grad[0],
                                             has_bias and ctx.needs_input_grad[1])
        dxs = []
        off = 0
        for k, x in enumerate(xs):
            K = x.shape[1]
            if ctx.needs_input_grad[8 + k]:
                # dX_k[M, K_k] = dZ[M, N] @ W_k with W_k presented as the NT "B" operand [K_k, N]
                wk = weight[off:off + K, :] if weight_is_kn else weight[:, off:off + K].t()
                dxs.append(gemm_nt([dZ], [wk], fp32_mode=GemmConfig.fp32_mode_bwd))
            else:
                dxs.append(None)
            off += K
        return (d_weight, d_bias, None, None, None, None, None, None, *dxs)


def _linear_grads(dY: torch.Tensor, y: Optional[torch.Tensor], xs: Sequence[torch.Tensor], relu, p: float,
                  weight_is_kn: bool, want_w: bool, want_b: bool):
    """Shared backward of y = dropout(act(sum_k x_k W_k + b)): returns (dZ, dW or None, db or None).  dZ is the gradient
    at the pre-activation (autograd of F.relu + nn.Dropout, drop_robust_gcn.py:64-100); dW is the fp32 master gradient,
    [K, N] for h_weights (``weight_is_kn``), [N, K] for nn.Linear."""
    dY = dY.contiguous()
    d_weight = d_bias = None
    bwd, cache = GemmConfig.fp32_mode_bwd, {}
    M, C = dY.shape
    fused = (relu is True and dY.dtype == torch.float32 and bwd == "tf32bc" and C % 8 == 0 and C <= 2048 and
             256 % (C // 8) == 0 and M > 0 and y.stride(0) % 4 == 0)
    if fused:
        # one pass: dZ, its split for the weight-gradient kernel (when dZ is the narrow operand) and the bias gradient
        lib = _lib.load()
        dZ = torch.empty_like(dY)
        # gemm_tn takes dZ as the narrow (pre-split) operand when it is narrower than x (ties: its second argument)
        split = want_w and all((x.shape[1] >= C) if weight_is_kn else (x.shape[1] > C) for x in xs)
        hi = torch.empty(M, C, dtype=torch.float32, device=dY.device) if split else None
        packed = torch.empty((M + 31) // 32 * 64, C, dtype=torch.bfloat16, device=dY.device) if split else None
        d_bias = torch.empty(C, dtype=torch.float32, device=dY.device) if want_b else None
        ws = torch.empty(lib.grl_dz_split_workspace_bytes(M, C) if want_b else 0, dtype=torch.uint8, device=dY.device)
        flags = _lib.EPI_RELU | (_lib.EPI_DROPOUT if p > 0.0 else 0)
        _lib.check(lib.grl_dz_split(dY.data_ptr(), dY.stride(0), y.data_ptr(), y.stride(0), dZ.data_ptr(), dZ.stride(0),
                                    _lib.ptr(hi), _lib.ptr(packed), C, M, C, flags, float(p), _lib.ptr(d_bias),
                                    _lib.ptr(ws) if want_b else 0, ws.numel(), _lib.stream_ptr(dY.device)),
                   "grl_dz_split")
        _lib.count_launch(2 if want_b else 1)
        if split:
            cache[(dZ.data_ptr(), dZ.stride(0), M, C)] = (hi, packed, dZ)
    elif (relu is True and dY.dtype == torch.bfloat16 and C % 8 == 0 and C <= 2048 and 256 % (C // 8) == 0 and M > 0 and
          y.stride(0) % 8 == 0 and dY.stride(0) % 8 == 0):
        # bf16: dZ and the bias gradient in one pass (grl_dz_colsum_bf16), the bias gradient summed in a fixed order
        lib = _lib.load()
        dZ = torch.empty_like(dY)
        d_bias = torch.empty(C, dtype=torch.float32, device=dY.device) if want_b else None
        ws = torch.empty(lib.grl_dz_split_workspace_bytes(M, C) if want_b else 0, dtype=torch.uint8, device=dY.device)
        flags = _lib.EPI_RELU | (_lib.EPI_DROPOUT if p > 0.0 else 0)
        _lib.check(lib.grl_dz_colsum_bf16(dY.data_ptr(), dY.stride(0), y.data_ptr(), y.stride(0), dZ.data_ptr(), dZ.stride(0),
                                          M, C, flags, float(p), _lib.ptr(d_bias), _lib.ptr(ws) if want_b else 0, ws.numel(),
                                          _lib.stream_ptr(dY.device)), "grl_dz_colsum_bf16")
        _lib.count_launch(2 if want_b else 1)
    elif relu or p > 0.0:
        dZ = torch.empty_like(dY)
        epilogue_bwd_(dY, y, dZ, relu, p)
    else:
        dZ = dY
    dZ = _tma_ready(dZ)          # a width like 53 is re-laid to a 16-byte pitch ONCE, for the dW and the dX products alike
    if want_w:
        parts = [gemm_tn(x, dZ, bwd, cache) if weight_is_kn else gemm_tn(dZ, x, bwd, cache) for x in xs]
        d_weight = parts[0] if len(parts) == 1 else torch.cat(parts, dim=0 if weight_is_kn else 1)
    if want_b and d_bias is None:
        d_bias = colsum(dZ)
    return dZ, d_weight, d_bias


def linear_act(xs: Sequence[torch.Tensor], weight: torch.Tensor, bias: Optional[torch.Tensor], relu: bool = False,
               p: float = 0.0, mask: Optional[torch.Tensor] = None, seed: int = 0, stream_id: int = 0,
               weight_is_kn: bool = False) -> torch.Tensor:
    return _LinearAct.apply(weight, bias, relu, float(p), mask, seed, stream_id, weight_is_kn, *xs)


class _SparseLinearAct(torch.autograd.Function):
    """y = dropout(relu(V @ W^T + b)) for V given as ``SparseNodeFeatures`` (features.py): heavy columns through a
    K = n_heavy tensor-core GEMM, light columns as a gather-sum over the rows of W^T (csrc/spmm.cu: fp32 FMA in CSR
    order); weight gradient = the same two pieces transposed.  No gradient flows to the features."""

    @staticmethod
    def forward(ctx, feats, weight, bias, relu, p, mask, seed, stream_id, cdt):
        lib = _lib.load()
        Nt, S, D = feats.num_nodes, weight.shape[0], weight.shape[1]
        if D != feats.shape[2]:
            raise _lib.GrlError(f"sparse features have {feats.shape[2]} columns, the layer expects {D}")
        Dh = int(feats.heavy_cols.numel())
        Wt = weight.detach().t().to(cdt).contiguous()                      # [D, S]: one row per feature column
        if Dh:
            hv = feats.heavy[:, :Dh] if cdt == torch.float32 else feats.heavy.to(cdt)[:, :Dh]
            z = gemm_nt([hv], [weight.detach()[:, feats.heavy_cols]], out_dtype=cdt)
        else:
            z = torch.empty(Nt, S, dtype=cdt, device=weight.device)
        if feats.nnz or not Dh:
            _lib.check(lib.grl_spmm_bwd(feats.rowptr.data_ptr(), feats.col.data_ptr(), feats.identity_perm().data_ptr(),
                                        feats.val.data_ptr(), Wt.data_ptr(), S, z.data_ptr(), S, _lib.dtype_code(z), 1, Nt, 1,
                                        S, 1 if Dh else 0, _lib.stream_ptr(z.device)), "grl_spmm_bwd(features)")
            _lib.count_launch(1)
        if bias is not None and (bias.dtype != torch.float32 or not bias.is_contiguous()):
            bias = bias.float().contiguous()
        epilogue_fwd_(z, bias, z, relu, p, mask, seed, stream_id)
        ctx.feats = feats
        ctx.save_for_backward(weight, z if relu else None)
        ctx.cfg = (relu, p, bias is not None)
        return z

    @staticmethod
    def backward(ctx, dY):
        weight, y = ctx.saved_tensors
        relu, p, has_bias = ctx.cfg
        feats = ctx.feats
        lib = _lib.load()
        dY = dY.contiguous()
        if relu or p > 0.0:
            dZ = torch.empty_like(dY)
            epilogue_bwd_(dY, y, dZ, relu, p)
        else:
            dZ = dY
        d_weight = d_bias = None
        if ctx.needs_input_grad[1]:
            S, D = weight.shape
            dZ32 = dZ if dZ.dtype == torch.float32 else dZ.float()          # fp32 master gradient
            feats.build_transpose()
            dWt = torch.empty(D, S, dtype=torch.float32, device=dZ.device)
            _lib.check(lib.grl_spmm_bwd(feats.colptr.data_ptr(), feats.rowidx.data_ptr(), feats.perm.data_ptr(),
                                        feats.val.data_ptr(), dZ32.data_ptr(), S, dWt.data_ptr(), S, _lib.dtype_code(dWt), 1, D, 1, S, 0,
                                        _lib.stream_ptr(dZ.device)), "grl_spmm_bwd(features^T)")
            _lib.count_launch(1)
            Dh = int(feats.heavy_cols.numel())
            if Dh:
                dWt[feats.heavy_cols] = gemm_tn(feats.heavy[:, :Dh], dZ32, GemmConfig.fp32_mode_bwd)     # [Dh, S]
            d_weight = dWt.t()
        if has_bias and ctx.needs_input_grad[2]:
            d_bias = colsum(dZ)
        return (None, d_weight, d_bias, None, None, None, None, None, None)


def sparse_linear_act(feats, weight: torch.Tensor, bias: Optional[torch.Tensor], relu: bool = False, p: float = 0.0,
                      mask: Optional[torch.Tensor] = None, seed: int = 0, stream_id: int = 0,
                      compute_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    return _SparseLinearAct.apply(feats, weight, bias, relu, float(p), mask, seed, stream_id, compute_dtype)


class _BatchNormAct(torch.autograd.Function):
    """leaky_relu(BatchNorm1d(x), slope) over the rows of x [M, C] (csrc/batchnorm.cu; fp32)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, momentum, eps, training, slope):
        lib = _lib.load()
        x = x.contiguous()
        M, C = x.shape
        y = torch.empty_like(x)
        mean = torch.empty(C, dtype=torch.float32, device=x.device)
        rstd = torch.empty(C, dtype=torch.float32, device=x.device)
        ws = torch.empty(lib.grl_bn_workspace_bytes(M, C), dtype=torch.uint8, device=x.device)
        _lib.check(lib.grl_bn_act_fwd(x.data_ptr(), x.stride(0), gamma.data_ptr(), beta.data_ptr(), _lib.ptr(running_mean),
                                      _lib.ptr(running_var), float(momentum), float(eps), int(training), float(slope),
                                      y.data_ptr(), y.stride(0), mean.data_ptr(), rstd.data_ptr(), M, C, ws.data_ptr(),
                                      ws.numel(), _lib.stream_ptr(x.device)), "grl_bn_act_fwd")
        _lib.count_launch(5 if training else 2)
        ctx.save_for_backward(x, y, gamma, mean, rstd)
        ctx.cfg = (bool(training), float(slope))
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, gamma, mean, rstd = ctx.saved_tensors
        training, slope = ctx.cfg
        lib = _lib.load()
        dy = dy.contiguous()
        M, C = x.shape
        dx = torch.empty_like(x)
        dgamma = torch.empty(C, dtype=torch.float32, device=x.device)
        dbeta = torch.empty(C, dtype=torch.float32, device=x.device)
        ws = torch.empty(lib.grl_bn_workspace_bytes(M, C), dtype=torch.uint8, device=x.device)
        _lib.check(lib.grl_bn_act_bwd(dy.data_ptr(), dy.stride(0), y.data_ptr(), y.stride(0), x.data_ptr(), x.stride(0),
                                      gamma.data_ptr(), mean.data_ptr(), rstd.data_ptr(), slope, int(training),
                                      dx.data_ptr(), dx.stride(0), dgamma.data_ptr(), dbeta.data_ptr(), M, C, ws.data_ptr(),
                                      ws.numel(), _lib.stream_ptr(x.device)), "grl_bn_act_bwd")
        _lib.count_launch(3)
        return dx, dgamma, dbeta, None, None, None, None, None, None


def batchnorm_act(x: torch.Tensor, bn: torch.nn.BatchNorm1d, slope: float, training: bool) -> torch.Tensor:
    """``F.leaky_relu(bn(x^T)^T, slope)`` for x [M, C] with ``bn`` an ``nn.BatchNorm1d`` used as the parameter / buffer
    container (GCNBlock / EmbeddingBlock, deep_rp_robust_gcn.py:30-62).  Training mode normalises with batch statistics
    and updates ``running_mean`` / ``running_var`` / ``num_batches_tracked`` like torch."""
    if x.dtype != torch.float32:
        raise _lib.GrlError("batchnorm_act runs in fp32")
    use_batch = training or bn.running_mean is None
    momentum = 0.0
    if training and bn.track_running_stats and bn.running_mean is not None:
        bn.num_batches_tracked += 1
        momentum = 1.0 / float(bn.num_batches_tracked) if bn.momentum is None else bn.momentum
    rm = bn.running_mean if (bn.track_running_stats and bn.running_mean is not None) else None
    rv = bn.running_var if rm is not None else None
    upd = training and rm is not None
    return _BatchNormAct.apply(x, bn.weight, bn.bias, rm if (upd or not use_batch) else None,
                               rv if (upd or not use_batch) else None, momentum, bn.eps, use_batch, slope)


def graph_conv(xs: Sequence[torch.Tensor], g: PackedGraphBatch, w_edge: torch.Tensor, h_weights: torch.Tensor,
               bias: Optional[torch.Tensor], relu: bool = False, p: float = 0.0, mask: Optional[torch.Tensor] = None,
               seed: int = 0, stream_id: int = 0) -> torch.Tensor:
    """GraphConv over a virtual concat of node-feature tensors: aggregate over the (L+1) typed adjacencies,
    then one projection with K = (L+1)*F (reference robust_gcn.py:45-51), epilogue fused."""
    if FusedGraphConv.wants(g, h_weights.shape[1]) and not (torch.is_grad_enabled() and (h_weights.requires_grad or any(x.requires_grad for x in xs))):
        out = graph_conv_fused(xs, g, w_edge, h_weights, bias, relu, p, mask, seed, stream_id)
        if out is not None:
            return out
    if GraphConvBackward.aggregate_first and h_weights.shape[1] <= sum(x.shape[1] for x in xs) and torch.is_grad_enabled() \
            and any(x.requires_grad for x in xs) and g.L + 1 <= 31:      # (the typed-transpose builder keeps a node's types one per lane)
        return _GraphConv.apply(g, w_edge, h_weights, bias, relu, float(p), mask, seed, stream_id, *xs)
    if not torch.is_grad_enabled() and len(xs) == 2:
        # no-grad layers over a virtual concat (gcn3 = GraphConv([g1, g2])): aggregate every input on its own -- in eval the
        # edge weights are the same object for all layers, so A' g1 of the previous layer (gcn2 = GraphConv([g1])) is reused
        # instead of being aggregated again -- and project with one K segment per input
        L1, F_total, C = g.L + 1, sum(x.shape[1] for x in xs), h_weights.shape[1]
        w3 = h_weights.view(L1, F_total, C)
        aggs, b_views, off = [], [], 0
        for x in xs:
            aggs.append(_aggregate_cached(g, w_edge, x))
            b_views.append(w3[:, off:off + x.shape[1], :].reshape(L1 * x.shape[1], C).t())
            off += x.shape[1]
        return gemm_nt(aggs, b_views, None, bias, relu, p, mask, seed, stream_id)
    if not torch.is_grad_enabled() and len(xs) == 1:
        return linear_act([_aggregate_cached(g, w_edge, xs[0])], h_weights, bias, relu, p, mask, seed, stream_id, weight_is_kn=True)
    new_v = spmm_aggregate(g, w_edge, xs)
    return linear_act([new_v], h_weights, bias, relu, p, mask, seed, stream_id, weight_is_kn=True)


def _aggregate_cached(g: PackedGraphBatch, w_edge: torch.Tensor, x: torch.Tensor) -> torch.Tensor:
    """A' x [B*N, (L+1) F] for a no-grad layer; the LAST aggregate is remembered on the weight tensor (keyed by the input's
    storage and version), which is what lets gcn3 reuse gcn2's A' g1 in eval"""
    key = (x.data_ptr(), x._version, tuple(x.shape), x.dtype)
    hit = getattr(w_edge, "_grl_agg", None)
    if hit is not None and hit[0] == key:
        return hit[1]
    y = torch.empty(g.num_nodes, (g.L + 1) * x.shape[1], dtype=x.dtype, device=x.device)
    spmm_fwd_(g, w_edge, x, y.view(g.num_rows, x.shape[1]))
    w_edge._grl_agg = (key, y, x)          # keeps x alive: the key is its address
    return y


class GraphConvBackward:
    """``aggregate_first``: the input gradient of GraphConv as aggregate-then-project on the TRANSPOSED typed graph,
        dX_k = sum_t (A_t^T dZ) W_{t,k}^T
    (ops._GraphConv) instead of autograd's project-then-aggregate dX_k = sum_t A_t^T (dZ W_{t,k}^T): the aggregation
    runs the forward slab kernel over [B*N, C] instead of the L2-gather transposed kernel over [B*N*(L+1), F], the
    projection becomes a long-K product (K = (L+1) C) instead of a short-K, epilogue-bound one, and the intermediate is
    [R, C] instead of [R, F].  Same mathematics in a different association: gradients agree to fp32 rounding."""
    aggregate_first = True


class _GraphConv(torch.autograd.Function):
    """GraphConv (robust_gcn.py:45-51) with its ReLU / dropout epilogue as ONE autograd node: forward = aggregation +
    projection exactly as ``spmm_aggregate`` + ``linear_act``; backward = dZ / dW / db as ``_LinearAct`` and the input
    gradients through the typed transpose of the batch (``GraphConvBackward``)."""

    @staticmethod
    def forward(ctx, g, w_edge, h_weights, bias, relu, p, mask, seed, stream_id, *xs):
        F_total = sum(x.shape[1] for x in xs)
        new_v = torch.empty(g.num_nodes, (g.L + 1) * F_total, dtype=xs[0].dtype, device=xs[0].device)
        nv = new_v.view(g.num_rows, F_total)
        off = 0
        for x in xs:
            spmm_fwd_(g, w_edge, x, nv[:, off:off + x.shape[1]])
            off += x.shape[1]
        z = gemm_nt([new_v], [h_weights.t()], None, bias, relu, p, mask, seed, stream_id)
        ctx.save_for_backward(h_weights, z if relu else None, new_v)
        ctx.g, ctx.w = g, w_edge
        ctx.widths = [x.shape[1] for x in xs]
        ctx.cfg = (relu, p, bias is not None)
        return z

    @staticmethod
    def backward(ctx, dY):
        h_weights, y, new_v = ctx.saved_tensors
        relu, p, has_bias = ctx.cfg
        g, w = ctx.g, ctx.w
        dZ, d_weight, d_bias = _linear_grads(dY, y, [new_v], relu, p, True, ctx.needs_input_grad[2],
                                             has_bias and ctx.needs_input_grad[3])
        grads: List[Optional[torch.Tensor]] = [None] * len(ctx.widths)
        if any(ctx.needs_input_grad[9:]):
            C = dZ.shape[1]
            gT = g.typed_transpose()
            wT = g.transposed_weights(w)
            U = torch.empty(g.num_nodes, (g.L + 1) * C, dtype=dZ.dtype, device=dZ.device)       # rows (j, t) of A_t^T dZ
            spmm_fwd_(gT, wT, dZ, U.view(g.num_rows, C))
            F_total = sum(ctx.widths)
            w3 = h_weights.detach().view(g.L + 1, F_total, C)
            off = 0
            for k, Fk in enumerate(ctx.widths):
                if ctx.needs_input_grad[9 + k]:
                    # B[f, t*C + c] = h_weights[t*F + off + f, c]: the NT "B" operand [F_k, (L+1) C] of dX_k = U B^T
                    bk = w3[:, off:off + Fk, :].permute(1, 0, 2).reshape(Fk, (g.L + 1) * C)
                    grads[k] = gemm_nt([U], [bk], fp32_mode=GemmConfig.fp32_mode_bwd)
                off += Fk
        return (None, None, d_weight, d_bias, None, None, None, None, None, *grads)


class FusedGraphConv:
    """GraphConv layers that need no gradient (eval, ``KVInference``, cfg5) can run as ONE kernel
    (csrc/graphconv_fused.cu): the aggregation of each (node tile, 32-column chunk, edge type) is produced straight into
    the operand tile of the tcgen05 projection, so ``new_V`` ((L+1) x the layer input) is neither written nor re-read
    (1.84 GB of HBM traffic per cfg2 layer).  Training keeps the unfused pair: dW = new_V^T dZ needs ``new_V`` again.

    ``enabled``: True = every supported layer, False = never, "auto" (default) = when the layer is at most ``max_waves``
    waves of CTA clusters.  Measured (profiles/r02_s58_fused_bench_16_gather_warps.jsonl): the aggregation producer is bound
    by its own dependent instruction stream, so a full-machine layer runs 1.38x the unfused pair's time (1.15 vs 0.83 ms
    at cfg2's gcn1) while small batches, which are launch-bound, gain (whole eval forward 64 x 100 nodes: 0.99 -> 0.78 ms;
    8 x 200: 0.94 -> 0.78 ms)."""
    enabled = "auto"
    max_waves = 1.0

    @classmethod
    def wants(cls, g: PackedGraphBatch, C: int) -> bool:
        if cls.enabled is True or cls.enabled is False:
            return cls.enabled
        ctas = g.B * ((g.N + 127) // 128) * (1 if C <= 128 else (2 if C <= 256 else 4))
        return ctas <= cls.max_waves * 148


def graph_conv_fused(xs: Sequence[torch.Tensor], g: PackedGraphBatch, w_edge: torch.Tensor, h_weights: torch.Tensor,
                     bias: Optional[torch.Tensor], relu=False, p: float = 0.0, mask: Optional[torch.Tensor] = None,
                     seed: int = 0, stream_id: int = 0) -> Optional[torch.Tensor]:
    """Forward-only fused GraphConv (grl_graphconv_fused_fwd).  Returns None for shapes / dtypes the kernel is not built
    for (the caller then runs spmm_aggregate + linear_act): fp32 inputs under the "tf32x3" forward mode, one or two
    inputs with widths that are multiples of 32, C <= 512."""
    if not 1 <= len(xs) <= 2 or any(x.dtype != torch.float32 for x in xs) or GemmConfig.fp32_mode != "tf32x3":
        return None
    F_total = sum(x.shape[1] for x in xs)
    K, C = h_weights.shape
    if K != (g.L + 1) * F_total or any(x.shape[1] % 32 for x in xs) or C > 512 or g.L + 2 > 32 or g.num_nodes == 0:
        return None
    xs = [_tma_ready(x) for x in xs]
    hi, lo = _prep_weight(h_weights.t(), torch.float32, _lib.GEMM_TF32X3)
    if bias is not None and (bias.dtype != torch.float32 or not bias.is_contiguous()):
        bias = bias.float().contiguous()
    M = g.num_nodes
    if mask is not None and (mask.dtype != torch.uint8 or mask.numel() != M * C or not mask.is_contiguous()):
        raise _lib.GrlError("dropout mask must be contiguous uint8 [M, C]")
    Y = torch.empty(M, C, dtype=torch.float32, device=xs[0].device)
    flags = (_lib.EPI_BIAS if bias is not None else 0) | _act_flag(relu) | (_lib.EPI_DROPOUT if p > 0.0 else 0)
    x1 = xs[1] if len(xs) == 2 else None
    rc = _lib.load().grl_graphconv_fused_fwd(
        g.rowptr.data_ptr(), g.col.data_ptr(), w_edge.data_ptr(), xs[0].data_ptr(), xs[0].stride(0), xs[0].shape[1],
        _lib.ptr(x1), x1.stride(0) if x1 is not None else 0, x1.shape[1] if x1 is not None else 0,
        hi.data_ptr(), lo.data_ptr(), hi.stride(0), _lib.ptr(bias), Y.data_ptr(), Y.stride(0), g.B, g.N, g.L, C,
        flags, float(p), _lib.ptr(mask), int(seed) & (2**64 - 1), int(stream_id), _lib.stream_ptr(Y.device))
    if rc == _lib.GRL_E_UNSUP:
        return None
    _lib.check(rc, "grl_graphconv_fused_fwd")
    _lib.count_launch(1)
    return Y


class AttentionConfig:
    """``tensor_cores``: NodeSelfAtten on the tcgen05 kernels of csrc/attention_tc.cu where they are instantiated --
    (fp32, H=128, d=16), (bf16, 128, 16), (bf16, 256, 32) -- and, for fp32, while the graph has at most
    ``max_nodes_fp32`` nodes (one TMEM accumulator over all keys: its truncating adds cost ~1.5e-8 per 16 keys and
    product, 3e-6 at 500 nodes).  Everything else runs the fp32 FFMA kernels of csrc/attention.cu."""
    tensor_cores = True
    max_nodes_fp32 = 2048


def _tc_pieces(dtype: torch.dtype, H: int, D: int):
    import ctypes
    a, b, c = ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(0)
    ok = _lib.load().grl_selfatten_tc_pieces(_DTYPE_CODE[dtype], H, D, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
    return (a.value, b.value, c.value) if ok == 1 else None


_DTYPE_CODE = {torch.float32: _lib.GRL_F32, torch.bfloat16: _lib.GRL_BF16}


def _split_pieces(x: torch.Tensor, np_: int) -> torch.Tensor:
    """[rows, C] -> bf16 [np, rows, C] (csrc/attention_tc.cu grl_attn_split); a contiguous bf16 x with np = 1 is itself"""
    rows, C = x.shape
    if np_ == 1 and x.dtype == torch.bfloat16 and x.is_contiguous():
        return x.view(1, rows, C)
    out = torch.empty(np_, rows, C, dtype=torch.bfloat16, device=x.device)
    _lib.check(_lib.load().grl_attn_split(x.data_ptr(), _row_major_2d(x, "x"), _lib.dtype_code(x), rows, C, 0, np_,
                                          out.data_ptr(), _lib.stream_ptr(x.device)), "grl_attn_split")
    _lib.count_launch(1)
    return out


class _SelfAttentionTC(torch.autograd.Function):
    """gamma * (softmax(f g^T) h) + V per graph on the tensor cores (csrc/attention_tc.cu); saves o, the row max and the
    row sum (3 floats x N per graph) instead of the N x N probabilities."""

    @staticmethod
    def forward(ctx, f, g, h, gamma, V, pieces):
        B, N, H = h.shape
        D = f.shape[2]
        np_f, nps_b, np_b = pieces
        Nt = B * N
        f2, g2, h2 = f.reshape(Nt, D), g.reshape(Nt, D), h.reshape(Nt, H)
        V = V.contiguous()
        gam = gamma.detach().float().contiguous()
        fp, gp, hp = _split_pieces(f2, np_f), _split_pieces(g2, np_f), _split_pieces(h2, np_f)
        out = torch.empty_like(V)
        need_grad = any(ctx.needs_input_grad)
        o = torch.empty(Nt, H, dtype=torch.float32, device=V.device) if need_grad else None
        m = torch.empty(Nt, dtype=torch.float32, device=V.device) if need_grad else None
        l = torch.empty(Nt, dtype=torch.float32, device=V.device) if need_grad else None
        _lib.check(_lib.load().grl_selfatten_tc_fwd(fp.data_ptr(), gp.data_ptr(), hp.data_ptr(), V.data_ptr(),
                                                    gam.data_ptr(), out.data_ptr(), _lib.ptr(o), _lib.ptr(m), _lib.ptr(l),
                                                    _lib.dtype_code(V), B, N, H, D, _lib.stream_ptr(V.device)),
                   "grl_selfatten_tc_fwd")
        _lib.count_launch(1)
        if need_grad:
            # the backward reads prefixes of the forward's pieces (the first k pieces of a 3-way split ARE the k-way split)
            ctx.save_for_backward(fp, gp, hp, gam, o, m, l)
            ctx.shape = (B, N, H, D, f.dtype)
            ctx.pieces = pieces
        return out

    @staticmethod
    def backward(ctx, dOut):
        fp, gp, hp, gam, o, m, l = ctx.saved_tensors
        B, N, H, D, dt = ctx.shape
        np_f, nps_b, np_b = ctx.pieces
        Nt = B * N
        lib = _lib.load()
        dOut = dOut.contiguous()
        dev = dOut.device
        # the kernels index pieces as [np][Nt, C]: with fewer pieces than the forward saved, the leading ones are used
        do_p = torch.empty(np_b, Nt, H, dtype=torch.bfloat16, device=dev)
        Dvec = torch.empty(Nt, dtype=torch.float32, device=dev)
        part = torch.empty(lib.grl_selfatten_tc_prep_blocks(Nt), H, dtype=torch.float32, device=dev)
        df = torch.empty(B, N, D, dtype=dt, device=dev)
        dg = torch.empty(B, N, D, dtype=dt, device=dev)
        dh = torch.empty(B, N, H, dtype=dt, device=dev)
        _lib.check(lib.grl_selfatten_tc_bwd(dOut.data_ptr(), o.data_ptr(), m.data_ptr(), l.data_ptr(), fp.data_ptr(),
                                            gp.data_ptr(), hp.data_ptr(), gam.data_ptr(), do_p.data_ptr(), Dvec.data_ptr(),
                                            part.data_ptr(), df.data_ptr(), dg.data_ptr(), dh.data_ptr(),
                                            _DTYPE_CODE[dt], B, N, H, D, _lib.stream_ptr(dev)), "grl_selfatten_tc_bwd")
        _lib.count_launch(3)
        dgamma = colsum(part)
        return df, dg, dh, dgamma, dOut, None


class _SelfAttention(torch.autograd.Function):
    """gamma * (softmax(f g^T) h) + V per graph (csrc/attention.cu).  f, g: [B, N, d]; h, V: [B, N, H]."""

    @staticmethod
    def forward(ctx, f, g, h, gamma, V):
        B, N, H = h.shape
        D = f.shape[2]
        f, g, h, V = f.contiguous(), g.contiguous(), h.contiguous(), V.contiguous()
        gam = gamma.detach().float().contiguous()
        out = torch.empty_like(V)
        need_grad = any(ctx.needs_input_grad)
        Np = (N + 63) // 64 * 64
        tiled = (H, D) in ((64, 16), (128, 16), (256, 32))      # csrc/attention.cu instantiations; else generic kernels
        o = torch.empty(B * N, H, dtype=torch.float32, device=V.device) if need_grad else None
        P = torch.empty(B, N, Np, dtype=torch.float32, device=V.device) if (need_grad or not tiled) else None
        _lib.check(_lib.load().grl_selfatten_fwd(f.data_ptr(), g.data_ptr(), h.data_ptr(), V.data_ptr(), gam.data_ptr(),
                                                 out.data_ptr(), _lib.ptr(o), _lib.ptr(P), _lib.dtype_code(V), B, N, H, D,
                                                 Np, _lib.stream_ptr(V.device)), "grl_selfatten_fwd")
        _lib.count_launch(1)
        if need_grad:
            ctx.save_for_backward(f, g, h, gam, o, P)
        return out

    @staticmethod
    def backward(ctx, dOut):
        f, g, h, gam, o, P = ctx.saved_tensors
        B, N, H = h.shape
        D = f.shape[2]
        Np = P.shape[2]
        dOut = dOut.contiguous()
        dS = torch.empty_like(P)
        df, dg, dh = torch.empty_like(f), torch.empty_like(g), torch.empty_like(h)
        n_qb = (N + 63) // 64
        part = torch.empty(B * n_qb, H, dtype=torch.float32, device=h.device)
        _lib.check(_lib.load().grl_selfatten_bwd(dOut.data_ptr(), P.data_ptr(), o.data_ptr(), f.data_ptr(), g.data_ptr(),
                                                 h.data_ptr(), gam.data_ptr(), dS.data_ptr(), df.data_ptr(),
                                                 dg.data_ptr(), dh.data_ptr(), part.data_ptr(), _lib.dtype_code(h), B, N,
                                                 H, D, Np, _lib.stream_ptr(h.device)), "grl_selfatten_bwd")
        _lib.count_launch(2)
        dgamma = colsum(part)                     # deterministic reduction of the per-block partial sums
        return df, dg, dh, dgamma, dOut


def self_attention(f_out: torch.Tensor, g_out: torch.Tensor, h_out: torch.Tensor, gamma: torch.Tensor,
                   V: torch.Tensor) -> torch.Tensor:
    """NodeSelfAtten core (reference robust_gcn.py:94-96): gamma * (softmax(f g^T) h) + V, per graph,
    unscaled and unmasked.  Tensor-core kernels (csrc/attention_tc.cu) where instantiated, else the tiled fp32 kernels of
    csrc/attention.cu; neither writes an N x N tensor in eval, the tensor-core path not in training either."""
    B, N, H = h_out.shape
    D = f_out.shape[2]
    if (AttentionConfig.tensor_cores and V.dtype in _DTYPE_CODE and f_out.dtype == V.dtype and g_out.dtype == V.dtype and
            h_out.dtype == V.dtype and (V.dtype == torch.bfloat16 or N <= AttentionConfig.max_nodes_fp32)):
        pieces = _tc_pieces(V.dtype, H, D)
        if pieces is not None:
            return _SelfAttentionTC.apply(f_out, g_out, h_out, gamma, V, pieces)
    return _SelfAttention.apply(f_out, g_out, h_out, gamma, V)
