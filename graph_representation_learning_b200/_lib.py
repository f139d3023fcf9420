"""ctypes binding of the C-ABI library ``lib/libgrl_b200.so`` (declared in ``include/grl_b200.h``).

There is no CPU or eager fallback: if the library is missing the import of any compute entry
point raises, and every non-zero return code becomes a ``GrlError`` carrying ``grl_last_error()``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import c_char_p, c_float, c_int, c_int64, c_size_t, c_uint32, c_uint64, c_void_p

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libgrl_b200.so")

GRL_F32, GRL_BF16, GRL_F16 = 0, 1, 2
EPI_BIAS, EPI_RELU, EPI_DROPOUT, EPI_LEAKY, EPI_LEAKY02 = 1, 2, 4, 8, 16
GRL_E_BADARG, GRL_E_ALIGN, GRL_E_UNSUP, GRL_E_WORKSPACE = -1, -2, -3, -4

_DTYPES = {torch.float32: GRL_F32, torch.bfloat16: GRL_BF16, torch.float16: GRL_F16}


class GrlError(RuntimeError):
    pass


# name -> (restype, argtypes); mirrors include/grl_b200.h one to one (tests check the two agree)
SIGNATURES = {
    "grl_abi_version": (c_int, []),
    "grl_last_error": (c_char_p, []),
    "grl_device_ok": (c_int, []),
    "grl_csr_count": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "grl_scan_workspace_bytes": (c_size_t, [c_int64]),
    "grl_scan_inclusive_i32": (c_int, [c_void_p, c_int64, c_void_p, c_size_t, c_void_p]),
    "grl_csr_fill": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "grl_coo_count": (c_int, [c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p]),
    "grl_coo_fill": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                             c_void_p, c_void_p]),
    "grl_csr_transpose_workspace_bytes": (c_size_t, [c_int64, c_int64]),
    "grl_csr_bitmap_bytes": (c_size_t, [c_int, c_int, c_int]),
    "grl_csr_count_bitmap": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "grl_csr_fill_bitmap": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                    c_void_p]),
    "grl_csr_clamp_rowptr": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p]),
    "grl_csr_transpose": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_int64, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_size_t, c_void_p]),
    "grl_csr_transpose_general": (c_int, [c_void_p, c_void_p, c_int64, c_int, c_int64, c_void_p, c_void_p, c_void_p,
                                          c_void_p, c_size_t, c_void_p]),
    "grl_dropedge_weights": (c_int, [c_void_p, c_void_p, c_int64, c_float, c_uint64, c_uint64, c_void_p, c_void_p]),
    "grl_dropedge_stream": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_float, c_uint64, c_uint64,
                                    c_void_p, c_void_p, c_void_p, c_void_p]),
    "grl_stream_from_weights": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_void_p, c_void_p,
                                        c_void_p]),
    "grl_typed_transpose": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p]),
    "grl_csc_typed_regroup": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                      c_void_p]),
    "grl_spmm_fwd_stream": (c_int, [c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int,
                                    c_int, c_int, c_void_p]),
    "grl_graphconv_fused_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_int64, c_int,
                                        c_void_p, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_int, c_int,
                                        c_int, c_uint32, c_float, c_void_p, c_uint64, c_uint64, c_void_p]),
    "grl_graphconv_fused_set_mode": (c_int, [c_int]),
    "grl_spmm_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int,
                             c_int, c_int, c_void_p]),
    "grl_spmm_set_tile_mode": (c_int, [c_int]),
    "grl_spmm_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_void_p, c_int64, c_int,
                             c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "grl_bias_act_dropout_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_int64, c_int,
                                         c_uint32, c_float, c_void_p, c_uint64, c_uint64, c_void_p]),
    "grl_bias_act_dropout_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64,
                                         c_int, c_uint32, c_float, c_void_p]),
    "grl_selfatten_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                  c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "grl_selfatten_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int,
                                  c_void_p]),
    "grl_selfatten_tc_pieces": (c_int, [c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "grl_attn_split": (c_int, [c_void_p, c_int64, c_int, c_int64, c_int, c_void_p, c_int, c_void_p, c_void_p]),
    "grl_selfatten_tc_fwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "grl_selfatten_tc_prep_blocks": (c_size_t, [c_int64]),
    "grl_selfatten_tc_bwd": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                     c_int, c_int, c_void_p]),
    "grl_reduce_workspace_bytes": (c_size_t, []),
    "grl_ce_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p,
                           c_void_p]),
    "grl_ce_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p, c_void_p, c_void_p,
                           c_void_p, c_int64, c_void_p]),
    "grl_l2_norm": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p]),
    "grl_l2_norm_static": (c_int, [c_void_p, c_int64, c_void_p, c_void_p]),
    "grl_adam_step": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int64, c_float, c_float, c_float, c_float,
                              c_float, c_int64, c_void_p, c_float, c_void_p]),
    "grl_colsum": (c_int, [c_void_p, c_int64, c_int, c_int64, c_int, c_void_p, c_void_p]),
    "grl_gemm_nt": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int,
                            c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int,
                            c_void_p, c_int64, c_int, c_int, c_int, c_int,
                            c_void_p, c_uint32, c_float, c_void_p, c_uint64, c_uint64, c_void_p]),
    "grl_gemm_x3_set_promo": (c_int, [c_int]),
    "grl_gemm_nt_bf16_set_persistent": (c_int, [c_int]),
    "grl_gemm_bc_set_multicast": (c_int, [c_int]),
    "grl_gemm_bc_set_pairs": (c_int, [c_int]),
    "grl_rng_set_offset_ptr": (c_int, [c_void_p]),
    "grl_gemm_bc_set_promo": (c_int, [c_int]),
    "grl_gemm_tn_workspace_bytes": (c_size_t, [c_int, c_int]),
    "grl_gemm_tn": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int64,
                            c_void_p, c_size_t, c_void_p]),
    "grl_gemm_tn_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int, c_int64,
                                 c_void_p, c_size_t, c_void_p]),
    "grl_split_tf32": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "grl_split_tf32_strided": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "grl_split_tf32_mix_strided": (c_int, [c_void_p, c_int64, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int,
                                           c_void_p]),
    "grl_dz_split_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "grl_dz_split": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int64,
                             c_int, c_uint32, c_float, c_void_p, c_void_p, c_size_t, c_void_p]),
    "grl_dz_colsum_bf16": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_uint32, c_float,
                                   c_void_p, c_void_p, c_size_t, c_void_p]),
    "grl_split_act_mn": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "grl_gemm_tn_bc": (c_int, [c_void_p, c_int64, c_int, c_void_p, c_int64, c_void_p, c_int64, c_int, c_int64, c_void_p,
                               c_int, c_void_p, c_size_t, c_void_p]),
    "grl_split_tf32_mix": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_int64, c_int, c_void_p]),
    "grl_bn_workspace_bytes": (c_size_t, [c_int64, c_int]),
    "grl_bn_act_fwd": (c_int, [c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_void_p, c_float, c_float, c_int, c_float,
                               c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_size_t, c_void_p]),
    "grl_bn_act_bwd": (c_int, [c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_int64, c_void_p, c_void_p, c_void_p, c_float,
                               c_int, c_void_p, c_int64, c_void_p, c_void_p, c_int64, c_int, c_void_p, c_size_t, c_void_p]),
}

GEMM_BF16, GEMM_TF32, GEMM_TF32X3, GEMM_TF32_BF16C = 0, 1, 2, 3

_lib = None


def load() -> ctypes.CDLL:
    """dlopen the library (once) and attach prototypes.  Raises if it was never built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise GrlError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C graph_representation_learning_b200/csrc`). There is no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header / library mismatch: fail loudly
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = load().grl_last_error()
        raise GrlError(f"{what} failed with code {rc}: {msg.decode() if msg else '?'}")


def dtype_code(t: torch.Tensor) -> int:
    try:
        return _DTYPES[t.dtype]
    except KeyError:
        raise GrlError(f"dtype {t.dtype} is not supported by the B200 kernels") from None


def ptr(t) -> int:
    return 0 if t is None else t.data_ptr()


def stream_ptr(device=None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str) -> None:
    if not t.is_cuda:
        raise GrlError(f"{name} must live on a CUDA device (got {t.device}); the B200 path has no CPU fallback")


# Launch counter: bench.py reports how many of OUR kernels ran inside the timed region.
class _Counter:
    launches = 0


def count_launch(n: int = 1) -> None:
    _Counter.launches += n


def launches() -> int:
    return _Counter.launches
