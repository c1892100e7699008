"""ctypes binding of ``libtopo_b200.so`` (C ABI declared in ``include/topo_b200.h``).

PyTorch is plumbing only: it owns device memory and the CUDA stream; every kernel on the hot path is
launched through the functions bound here.  Loading fails loudly if the library is missing — there is
no PyTorch/CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("TOPO_LIB") or os.path.join(HERE, "libtopo_b200.so")   # TOPO_LIB: another build of the same library (A/B runs)

ACT_NONE, ACT_SILU, ACT_GEGLU, ACT_GELU, ACT_RELU = 0, 1, 2, 3, 4

# every symbol include/topo_b200.h declares (tests check the library exports all of them)
SYMBOLS = [
    "topo_create", "topo_destroy", "topo_last_error", "topo_last_kernel", "topo_version", "topo_lora_linear", "topo_conv3x3",
    "topo_attention", "topo_attention_lcol", "topo_attention_varlen", "topo_embed_sum", "topo_groupnorm_silu", "topo_groupnorm_ws_floats", "topo_layernorm",
    "topo_zeroconv_residual", "topo_timestep_embedding", "topo_ddim_step", "topo_ddim_step_x0", "topo_conv_direct",
    "topo_nchw_f32_to_nhwc_bf16", "topo_nhwc_to_nchw_f32", "topo_upsample2x", "topo_im2col3x3_s2", "topo_im2col3x3_s2_pad", "topo_tile_metrics", "topo_im2col3x3_s1_k64",
    "topo_add_bf16", "topo_tile_to_u8", "topo_tile_to_u8_round", "topo_softmax_rows", "topo_geglu_tile",
    "topo_groupnorm_silu_static", "topo_groupnorm_static_max_floats",
]


class TopoLinear(C.Structure):
    """Mirror of ``topo_linear_t``."""
    _fields_ = [
        ("W", C.c_void_p), ("bias", C.c_void_p), ("A", C.c_void_p), ("B", C.c_void_p),
        ("N", C.c_int32), ("r_pad", C.c_int32), ("s", C.c_float), ("alpha", C.c_float),
        ("out", C.c_void_p), ("out_ld", C.c_int32), ("out_col", C.c_int32), ("out_fp32", C.c_int32),
        ("res", C.c_void_p), ("res_ld", C.c_int32), ("res_col", C.c_int32),
        ("rowbias", C.c_void_p), ("rowbias_ld", C.c_int32), ("rows_per_group", C.c_int32),
        ("act", C.c_int32),
    ]


class TopoError(RuntimeError):
    pass


_lib: Optional[C.CDLL] = None
_handles: Dict[int, C.c_void_p] = {}


def load_library(path: str = LIB_PATH) -> C.CDLL:
    """dlopen the in-tree library; raise (never fall back) when it is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(path):
        raise TopoError(
            f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). The B200 path has no PyTorch/CPU fallback.")
    lib = C.CDLL(path)
    lib.topo_last_error.restype = C.c_char_p
    lib.topo_last_kernel.restype = C.c_char_p
    lib.topo_version.restype = C.c_int
    lib.topo_create.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    lib.topo_destroy.argtypes = [C.c_void_p]
    vp, i32, f32, i64 = C.c_void_p, C.c_int, C.c_float, C.c_int64
    lib.topo_lora_linear.argtypes = [vp, vp, i32, i32, i32, C.POINTER(TopoLinear), i32, vp]
    lib.topo_conv3x3.argtypes = [vp, vp, i32, i32, i32, i32, i32, C.POINTER(TopoLinear), vp]
    lib.topo_attention.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp]
    lib.topo_attention_lcol.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, f32, i32, vp]
    lib.topo_attention_varlen.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, f32, vp, vp]
    lib.topo_embed_sum.argtypes = [vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp]
    lib.topo_groupnorm_silu.argtypes = [vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, vp, vp]
    lib.topo_groupnorm_ws_floats.argtypes = [i32, i32]
    lib.topo_groupnorm_silu_static.argtypes = [vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, f32, i32, vp]
    lib.topo_groupnorm_static_max_floats.argtypes = []
    lib.topo_geglu_tile.argtypes = [i32]
    lib.topo_layernorm.argtypes = [vp, vp, i32, vp, vp, vp, i32, i32, i32, f32, vp]
    lib.topo_zeroconv_residual.argtypes = [vp, vp, i32, vp, vp, f32, vp, i32, vp, i32, i32, i32, i32, vp]
    lib.topo_timestep_embedding.argtypes = [vp, vp, vp, i32, i32, vp]
    lib.topo_ddim_step.argtypes = [vp, vp, vp, vp, i64, i32, f32, f32, f32, f32, vp, vp]
    lib.topo_ddim_step_x0.argtypes = [vp, vp, vp, vp, vp, i64, i32, f32, f32, f32, f32, vp, vp]
    lib.topo_conv_direct.argtypes = [vp, vp, i32, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, i32, i32, vp,
                                     i32, i32, vp]
    lib.topo_nchw_f32_to_nhwc_bf16.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp]
    lib.topo_nhwc_to_nchw_f32.argtypes = [vp, vp, i32, i32, vp, i32, i32, i32, vp]
    lib.topo_upsample2x.argtypes = [vp, vp, i32, vp, i32, i32, i32, i32, i32, vp]
    lib.topo_im2col3x3_s2.argtypes = [vp, vp, i32, vp, i32, i32, i32, i32, vp]
    lib.topo_tile_metrics.argtypes = [vp, vp, vp, i32, i32, i32, i64, vp, i32, vp, vp, vp]
    lib.topo_im2col3x3_s2_pad.argtypes = [vp, vp, i32, vp, i32, i32, i32, i32, i32, vp]
    lib.topo_im2col3x3_s1_k64.argtypes = [vp, vp, i32, vp, i32, i32, i32, i32, vp]
    lib.topo_add_bf16.argtypes = [vp, vp, i32, vp, i32, vp, i32, i32, i32, vp]
    lib.topo_tile_to_u8.argtypes = [vp, vp, i32, i32, vp, i64, i32, vp]
    lib.topo_tile_to_u8_round.argtypes = [vp, vp, i32, i32, vp, i64, i32, vp]
    lib.topo_softmax_rows.argtypes = [vp, vp, i32, vp, i32, i32, i32, f32, vp]
    _lib = lib
    return lib


def last_error() -> str:
    return load_library().topo_last_error().decode()


def last_kernel() -> str:
    return load_library().topo_last_kernel().decode()


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        raise TopoError(f"{what or 'libtopo_b200'} failed (code {rc}): {last_error()}")


def handle(device: Optional[int] = None) -> C.c_void_p:
    """One handle per (process, device)."""
    lib = load_library()
    if device is None:
        if not torch.cuda.is_available():
            raise TopoError("no CUDA device: the B200 path has no CPU fallback")
        device = torch.cuda.current_device()
    h = _handles.get(device)
    if h is None:
        h = C.c_void_p()
        check(lib.topo_create(C.byref(h), device), "topo_create")
        _handles[device] = h
    return h


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def make_linear(W, out, N, *, bias=None, A=None, B=None, r_pad=0, s=1.0, alpha=1.0, out_ld=None, out_col=0,
                out_fp32=False, res=None, res_ld=0, res_col=0, rowbias=None, rowbias_ld=0, rows_per_group=1,
                act=ACT_NONE) -> TopoLinear:
    """Fill a ``topo_linear_t`` from torch tensors (pointers only; the caller keeps them alive)."""
    seg = TopoLinear()
    seg.W, seg.bias, seg.A, seg.B = ptr(W), ptr(bias), ptr(A), ptr(B)
    seg.N, seg.r_pad, seg.s, seg.alpha = N, r_pad, s, alpha
    seg.out = ptr(out)
    seg.out_ld = out_ld if out_ld is not None else (N // 2 if act == ACT_GEGLU else N)
    seg.out_col, seg.out_fp32 = out_col, int(out_fp32)
    seg.res, seg.res_ld, seg.res_col = ptr(res), res_ld, res_col
    seg.rowbias, seg.rowbias_ld, seg.rows_per_group = ptr(rowbias), rowbias_ld, rows_per_group
    seg.act = act
    return seg
