"""Tensor-level wrappers over the C ABI (one function per ``topo_*`` entry point) plus the stand-alone
module forwards of the boundary rows a1 (``LoRALinear``) and a5/u6 (``CrossAttention``).

All tensors are CUDA tensors; activations bf16 row-major.  Nothing here computes with PyTorch ops beyond
allocation, dtype casts and zero padding of weights — the arithmetic is in ``csrc/*.cu``.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch

from . import capi
from .capi import ACT_GEGLU, ACT_GELU, ACT_NONE, ACT_RELU, ACT_SILU, TopoLinear, make_linear

BF16 = torch.bfloat16


def _cuda(t: torch.Tensor, what: str) -> None:
    if not t.is_cuda:
        raise capi.TopoError(f"{what}: expected a CUDA tensor — the B200 path has no CPU fallback "
                             "(the fp32 checker lives in oracle/ and is test infrastructure)")


# ----------------------------------------------------------------------------------------------
# weight preparation (host side, once per checkpoint / surgery)
# ----------------------------------------------------------------------------------------------
def lora_rpad(r: int) -> int:
    if r <= 0:
        return 0
    if r <= 16:
        return 16
    if r <= 32:
        return 32
    raise ValueError(f"LoRA rank {r} > 32 is not supported by the fused kernel")


def prep_lora(A: torch.Tensor, B: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor, int]:
    """Zero-pad LoRA down ``A [r,K]`` / up ``B [N,r]`` to the UMMA-legal rank (16 or 32), bf16."""
    r = A.shape[0]
    rp = lora_rpad(r)
    Ap = torch.zeros(rp, A.shape[1], dtype=BF16, device=A.device)
    Ap[:r] = A.to(BF16)
    Bp = torch.zeros(B.shape[0], rp, dtype=BF16, device=B.device)
    Bp[:, :r] = B.to(BF16)
    return Ap.contiguous(), Bp.contiguous(), rp


def head_pad(d: int) -> int:
    """Head dim padded to the next size the attention kernel is instantiated for."""
    for dp in (48, 64, 80, 160):
        if d <= dp and (dp - d) < 16:
            return dp
    raise ValueError(f"head dim {d} unsupported (kernels are instantiated for 40->48, 80, 160)")


def pad_head_rows(W: torch.Tensor, heads: int, d: int, dp: int) -> torch.Tensor:
    """[heads*d, K] -> [heads*dp, K] with zero rows, so q/k/v come out of the GEMM already padded."""
    if d == dp:
        return W.contiguous()
    out = torch.zeros(heads, dp, W.shape[1], dtype=W.dtype, device=W.device)
    out[:, :d] = W.reshape(heads, d, -1)
    return out.reshape(heads * dp, -1).contiguous()


def pad_head_cols(W: torch.Tensor, heads: int, d: int, dp: int) -> torch.Tensor:
    """[N, heads*d] -> [N, heads*dp] with zero columns (to_out consumes the padded attention output)."""
    if d == dp:
        return W.contiguous()
    out = torch.zeros(W.shape[0], heads, dp, dtype=W.dtype, device=W.device)
    out[:, :, :d] = W.reshape(W.shape[0], heads, d)
    return out.reshape(W.shape[0], heads * dp).contiguous()


def geglu_index(n_rows: int, device) -> torch.Tensor:
    """Row permutation for GEGLU ``proj`` ([2*inner, K]: first half values, second half gates): every N tile of the
    kernel (``topo_geglu_tile(N)`` = 128 or 256 rows) holds tile/2 value rows followed by their tile/2 gate rows."""
    tile = capi.load_library().topo_geglu_tile(n_rows)
    inner, half = n_rows // 2, tile // 2
    assert inner % half == 0
    idx = []
    for t in range(inner // half):
        idx += list(range(t * half, (t + 1) * half)) + list(range(inner + t * half, inner + (t + 1) * half))
    return torch.tensor(idx, device=device)


def geglu_interleave(W: torch.Tensor, bias: Optional[torch.Tensor], tile: Optional[int] = None):
    """Apply ``geglu_index`` to the weight rows (and bias) of a GEGLU projection."""
    if tile is None:
        idx = geglu_index(W.shape[0], W.device)
        return W[idx].contiguous(), (None if bias is None else bias[idx].contiguous())
    inner = W.shape[0] // 2
    half = tile // 2
    assert inner % half == 0
    idx = []
    for t in range(inner // half):
        idx += list(range(t * half, (t + 1) * half))
        idx += list(range(inner + t * half, inner + (t + 1) * half))
    idx = torch.tensor(idx, device=W.device)
    return W[idx].contiguous(), (None if bias is None else bias[idx].contiguous())


def conv3x3_weight(w: torch.Tensor) -> torch.Tensor:
    """[Cout, Cin, 3, 3] -> [Cout, 9*Cin] tap-major (column = (ky*3+kx)*Cin + ci), bf16."""
    return w.permute(0, 2, 3, 1).reshape(w.shape[0], -1).to(BF16).contiguous()


def conv_direct_weight(w: torch.Tensor) -> torch.Tensor:
    """[Cout, Cin, k, k] -> [k*k*Cin, Cout] (Cout contiguous) for the narrow direct convolution."""
    return w.permute(2, 3, 1, 0).reshape(-1, w.shape[0]).to(BF16).contiguous()


# ----------------------------------------------------------------------------------------------
# raw op wrappers
# ----------------------------------------------------------------------------------------------
def lora_linear_raw(x: torch.Tensor, M: int, K: int, x_ld: int, segs: Sequence[TopoLinear]) -> None:
    arr = (TopoLinear * len(segs))(*segs)
    capi.check(capi.load_library().topo_lora_linear(capi.handle(), x.data_ptr(), M, K, x_ld, arr, len(segs),
                                                    capi.stream()), "topo_lora_linear")


def conv3x3_raw(x: torch.Tensor, B: int, H: int, W: int, Cc: int, x_ld: int, seg: TopoLinear) -> None:
    capi.check(capi.load_library().topo_conv3x3(capi.handle(), x.data_ptr(), B, H, W, Cc, x_ld, seg, capi.stream()),
               "topo_conv3x3")


def linear(x: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor] = None, A=None, Bm=None, s: float = 1.0,
           res: Optional[torch.Tensor] = None, act: int = ACT_NONE, alpha: float = 1.0, out_fp32=False):
    """out = alpha*(x W^T + s (x A^T) Bm^T + bias) + res on 2-D bf16 ``x [M,K]``; returns ``[M,N]``."""
    _cuda(x, "linear")
    M, K = x.shape
    N = W.shape[0]
    keep = []
    rp = 0
    if A is not None:
        A, Bm, rp = prep_lora(A, Bm)
        keep += [A, Bm]
    Wb = W.to(BF16).contiguous()
    if act == ACT_GEGLU:
        Wb, bias = geglu_interleave(Wb, bias)
    bias = None if bias is None else bias.float().contiguous()
    n_out = N // 2 if act == ACT_GEGLU else N
    out = torch.empty(M, n_out, dtype=torch.float32 if out_fp32 else BF16, device=x.device)
    seg = make_linear(Wb, out, N, bias=bias, A=A, B=Bm, r_pad=rp, s=s, alpha=alpha, out_fp32=out_fp32,
                      res=res, res_ld=0 if res is None else res.stride(0), act=act)
    lora_linear_raw(x, M, K, x.stride(0), [seg])
    return out


def conv3x3(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, res=None, rowbias=None,
            act: int = ACT_NONE, out_fp32=False):
    """3x3/s1/p1 conv on NHWC bf16 ``x [B,H,W,C]`` with ``w [Cout,Cin,3,3]``; returns NHWC ``[B,H,W,Cout]``."""
    _cuda(x, "conv3x3")
    Bn, H, Wd, Cc = x.shape
    Cout = w.shape[0]
    wk = conv3x3_weight(w)
    bias = None if bias is None else bias.float().contiguous()
    out = torch.empty(Bn, H, Wd, Cout, dtype=torch.float32 if out_fp32 else BF16, device=x.device)
    seg = make_linear(wk, out, Cout, bias=bias, out_fp32=out_fp32, res=res, res_ld=0 if res is None else res.shape[-1],
                      rowbias=rowbias, rowbias_ld=0 if rowbias is None else rowbias.stride(0),
                      rows_per_group=H * Wd, act=act)
    conv3x3_raw(x, Bn, H, Wd, Cc, Cc, seg)
    return out


def attention(q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, heads: int, dp: int, scale: float,
              kv_len: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None, l_col: int = -1) -> torch.Tensor:
    """q [B,Nq,heads*dp], k/v [B,Nk,heads*dp] (bf16, padded head columns zero) -> o [B,Nq,heads*dp].
    ``kv_len`` (int32 [B] on the device): keys at or beyond the sample's length are masked (right-padded prompts).
    ``l_col >= 0``: v holds 1.0 in padded column ``l_col`` of every head and the row sum is read from the P V product."""
    _cuda(q, "attention")
    Bn, Nq, ld = q.shape
    Nk = k.shape[1]
    o = torch.empty(Bn, Nq, heads * dp, dtype=BF16, device=q.device) if out is None else out
    if o.shape != (Bn, Nq, heads * dp) or o.dtype != BF16 or not o.is_contiguous():
        raise capi.TopoError(f"attention: out must be a contiguous bf16 [{Bn},{Nq},{heads * dp}] tensor")
    if kv_len is not None:
        kv_len = kv_len.to(device=q.device, dtype=torch.int32).contiguous()
        capi.check(capi.load_library().topo_attention_varlen(
            capi.handle(), q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), Bn, heads, Nq, Nk, dp,
            q.stride(1), k.stride(1), v.stride(1), o.stride(1), scale, kv_len.data_ptr(), capi.stream()),
            "topo_attention_varlen")
        return o
    if l_col >= 0:
        capi.check(capi.load_library().topo_attention_lcol(
            capi.handle(), q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), Bn, heads, Nq, Nk, dp,
            q.stride(1), k.stride(1), v.stride(1), o.stride(1), scale, l_col, capi.stream()), "topo_attention_lcol")
        return o
    capi.check(capi.load_library().topo_attention(
        capi.handle(), q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), Bn, heads, Nq, Nk, dp,
        q.stride(1), k.stride(1), v.stride(1), o.stride(1), scale, capi.stream()), "topo_attention")
    return o


_gn_ws = {}
# Which of the (at most two) concurrent kernel streams of a step is issuing: 0 = main / UNet, 1 = the ControlNet trunk on
# the engine's side stream.  The workspace is keyed on this slot, not on the stream handle, so the warm-up pass that
# precedes a CUDA-graph capture (a different stream) creates the very buffers the capture uses: no process-global
# tensor is ever allocated inside a graph's private memory pool.
GN_SLOT = [0]


def groupnorm_ws(B: int, G: int, device) -> torch.Tensor:
    n = capi.load_library().topo_groupnorm_ws_floats(B, G)
    key = (str(device), n, GN_SLOT[0])
    if key not in _gn_ws:
        if torch.cuda.is_current_stream_capturing():
            raise capi.TopoError("groupnorm workspace requested for the first time inside a CUDA-graph capture; run the "
                                 "same shapes once eagerly first (GraphedModel's warm-up does)")
        _gn_ws[key] = torch.zeros(n, dtype=torch.float32, device=device)  # the fused kernel's counters start at zero
    return _gn_ws[key]


def groupnorm_silu(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, G: int, eps: float, silu: bool,
                   out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x NHWC bf16 [B, ..., C] (contiguous pixels) -> same shape."""
    _cuda(x, "groupnorm_silu")
    Bn, Cc = x.shape[0], x.shape[-1]
    HW = x.numel() // (Bn * Cc)
    y = torch.empty_like(x) if out is None else out
    ws = groupnorm_ws(Bn, G, x.device)
    capi.check(capi.load_library().topo_groupnorm_silu(
        capi.handle(), x.data_ptr(), Cc, gamma.data_ptr(), beta.data_ptr(), y.data_ptr(), Cc, Bn, HW, Cc, G, eps,
        int(silu), ws.data_ptr(), capi.stream()), "topo_groupnorm_silu")
    return y


def layernorm(x: torch.Tensor, gamma: torch.Tensor, beta: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    _cuda(x, "layernorm")
    Cc = x.shape[-1]
    M = x.numel() // Cc
    y = torch.empty_like(x)
    capi.check(capi.load_library().topo_layernorm(capi.handle(), x.data_ptr(), Cc, gamma.data_ptr(), beta.data_ptr(),
                                                  y.data_ptr(), Cc, M, Cc, eps, capi.stream()), "topo_layernorm")
    return y


def zeroconv_residual(hc: torch.Tensor, Wz: torch.Tensor, bz: torch.Tensor, scale: float, skip: torch.Tensor,
                      out: torch.Tensor, out_off: int) -> None:
    """out[:, off:off+C] = skip + scale*(hc Wz^T + bz);  hc/skip [M,C] bf16, out [M,out_ld] bf16."""
    M, Cc = hc.shape
    capi.check(capi.load_library().topo_zeroconv_residual(
        capi.handle(), hc.data_ptr(), hc.stride(0), Wz.data_ptr(), bz.data_ptr(), scale, skip.data_ptr(),
        skip.stride(0), out.data_ptr(), M, Cc, out.stride(0), out_off, capi.stream()), "topo_zeroconv_residual")


def timestep_embedding(t: torch.Tensor, dim: int) -> torch.Tensor:
    t = t.float().contiguous()
    emb = torch.empty(t.shape[0], dim, dtype=BF16, device=t.device)
    capi.check(capi.load_library().topo_timestep_embedding(capi.handle(), t.data_ptr(), emb.data_ptr(), t.shape[0], dim,
                                                           capi.stream()), "topo_timestep_embedding")
    return emb


def ddim_step(x: torch.Tensor, eps: torch.Tensor, cfg: bool, w: float, a_t: float, a_prev: float, sigma: float = 0.0,
              noise: Optional[torch.Tensor] = None, out: Optional[torch.Tensor] = None,
              pred_x0: Optional[torch.Tensor] = None) -> torch.Tensor:
    """x' of one DDIM step (CFG combine fused); ``pred_x0`` (fp32, x-shaped) additionally receives the predicted x0."""
    xo = torch.empty_like(x) if out is None else out
    if pred_x0 is not None:
        capi.check(capi.load_library().topo_ddim_step_x0(
            capi.handle(), x.data_ptr(), eps.data_ptr(), xo.data_ptr(), pred_x0.data_ptr(), x.numel(), int(cfg), w, a_t,
            a_prev, sigma, capi.ptr(noise), capi.stream()), "topo_ddim_step_x0")
        return xo
    capi.check(capi.load_library().topo_ddim_step(
        capi.handle(), x.data_ptr(), eps.data_ptr(), xo.data_ptr(), x.numel(), int(cfg), w, a_t, a_prev, sigma,
        capi.ptr(noise), capi.stream()), "topo_ddim_step")
    return xo


def conv_direct(x: torch.Tensor, w_t: torch.Tensor, bias: Optional[torch.Tensor], Cin: int, Cout: int, ksize: int,
                stride: int, act: int = ACT_NONE, add: Optional[torch.Tensor] = None, out_fp32=False,
                out: Optional[torch.Tensor] = None, pad_lo: int = -1) -> torch.Tensor:
    """x NHWC bf16 [B,H,W,Cin]; w_t from ``conv_direct_weight``; returns NHWC [B,Ho,Wo,Cout].  ``pad_lo=0`` with
    stride 2 is diffusers' ``Downsample2D(padding=0)`` (zero row/column appended after the image only)."""
    Bn, H, Wd, _ = x.shape
    pad = ksize // 2
    Ho, Wo = (H + 2 * pad - ksize) // stride + 1, (Wd + 2 * pad - ksize) // stride + 1
    if out is None:
        out = torch.empty(Bn, Ho, Wo, Cout, dtype=torch.float32 if out_fp32 else BF16, device=x.device)
    capi.check(capi.load_library().topo_conv_direct(
        capi.handle(), x.data_ptr(), x.shape[-1], w_t.data_ptr(), capi.ptr(bias), out.data_ptr(), out.shape[-1], Bn, H,
        Wd, Cin, Cout, ksize, stride, pad_lo, act, capi.ptr(add), 0 if add is None else add.shape[-1], int(out_fp32),
        capi.stream()), "topo_conv_direct")
    return out


def embed_sum(ids: torch.Tensor, word: torch.Tensor, pos: torch.Tensor, type0: torch.Tensor) -> torch.Tensor:
    """BERT embeddings ``word[ids] + pos[t] + type[0]`` -> bf16 [B, T, C] (fp32 tables, int ids on the device)."""
    _cuda(ids, "embed_sum")
    Bn, T = ids.shape
    Cc = word.shape[1]
    ids = ids.to(torch.int32).contiguous()
    y = torch.empty(Bn, T, Cc, dtype=BF16, device=ids.device)
    capi.check(capi.load_library().topo_embed_sum(capi.handle(), ids.data_ptr(), word.data_ptr(), pos.data_ptr(),
                                                  type0.data_ptr(), y.data_ptr(), Bn, T, Cc, word.shape[0],
                                                  capi.stream()), "topo_embed_sum")
    return y


def nchw_to_nhwc(x: torch.Tensor) -> torch.Tensor:
    """fp32 NCHW -> bf16 NHWC."""
    Bn, Cc, H, Wd = x.shape
    x = x.float().contiguous()
    y = torch.empty(Bn, H, Wd, Cc, dtype=BF16, device=x.device)
    capi.check(capi.load_library().topo_nchw_f32_to_nhwc_bf16(capi.handle(), x.data_ptr(), y.data_ptr(), Bn, Cc, H * Wd,
                                                              Cc, capi.stream()), "topo_nchw_f32_to_nhwc_bf16")
    return y


def nhwc_to_nchw(x: torch.Tensor) -> torch.Tensor:
    """bf16/fp32 NHWC -> fp32 NCHW."""
    Bn, H, Wd, Cc = x.shape
    y = torch.empty(Bn, Cc, H, Wd, dtype=torch.float32, device=x.device)
    capi.check(capi.load_library().topo_nhwc_to_nchw_f32(capi.handle(), x.data_ptr(), int(x.dtype == torch.float32), x.stride(-2),
                                                         y.data_ptr(), Bn, Cc, H * Wd, capi.stream()),
               "topo_nhwc_to_nchw_f32")
    return y


def upsample2x(x: torch.Tensor) -> torch.Tensor:
    Bn, H, Wd, Cc = x.shape
    y = torch.empty(Bn, 2 * H, 2 * Wd, Cc, dtype=BF16, device=x.device)
    capi.check(capi.load_library().topo_upsample2x(capi.handle(), x.data_ptr(), Cc, y.data_ptr(), Cc, Bn, H, Wd, Cc,
                                                   capi.stream()), "topo_upsample2x")
    return y


def im2col3x3_s1_k64(x: torch.Tensor) -> torch.Tensor:
    """NHWC bf16 [B,H,W,C] with 9*C <= 64 -> [B*H*W, 64] (tap-major columns, zero padded): the narrow input convs as
    one-K-block GEMMs."""
    _cuda(x, "im2col3x3_s1_k64")
    Bn, H, Wd, Cc = x.shape
    y = torch.empty(Bn * H * Wd, 64, dtype=BF16, device=x.device)
    capi.check(capi.load_library().topo_im2col3x3_s1_k64(capi.handle(), x.data_ptr(), x.stride(2), y.data_ptr(), Bn, H, Wd,
                                                         Cc, capi.stream()), "topo_im2col3x3_s1_k64")
    return y


def im2col3x3_s2(x: torch.Tensor, pad_lo: int = 1) -> torch.Tensor:
    """NHWC bf16 -> [B*(H/2)*(W/2), 9*C] for a 3x3 stride-2 conv.  ``pad_lo=1``: symmetric padding 1 (UNet Downsample);
    ``pad_lo=0``: the VAE encoder's pad-(0,1,0,1) form."""
    Bn, H, Wd, Cc = x.shape
    y = torch.empty(Bn * (H // 2) * (Wd // 2), 9 * Cc, dtype=BF16, device=x.device)
    capi.check(capi.load_library().topo_im2col3x3_s2_pad(capi.handle(), x.data_ptr(), Cc, y.data_ptr(), Bn, H, Wd, Cc,
                                                         pad_lo, capi.stream()), "topo_im2col3x3_s2_pad")
    return y


def add(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    Cc = a.shape[-1]
    y = torch.empty_like(a)
    capi.check(capi.load_library().topo_add_bf16(capi.handle(), a.data_ptr(), Cc, b.data_ptr(), Cc, y.data_ptr(), Cc,
                                                 a.numel() // Cc, Cc, capi.stream()), "topo_add_bf16")
    return y


def tile_to_u8(x: torch.Tensor, round_half_even: bool = False) -> torch.Tensor:
    """NHWC [-1,1] (bf16 or fp32) -> uint8 NHWC.  Default: ``x*127.5+127.5``, clip, truncate
    (BaseModel/evaluateCN.py:101-103); ``round_half_even``: ``((x+1)*127.5).clamp(0,255).round()``
    (LoRA_modified/quantitative_metrics.py:206-207)."""
    _cuda(x, "tile_to_u8")
    Cc = x.shape[-1]
    ld = x.stride(-2)  # a channel slice of a wider pixel-major buffer is fine (pixels must be evenly strided)
    y = torch.empty(x.shape, dtype=torch.uint8, device=x.device)
    lib = capi.load_library()
    fn = lib.topo_tile_to_u8_round if round_half_even else lib.topo_tile_to_u8
    capi.check(fn(capi.handle(), x.data_ptr(), int(x.dtype == torch.float32), ld, y.data_ptr(), x.numel() // Cc, Cc,
                  capi.stream()), "topo_tile_to_u8")
    return y


def tile_metrics(pred: torch.Tensor, gt: torch.Tensor, lut: torch.Tensor, bins: int, planar: bool):
    """One pass over T tile pairs holding integral 0..255 values (uint8 or fp32; ``[T,P,3]`` or planar ``[T,3,P]``):
    returns ``(sse int64 [T], hist int32 [T,2,3,bins])`` — exact counts."""
    _cuda(pred, "tile_metrics")
    if pred.shape != gt.shape or pred.dtype != gt.dtype or pred.dtype not in (torch.uint8, torch.float32):
        raise capi.TopoError(f"tile_metrics: pred/gt must be same-shape uint8 or float32 tensors, got {pred.shape} "
                             f"{pred.dtype} / {gt.shape} {gt.dtype}")
    pred, gt = pred.contiguous(), gt.contiguous()
    T = pred.shape[0]
    n_pixels = pred[0].numel() // 3 if T else 1
    sse = torch.zeros(T, dtype=torch.int64, device=pred.device)
    hist = torch.zeros(T, 2, 3, bins, dtype=torch.int32, device=pred.device)
    capi.check(capi.load_library().topo_tile_metrics(
        capi.handle(), pred.data_ptr(), gt.data_ptr(), int(pred.dtype == torch.float32), int(planar), T, n_pixels,
        lut.data_ptr(), bins, sse.data_ptr(), hist.data_ptr(), capi.stream()), "topo_tile_metrics")
    return sse, hist


def softmax_rows(x: torch.Tensor, scale: float) -> torch.Tensor:
    M, N = x.shape
    y = torch.empty_like(x)
    capi.check(capi.load_library().topo_softmax_rows(capi.handle(), x.data_ptr(), x.stride(0), y.data_ptr(), y.stride(0),
                                                     M, N, scale, capi.stream()), "topo_softmax_rows")
    return y


# ----------------------------------------------------------------------------------------------
# stand-alone module forwards (boundary rows a1, a5/u6)
# ----------------------------------------------------------------------------------------------
def _proj_params(mod):
    """(W, bias, A, B, s) of a projection module: nn.Linear, LoRALinear-like (base/A/B/scaling, the
    reference's LoRA_modified/LoRA_utils.py:20-42) or LoRACompatibleLinear (lora_layer.down/up, scale)."""
    if hasattr(mod, "base") and hasattr(mod, "A") and hasattr(mod, "B"):
        return mod.base.weight, mod.base.bias, mod.A.weight, mod.B.weight, float(getattr(mod, "scaling", 2.0))
    if getattr(mod, "lora_layer", None) is not None:
        return mod.weight, mod.bias, mod.lora_layer.down.weight, mod.lora_layer.up.weight, float(getattr(mod, "scale", 1.0))
    if isinstance(mod, torch.nn.Linear):
        return mod.weight, mod.bias, None, None, 1.0
    raise capi.TopoError(
        f"unsupported projection module {type(mod).__name__}: the B200 path accepts nn.Linear, LoRALinear "
        "(base/A/B/scaling) and LoRACompatibleLinear (lora_layer); there is no eager fallback")


def lora_linear_module(mod, x: torch.Tensor) -> torch.Tensor:
    """y = base(x) + B(A(x)) * scaling on the fused kernel; x [..., in] any float dtype on CUDA."""
    _cuda(x, "LoRALinear.forward")
    W, b, A, Bm, s = _proj_params(mod)
    x2 = x.reshape(-1, x.shape[-1]).to(BF16).contiguous()
    y = linear(x2, W, b, A, Bm, s)
    return y.reshape(*x.shape[:-1], W.shape[0]).to(x.dtype)


def cross_attention_module(attn, x: torch.Tensor, context: Optional[torch.Tensor], mask=None) -> torch.Tensor:
    """CrossAttention.forward on the kernels: projections through ``attn.to_q/to_k/to_v/to_out[0]`` as
    *modules* (so any replacement changes the result, LoRA_utils.py:50-58), core through topo_attention."""
    _cuda(x, "CrossAttention.forward")
    if mask is not None:
        raise capi.TopoError("attention masks are not on the reference's path (always None) and are not implemented")
    ctx = x if context is None else context
    h, d = attn.heads, attn.dim_head
    dp = head_pad(d)
    Bn, Nq, _ = x.shape
    Nk = ctx.shape[1]

    def proj(mod, inp):
        W, b, A, Bm, s = _proj_params(mod)
        inp2 = inp.reshape(-1, inp.shape[-1]).to(BF16).contiguous()
        Wp = pad_head_rows(W, h, d, dp)
        Bp = None if Bm is None else pad_head_rows(Bm, h, d, dp)
        bp = None if b is None else pad_head_rows(b[:, None], h, d, dp)[:, 0]
        return linear(inp2, Wp, bp, A, Bp, s)

    q = proj(attn.to_q, x).reshape(Bn, Nq, h * dp)
    k = proj(attn.to_k, ctx).reshape(Bn, Nk, h * dp)
    v = proj(attn.to_v, ctx).reshape(Bn, Nk, h * dp)
    o = attention(q, k, v, h, dp, d ** -0.5)
    W, b, A, Bm, s = _proj_params(attn.to_out[0])
    Wp = pad_head_cols(W, h, d, dp)
    Ap = None if A is None else pad_head_cols(A, h, d, dp)
    y = linear(o.reshape(Bn * Nq, h * dp), Wp, b, Ap, Bm, s)
    return y.reshape(Bn, Nq, -1).to(x.dtype)


def control_branch(attn, hidden: torch.Tensor, scale: float = 1.0, add_to: Optional[torch.Tensor] = None):
    """``scale * to_control(control_states)`` of CrossAttention_LoRA_Control (LoRA_modified/attention_lora.py:61-80)
    as two skinny GEMMs (rank zero-padded to one 64-wide K block); with ``add_to`` the sum is fused into the second
    GEMM's residual epilogue and only applied when the shapes agree, as the reference does (:77-80)."""
    c = attn.control_states
    _cuda(c, "to_control")
    if c.ndim == 4 and hidden.ndim == 3:
        Bn, Cc, H, Wd = c.shape
        c = nchw_to_nhwc(c).reshape(Bn, H * Wd, Cc)
    down, up = attn.to_control[0].weight, attn.to_control[1].weight  # [rc, C], [C, rc]
    rc, Cc = down.shape
    if c.shape[-1] != Cc:   # the reference's nn.Linear raises here too (only batch / spatial mismatches are ignored)
        raise capi.TopoError(f"to_control: control states have {c.shape[-1]} channels, the branch expects {Cc}")
    kp = ((rc + 63) // 64) * 64
    dpad = torch.zeros(kp, Cc, dtype=BF16, device=c.device)
    dpad[:rc] = down.to(BF16)
    upad = torch.zeros(up.shape[0], kp, dtype=BF16, device=c.device)
    upad[:, :rc] = up.to(BF16)
    x2 = c.reshape(-1, Cc).to(BF16).contiguous()
    t = linear(x2, dpad)
    out_shape = (*c.shape[:-1], up.shape[0])
    if add_to is not None and tuple(add_to.shape) != out_shape:
        return add_to
    res = None if add_to is None else add_to.reshape(-1, up.shape[0]).to(BF16).contiguous()
    y = linear(t, upad, res=res, alpha=scale)
    return y.reshape(out_shape).to(hidden.dtype)
