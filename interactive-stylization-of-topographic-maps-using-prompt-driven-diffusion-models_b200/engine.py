"""TopoEngine — the execution plan of ``ControlLDM.apply_model`` / ``decode_first_stage`` on the sm_100a kernels.

The module tree in ``ldm.py`` only holds parameters; this file walks it once (after every weight load / LoRA
surgery), packs the weights into the layouts the kernels want (bf16, K-major, padded heads, GEGLU-interleaved,
tap-major conv filters) and then executes a forward as a flat sequence of C-ABI calls.  PyTorch is used for
device memory (``torch.empty``), dtype casts of inputs and the CUDA stream — no arithmetic.

Data layout in HBM: activations are NHWC bf16, i.e. token matrices ``[B*H*W, C]`` (row-major, C contiguous);
the residual stream, attention q/k/v (heads padded 40→48) and the decoder concat buffers ``[h | skip]`` are all
token matrices, so "rearrange b c h w -> b (h w) c" and ``torch.cat`` cost nothing: producers write straight
into a column slice (``out_ld`` / ``out_col`` of ``topo_linear_t``).

Fusions (BASELINE.json north_star a–e):
 (a) q/k/v in one launch (3 segments, one read of x), LoRA delta folded into the same tcgen05 GEMM
 (b)/(c) flash attention on TMEM; cross-attention K/V are step-invariant and hoisted (``prepare``)
 (d) GroupNorm+SiLU one op
 (e) ControlNet zero-conv + ``hs.pop() + control.pop()`` + ``torch.cat`` = one GEMM epilogue writing the skip
     half of the UNet decoder's concat buffer; ``h += control.pop()`` at the mid block likewise.
Upstream semantics restated in SURVEY.md Appendix A; call sites: ``vis_metrics.py:437`` (sampling),
``…_ctrlora.py:388`` / ``vis_metrics.py:112`` (apply_model), ``vis_metrics.py:450`` (decode).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn as nn

from . import capi, ops
from .capi import ACT_GEGLU, ACT_NONE, ACT_SILU, make_linear
from .ops import BF16

LAUNCHES = [0]  # number of libtopo_b200 kernels launched (bench.py reports it as gpu_launches)


def _count(n: int = 1) -> None:
    LAUNCHES[0] += n


class _Prof:
    """Optional per-launch CUDA-event bracketing (bench.py's roofline leg): records (kind, flops, bytes, ev0, ev1)
    for the tensor-core / attention / norm launches on the stream they are launched on."""
    on = False
    recs: list = []


PROF = _Prof()


def _rec(kind: str, flops: float, nbytes: float, fn, tag: str = ""):
    if not PROF.on:
        return fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = fn()
    e1.record()
    PROF.recs.append((kind, flops, nbytes, e0, e1, tag, capi.last_kernel()))
    return r


def prof_summary(by: str = "kind") -> Dict[str, Dict[str, float]]:
    """``by='kind'``: op class -> {launches, ms, flops, bytes}; ``by='kernel'``: the same keyed by the kernel template
    instance the C side launched (the names of an ncu launch list).  Call after torch.cuda.synchronize()."""
    out: Dict[str, Dict[str, float]] = {}
    for kind, fl, nb, e0, e1, _tag, kern in PROF.recs:
        key = kind if by == "kind" else (kern or kind)
        d = out.setdefault(key, dict(launches=0, ms=0.0, flops=0.0, bytes=0.0, kinds=set()))
        d["kinds"].add(kind)
        d["launches"] += 1
        d["ms"] += e0.elapsed_time(e1)
        d["flops"] += fl
        d["bytes"] += nb
    return out


# ----------------------------------------------------------------------------------------------
# packed weights
# ----------------------------------------------------------------------------------------------
@dataclass
class Lin:
    """One projection ``y = x W^T (+b) + s (x A^T) B^T`` packed for topo_lora_linear."""
    W: torch.Tensor                       # bf16 [N, K]
    bias: Optional[torch.Tensor]          # fp32 [N]
    A: Optional[torch.Tensor] = None      # bf16 [rp, K]
    B: Optional[torch.Tensor] = None      # bf16 [N, rp]
    rp: int = 0
    s: float = 1.0
    r: int = 0                            # true LoRA rank (algorithmic FLOPs are counted on r, not on the padded rp)

    @property
    def N(self) -> int:
        return self.W.shape[0]

    @property
    def K(self) -> int:
        return self.W.shape[1]


def _dev(t: Optional[torch.Tensor], device, dtype) -> Optional[torch.Tensor]:
    return None if t is None else t.detach().to(device=device, dtype=dtype).contiguous()


def pack_linear(mod: nn.Module, device, heads: int = 0, d: int = 0, dp: int = 0, pad: str = "") -> Lin:
    """Pack an ``nn.Linear`` / ``LoRALinear`` / ``LoRACompatibleLinear``.  ``pad='rows'`` zero-pads the output
    features per head (q/k/v: 40→48), ``pad='cols'`` the input features (to_out consumes the padded layout)."""
    W, b, A, Bm, s = ops._proj_params(mod)
    W = W.detach().to(device)
    if A is not None:
        A, Bm = A.detach().to(device), Bm.detach().to(device)
    if b is not None:
        b = b.detach().to(device)
    if pad == "rows" and dp != d:
        W = ops.pad_head_rows(W, heads, d, dp)
        Bm = None if Bm is None else ops.pad_head_rows(Bm, heads, d, dp)
        b = None if b is None else ops.pad_head_rows(b[:, None], heads, d, dp)[:, 0]
    elif pad == "cols" and dp != d:
        W = ops.pad_head_cols(W, heads, d, dp)
        A = None if A is None else ops.pad_head_cols(A, heads, d, dp)
    rp = r = 0
    if A is not None:
        r = A.shape[0]
        A, Bm, rp = ops.prep_lora(A, Bm)
    return Lin(W.to(BF16).contiguous(), _dev(b, device, torch.float32), A, Bm, rp, s, r)


def _lcol_enabled() -> bool:
    import os
    return os.environ.get("TOPO_ATTN_LCOL", "1") != "0"


def pack_conv1x1(conv: nn.Conv2d, device) -> Lin:
    W = conv.weight.detach().to(device).reshape(conv.out_channels, conv.in_channels)
    return Lin(W.to(BF16).contiguous(), _dev(conv.bias, device, torch.float32))


def pack_conv3x3(conv: nn.Conv2d, device) -> Lin:
    return Lin(ops.conv3x3_weight(conv.weight.detach().to(device)), _dev(conv.bias, device, torch.float32))


def pack_geglu(mod: nn.Module, device) -> Lin:
    """GEGLU ``proj`` with rows interleaved per kernel N tile (values | gates) for the fused epilogue."""
    lin = pack_linear(mod, device)
    idx = ops.geglu_index(lin.N, device)
    lin.W = lin.W[idx].contiguous()
    lin.bias = None if lin.bias is None else lin.bias[idx].contiguous()
    lin.B = None if lin.B is None else lin.B[idx].contiguous()
    return lin


@dataclass
class Norm:
    gamma: torch.Tensor
    beta: torch.Tensor
    eps: float
    groups: int = 0


def pack_norm(m: nn.Module, device) -> Norm:
    return Norm(_dev(m.weight, device, torch.float32), _dev(m.bias, device, torch.float32), float(m.eps),
                int(getattr(m, "num_groups", 0)))


# ----------------------------------------------------------------------------------------------
# launch helpers (each = exactly one topo_* call)
# ----------------------------------------------------------------------------------------------
def gemm(x: torch.Tensor, lin: Lin, out: Optional[torch.Tensor] = None, out_col: int = 0,
         res: Optional[torch.Tensor] = None, res_col: int = 0, act: int = ACT_NONE, alpha: float = 1.0,
         rowbias: Optional[torch.Tensor] = None, rows_per_group: int = 1, out_fp32: bool = False) -> torch.Tensor:
    """x [M, K] (row stride arbitrary) → out[:, out_col : out_col+N] ; returns ``out`` (allocated if None)."""
    M = x.shape[0]
    n_out = lin.N // 2 if act == ACT_GEGLU else lin.N
    if out is None:
        out = torch.empty(M, n_out, dtype=torch.float32 if out_fp32 else BF16, device=x.device)
    seg = make_linear(lin.W, out, lin.N, bias=lin.bias, A=lin.A, B=lin.B, r_pad=lin.rp, s=lin.s, alpha=alpha,
                      out_ld=out.stride(0), out_col=out_col, out_fp32=out_fp32, res=res,
                      res_ld=0 if res is None else res.stride(0), res_col=res_col, rowbias=rowbias,
                      rowbias_ld=0 if rowbias is None else rowbias.stride(0), rows_per_group=rows_per_group, act=act)
    fl = 2.0 * M * lin.N * lin.K + 2.0 * M * lin.r * (lin.K + lin.N)
    _rec("gemm_lora" if lin.r else "gemm", fl, 2.0 * (M * lin.K + lin.N * lin.K + M * n_out),
         lambda: ops.lora_linear_raw(x, M, lin.K, x.stride(0), [seg]), f"M{M} N{lin.N} K{lin.K} act{act} res{int(res is not None)}")
    _count()
    return out


def gemm_multi(x: torch.Tensor, lins: Sequence[Lin], out: torch.Tensor, cols: Sequence[int]) -> None:
    """Several projections of the same x in ONE launch when their padded LoRA ranks agree, else one launch each."""
    M = x.shape[0]
    if len({l.rp for l in lins}) == 1 and len({l.K for l in lins}) == 1 and len(lins) <= 3:
        segs = [make_linear(l.W, out, l.N, bias=l.bias, A=l.A, B=l.B, r_pad=l.rp, s=l.s, out_ld=out.stride(0),
                            out_col=c) for l, c in zip(lins, cols)]
        K = lins[0].K
        fl = sum(2.0 * M * l.N * K + 2.0 * M * l.r * (K + l.N) for l in lins)
        nb = 2.0 * (M * K + sum(l.N * K + M * l.N for l in lins))
        _rec("gemm_lora" if any(l.r for l in lins) else "gemm", fl, nb,
             lambda: ops.lora_linear_raw(x, M, K, x.stride(0), segs), f"M{M} N{'+'.join(str(l.N) for l in lins)} K{K} multi")
        _count()
    else:
        for l, c in zip(lins, cols):
            gemm(x, l, out, c)


def conv3(x: torch.Tensor, B: int, H: int, W: int, lin: Lin, out: Optional[torch.Tensor] = None, out_col: int = 0,
          res: Optional[torch.Tensor] = None, res_col: int = 0, rowbias: Optional[torch.Tensor] = None,
          out_fp32: bool = False) -> torch.Tensor:
    """3x3/s1/p1 implicit-GEMM conv on the token matrix x [B*H*W, C]."""
    C = x.shape[1]
    if out is None:
        out = torch.empty(B * H * W, lin.N, dtype=torch.float32 if out_fp32 else BF16, device=x.device)
    seg = make_linear(lin.W, out, lin.N, bias=lin.bias, out_ld=out.stride(0), out_col=out_col, out_fp32=out_fp32,
                      res=res, res_ld=0 if res is None else res.stride(0), res_col=res_col, rowbias=rowbias,
                      rowbias_ld=0 if rowbias is None else rowbias.stride(0), rows_per_group=H * W)
    M = B * H * W
    _rec("conv3x3", 2.0 * M * lin.N * 9 * C, 2.0 * (M * C + lin.N * 9 * C + M * lin.N),
         lambda: ops.conv3x3_raw(x, B, H, W, C, x.stride(0), seg), f"B{B} H{H} C{C} N{lin.N} res{int(res is not None)}")
    _count()
    return out


def gn(x: torch.Tensor, B: int, n: Norm, silu: bool) -> torch.Tensor:
    y = _rec("groupnorm_silu", 0.0, 4.0 * x.numel(),
             lambda: ops.groupnorm_silu(x.view(B, -1, x.shape[1]), n.gamma, n.beta, n.groups, n.eps, silu))
    _count(2)
    return y.view(x.shape)


def ln(x: torch.Tensor, n: Norm) -> torch.Tensor:
    _count()
    return _rec("layernorm", 0.0, 4.0 * x.numel(), lambda: ops.layernorm(x, n.gamma, n.beta, n.eps))


# ----------------------------------------------------------------------------------------------
# packed blocks
# ----------------------------------------------------------------------------------------------
class PRes:
    """ResBlock (upstream openaimodel.py).  ``emb_col`` = this block's column offset in the
    per-net batched time-embedding projection (all ``emb_layers.1`` of a net run as ONE GEMM)."""

    def __init__(self, rb, device):
        self.cin, self.cout = rb.channels, rb.out_channels
        self.n1 = pack_norm(rb.in_layers[0], device)
        self.c1 = pack_conv3x3(rb.in_layers[2], device)
        self.n2 = pack_norm(rb.out_layers[0], device)
        self.c2 = pack_conv3x3(rb.out_layers[3], device)
        self.skip = None if isinstance(rb.skip_connection, nn.Identity) else pack_conv1x1(rb.skip_connection, device)
        self.emb_mod = rb.emb_layers[1]
        self.emb_col = 0


class PAttn:
    def __init__(self, attn, device):
        self.heads, self.d = attn.heads, attn.dim_head
        self.dp = ops.head_pad(self.d)
        self.hdp = self.heads * self.dp
        self.scale = self.d ** -0.5
        self.is_self = attn.context_dim is None
        kw = dict(heads=self.heads, d=self.d, dp=self.dp)
        self.q = pack_linear(attn.to_q, device, pad="rows", **kw)
        self.k = pack_linear(attn.to_k, device, pad="rows", **kw)
        self.v = pack_linear(attn.to_v, device, pad="rows", **kw)
        self.o = pack_linear(attn.to_out[0], device, pad="cols", **kw)
        self.module = attn  # for the optional to_control branch (CrossAttention_LoRA_Control)
        # d = 40 self-attention (N up to 4096, the most expensive kernel of the step): the zero padding column 40 of every V
        # head is set to 1.0 through the projection's bias, so the softmax row sum falls out of the P V UMMA (O[:, 40]) and
        # the softmax warps do not accumulate it (ops.attention l_col).  TOPO_ATTN_LCOL=0 disables (A/B runs).
        self.l_col = -1
        if self.is_self and self.d == 40 and self.dp == 48 and _lcol_enabled():
            bv = torch.zeros(self.hdp, dtype=torch.float32, device=device) if self.v.bias is None else self.v.bias.clone()
            bv[torch.arange(self.heads, device=device) * self.dp + self.d] = 1.0
            self.v.bias = bv
            self.l_col = self.d


class PBlock:
    def __init__(self, blk, device):
        self.n1, self.n2, self.n3 = (pack_norm(m, device) for m in (blk.norm1, blk.norm2, blk.norm3))
        self.attn1, self.attn2 = PAttn(blk.attn1, device), PAttn(blk.attn2, device)
        self.ff0 = pack_geglu(blk.ff.net[0].proj, device)
        self.ff2 = pack_linear(blk.ff.net[2], device)


class PST:
    def __init__(self, st, device):
        self.norm = pack_norm(st.norm, device)
        self.proj_in = pack_conv1x1(st.proj_in, device)
        self.blocks = [PBlock(b, device) for b in st.transformer_blocks]
        self.proj_out = pack_conv1x1(st.proj_out, device)


class PDown:
    def __init__(self, m, device):
        self.conv = pack_conv3x3(m.op, device)  # [Cout, 9C] tap-major, consumed after im2col (stride 2)


class PUp:
    def __init__(self, m, device):
        self.conv = pack_conv3x3(m.conv, device)


class PConvIn:
    def __init__(self, conv, device):
        self.w = ops.conv_direct_weight(conv.weight.detach().to(device))
        self.b = _dev(conv.bias, device, torch.float32)
        self.cin, self.cout, self.k, self.stride = conv.in_channels, conv.out_channels, conv.kernel_size[0], conv.stride[0]
        # narrow 3x3/s1 input convs (4-channel latents): im2col into ONE 64-wide K block + tensor-core GEMM
        self.gemm = None
        if self.k == 3 and self.stride == 1 and 9 * self.cin <= 64:
            w = torch.zeros(self.cout, 64, device=device)
            w[:, :9 * self.cin] = conv.weight.detach().to(device).float().permute(0, 2, 3, 1).reshape(self.cout, -1)
            self.gemm = Lin(w.to(BF16).contiguous(), self.b)


def _pack_layer(m, device):
    from . import ldm
    if isinstance(m, ldm.ResBlock):
        return PRes(m, device)
    if isinstance(m, ldm.SpatialTransformer):
        return PST(m, device)
    if isinstance(m, ldm.Downsample):
        return PDown(m, device)
    if isinstance(m, ldm.Upsample):
        return PUp(m, device)
    if isinstance(m, nn.Conv2d):
        return PConvIn(m, device)
    raise capi.TopoError(f"engine: unsupported layer {type(m).__name__}")


class PNet:
    """Packed UNet or ControlNet: time MLP, block lists, and the batched ResBlock time-embedding projection."""

    def __init__(self, net, device, control: bool):
        self.mc = net.model_channels
        self.t0 = pack_linear(net.time_embed[0], device)
        self.t2 = pack_linear(net.time_embed[2], device)
        self.input_blocks = [[_pack_layer(m, device) for m in blk] for blk in net.input_blocks]
        self.middle = [_pack_layer(m, device) for m in net.middle_block]
        self.output_blocks = [] if control else [[_pack_layer(m, device) for m in blk] for blk in net.output_blocks]
        res = [l for blk in self.input_blocks + [self.middle] + self.output_blocks for l in blk if isinstance(l, PRes)]
        Ws, bs, col = [], [], 0
        for r in res:
            r.emb_col = col
            Ws.append(r.emb_mod.weight.detach().to(device))
            bs.append(r.emb_mod.bias.detach().to(device))
            col += r.cout
        self.emb_all = Lin(torch.cat(Ws).to(BF16).contiguous(), torch.cat(bs).float().contiguous())
        if control:
            self.zero = [pack_conv1x1(z[0], device) for z in net.zero_convs]
            self.mid_out = pack_conv1x1(net.middle_block_out[0], device)
            hb = [m for m in net.input_hint_block if isinstance(m, nn.Conv2d)]
            self.hint = [PConvIn(c, device) for c in hb]
            # the last hint conv (256 -> model_channels at latent resolution, 41 % of the block's FLOPs) runs on the
            # tensor-core implicit-GEMM kernel when its input width allows (C % 64 == 0)
            last = hb[-1]
            self.hint_last = pack_conv3x3(last, device) if (last.in_channels % 64 == 0 and last.kernel_size[0] == 3
                                                            and last.stride[0] == 1) else None
        else:
            self.out_norm = pack_norm(net.out[0], device)
            self.out_conv = pack_conv3x3(net.out[2], device)

    def sts(self) -> List[PST]:
        return [l for blk in self.input_blocks + [self.middle] + self.output_blocks for l in blk if isinstance(l, PST)]


# ----------------------------------------------------------------------------------------------
# block forwards.  ``dst = (buffer, col)`` makes the block's LAST kernel write its result into a column slice
# of ``buffer`` (a decoder concat buffer); the function then returns ``buffer``.
# ----------------------------------------------------------------------------------------------
Dst = Optional[Tuple[torch.Tensor, int]]


def run_res(r: PRes, x: torch.Tensor, B: int, H: int, W: int, emb: torch.Tensor, dst: Dst = None) -> torch.Tensor:
    rb = emb[:, r.emb_col:r.emb_col + r.cout]
    h = gn(x, B, r.n1, True)
    h = conv3(h, B, H, W, r.c1, rowbias=rb)            # + Linear(SiLU(emb))[:, :, None, None] in the epilogue
    h = gn(h, B, r.n2, True)
    sk = x if r.skip is None else gemm(x, r.skip)
    out, col = dst if dst is not None else (None, 0)
    return conv3(h, B, H, W, r.c2, out=out, out_col=col, res=sk)  # + skip(x) in the epilogue


def run_attn(a: PAttn, n: torch.Tensor, h: torch.Tensor, B: int, kv) -> torch.Tensor:
    """h + to_out(softmax(q k^T d^-1/2) v); ``n`` = LayerNorm(h).  kv None → self-attention (fused q/k/v launch)."""
    M = n.shape[0]
    N = M // B
    hdp = a.hdp
    if kv is None:
        qkv = torch.empty(M, 3 * hdp, dtype=BF16, device=n.device)
        gemm_multi(n, [a.q, a.k, a.v], qkv, [0, hdp, 2 * hdp])
        t = qkv.view(B, N, 3 * hdp)
        q, k, v = t[:, :, :hdp], t[:, :, hdp:2 * hdp], t[:, :, 2 * hdp:]
    else:
        q = gemm(n, a.q).view(B, N, hdp)
        k, v = kv
    Nk = k.shape[1]
    o = _rec("attn_self" if kv is None else "attn_cross", 4.0 * B * a.heads * N * Nk * a.d,
             2.0 * (2 * M * hdp + 2 * B * Nk * hdp),
             lambda: ops.attention(q, k, v, a.heads, a.dp, a.scale, l_col=a.l_col if (kv is None and Nk > 80) else -1),
             f"B{B} N{N} Nk{Nk} d{a.d}")
    _count()
    out = gemm(o.view(M, hdp), a.o, res=h)
    if getattr(a.module, "control_states", None) is not None:  # CrossAttention_LoRA_Control (attention_lora.py:73-81)
        out = ops.control_branch(a.module, n.view(B, N, -1), 1.0, add_to=out.view(B, N, -1)).reshape(M, -1)
        _count(2)
    return out


def run_block(b: PBlock, h: torch.Tensor, B: int, kv) -> torch.Tensor:
    h = run_attn(b.attn1, ln(h, b.n1), h, B, None)
    h = run_attn(b.attn2, ln(h, b.n2), h, B, kv)
    g = gemm(ln(h, b.n3), b.ff0, act=ACT_GEGLU)
    return gemm(g, b.ff2, res=h)


def run_st(st: PST, x: torch.Tensor, B: int, kvs, dst: Dst = None) -> torch.Tensor:
    h = gemm(gn(x, B, st.norm, False), st.proj_in)
    for blk, kv in zip(st.blocks, kvs):
        h = run_block(blk, h, B, kv)
    out, col = dst if dst is not None else (None, 0)
    return gemm(h, st.proj_out, out=out, out_col=col, res=x)


def run_down(d: PDown, x: torch.Tensor, B: int, H: int, W: int, dst: Dst = None) -> torch.Tensor:
    cols = ops.im2col3x3_s2(x.view(B, H, W, -1))
    _count()
    out, col = dst if dst is not None else (None, 0)
    return gemm(cols, d.conv, out=out, out_col=col)


def run_up(u: PUp, x: torch.Tensor, B: int, H: int, W: int, dst: Dst = None) -> torch.Tensor:
    y = ops.upsample2x(x.view(B, H, W, -1))
    _count()
    out, col = dst if dst is not None else (None, 0)
    return conv3(y.view(B * 4 * H * W, -1), B, 2 * H, 2 * W, u.conv, out=out, out_col=col)


def run_convin(c: PConvIn, x4: torch.Tensor, act: int = ACT_NONE, add: Optional[torch.Tensor] = None,
               out_fp32: bool = False) -> torch.Tensor:
    _count()
    Bn, H, W, _ = x4.shape
    if c.gemm is not None and act == ACT_NONE:
        cols = ops.im2col3x3_s1_k64(x4)
        _count()
        res = None if add is None else add.reshape(Bn * H * W, -1)
        return gemm(cols, c.gemm, res=res, out_fp32=out_fp32).view(Bn, H, W, -1)
    Ho, Wo = H // c.stride, W // c.stride
    return _rec("conv_direct", 2.0 * Bn * Ho * Wo * c.cout * c.k * c.k * c.cin,
                2.0 * (x4.numel() + Bn * Ho * Wo * c.cout * (2 if out_fp32 else 1)),
                lambda: ops.conv_direct(x4, c.w, c.b, c.cin, c.cout, c.k, c.stride, act=act, add=add, out_fp32=out_fp32),
                f"B{Bn} H{H} C{c.cin} N{c.cout} k{c.k} s{c.stride}")


def run_layers(layers, h: torch.Tensor, B: int, H: int, W: int, emb, kv_iter, dst: Dst = None):
    for i, l in enumerate(layers):
        d = dst if i == len(layers) - 1 else None
        if isinstance(l, PRes):
            h = run_res(l, h, B, H, W, emb, d)
        elif isinstance(l, PST):
            h = run_st(l, h, B, next(kv_iter), d)
        elif isinstance(l, PDown):
            h = run_down(l, h, B, H, W, d)
            H, W = H // 2, W // 2
        elif isinstance(l, PUp):
            h = run_up(l, h, B, H, W, d)
            H, W = 2 * H, 2 * W
        else:
            raise capi.TopoError(f"engine: unexpected layer {type(l).__name__}")
    return h, H, W


def _two_streams() -> bool:
    import os
    return os.environ.get("TOPO_TWO_STREAMS", "1") != "0"


def _shareable(layers) -> bool:
    return (len(layers) == 2 and isinstance(layers[0], PRes) and isinstance(layers[1], PST) and len(layers[1].blocks) >= 1
            and getattr(layers[1].blocks[0].attn1.module, "control_states", None) is None)


def run_first_block_shared(layers, h_full: torch.Tensor, Bs: int, rep: int, H: int, W: int, emb: torch.Tensor, kv_iter):
    """``input_blocks[1]`` = [ResBlock, SpatialTransformer] when the batch is ``rep`` identical copies of ``Bs`` latents
    that differ only in their cross-attention context (the batched CFG call: ``x_in = cat([x] * 2)``, ``t_in = cat([t] *
    2)``, ldm ``p_sample_ddim``).  Nothing before the first cross-attention sees the context, so the ResBlock, GroupNorm,
    ``proj_in``, LayerNorm, q/k/v projection, the 4096-token SELF-attention and its out-projection run once on ``Bs``
    samples; their result is repeated along the batch and the rest of the block runs on all ``rep * Bs``.  Bit-identical
    to the duplicated computation (every kernel is deterministic and per-sample)."""
    res, st = layers
    M = Bs * H * W
    r = run_res(res, h_full[:M], Bs, H, W, emb[:Bs])
    kvs = next(kv_iter)
    h = gemm(gn(r, Bs, st.norm, False), st.proj_in)
    b0 = st.blocks[0]
    h = run_attn(b0.attn1, ln(h, b0.n1), h, Bs, None)
    Be = Bs * rep
    a2 = b0.attn2
    if getattr(a2.module, "control_states", None) is None:
        # The context enters through K/V only: LayerNorm and the q projection of the first CROSS-attention are shared
        # too, and the copies differ from here on.  Nothing is repeated in memory: the attention and its out-projection
        # run once per copy on the shared q / shared residual and write their slice of the full-batch tensors.
        N, hdp = H * W, a2.hdp
        q = gemm(ln(h, b0.n2), a2.q).view(Bs, N, hdp)
        k, v = kvs[0]
        Nk = k.shape[1]
        o = torch.empty(Be, N, hdp, dtype=BF16, device=h.device)
        hf = torch.empty(Be * N, h.shape[1], dtype=BF16, device=h.device)
        for i in range(rep):
            sl = slice(i * Bs, (i + 1) * Bs)
            _rec("attn_cross", 4.0 * Bs * a2.heads * N * Nk * a2.d, 2.0 * (2 * M * hdp + 2 * Bs * Nk * hdp),
                 lambda: ops.attention(q, k[sl], v[sl], a2.heads, a2.dp, a2.scale, out=o[sl]), f"B{Bs} N{N} Nk{Nk} d{a2.d}")
            _count()
            gemm(o[sl].view(M, hdp), a2.o, out=hf[i * M:(i + 1) * M], res=h)
        h = hf
    else:
        h = h.repeat(rep, 1)
        _count()
        h = run_attn(a2, ln(h, b0.n2), h, Be, kvs[0])
    g = gemm(ln(h, b0.n3), b0.ff0, act=ACT_GEGLU)
    h = gemm(g, b0.ff2, res=h)
    for blk, kv in zip(st.blocks[1:], kvs[1:]):
        h = run_block(blk, h, Be, kv)
    out = torch.empty(Be * H * W, st.proj_out.N, dtype=BF16, device=h.device)
    for i in range(rep):  # the block's residual r is shared as well: one GEMM per copy instead of r.repeat
        gemm(h[i * M:(i + 1) * M], st.proj_out, out=out[i * M:(i + 1) * M], res=r)
    return out


def _block_out_channels(layers) -> int:
    return [l.cout for l in layers if isinstance(l, PRes)][-1]


@dataclass
class Prepared:
    """Step-invariant work of one sampling run (hoisted out of the DDIM loop): cross-attention K/V of every
    transformer block for this context, and the ControlNet hint-block feature map for this hint."""
    ctx_key: Tuple
    unet_kv: list
    cnet_kv: list
    hint_feat: Optional[torch.Tensor]
    hint_rep: int = 1                     # the hint feature is this many copies of its first Be / hint_rep samples


def fingerprint(model: nn.Module) -> Tuple:
    """Changes whenever weights are written in place through autograd-visible ops (``load_state_dict``, ``p.copy_()``
    under ``no_grad`` bump ``_version``) or modules are swapped.  Writes through ``p.data`` are invisible to it: after
    such an edit call ``model.invalidate_engine()``."""
    nv = np_ = 0
    dev = None
    for p in model.parameters():
        nv += p._version
        np_ += 1
        dev = p.device
    return (np_, nv, sum(1 for _ in model.modules()), str(dev), getattr(model, "_topo_surgery_epoch", 0))


class TopoEngine:
    def __init__(self, model):
        dev = next(model.parameters()).device
        if dev.type != "cuda":
            raise capi.TopoError(
                "TopoEngine: the model is on %s; move it to a B200 (`model.to('cuda')`). There is no CPU forward in "
                "this package (the fp32 checker lives in oracle/ and is test infrastructure)." % dev)
        capi.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        self.device = dev
        self.fp = fingerprint(model)
        self.scale_factor = float(model.scale_factor)
        with torch.no_grad():
            self.unet = PNet(model.model.diffusion_model, dev, control=False)
            self.cnet = PNet(model.control_model, dev, control=True)
            self.vae = None if model.first_stage_model is None else PVae(model.first_stage_model, dev, self.scale_factor)

    def stale(self, model) -> bool:
        return fingerprint(model) != self.fp

    # ---- step-invariant work -----------------------------------------------------------------
    def _kv(self, net: PNet, ctx2: torch.Tensor, Be: int):
        out = []
        for st in net.sts():
            per = []
            for blk in st.blocks:
                a = blk.attn2
                kv = torch.empty(ctx2.shape[0], 2 * a.hdp, dtype=BF16, device=ctx2.device)
                gemm_multi(ctx2, [a.k, a.v], kv, [0, a.hdp])
                t = kv.view(Be, -1, 2 * a.hdp)
                per.append((t[:, :, :a.hdp], t[:, :, a.hdp:]))
            out.append(per)
        return out

    def _hint(self, hint: torch.Tensor) -> torch.Tensor:
        """input_hint_block: 8 convs with SiLU between → [Be, H/8, W/8, mc] bf16."""
        g = ops.nchw_to_nhwc(hint)
        _count()
        convs = self.cnet.hint
        for i, c in enumerate(convs[:-1]):
            g = run_convin(c, g, act=ACT_SILU)
        if self.cnet.hint_last is not None:
            Bh, Hh, Wh, Ch = g.shape
            if (Hh * Wh) % 128 == 0 or 128 % (Hh * Wh) == 0:
                return conv3(g.view(Bh * Hh * Wh, Ch), Bh, Hh, Wh, self.cnet.hint_last).view(Bh, Hh, Wh, -1)
        return run_convin(convs[-1], g, act=ACT_NONE)

    @torch.no_grad()
    def prepare(self, ctx: torch.Tensor, hint: Optional[torch.Tensor]) -> Prepared:
        """``hint`` may hold fewer samples than ``ctx`` (Be = k * Bh): the batched CFG call passes the SAME control image
        for the unconditional and the conditional half (vis_metrics.py:427-428), so the hint block runs once per tile
        and its feature map is repeated along the batch."""
        Be = ctx.shape[0]
        ctx2 = ctx.reshape(-1, ctx.shape[-1]).to(BF16).contiguous()
        key = (ctx.data_ptr(), tuple(ctx.shape), None if hint is None else (hint.data_ptr(), tuple(hint.shape)))
        feat, hint_rep = None, 1
        if hint is not None:
            feat = self._hint(hint)
            if hint.shape[0] != Be:
                if Be % hint.shape[0]:
                    raise capi.TopoError(f"prepare: hint batch {hint.shape[0]} does not divide the context batch {Be}")
                hint_rep = Be // hint.shape[0]
                feat = feat.repeat(hint_rep, 1, 1, 1)
        return Prepared(key, self._kv(self.unet, ctx2, Be), self._kv(self.cnet, ctx2, Be), feat, hint_rep)

    def _time(self, net: PNet, t: torch.Tensor) -> torch.Tensor:
        e = ops.timestep_embedding(t, net.mc)
        e = gemm(e, net.t0, act=ACT_SILU)
        e = gemm(e, net.t2, act=ACT_SILU)             # SiLU(emb): every ResBlock starts its emb_layers with SiLU
        _count()
        return gemm(e, net.emb_all, out_fp32=True)    # all ResBlock Linear(1280→Cout) in one GEMM, fp32 [Be, ΣCout]

    # ---- the hot path --------------------------------------------------------------------------
    @torch.no_grad()
    def apply_model(self, x: torch.Tensor, t: torch.Tensor, ctx: torch.Tensor, hint: Optional[torch.Tensor],
                    control_scales: Sequence[float], only_mid_control: bool = False,
                    prepared: Optional[Prepared] = None, x_rep: int = 1) -> torch.Tensor:
        """eps [Be,4,H,W] fp32 = UNet(x, t, ctx, control = scales * ControlNet(x, hint, t, ctx))  (SURVEY.md u1-u3).
        ``x_rep = k`` is the caller's promise that ``x`` and ``t`` are k identical copies of their first ``Be / k``
        entries (the batched CFG call): the context-free prefix of both networks then runs once per latent
        (``run_first_block_shared``)."""
        if not x.is_cuda:
            raise capi.TopoError("apply_model: x must be a CUDA tensor (no CPU fallback)")
        P = prepared if prepared is not None else self.prepare(ctx, hint)
        Be, _, H0, W0 = x.shape
        tf = t.to(device=x.device, dtype=torch.float32).contiguous()
        U, Cn = self.unet, self.cnet
        x4 = ops.nchw_to_nhwc(x)
        _count()

        # -- ControlNet trunk: independent of the UNet encoder until the zero-convs add its outputs to the UNet skips, so it
        #    is issued on a second stream (a fork/join inside the captured graph): the persistent kernels of the two
        #    networks fill each other's tails and the SM-starved 8x8 / 16x16 levels.  TOPO_TWO_STREAMS=0: one stream.
        use_control = P.hint_feat is not None

        def cnet_trunk():
            embc = self._time(Cn, tf)
            kvc = iter(P.cnet_kv)
            hc = run_convin(Cn.input_blocks[0][0], x4, add=P.hint_feat).view(Be * H0 * W0, -1)
            Hc, Wc = H0, W0
            outs = []
            for i, blk in enumerate(Cn.input_blocks):
                if i == 1 and x_rep > 1 and P.hint_rep == x_rep and _shareable(blk):   # same latents AND same hints
                    hc = run_first_block_shared(blk, hc, Be // x_rep, x_rep, Hc, Wc, embc, kvc)
                elif i:
                    hc, Hc, Wc = run_layers(blk, hc, Be, Hc, Wc, embc, kvc)
                outs.append(hc)
            hm, _, _ = run_layers(Cn.middle, hc, Be, Hc, Wc, embc, kvc)
            return outs, hm

        side, trunk = None, None
        if use_control:
            if _two_streams() and not PROF.on:
                side = self.__dict__.get("_side")
                if side is None:
                    side = self.__dict__["_side"] = torch.cuda.Stream(device=x.device)
                side.wait_stream(torch.cuda.current_stream(x.device))
                ops.GN_SLOT[0] = 1          # the trunk's GroupNorms use their own statistics workspace
                try:
                    with torch.cuda.stream(side):
                        trunk = cnet_trunk()
                finally:
                    ops.GN_SLOT[0] = 0
            else:
                trunk = cnet_trunk()

        # -- UNet encoder + middle
        emb = self._time(U, tf)
        kv = iter(P.unet_kv)
        h = run_convin(U.input_blocks[0][0], x4).view(Be * H0 * W0, -1)
        H, W = H0, W0
        hs, dims = [h], [(H, W)]
        if x_rep > 1 and Be % x_rep:
            raise capi.TopoError(f"apply_model: x_rep={x_rep} does not divide the batch {Be}")
        for bi, blk in enumerate(U.input_blocks[1:]):
            if bi == 0 and x_rep > 1 and _shareable(blk):
                h = run_first_block_shared(blk, h, Be // x_rep, x_rep, H, W, emb, kv)
            else:
                h, H, W = run_layers(blk, h, Be, H, W, emb, kv)
            hs.append(h)
            dims.append((H, W))
        h_mid, Hm, Wm = run_layers(U.middle, h, Be, H, W, emb, kv)

        # -- decoder concat buffers [h | skip]; the skip halves are filled by the ControlNet zero-conv epilogues
        nb = len(U.output_blocks)
        ch_in = [h_mid.shape[1]] + [_block_out_channels(b) for b in U.output_blocks[:-1]]
        cat = [torch.empty(hs[nb - 1 - j].shape[0], ch_in[j] + hs[nb - 1 - j].shape[1], dtype=BF16, device=x.device)
               for j in range(nb)]

        if use_control:
            if side is not None:
                torch.cuda.current_stream(x.device).wait_stream(side)
            hcs, hc_mid = trunk
            for i, hc in enumerate(hcs):
                j = nb - 1 - i
                if only_mid_control:
                    cat[j][:, ch_in[j]:].copy_(hs[i])
                else:  # (e) skip + scale*zero_conv(hc) straight into the skip half of concat buffer j
                    gemm(hc, Cn.zero[i], out=cat[j], out_col=ch_in[j], res=hs[i], alpha=float(control_scales[i]))
            gemm(hc_mid, Cn.mid_out, out=cat[0], out_col=0, res=h_mid, alpha=float(control_scales[len(Cn.zero)]))
        else:
            for j in range(nb):
                cat[j][:, ch_in[j]:].copy_(hs[nb - 1 - j])
            cat[0][:, :ch_in[0]].copy_(h_mid)

        # -- UNet decoder: block j reads cat[j] and writes the h half of cat[j+1]
        H, W = Hm, Wm
        h = cat[0]
        for j, blk in enumerate(U.output_blocks):
            dst = (cat[j + 1], 0) if j + 1 < nb else None
            h, H, W = run_layers(blk, h, Be, H, W, emb, kv, dst)
        h = gn(h, Be, U.out_norm, True)
        e = conv3(h, Be, H, W, U.out_conv, out_fp32=True)
        _count()
        return ops.nhwc_to_nchw(e.view(Be, H, W, -1))

    @torch.no_grad()
    def decode_first_stage(self, z: torch.Tensor) -> torch.Tensor:
        if self.vae is None:
            raise capi.TopoError("decode_first_stage: model was built with with_vae=False")
        return self.vae.decode(z)

    @torch.no_grad()
    def encode_moments(self, x: torch.Tensor) -> torch.Tensor:
        """Posterior moments of ``first_stage_model.encode`` (``vis_metrics.py:86``)."""
        if self.vae is None or not hasattr(self.vae, "e_conv_in"):
            raise capi.TopoError("encode_first_stage: model was built without the VAE encoder")
        return self.vae.encode_moments(x)

    def decode_to_u8(self, z: torch.Tensor) -> torch.Tensor:
        """decode_first_stage + uint8 HWC conversion on the device → [B, 8H, 8W, 3] uint8."""
        if self.vae is None:
            raise capi.TopoError("decode_to_u8: model was built with with_vae=False")
        return self.vae.decode(z, u8=True)


# ----------------------------------------------------------------------------------------------
# VAE decoder (AutoencoderKL.decode)
# ----------------------------------------------------------------------------------------------
class PVRes:
    def __init__(self, rb, device):
        self.n1, self.c1 = pack_norm(rb.norm1, device), pack_conv3x3(rb.conv1, device)
        self.n2, self.c2 = pack_norm(rb.norm2, device), pack_conv3x3(rb.conv2, device)
        self.nin = pack_conv1x1(rb.nin_shortcut, device) if hasattr(rb, "nin_shortcut") else None

    def run(self, x, B, H, W):
        h = conv3(gn(x, B, self.n1, True), B, H, W, self.c1)
        h = gn(h, B, self.n2, True)
        sk = x if self.nin is None else gemm(x, self.nin)
        return conv3(h, B, H, W, self.c2, res=sk)


class PVAttn:
    """Single-head attention over H*W tokens with d = C (512): scores are materialised per tile
    (S = q k^T as a GEMM, row softmax, P v as a GEMM against v^T); the v bias is added after P·v, which is
    exact because softmax rows sum to one."""

    def __init__(self, at, device):
        self.norm = pack_norm(at.norm, device)
        self.q, self.k = pack_conv1x1(at.q, device), pack_conv1x1(at.k, device)
        v = pack_conv1x1(at.v, device)
        self.vW, self.vb = v.W, v.bias
        self.proj = pack_conv1x1(at.proj_out, device)
        self.C = self.vW.shape[0]

    def run(self, x, B, H, W):
        N, C = H * W, self.C
        hn = gn(x, B, self.norm, False)
        q, k = gemm(hn, self.q), gemm(hn, self.k)
        o = torch.empty(B * N, C, dtype=BF16, device=x.device)
        for b in range(B):
            sl = slice(b * N, (b + 1) * N)
            vT = gemm(self.vW, Lin(hn[sl], None))                       # [C, N] = Wv hn^T  (v without bias, transposed)
            s = gemm(q[sl], Lin(k[sl], None))                           # [N, N]
            p = _rec("softmax_rows", 0.0, 4.0 * s.numel(), lambda: ops.softmax_rows(s, C ** -0.5), f"N{N}")
            _count()
            gemm(p, Lin(vT, self.vb), out=o[sl])
        return gemm(o, self.proj, res=x)


class PVae:
    def __init__(self, vae, device, scale_factor: float):
        dec = vae.decoder
        pq = vae.post_quant_conv
        # z / scale_factor folded into post_quant_conv: W (z/s) + b = (W/s) z + b
        w = pq.weight.detach().to(device).float() / scale_factor
        self.pq_w, self.pq_b = ops.conv_direct_weight(w), _dev(pq.bias, device, torch.float32)
        self.zc = pq.in_channels
        self.conv_in = PConvIn(dec.conv_in, device)
        self.mid1, self.attn, self.mid2 = PVRes(dec.mid.block_1, device), PVAttn(dec.mid.attn_1, device), PVRes(dec.mid.block_2, device)
        self.levels = []
        for lvl in reversed(list(dec.up)):
            blocks = [PVRes(b, device) for b in lvl.block]
            up = pack_conv3x3(lvl.upsample.conv, device) if hasattr(lvl, "upsample") else None
            self.levels.append((blocks, up))
        self.norm_out = pack_norm(dec.norm_out, device)
        self.conv_out = PConvIn(dec.conv_out, device)
        # conv_out (128 -> 3 at 512x512) as a tensor-core implicit GEMM with the 3 output channels zero-padded to the
        # smallest UMMA N (16): the direct kernel spent 31.8 ms of a 87 ms decode here (profiles/r01d_stage.txt)
        co = dec.conv_out
        self.conv_out_tc = None
        if co.in_channels % 64 == 0 and co.kernel_size[0] == 3 and co.out_channels <= 16:
            w = torch.zeros(16, co.in_channels, 3, 3, device=device)
            w[:co.out_channels] = co.weight.detach().to(device).float()
            b = torch.zeros(16, device=device)
            if co.bias is not None:
                b[:co.out_channels] = co.bias.detach().to(device).float()
            self.conv_out_tc = Lin(ops.conv3x3_weight(w), b)
            self.out_ch = co.out_channels
        if hasattr(vae, "encoder"):
            self._pack_encoder(vae, device)

    def _pack_encoder(self, vae, device):
        enc = vae.encoder
        self.e_conv_in = PConvIn(enc.conv_in, device)
        self.e_levels = []
        for lvl in enc.down:
            blocks = [PVRes(b, device) for b in lvl.block]
            down = pack_conv3x3(lvl.downsample.conv, device) if hasattr(lvl, "downsample") else None
            self.e_levels.append((blocks, down))
        self.e_mid1, self.e_attn, self.e_mid2 = (PVRes(enc.mid.block_1, device), PVAttn(enc.mid.attn_1, device),
                                                 PVRes(enc.mid.block_2, device))
        self.e_norm_out = pack_norm(enc.norm_out, device)
        # conv_out (512 -> 2z) then quant_conv (1x1, 2z -> 2z) composed into ONE 3x3 conv: Wq (Wc * x + bc) + bq,
        # output channels zero-padded to the smallest UMMA N (16)
        co, qc = enc.conv_out, vae.quant_conv
        Wq = qc.weight.detach().to(device).float().view(qc.out_channels, qc.in_channels)
        w = torch.einsum("om,mikl->oikl", Wq, co.weight.detach().to(device).float())
        b = Wq @ co.bias.detach().to(device).float() + qc.bias.detach().to(device).float()
        self.mom_ch = qc.out_channels
        if self.mom_ch > 16 or co.in_channels % 64 != 0:
            raise capi.TopoError("VAE encoder: conv_out wider than 16 moments or Cin not a multiple of 64 is not implemented")
        wp = torch.zeros(16, co.in_channels, 3, 3, device=device)
        wp[:self.mom_ch] = w
        bp = torch.zeros(16, device=device)
        bp[:self.mom_ch] = b
        self.e_conv_out = Lin(ops.conv3x3_weight(wp), bp)

    def encode_moments(self, x: torch.Tensor) -> torch.Tensor:
        """``quant_conv(encoder(x))`` -> fp32 NCHW ``[B, 2z, H/8, W/8]`` (mean | logvar), x ``[B,3,H,W]`` fp32."""
        B, _, H, W = x.shape
        if H % 8 or W % 8:
            raise capi.TopoError("VAE encode: H and W must be multiples of 8")
        h = run_convin(self.e_conv_in, ops.nchw_to_nhwc(x.float()))
        _count()
        h = h.view(B * H * W, -1)
        for blocks, down in self.e_levels:
            for rb in blocks:
                h = rb.run(h, B, H, W)
            if down is not None:   # pad (0,1,0,1) + stride-2 conv: im2col with no leading padding, then a GEMM
                cols = ops.im2col3x3_s2(h.view(B, H, W, -1), pad_lo=0)
                _count()
                H, W = H // 2, W // 2
                h = gemm(cols, down)
        h = self.e_mid1.run(h, B, H, W)
        h = self.e_attn.run(h, B, H, W)
        h = self.e_mid2.run(h, B, H, W)
        h = gn(h, B, self.e_norm_out, True)
        if W % 8 != 0:
            raise capi.TopoError("VAE encode: latent width must be a multiple of 8 for the tensor-core conv_out")
        m = conv3(h, B, H, W, self.e_conv_out, out_fp32=True).view(B, H, W, 16)[..., :self.mom_ch]
        return ops.nhwc_to_nchw(m)

    def decode(self, z: torch.Tensor, u8: bool = False) -> torch.Tensor:
        B, _, H, W = z.shape
        h = ops.nchw_to_nhwc(z.float())
        h = ops.conv_direct(h, self.pq_w, self.pq_b, self.zc, self.zc, 1, 1)
        h = run_convin(self.conv_in, h)
        _count(2)
        h = h.view(B * H * W, -1)
        h = self.mid1.run(h, B, H, W)
        h = self.attn.run(h, B, H, W)
        h = self.mid2.run(h, B, H, W)
        for blocks, up in self.levels:
            for rb in blocks:
                h = rb.run(h, B, H, W)
            if up is not None:
                y = _rec("upsample2x", 0.0, 2.0 * 5 * h.numel(), lambda: ops.upsample2x(h.view(B, H, W, -1)), f"B{B} H{H} C{h.shape[1]}")
                _count()
                H, W = 2 * H, 2 * W
                h = conv3(y.view(B * H * W, -1), B, H, W, up)
        h = gn(h, B, self.norm_out, True)
        if self.conv_out_tc is not None and W % 128 == 0:
            img = conv3(h, B, H, W, self.conv_out_tc, out_fp32=True).view(B, H, W, 16)[..., :self.out_ch]
        else:
            img = run_convin(self.conv_out, h.view(B, H, W, -1), out_fp32=True)
        _count()
        if u8:  # evaluateCN.py:101-103: *127.5+127.5, clip, truncate → uint8 HWC, fused before the gather
            return _rec("tile_to_u8", 0.0, 5.0 * img.numel(), lambda: ops.tile_to_u8(img))
        return ops.nhwc_to_nchw(img)


# ----------------------------------------------------------------------------------------------
# CUDA-graph replay (one apply_model = ~550 launches; captured once per shape, replayed per DDIM step)
# ----------------------------------------------------------------------------------------------
class GraphedModel:
    """Static-buffer CUDA graphs of ``prepare`` (per tile batch) and ``apply_model`` (per DDIM step).

    ``t`` lives in a device buffer read by the timestep-embedding kernel, so ONE step graph serves every timestep.
    The kernels are captured on torch's capture stream through the C ABI's ``stream`` argument; TMA descriptors are
    encoded on the host at capture time and baked into the graph as kernel parameters.  Control states injected into
    ``CrossAttention_LoRA_Control`` modules (``attention_lora.py:57-59``) are read through static copies that
    ``set_cond`` refreshes, so a later ``inject_control_states`` with same-shaped tensors is honoured by the replay."""

    def __init__(self, eng: TopoEngine, Be: int, lat_hw: Tuple[int, int], ctx_shape: Tuple[int, int],
                 hint_hw: Optional[Tuple[int, int]], control_scales: Sequence[float], only_mid_control: bool,
                 hint_batch: Optional[int] = None, x_rep: int = 1):
        dev = eng.device
        self.eng = eng
        self.x = torch.zeros(Be, 4, *lat_hw, device=dev)
        self.t = torch.zeros(Be, device=dev)
        self.ctx = torch.zeros(Be, *ctx_shape, device=dev)
        # hint_batch < Be: the CFG halves share their control images (see TopoEngine.prepare)
        self.hint = None if hint_hw is None else torch.zeros(hint_batch or Be, 3, *hint_hw, device=dev)
        scales = [float(s) for s in control_scales]
        self.ctrl = [(m, m.control_states.detach().to(dev).clone()) for m in eng.control_modules()
                     if m.control_states is not None]
        live = [m.control_states for m, _ in self.ctrl]
        for m, buf in self.ctrl:
            m.control_states = buf
        try:
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):  # warm-up outside capture: function attributes, workspaces, lazy inits
                P = eng.prepare(self.ctx, self.hint)
                eng.apply_model(self.x, self.t, self.ctx, self.hint, scales, only_mid_control, prepared=P, x_rep=x_rep)
            torch.cuda.current_stream(dev).wait_stream(side)
            torch.cuda.synchronize(dev)
            n0 = LAUNCHES[0]
            self.g_prep = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_prep):
                self.P = eng.prepare(self.ctx, self.hint)
            self.prep_launches = LAUNCHES[0] - n0
            n0 = LAUNCHES[0]
            self.g_step = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_step, pool=self.g_prep.pool()):
                self.eps = eng.apply_model(self.x, self.t, self.ctx, self.hint, scales, only_mid_control, prepared=self.P,
                                           x_rep=x_rep)
            self.step_launches = LAUNCHES[0] - n0
        finally:
            for (m, _), cs in zip(self.ctrl, live):
                m.control_states = cs

    def set_cond(self, ctx: torch.Tensor, hint: Optional[torch.Tensor]) -> None:
        self.ctx.copy_(ctx, non_blocking=True)
        if self.hint is not None:
            self.hint.copy_(hint, non_blocking=True)
        for m, buf in self.ctrl:
            cs = m.control_states
            if cs is None or tuple(cs.shape) != tuple(buf.shape):
                raise capi.TopoError("GraphedModel.set_cond: control_states changed shape or were removed since this graph "
                                     "was captured; fetch the graph again through TopoEngine.graphed()")
            buf.copy_(cs, non_blocking=True)
        self.g_prep.replay()
        _count(self.prep_launches)

    def step(self) -> torch.Tensor:
        """Replay apply_model on the current contents of ``self.x`` / ``self.t``; returns the static eps buffer."""
        self.g_step.replay()
        _count(self.step_launches)
        return self.eps


def _control_modules(self: TopoEngine):
    """Attention modules that can carry injected control states (``CrossAttention_LoRA_Control``), in execution order."""
    out = []
    for net in (self.unet, self.cnet):
        for st in net.sts():
            for blk in st.blocks:
                for a in (blk.attn1, blk.attn2):
                    if hasattr(a.module, "control_states"):
                        out.append(a.module)
    return out


def _graphed(self: TopoEngine, Be, lat_hw, ctx_shape, hint_hw, control_scales, only_mid_control,
             hint_batch=None, x_rep=1, branch: int = 0) -> GraphedModel:
    """Cached step graph.  ``branch`` separates graphs of identical shape that must own distinct static buffers (the two
    apply_model calls of ``cldm/ddim_hacked``: 0 = unconditional, 1 = conditional); the key also carries which attention
    modules hold control states and their shapes, so injecting / removing control states selects another capture."""
    ctrl_sig = tuple((i, tuple(m.control_states.shape)) for i, m in enumerate(self.control_modules())
                     if m.control_states is not None)
    key = (Be, hint_batch or Be, tuple(lat_hw), tuple(ctx_shape), None if hint_hw is None else tuple(hint_hw),
           tuple(float(s) for s in control_scales), bool(only_mid_control), int(x_rep), int(branch), ctrl_sig)
    cache = self.__dict__.setdefault("_graphs", {})
    if key not in cache:
        cache[key] = GraphedModel(self, Be, lat_hw, ctx_shape, hint_hw, control_scales, only_mid_control, hint_batch,
                                  x_rep)
    return cache[key]


TopoEngine.control_modules = _control_modules
TopoEngine.graphed = _graphed
