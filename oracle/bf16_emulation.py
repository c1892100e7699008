"""ORACLE (test infrastructure) — bf16 rounding emulation inside the fp32 oracle.

``with rounding(a=.., r=.., o=..):`` makes ``oracle.ldm_oracle`` round to bf16 (and back to fp32) at one class of
places at a time, so that a test or tool can ask "how far from the fp32 reference does an IDEAL bf16 pipeline land?"
and compare the CUDA kernels with that instead of with an arbitrary tolerance:

  a  operands: the input of every linear / conv and q, k, v, P of every attention (what a bf16 tensor-core GEMM reads)
  r  residual stream: every residual SUM (ResBlock output, the three adds of a transformer block, SpatialTransformer
     output)
  o  every linear / conv / norm OUTPUT (what a bf16 activation store does)
Weights are rounded by passing ``round_weights(sd)``.  Used by tests/test_gpu_fullsize.py and tools/parity_ablation.py;
never imported by the product package.
"""
import contextlib

import torch
import torch.nn.functional as F

from . import ldm_oracle as O


def rb(x):
    return x.to(torch.bfloat16).to(torch.float32)


def round_weights(sd):
    return {k: (rb(v) if v.ndim >= 2 else v) for k, v in sd.items()}


class _Pol:
    a = r = o = False


_orig = {}


def _lora_linear(sd, p, x, scales=None):
    y = _orig["lora_linear"](sd, p, rb(x) if _Pol.a else x, scales)
    return rb(y) if _Pol.o else y


def _conv(sd, p, x, stride=1, padding=0):
    y = _orig["conv"](sd, p, rb(x) if _Pol.a else x, stride, padding)
    return rb(y) if _Pol.o else y


def _group_norm(sd, p, x, eps, groups=None):
    y = _orig["group_norm"](sd, p, x, eps, groups)
    return rb(y) if _Pol.o else y


def _attention_core(q, k, v, heads):
    if not _Pol.a:
        return _orig["attention_core"](q, k, v, heads)
    b, n, c = q.shape
    d = c // heads
    q, k, v = (rb(t).reshape(b, -1, heads, d).permute(0, 2, 1, 3).reshape(b * heads, -1, d) for t in (q, k, v))
    sim = torch.einsum("bid,bjd->bij", q, k) * (d ** -0.5)
    sim = sim - sim.amax(-1, keepdim=True)
    e = sim.exp()
    out = torch.einsum("bij,bjd->bid", rb(e), v) / e.sum(-1, keepdim=True)   # P stored as bf16, normalised by the fp32 sum
    return out.reshape(b, heads, n, d).permute(0, 2, 1, 3).reshape(b, n, c)


def _rr(x):
    return rb(x) if _Pol.r else x


def _basic_transformer_block(sd, p, x, ctx, heads, scales=None, ctrl=None):
    c = x.shape[-1]
    ctrl = ctrl or {}

    def ln(n, t):
        y = F.layer_norm(t, (c,), sd[f"{p}.{n}.weight"], sd[f"{p}.{n}.bias"], 1e-5)
        return rb(y) if _Pol.o else y
    x = _rr(O.lora_control_attention(sd, p + ".attn1", ln("norm1", x), None, heads, ctrl.get(p + ".attn1"), 1.0, scales) + x)
    x = _rr(O.lora_control_attention(sd, p + ".attn2", ln("norm2", x), ctx, heads, ctrl.get(p + ".attn2"), 1.0, scales) + x)
    h = O.lora_linear(sd, p + ".ff.net.0.proj", ln("norm3", x), scales)
    a, g = h.chunk(2, dim=-1)
    return _rr(O.lora_linear(sd, p + ".ff.net.2", a * F.gelu(g), scales) + x)


def _spatial_transformer(sd, p, x, ctx, heads, scales=None, ctrl=None):
    b, c, h, w = x.shape
    x_in = x
    x = O.group_norm(sd, p + ".norm", x, 1e-6)
    x = O.conv(sd, p + ".proj_in", x)
    x = x.permute(0, 2, 3, 1).reshape(b, h * w, -1)
    i = 0
    while f"{p}.transformer_blocks.{i}.norm1.weight" in sd:
        x = O.basic_transformer_block(sd, f"{p}.transformer_blocks.{i}", x, ctx, heads, scales, ctrl)
        i += 1
    x = x.reshape(b, h, w, -1).permute(0, 3, 1, 2)
    return _rr(O.conv(sd, p + ".proj_out", x) + x_in)


def _res_block(sd, p, x, emb):
    return _rr(_orig["res_block"](sd, p, x, emb))


_PATCH = dict(lora_linear=_lora_linear, conv=_conv, group_norm=_group_norm, attention_core=_attention_core,
              basic_transformer_block=_basic_transformer_block, spatial_transformer=_spatial_transformer,
              res_block=_res_block)


@contextlib.contextmanager
def rounding(a=False, r=False, o=False):
    """Patch the oracle's primitives for the duration of the block (not re-entrant, not thread-safe: test code)."""
    assert not _orig, "rounding() is not re-entrant"
    for k in _PATCH:
        _orig[k] = getattr(O, k)
    _Pol.a, _Pol.r, _Pol.o = bool(a), bool(r), bool(o)
    try:
        for k, f in _PATCH.items():
            setattr(O, k, f)
        yield
    finally:
        for k, f in _orig.items():
            setattr(O, k, f)
        _orig.clear()
        _Pol.a = _Pol.r = _Pol.o = False
