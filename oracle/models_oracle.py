"""TEST INFRASTRUCTURE — fp32 PyTorch restatement of ``LoRA_modified/models.py`` (processor algebra + ControlLoRA
pyramid), the checker for ``topo_b200/models.py``.  Never imported by the product package.

PARITY PINNED on the reference's own processor code (round 2): ``tests/golden/make_golden_models.py`` runs
``/root/reference/LoRA_modified/models.py`` itself — nine processor configurations and the ControlLoRA pyramid — and
``tests/test_oracle_golden.py::test_models_oracle_matches_reference_models_py`` holds this file to those outputs at
2e-5.  The reference module imports ``diffusers`` (0.12-0.14 API, ``models.py:7-12``), which is not installed here, so the
few diffusers primitives it touches (``LoRALinearLayer``, the ``CrossAttention`` host module, ``Downsample2D``, the mixins)
are STAND-INS restating diffusers 0.12 in the generator; everything the reference wrote on top of them (which tensor
feeds which LoRA, skip flags, control-state entry points, pre/post stacks, concat / self-add / batch repeat, the pyramid
wiring) is executed as is.  The functions below restate the reference line by line in functional form over the SAME
parameter containers the CUDA path uses (anything with ``to_q/to_k/to_v/to_out`` Linears and ``to_*_lora.{down,up}``
pairs); citations are ``LoRA_modified/models.py:<line>``.  Attention is the explicit ``softmax(q k^T * scale) v`` of
``get_attention_scores`` + ``torch.bmm`` (:143-144).
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn.functional as F


def _lora(layer, x):                      # diffusers LoRALinearLayer.forward: up(down(x)), no scaling inside
    return F.linear(F.linear(x, layer.down.weight.float()), layer.up.weight.float())


def _lin(m, x):
    return F.linear(x, m.weight.float(), None if m.bias is None else m.bias.float())


def _heads(t, h):                         # head_to_batch_dim
    b, n, c = t.shape
    return t.view(b, n, h, c // h).transpose(1, 2).reshape(b * h, n, c // h)


def _unheads(t, h):                       # batch_to_head_dim
    bh, n, d = t.shape
    return t.view(bh // h, h, n, d).transpose(1, 2).reshape(bh // h, n, h * d)


def attention_core(attn, q, k, v):
    h = attn.heads
    qh, kh, vh = _heads(q, h), _heads(k, h), _heads(v, h)
    p = torch.softmax(torch.bmm(qh, kh.transpose(1, 2)) * attn.scale, dim=-1)   # :143
    return _unheads(torch.bmm(p, vh), h)                                          # :144-145


def process_control_states(proc, hidden, scale=1.0, is_out=False):
    """:201-220 (v1) / :338-357 (V2).  Caches the flattened map back on ``proc`` as the reference does."""
    cs = proc.control_states.to(hidden.dtype)
    if hidden.ndim == 3 and cs.ndim == 4:                                         # :203-206
        b, _, hh, ww = cs.shape
        cs = cs.permute(0, 2, 3, 1).reshape(b, hh * ww, -1)
        proc.control_states = cs
    inp = cs
    if proc.concat_hidden:                                                        # :208-214
        b1, b2 = cs.shape[0], hidden.shape[0]
        if b1 != b2:
            cs = cs[:, None].repeat(1, b2 // b1, *([1] * (cs.ndim - 1))).view(-1, *cs.shape[1:])
        inp = torch.cat([hidden, cs], -1)
    layer = proc.to_control_out if is_out else proc.to_control
    out = scale * _lora(layer, inp)                                               # :215
    return cs + out if proc.control_self_add else out                             # :216-219


def _chain(proc) -> List:
    return list(getattr(proc, "pre_loras", [])) + [proc] + list(getattr(proc, "post_loras", []))


def _is_v1(p):
    return hasattr(p, "to_control") and not hasattr(p, "to_control_out")


def _is_v2(p):
    return hasattr(p, "to_control_out")


def processor_call(attn, proc, hidden, enc=None, scale=1.0):
    """One attention layer under ``proc`` (LoRACrossAttnProcessor :118-152, ControlLoRACrossAttnProcessor :222-287,
    ControlLoRACrossAttnProcessorV2 :357-431).  fp32 in, fp32 out."""
    hidden = hidden.float()
    v1, v2 = _is_v1(proc), _is_v2(proc)
    chain = _chain(proc)
    if v2:                                                                        # :366-372
        for p in chain:
            if _is_v2(p):
                hidden = hidden + process_control_states(p, hidden, scale)
    q = _lin(attn.to_q, hidden)
    for p in chain:                                                               # :232-243 / :374-380 / :125-126
        lin = q if p.post_add else hidden
        if v1 and _is_v1(p):
            lin = lin + process_control_states(p, hidden, scale)
        q = q + scale * _lora(p.to_q_lora, lin)
    e = hidden if enc is None else enc.float()
    k = _lin(attn.to_k, e)
    for p in chain:                                                               # :249-256
        if not p.key_states_skipped:
            k = k + scale * _lora(p.to_k_lora, k if p.post_add else e)
    v = _lin(attn.to_v, e)
    for p in chain:                                                               # :257-265: stacks add WITHOUT scale
        if not p.value_states_skipped:
            v = v + (scale if p is proc else 1.0) * _lora(p.to_v_lora, v if p.post_add else e)
    o = attention_core(attn, q, k, v)
    if v2:                                                                        # :412-418
        for p in chain:
            if _is_v2(p):
                o = o + process_control_states(p, o, scale, is_out=True)
    out = _lin(attn.to_out[0], o)
    own_always = v1 or v2                                                         # :279 / :422 (no skip test on self)
    for p in chain:
        if (p is proc and own_always) or not p.output_states_skipped:
            out = out + scale * _lora(p.to_out_lora, out if p.post_add else o)
    return out                                                                    # to_out[1] = Dropout (eval: identity)


# ----------------------------------------------------------------------------------------------------------------
# ControlLoRA pyramid (:434-833)
# ----------------------------------------------------------------------------------------------------------------
def _gn(n, x):
    return F.group_norm(x, n.num_groups, n.weight.float(), n.bias.float(), n.eps)


def _conv(c, x):
    return F.conv2d(x, c.weight.float(), None if c.bias is None else c.bias.float(), c.stride, c.padding)


def conv_block(blk, x):                   # ConvBlock2D.forward with temb=None (:513-547)
    h = _conv(blk.conv1, F.silu(_gn(blk.norm1, x)))
    return F.silu(_gn(blk.norm2, h))


def down_encoder_block(blk, x):           # SimpleDownEncoderBlock2D.forward (:601-610)
    for c in blk.convnets:
        x = conv_block(c, x)
    if blk.downsamplers is not None:
        for d in blk.downsamplers:        # diffusers Downsample2D(use_conv, padding=0): pad (0,1,0,1) then stride-2 conv
            if d.padding == 0:
                x = F.pad(x, (0, 1, 0, 1))
            x = F.conv2d(x, d.conv.weight.float(), d.conv.bias.float(), 2, 0 if d.padding == 0 else d.padding)
    return x


def _block(b, x):
    if isinstance(b, torch.nn.Identity):
        return x
    if isinstance(b, torch.nn.Sequential):
        for m in b:
            x = down_encoder_block(m, x)
        return x
    return down_encoder_block(b, x)


def control_lora_forward(model, x, inject: bool = False):
    """ControlLoRA.forward(x, return_dict=False) (:800-833): tuple of control maps [B, C_i, H_i, W_i] fp32."""
    h = _conv(model.conv_in, x.float())
    outs = []
    for down, pre, layers in zip(model.down_blocks, model.pre_lora_layers, model.lora_layers):
        h = _block(down, h)
        cs = _block(pre, h)
        if inject:
            for l in layers:
                l.control_states = cs
        outs.append(cs)
    return tuple(outs)
