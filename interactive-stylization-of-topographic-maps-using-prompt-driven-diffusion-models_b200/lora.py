"""The ``LoRA_modified`` model-surgery API, kept name-for-name, on top of the fused sm_100a projection kernel.

Reference surface reproduced here (paths relative to /root/reference):

* ``LoRALinear``, ``lora_qkv``, ``ensure_lora_to_device``, ``disable_bert_dropout`` — ``LoRA_modified/LoRA_utils.py:6-69``
* ``lora_attn_qkvo`` — ``TextEncoder_Finetuning/Training_sdfusion_bert_lora_qkvo.py:149-187`` (scaling ``(alpha or 4r)/r``)
* ``CrossAttention_LoRA_Control``, ``ControlLoRA``, ``inject_control_lora_into_ldm``, ``freeze_all_but_lora``,
  ``get_lora_parameters`` — ``LoRA_modified/attention_lora.py:13-157``

State-dict keys are the reference's (``*.to_q.base.weight``, ``*.to_q.A.weight``, ``*.to_q.B.weight``), so the
``".A."/".B."`` checkpoint filter of ``multiseg_19prompts_batch1_shuffle_addlrSD.py:539-543`` keeps working.

Forward passes run on the CUDA kernels only (``y = x W^T + s (x A^T) B^T`` in one tcgen05 GEMM, the low-rank delta
accumulated in TMEM); on a CPU tensor they raise — the fp32 checker lives in ``oracle/``.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional

import torch
import torch.nn as nn

from .ldm import CrossAttention


def disable_bert_dropout(hf_bert) -> None:
    """Zero every dropout probability of a HuggingFace BERT (config fields and ``nn.Dropout`` instances)."""
    cfg = getattr(hf_bert, "config", None)
    if cfg is not None:
        cfg.hidden_dropout_prob = 0.0
        cfg.attention_probs_dropout_prob = 0.0
    for mod in hf_bert.modules():
        if isinstance(mod, nn.Dropout):
            mod.p = 0.0


class LoRALinear(nn.Module):
    """Frozen ``base`` Linear plus a rank-``r`` update: ``base(x) + B(A(x)) * scaling``.

    ``scaling = (alpha or default_alpha_mult * r) / r`` — 2.0 with the ``LoRA_utils`` default, 4.0 in the qkvo script.
    ``A`` is kaiming-uniform(a=sqrt 5), ``B`` is zero, both created on ``base``'s device/dtype, so a freshly wrapped
    model computes exactly what the base model does.
    """

    def __init__(self, base: nn.Linear, r: int = 16, alpha: Optional[float] = None, default_alpha_mult: int = 2):
        super().__init__()
        self.base = base
        for p in base.parameters():
            p.requires_grad = False
        w = base.weight
        self.A = nn.Linear(base.in_features, r, bias=False).to(device=w.device, dtype=w.dtype)
        self.B = nn.Linear(r, base.out_features, bias=False).to(device=w.device, dtype=w.dtype)
        nn.init.kaiming_uniform_(self.A.weight, a=math.sqrt(5))
        nn.init.zeros_(self.B.weight)
        self.scaling = (alpha or (default_alpha_mult * r)) / r

    # nn.Linear look-alike attributes some callers read
    @property
    def in_features(self) -> int:
        return self.base.in_features

    @property
    def out_features(self) -> int:
        return self.base.out_features

    def forward(self, x):
        from .ops import lora_linear_module
        return lora_linear_module(self, x)


def _wrap(parent: nn.Module, attr: str, r: int, counts: Dict[str, int], key: str, mult: int = 2) -> None:
    cur = getattr(parent, attr, None)
    if isinstance(cur, nn.Linear):
        setattr(parent, attr, LoRALinear(cur, r=r, default_alpha_mult=mult))
        counts[key] += 1


def lora_qkv(unet: nn.Module, r_q: int = 4, r_kv: int = 12) -> Dict[str, int]:
    """Wrap every ``to_q`` (rank ``r_q``) and ``to_k``/``to_v`` (rank ``r_kv``) that is an ``nn.Linear`` — both the
    self- and the cross-attention of every transformer block — in place; returns the wrap counts."""
    counts = {"q": 0, "k": 0, "v": 0}
    for mod in list(unet.modules()):
        _wrap(mod, "to_q", r_q, counts, "q")
        _wrap(mod, "to_k", r_kv, counts, "k")
        _wrap(mod, "to_v", r_kv, counts, "v")
    _invalidate(unet)
    return counts


def lora_attn_qkvo(unet: nn.Module, r: int = 8) -> int:
    """q, k, v and ``to_out[0]`` of every attention, scaling 4.0 (``Training_sdfusion_bert_lora_qkvo.py:167,175-187``)."""
    counts = {"n": 0}
    for mod in list(unet.modules()):
        if not all(hasattr(mod, a) for a in ("to_q", "to_k", "to_v")):
            continue
        for a in ("to_q", "to_k", "to_v"):
            _wrap(mod, a, r, counts, "n", mult=4)
        out = getattr(mod, "to_out", None)
        if isinstance(out, (nn.Sequential, nn.ModuleList)) and isinstance(out[0], nn.Linear):
            out[0] = LoRALinear(out[0], r=r, default_alpha_mult=4)
            counts["n"] += 1
    _invalidate(unet)
    return counts["n"]


def ensure_lora_to_device(unet: nn.Module, device=None) -> None:
    """Move the A/B of wrapped ``to_k``/``to_v`` next to their base weights.  (The reference ignores ``to_q`` and the
    ``device`` argument, ``LoRA_utils.py:62-69``; here ``to_q`` and ``to_out[0]`` are handled too — a strict superset.)"""
    for mod in unet.modules():
        cands = [getattr(mod, a, None) for a in ("to_q", "to_k", "to_v")]
        out = getattr(mod, "to_out", None)
        if isinstance(out, (nn.Sequential, nn.ModuleList)) and len(out):
            cands.append(out[0])
        for c in cands:
            if isinstance(c, LoRALinear):
                w = c.base.weight
                c.A.to(device=w.device, dtype=w.dtype)
                c.B.to(device=w.device, dtype=w.dtype)


def _invalidate(root: nn.Module) -> None:
    """Surgery changes the execution plan: drop any cached engine hanging off the owning ControlLDM."""
    for m in root.modules():
        if hasattr(m, "_engine"):
            m._engine = None
    setattr(root, "_topo_surgery_epoch", getattr(root, "_topo_surgery_epoch", 0) + 1)


# ------------------------------------------------------------------------------------------------
# "ControlLoRA for ldm" (attention_lora.py)
# ------------------------------------------------------------------------------------------------
class CrossAttention_LoRA_Control(CrossAttention):
    """CrossAttention whose q/k/v/out are ``LoRALinear`` plus a low-rank ``to_control`` branch that adds a projected
    control feature map to the attention output (``attention_lora.py:35-81``)."""

    def __init__(self, *args, lora_rank: int = 4, control_rank: Optional[int] = None, **kwargs):
        super().__init__(*args, **kwargs)
        self.to_q = LoRALinear(self.to_q, r=lora_rank)
        self.to_k = LoRALinear(self.to_k, r=lora_rank)
        self.to_v = LoRALinear(self.to_v, r=lora_rank)
        # upstream keeps to_out as [Linear, Dropout]; the reference wraps element 0 when it is indexable
        self.to_out[0] = LoRALinear(self.to_out[0], r=lora_rank)
        control_rank = control_rank or lora_rank
        hidden = self.to_q.base.out_features
        self.to_control = nn.Sequential(nn.Linear(hidden, control_rank, bias=False),
                                        nn.Linear(control_rank, hidden, bias=False))
        nn.init.zeros_(self.to_control[1].weight)
        self.control_states = None

    def inject_control_states(self, control_feat) -> None:
        """Stores an alias of the caller's tensor (stateful, as in the reference :57-59)."""
        self.control_states = control_feat

    def process_control_states(self, hidden_states, scale: float = 1.0):
        if self.control_states is None:
            return 0
        from .ops import control_branch
        return control_branch(self, hidden_states, scale)

    def forward(self, x, context=None, mask=None, scale: float = 1.0):
        out = super().forward(x, context=context, mask=mask)
        if self.control_states is not None:
            from .ops import control_branch
            out = control_branch(self, x, scale, add_to=out)
        return out


class ControlLoRA(nn.Module):
    """Three-conv hint encoder producing control states (``attention_lora.py:87-101``): 3→64 (s1) →128 (s2) →256 (s2),
    SiLU after each.  Runs on the direct-convolution kernel; input NCHW float on CUDA, output NCHW float32."""

    def __init__(self, in_channels: int = 3, base_channels: int = 64):
        super().__init__()
        c = base_channels
        self.encoder = nn.Sequential(
            nn.Conv2d(in_channels, c, 3, 1, 1), nn.SiLU(),
            nn.Conv2d(c, c * 2, 3, 2, 1), nn.SiLU(),
            nn.Conv2d(c * 2, c * 4, 3, 2, 1), nn.SiLU())

    def forward(self, x):
        from . import ops
        h = ops.nchw_to_nhwc(x)
        for i in (0, 2, 4):
            conv = self.encoder[i]
            h = ops.conv_direct(h, ops.conv_direct_weight(conv.weight), conv.bias.float().contiguous(),
                                conv.in_channels, conv.out_channels, 3, conv.stride[0], act=ops.ACT_SILU)
        return ops.nhwc_to_nchw(h)


def inject_control_lora_into_ldm(unet: nn.Module, rank: int = 4, control_rank: Optional[int] = None) -> int:
    """Replace every plain ``CrossAttention`` under ``unet`` by ``CrossAttention_LoRA_Control`` carrying the same base
    weights; returns the number of replacements (``attention_lora.py:107-137``)."""
    targets = [(n, m) for n, m in unet.named_modules()
               if isinstance(m, CrossAttention) and not isinstance(m, CrossAttention_LoRA_Control)]
    for name, old in targets:
        new = CrossAttention_LoRA_Control(query_dim=old.query_dim, context_dim=old.context_dim, heads=old.heads,
                                          dim_head=old.dim_head, dropout=old.dropout,
                                          bias=old.to_q.bias is not None, lora_rank=rank, control_rank=control_rank)
        new.to(device=old.to_q.weight.device, dtype=old.to_q.weight.dtype)
        for a in ("to_q", "to_k", "to_v"):
            getattr(new, a).base.load_state_dict(getattr(old, a).state_dict())
        new.to_out[0].base.load_state_dict(old.to_out[0].state_dict())
        parent = unet
        *path, leaf = name.split(".")
        for p in path:
            parent = getattr(parent, p)
        setattr(parent, leaf, new)
    _invalidate(unet)
    print(f"[inject_control_lora_into_ldm] Replaced {len(targets)} CrossAttention layers with LoRA-Control versions.")
    return len(targets)


def freeze_all_but_lora(model: nn.Module) -> int:
    for p in model.parameters():
        p.requires_grad = False
    for m in model.modules():
        if isinstance(m, LoRALinear) or hasattr(m, "to_control"):
            for p in m.parameters():
                p.requires_grad = True
    n_train = sum(p.numel() for p in model.parameters() if p.requires_grad)
    print(f"[freeze_all_but_lora] Trainable params: {n_train:,}")
    return n_train


def get_lora_parameters(model: nn.Module) -> List[nn.Parameter]:
    return [p for p in model.parameters() if p.requires_grad]
