"""Host-side mirror of ``LoRA_modified/models.py`` (the diffusers-processor form of the LoRA attention and the
``ControlLoRA`` conv pyramid) on the sm_100a kernels.  Names, constructor arguments, state-dict keys and the
stateful ``inject_control_states`` aliasing follow the reference (paths relative to /root/reference):

* ``LoRACrossAttnProcessor``            — ``LoRA_modified/models.py:72-152``
* ``ControlLoRACrossAttnProcessor``     — ``:155-287``  (``process_control_states`` :201-220)
* ``ControlLoRACrossAttnProcessorV2``   — ``:292-431``
* ``ConvBlock2D`` / ``SimpleDownEncoderBlock2D`` / ``get_down_block`` / ``ControlLoRA`` — ``:15-69, 434-833``
* ``CrossAttention`` (the diffusers 0.12-0.14 host class the processors are written against; diffusers itself is not a
  dependency here) with ``to_q/to_k/to_v/to_out``, ``head_to_batch_dim``, ``batch_to_head_dim``,
  ``get_attention_scores``, ``prepare_attention_mask``, ``set_processor``.

Arithmetic: every projection ``W x + sum_i s_i up_i(down_i(in_i))`` is ONE tcgen05 GEMM when the LoRA inputs are the
projection's own input (all ranks concatenated into the fused low-rank side product), plus a pair of skinny GEMMs for
each LoRA fed by something else (``post_add`` stacks, the control-state sums); ``softmax(q k^T) v`` is the flash
kernel, so the ``[B*h, N, N]`` probabilities the reference materialises (1 GiB fp32 at N = 4096) never exist.
q/k/v live in the head-padded layout (40 -> 48) end to end: the pads are zero rows/columns of the packed weights.
There is no CPU path: CPU tensors raise ``TopoError`` (the fp32 checker is ``oracle/models_oracle.py``).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple, Union

import torch
import torch.nn as nn

from . import capi, ops
from .ldm import LoRALinearLayer
from .ops import BF16

__all__ = ["LoRALinearLayer", "CrossAttention", "LoRACrossAttnProcessor", "ControlLoRACrossAttnProcessor",
           "ControlLoRACrossAttnProcessorV2", "Downsample2D", "ConvBlock2D", "SimpleDownEncoderBlock2D",
           "get_down_block", "ControlLoRAOutput", "ControlLoRA"]


# ------------------------------------------------------------------------------------------------------------
# small kernel-level helpers
# ------------------------------------------------------------------------------------------------------------
def _rows(x: torch.Tensor) -> torch.Tensor:
    """[..., C] any float dtype on CUDA -> contiguous bf16 [M, C]."""
    ops._cuda(x, "models")
    return x.reshape(-1, x.shape[-1]).to(BF16).contiguous()


def _pad_rank(down: torch.Tensor, up: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Zero-pad the rank of a LoRA pair to one 64-wide K block: down [r,K] -> [64k,K], up [N,r] -> [N,64k]."""
    r = down.shape[0]
    kp = ((r + 63) // 64) * 64
    d = torch.zeros(kp, down.shape[1], dtype=BF16, device=down.device)
    d[:r] = down.to(BF16)
    u = torch.zeros(up.shape[0], kp, dtype=BF16, device=up.device)
    u[:, :r] = up.to(BF16)
    return d, u


def lora_delta(x2: torch.Tensor, down: torch.Tensor, up: torch.Tensor, s: float,
               res: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``res + s * (x2 down^T) up^T`` as two skinny GEMMs (the second one adds ``res`` in its epilogue)."""
    d, u = _pad_rank(down, up)
    t = ops.linear(x2, d)
    return ops.linear(t, u, res=res, alpha=float(s))


class _HeadPad:
    """Padding of the per-head feature dimension (d -> dp) applied to packed weights, never to activations."""

    def __init__(self, heads: int, d: int):
        self.h, self.d, self.dp = heads, d, ops.head_pad(d)

    def rows(self, W: torch.Tensor) -> torch.Tensor:          # [h*d, K] -> [h*dp, K]
        return ops.pad_head_rows(W, self.h, self.d, self.dp)

    def cols(self, W: torch.Tensor) -> torch.Tensor:          # [N, h*d] -> [N, h*dp]
        return ops.pad_head_cols(W, self.h, self.d, self.dp)

    def vec(self, b: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
        return None if b is None else ops.pad_head_rows(b[:, None], self.h, self.d, self.dp)[:, 0]


def _fused_proj(x2: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor],
                loras: Sequence[Tuple[torch.Tensor, torch.Tensor, float]]) -> torch.Tensor:
    """``x2 W^T + bias + sum_i s_i (x2 down_i^T) up_i^T`` with every LoRA sharing the GEMM's own input: the ranks are
    concatenated (scales folded into ``up``) and ride the fused low-rank side product of the tcgen05 kernel; a total
    rank above 32 (the kernel's widest side product) spills the remainder to skinny GEMMs."""
    fused, spill, r = [], [], 0
    for down, up, s in loras:
        if r + down.shape[0] <= 32:
            fused.append((down, up, s))
            r += down.shape[0]
        else:
            spill.append((down, up, s))
    A = Bm = None
    if fused:
        A = torch.cat([d.float() for d, _, _ in fused], 0)
        Bm = torch.cat([u.float() * float(s) for _, u, s in fused], 1)
    y = ops.linear(x2, W, bias, A, Bm, 1.0)
    for down, up, s in spill:
        y = lora_delta(x2, down, up, s, res=y)
    return y


# ------------------------------------------------------------------------------------------------------------
# diffusers-style attention host (diffusers.models.cross_attention.CrossAttention, imported at models.py:12)
# ------------------------------------------------------------------------------------------------------------
class CrossAttention(nn.Module):
    """Parameter container + processor protocol of diffusers' ``CrossAttention``: ``to_q/to_k/to_v`` bias-free
    Linears (``bias=False`` default), ``to_out = [Linear(inner, query_dim), Dropout]``, ``heads``, ``scale``.
    ``forward`` hands over to ``self.processor(self, hidden_states, encoder_hidden_states=, attention_mask=)``."""

    def __init__(self, query_dim: int, cross_attention_dim: Optional[int] = None, heads: int = 8, dim_head: int = 64,
                 dropout: float = 0.0, bias: bool = False, upcast_attention: bool = False, processor=None):
        super().__init__()
        inner = heads * dim_head
        ctx = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.query_dim, self.cross_attention_dim, self.heads, self.dim_head = query_dim, cross_attention_dim, heads, dim_head
        self.inner_dim = inner
        self.scale = dim_head ** -0.5
        self.upcast_attention = upcast_attention
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(ctx, inner, bias=bias)
        self.to_v = nn.Linear(ctx, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(dropout)])
        self.processor = processor if processor is not None else CrossAttnProcessor()

    def set_processor(self, processor) -> None:
        # diffusers pops a replaced nn.Module processor out of _modules; same here so state_dict() does not keep it
        if isinstance(getattr(self, "processor", None), nn.Module) and not isinstance(processor, nn.Module):
            self._modules.pop("processor", None)
        self.processor = processor

    def forward(self, hidden_states, encoder_hidden_states=None, attention_mask=None, **cross_attention_kwargs):
        return self.processor(self, hidden_states, encoder_hidden_states=encoder_hidden_states,
                              attention_mask=attention_mask, **cross_attention_kwargs)

    # -- tensor-layout helpers of the protocol (pure views; used by reference processors, models.py:127-141)
    def head_to_batch_dim(self, t: torch.Tensor) -> torch.Tensor:
        b, n, c = t.shape
        h = self.heads
        return t.reshape(b, n, h, c // h).permute(0, 2, 1, 3).reshape(b * h, n, c // h)

    def batch_to_head_dim(self, t: torch.Tensor) -> torch.Tensor:
        bh, n, d = t.shape
        h = self.heads
        return t.reshape(bh // h, h, n, d).permute(0, 2, 1, 3).reshape(bh // h, n, h * d)

    def prepare_attention_mask(self, attention_mask, target_length, batch_size=None):
        if attention_mask is None:
            return None
        raise capi.TopoError("attention masks are not on the reference's path (always None, models.py:122-123) and are "
                             "not implemented by the flash kernel")

    def get_attention_scores(self, query, key, attention_mask=None) -> torch.Tensor:
        """softmax(q k^T * scale) for ``[B*h, N, d]`` operands, MATERIALISED (GEMM + row-softmax kernels, one launch
        pair per batch-head).  Exists for protocol completeness; the processors below never call it."""
        if attention_mask is not None:
            self.prepare_attention_mask(attention_mask, query.shape[1])
        ops._cuda(query, "get_attention_scores")
        d = query.shape[-1]
        kp = ((d + 63) // 64) * 64   # the GEMM's K granularity; zero columns do not change q k^T
        q = torch.zeros(*query.shape[:-1], kp, dtype=BF16, device=query.device)
        k = torch.zeros(*key.shape[:-1], kp, dtype=BF16, device=key.device)
        q[..., :d], k[..., :d] = query, key
        out = []
        for i in range(q.shape[0]):
            s = ops.linear(q[i], k[i], out_fp32=False)
            out.append(ops.softmax_rows(s, float(self.scale)))
        return torch.stack(out).to(query.dtype)

    # -- the fused core shared by every processor: padded q/k/v -> padded o
    def _core(self, q: torch.Tensor, k: torch.Tensor, v: torch.Tensor, B: int) -> torch.Tensor:
        hp = _HeadPad(self.heads, self.dim_head)
        Nq, Nk = q.shape[0] // B, k.shape[0] // B
        o = ops.attention(q.view(B, Nq, -1), k.view(B, Nk, -1), v.view(B, Nk, -1), self.heads, hp.dp, float(self.scale))
        return o.view(B * Nq, -1)


class CrossAttnProcessor:
    """diffusers' default processor (no LoRA): q/k/v projections, flash core, ``to_out``."""

    def __call__(self, attn: CrossAttention, hidden_states, encoder_hidden_states=None, attention_mask=None, scale=1.0):
        attn.prepare_attention_mask(attention_mask, hidden_states.shape[1])
        B = hidden_states.shape[0]
        hp = _HeadPad(attn.heads, attn.dim_head)
        h2 = _rows(hidden_states)
        e2 = h2 if encoder_hidden_states is None else _rows(encoder_hidden_states)
        q = ops.linear(h2, hp.rows(attn.to_q.weight), hp.vec(attn.to_q.bias))
        k = ops.linear(e2, hp.rows(attn.to_k.weight), hp.vec(attn.to_k.bias))
        v = ops.linear(e2, hp.rows(attn.to_v.weight), hp.vec(attn.to_v.bias))
        o = attn._core(q, k, v, B)
        y = ops.linear(o, hp.cols(attn.to_out[0].weight), attn.to_out[0].bias)
        return y.view(B, -1, y.shape[-1]).to(hidden_states.dtype)


# ------------------------------------------------------------------------------------------------------------
# LoRA processors
# ------------------------------------------------------------------------------------------------------------
def _apply_stack(x2: torch.Tensor, W: torch.Tensor, bias: Optional[torch.Tensor], entries, pad_out=None,
                 pad_in=None, pad_run=None) -> torch.Tensor:
    """Running value of one projection through an ordered LoRA stack.

    ``entries``: ``(layer, scale, post_add, extra)`` in the reference's order.  ``post_add=False`` entries read the
    projection's own input (order-free, fused into the base GEMM while it has not been launched yet);
    ``post_add=True`` entries read the running value (``Lq(query)``, models.py:126) and force it to be materialised.
    ``extra`` (optional bf16 rows) is added to the entry's input — by linearity ``L(in + extra) = L(in) + L(extra)``,
    so it costs one more skinny pair.  ``pad_out`` pads a LoRA ``up`` the way the base weight's rows are padded,
    ``pad_in`` pads the columns of a ``down`` that reads the (head-padded) input, ``pad_run`` those of a ``down`` that
    reads the (head-padded) running value."""
    pad_out = pad_out or (lambda w: w)
    pad_in = pad_in or (lambda w: w)
    pad_run = pad_run or (lambda w: w)
    fused, y = [], None
    extras = []

    def flush():
        nonlocal y, fused
        if y is None:
            y = _fused_proj(x2, W, bias, fused)
        else:
            for d, u, s in fused:
                y = lora_delta(x2, d, u, s, res=y)
        fused = []
        for src, d, u, s in extras:
            y = lora_delta(src, d, u, s, res=y)
        extras.clear()

    for layer, s, post_add, extra in entries:
        down, up = layer.down.weight, pad_out(layer.up.weight)
        if not post_add:
            fused.append((pad_in(down), up, s))
            if extra is not None:
                extras.append((extra, down, up, s))
        else:
            flush()
            y = lora_delta(y, pad_run(down), up, s, res=y)
            if extra is not None:
                y = lora_delta(extra, down, up, s, res=y)
    flush()
    return y


class LoRACrossAttnProcessor(nn.Module):
    """``LoRA_modified/models.py:72-152``: rank-``rank`` LoRA on q, k, v and ``to_out[0]`` of a diffusers attention.
    ``post_add=True`` feeds the *projected* tensor to the LoRA (``Lq(Wq h)``); k / v / out LoRAs can be left out
    (constructor) or muted (``skip_*_states``)."""

    def __init__(self, hidden_size, cross_attention_dim=None, rank=4, post_add=False, key_states_skipped=False,
                 value_states_skipped=False, output_states_skipped=False):
        super().__init__()
        self.hidden_size, self.cross_attention_dim, self.rank, self.post_add = hidden_size, cross_attention_dim, rank, post_add
        kv_in = hidden_size if post_add else (cross_attention_dim or hidden_size)
        self.to_q_lora = LoRALinearLayer(hidden_size, hidden_size, rank)
        if not key_states_skipped:
            self.to_k_lora = LoRALinearLayer(kv_in, hidden_size, rank)
        if not value_states_skipped:
            self.to_v_lora = LoRALinearLayer(kv_in, hidden_size, rank)
        if not output_states_skipped:
            self.to_out_lora = LoRALinearLayer(hidden_size, hidden_size, rank)
        self.key_states_skipped: bool = key_states_skipped
        self.value_states_skipped: bool = value_states_skipped
        self.output_states_skipped: bool = output_states_skipped

    def skip_key_states(self, is_skipped: bool = True):
        if is_skipped is False:
            assert hasattr(self, "to_k_lora")
        self.key_states_skipped = is_skipped

    def skip_value_states(self, is_skipped: bool = True):
        if is_skipped is False:
            assert hasattr(self, "to_q_lora")  # sic: the reference asserts on to_q_lora (models.py:110)
        self.value_states_skipped = is_skipped

    def skip_output_states(self, is_skipped: bool = True):
        if is_skipped is False:
            assert hasattr(self, "to_out_lora")
        self.output_states_skipped = is_skipped

    # ---- stacks: which LoRA layers act on q / k / v / out, in the reference's order -------------------------
    def _chain(self) -> List["LoRACrossAttnProcessor"]:
        return list(getattr(self, "pre_loras", [])) + [self] + list(getattr(self, "post_loras", []))

    def _q_entries(self, h2, scale, attn):
        return [(self.to_q_lora, scale, self.post_add, None)]

    def _k_entries(self, scale):
        return [(p.to_k_lora, scale, p.post_add, None) for p in self._chain() if not p.key_states_skipped]

    def _v_entries(self, scale):
        # pre / post stacks add their value LoRA WITHOUT `scale` (models.py:260,265 — kept as is)
        return [(p.to_v_lora, scale if p is self else 1.0, p.post_add, None) for p in self._chain()
                if not p.value_states_skipped]

    def _out_entries(self, scale):
        # the owning processor's to_out_lora is applied unconditionally in the Control variants (models.py:279,422)
        own_always = not type(self) is LoRACrossAttnProcessor
        return [(p.to_out_lora, scale, p.post_add, None) for p in self._chain()
                if (p is self and own_always) or not p.output_states_skipped]

    def _pre_hidden(self, h2, scale):    # V2 hook: control added to the hidden states before q
        return h2

    def _pre_out(self, o2, scale, hp):   # V2 hook: control added to the attention output before to_out
        return o2

    def __call__(self, attn: CrossAttention, hidden_states, encoder_hidden_states=None, attention_mask=None, scale=1.0):
        attn.prepare_attention_mask(attention_mask, hidden_states.shape[1])
        B = hidden_states.shape[0]
        hp = _HeadPad(attn.heads, attn.dim_head)
        self._dtype = hidden_states.dtype
        h2 = self._pre_hidden(_rows(hidden_states), scale)
        e2 = h2 if encoder_hidden_states is None else _rows(encoder_hidden_states)
        qkv = dict(pad_out=hp.rows, pad_run=hp.cols)
        q = _apply_stack(h2, hp.rows(attn.to_q.weight), hp.vec(attn.to_q.bias), self._q_entries(h2, scale, attn), **qkv)
        k = _apply_stack(e2, hp.rows(attn.to_k.weight), hp.vec(attn.to_k.bias), self._k_entries(scale), **qkv)
        v = _apply_stack(e2, hp.rows(attn.to_v.weight), hp.vec(attn.to_v.bias), self._v_entries(scale), **qkv)
        o = self._pre_out(attn._core(q, k, v, B), scale, hp)
        y = _apply_stack(o, hp.cols(attn.to_out[0].weight), attn.to_out[0].bias, self._out_entries(scale), pad_in=hp.cols)
        return y.view(B, -1, y.shape[-1]).to(hidden_states.dtype)   # to_out[1] is Dropout: identity at inference


class _ControlMixin:
    """Control-state plumbing shared by the two Control processors (models.py:191-220, 328-357)."""

    def inject_pre_lora(self, lora_layer):
        self.pre_loras.append(lora_layer)

    def inject_post_lora(self, lora_layer):
        self.post_loras.append(lora_layer)

    def inject_control_states(self, control_states):
        self.control_states = control_states   # an ALIAS of the caller's tensor, as in the reference (:198-199)

    def _control_rows(self, n_rows: int, like: torch.Tensor) -> torch.Tensor:
        """bf16 rows [M, Cc] of the injected control states.  A BCHW map is flattened to [B, HW, C] once and cached
        back on the processor (reference :203-206).  When the control batch divides the hidden batch the rows are
        repeated per sample (reference :209-212, there only under ``concat_hidden``; a plain broadcast otherwise)."""
        cs = self.control_states
        ops._cuda(cs, "control_states")
        if cs.ndim == 4:
            Bc, Cc, H, W = cs.shape
            cs = ops.nchw_to_nhwc(cs.float()).view(Bc, H * W, Cc).to(cs.dtype)
            self.control_states = cs
        rows = cs.reshape(-1, cs.shape[-1]).to(BF16).contiguous()
        if rows.shape[0] != n_rows:
            if n_rows % rows.shape[0]:
                raise capi.TopoError(f"control states with {rows.shape[0]} rows cannot serve {n_rows} hidden rows")
            b1 = cs.shape[0]
            rep = n_rows // rows.shape[0]
            rows = rows.view(b1, 1, -1, rows.shape[-1]).expand(b1, rep, -1, rows.shape[-1]).reshape(n_rows, -1).contiguous()
        return rows

    def _to_control(self, layer: LoRALinearLayer, h2: torch.Tensor, cs2: torch.Tensor, scale: float) -> torch.Tensor:
        """``scale * layer(cat([h, cs]) if concat_hidden else cs)`` (+ cs when ``control_self_add``) without the
        concatenation: the down projection is split column-wise into its hidden and control halves."""
        down, up = layer.down.weight, layer.up.weight
        if self.concat_hidden:
            C = h2.shape[1]
            dh, _ = _pad_rank(down[:, :C], up)
            dc, u = _pad_rank(down[:, C:], up)
            t = ops.linear(cs2, dc, res=ops.linear(h2, dh))
        else:
            dc, u = _pad_rank(down, up)
            t = ops.linear(cs2, dc)
        return ops.linear(t, u, res=cs2 if self.control_self_add else None, alpha=float(scale))


class ControlLoRACrossAttnProcessor(_ControlMixin, LoRACrossAttnProcessor):
    """``LoRA_modified/models.py:155-287``: the control map of this level goes through the low-rank ``to_control`` and
    is ADDED TO THE INPUT OF THE q LoRA: ``q += s * Lq(h + ctrl')``; stacks of ``pre_loras`` / ``post_loras`` apply
    around the processor's own LoRAs on q, k, v and out."""

    def __init__(self, hidden_size, cross_attention_dim=None, rank=4, control_rank=None, post_add=False,
                 concat_hidden=False, control_channels=None, control_self_add=True, key_states_skipped=False,
                 value_states_skipped=False, output_states_skipped=False, **kwargs):
        super().__init__(hidden_size, cross_attention_dim, rank, post_add=post_add, key_states_skipped=key_states_skipped,
                         value_states_skipped=value_states_skipped, output_states_skipped=output_states_skipped)
        control_rank = rank if control_rank is None else control_rank
        control_channels = hidden_size if control_channels is None else control_channels
        self.concat_hidden = concat_hidden
        # reference :182 tests `control_channels is None` AFTER defaulting it two lines above, so the flag is always
        # False there; reproduced (the argument is accepted and ignored exactly as upstream does)
        self.control_self_add = False
        self.control_states: Optional[torch.Tensor] = None
        self.to_control = LoRALinearLayer(control_channels + (hidden_size if concat_hidden else 0), hidden_size, control_rank)
        self.pre_loras: List[LoRACrossAttnProcessor] = []
        self.post_loras: List[LoRACrossAttnProcessor] = []

    def process_control_states(self, hidden_states, scale=1.0):
        """Public form (reference :201-220): returns ``[B, N, C]`` in ``hidden_states``' dtype."""
        h2 = _rows(hidden_states)
        y = self._to_control(self.to_control, h2, self._control_rows(h2.shape[0], hidden_states), scale)
        return y.view(*hidden_states.shape[:-1], y.shape[-1]).to(hidden_states.dtype)

    def _ctrl(self, h2, scale):
        assert self.control_states is not None
        return self._to_control(self.to_control, h2, self._control_rows(h2.shape[0], h2), scale)

    def _q_entries(self, h2, scale, attn):
        out = []
        for p in self._chain():
            extra = p._ctrl(h2, scale) if isinstance(p, ControlLoRACrossAttnProcessor) else None
            out.append((p.to_q_lora, scale, p.post_add, extra))
        return out


class ControlLoRACrossAttnProcessorV2(_ControlMixin, LoRACrossAttnProcessor):
    """``LoRA_modified/models.py:292-431``: control enters the HIDDEN STATES (``h += s * Lc([h | ctrl])`` before the q
    projection) and again the attention output before ``to_out`` through a second low-rank map ``to_control_out``;
    k / v LoRAs are absent by construction."""

    def __init__(self, hidden_size, cross_attention_dim=None, rank=4, control_rank=None, control_channels=None, **kwargs):
        super().__init__(hidden_size, cross_attention_dim, rank, post_add=False, key_states_skipped=True,
                         value_states_skipped=True, output_states_skipped=False)
        control_rank = rank if control_rank is None else control_rank
        control_channels = hidden_size if control_channels is None else control_channels
        self.concat_hidden = True
        self.control_self_add = False
        self.control_states: Optional[torch.Tensor] = None
        self.to_control = LoRALinearLayer(hidden_size + control_channels, hidden_size, control_rank)
        self.to_control_out = LoRALinearLayer(hidden_size + control_channels, hidden_size, control_rank)
        self.pre_loras: List[LoRACrossAttnProcessor] = []
        self.post_loras: List[LoRACrossAttnProcessor] = []

    def process_control_states(self, hidden_states, scale=1.0, is_out=False):
        h2 = _rows(hidden_states)
        layer = self.to_control_out if is_out else self.to_control
        y = self._to_control(layer, h2, self._control_rows(h2.shape[0], hidden_states), scale)
        return y.view(*hidden_states.shape[:-1], y.shape[-1]).to(hidden_states.dtype)

    def _q_entries(self, h2, scale, attn):
        return [(p.to_q_lora, scale, p.post_add, None) for p in self._chain()]

    def _v2_chain(self):
        return [p for p in self._chain() if isinstance(p, ControlLoRACrossAttnProcessorV2)]

    def _pre_hidden(self, h2, scale):
        assert self.control_states is not None
        for p in self._v2_chain():   # each term sees the hidden states already updated by the previous ones (:366-372)
            h2 = ops.add(h2, p._to_control(p.to_control, h2, p._control_rows(h2.shape[0], h2), scale))
        return h2

    def _pre_out(self, o2, scale, hp):
        # the attention output is head-padded here; the control maps read / write the unpadded hidden layout, so the
        # pads are dropped for this (rare, V2-only) step and restored by to_out's unpadded weight below
        if hp.dp != hp.d:
            o2 = o2.view(o2.shape[0], hp.h, hp.dp)[:, :, :hp.d].reshape(o2.shape[0], hp.h * hp.d).contiguous()
        for p in self._v2_chain():
            o2 = ops.add(o2, p._to_control(p.to_control_out, o2, p._control_rows(o2.shape[0], o2), scale))
        if hp.dp != hp.d:
            pad = torch.zeros(o2.shape[0], hp.h, hp.dp, dtype=o2.dtype, device=o2.device)
            pad[:, :, :hp.d] = o2.view(o2.shape[0], hp.h, hp.d)
            o2 = pad.view(o2.shape[0], hp.h * hp.dp)
        return o2


# ------------------------------------------------------------------------------------------------------------
# ControlLoRA conv pyramid (models.py:434-833)
# ------------------------------------------------------------------------------------------------------------
def _conv_nhwc(x: torch.Tensor, conv: nn.Conv2d, pad_lo: int = -1) -> torch.Tensor:
    """NHWC bf16 [B,H,W,Cin] -> [B,Ho,Wo,Cout]: 1x1 as a GEMM, 3x3/s1 with Cin % 64 == 0 as the tensor-core implicit
    GEMM, everything else (3 or 32 input channels, the stride-2 downsamplers) on the direct kernel."""
    B, H, W, C = x.shape
    k, st = conv.kernel_size[0], conv.stride[0]
    if k == 1 and st == 1:
        y = ops.linear(x.reshape(-1, C), conv.weight.reshape(conv.out_channels, C), conv.bias)
        return y.view(B, H, W, -1)
    if k == 3 and st == 1 and C % 64 == 0 and ((H * W) % 128 == 0) and (W % 128 == 0 or 128 % W == 0):
        return ops.conv3x3(x, conv.weight, conv.bias)
    b = None if conv.bias is None else conv.bias.float().contiguous()
    return ops.conv_direct(x, ops.conv_direct_weight(conv.weight), b, C, conv.out_channels, k, st, pad_lo=pad_lo)


class Downsample2D(nn.Module):
    """diffusers ``Downsample2D(channels, use_conv=True, out_channels=, padding=, name="op")``: stride-2 3x3 conv;
    ``padding=0`` pads one zero row/column AFTER the image (``F.pad(x, (0,1,0,1))``) — the form the pyramid uses
    (``downsample_padding=0``, models.py:709,758)."""

    def __init__(self, channels, use_conv=True, out_channels=None, padding=1, name="conv"):
        super().__init__()
        if not use_conv:
            raise capi.TopoError("Downsample2D(use_conv=False) (average pooling) is not on the reference's path")
        self.channels, self.out_channels, self.padding, self.name = channels, out_channels or channels, padding, name
        self.conv = nn.Conv2d(channels, self.out_channels, 3, stride=2, padding=padding)  # key "...downsamplers.0.conv.*"

    def _run(self, x: torch.Tensor) -> torch.Tensor:
        return _conv_nhwc(x, self.conv, pad_lo=0 if self.padding == 0 else -1)


class ConvBlock2D(nn.Module):
    """``GroupNorm -> SiLU -> conv(k) -> GroupNorm -> SiLU`` (models.py:434-547 with ``temb_channels=None``: the only
    form the pyramid instantiates; ``up`` / ``down`` resampling inside the block is never requested there)."""

    def __init__(self, *, in_channels, out_channels=None, conv_kernel_size=3, dropout=0.0, temb_channels=512, groups=32,
                 groups_out=None, pre_norm=True, eps=1e-6, non_linearity="swish", time_embedding_norm="default",
                 kernel=None, output_scale_factor=1.0, up=False, down=False):
        super().__init__()
        if up or down:
            raise capi.TopoError("ConvBlock2D(up/down=True) is not on the reference's path")
        if non_linearity not in ("swish", "silu"):
            raise capi.TopoError(f"ConvBlock2D: activation {non_linearity!r} is not on the reference's path (silu only)")
        out_channels = in_channels if out_channels is None else out_channels
        self.in_channels, self.out_channels = in_channels, out_channels
        self.pre_norm, self.time_embedding_norm, self.output_scale_factor = True, time_embedding_norm, output_scale_factor
        groups_out = groups if groups_out is None else groups_out
        self.norm1 = nn.GroupNorm(groups, in_channels, eps=eps, affine=True)
        self.conv1 = nn.Conv2d(in_channels, out_channels, conv_kernel_size, stride=1, padding=conv_kernel_size // 2)
        self.time_emb_proj = nn.Linear(temb_channels, out_channels) if temb_channels is not None else None
        self.norm2 = nn.GroupNorm(groups_out, out_channels, eps=eps, affine=True)
        self.dropout = nn.Dropout(dropout)

    @staticmethod
    def _gn_silu(x: torch.Tensor, n: nn.GroupNorm) -> torch.Tensor:
        return ops.groupnorm_silu(x, n.weight.float().contiguous(), n.bias.float().contiguous(), n.num_groups, float(n.eps), True)

    def _run(self, x: torch.Tensor) -> torch.Tensor:
        h = _conv_nhwc(self._gn_silu(x, self.norm1), self.conv1)
        return self._gn_silu(h, self.norm2)


class SimpleDownEncoderBlock2D(nn.Module):
    """``num_layers`` ConvBlock2D (no time embedding) + optional stride-2 ``Downsample2D`` (models.py:550-610)."""

    def __init__(self, in_channels: int, out_channels: int, dropout: float = 0.0, num_layers: int = 1,
                 convnet_eps: float = 1e-6, convnet_time_scale_shift: str = "default", convnet_act_fn: str = "swish",
                 convnet_groups: int = 32, convnet_pre_norm: bool = True, convnet_kernel_size: int = 3,
                 output_scale_factor=1.0, add_downsample=True, downsample_padding=1):
        super().__init__()
        convnets = []
        for i in range(num_layers):
            cin = in_channels if i == 0 else out_channels
            convnets.append(ConvBlock2D(in_channels=cin, out_channels=out_channels, temb_channels=None, eps=convnet_eps,
                                        groups=convnet_groups, dropout=dropout,
                                        time_embedding_norm=convnet_time_scale_shift, non_linearity=convnet_act_fn,
                                        output_scale_factor=output_scale_factor, pre_norm=convnet_pre_norm,
                                        conv_kernel_size=convnet_kernel_size))
        cin = in_channels if num_layers == 0 else out_channels
        self.convnets = nn.ModuleList(convnets)
        self.downsamplers = nn.ModuleList([Downsample2D(cin, use_conv=True, out_channels=out_channels,
                                                        padding=downsample_padding, name="op")]) if add_downsample else None

    def _run(self, x: torch.Tensor) -> torch.Tensor:
        for c in self.convnets:
            x = c._run(x)
        if self.downsamplers is not None:
            for d in self.downsamplers:
                x = d._run(x)
        return x


def get_down_block(down_block_type, num_layers, in_channels, out_channels, temb_channels, add_downsample, resnet_eps,
                   resnet_act_fn, attn_num_head_channels, resnet_groups=None, cross_attention_dim=None,
                   downsample_padding=None, dual_cross_attention=False, use_linear_projection=False,
                   only_cross_attention=False, upcast_attention=False, resnet_time_scale_shift="default",
                   resnet_kernel_size=3):
    """models.py:15-69.  Only ``SimpleDownEncoderBlock2D`` exists here; the reference forwards every other type to
    diffusers' own ``get_down_block`` (UNet blocks that the pyramid never asks for)."""
    t = down_block_type[7:] if down_block_type.startswith("UNetRes") else down_block_type
    if t != "SimpleDownEncoderBlock2D":
        raise capi.TopoError(f"get_down_block: {down_block_type!r} is a diffusers UNet block outside the ControlLoRA path")
    return SimpleDownEncoderBlock2D(num_layers=num_layers, in_channels=in_channels, out_channels=out_channels,
                                    add_downsample=add_downsample, convnet_eps=resnet_eps, convnet_act_fn=resnet_act_fn,
                                    convnet_groups=resnet_groups, downsample_padding=downsample_padding,
                                    convnet_time_scale_shift=resnet_time_scale_shift,
                                    convnet_kernel_size=resnet_kernel_size)


@dataclass
class ControlLoRAOutput:
    control_states: Tuple[torch.Tensor, ...]


class ControlLoRA(nn.Module):
    """``LoRA_modified/models.py:618-833``: conv pyramid on the 3x512x512 control image.

    ``conv_in 3->32`` -> four encoder blocks (32, 64, 128, 256; stride-2 ``Downsample2D(padding=0)`` between) ->
    256 ch @ 64x64; per level i a ``pre_lora_layer`` (1x1 ConvBlock2D 256 -> {320, 640, 1280, 1280}) gives the control
    map that is injected into that level's processors (``lora_layers[i]``), then a stride-2 block (256 -> 256) feeds
    the next level: control maps 320@64^2, 640@32^2, 1280@16^2, 1280@8^2.  ``forward(x, return_dict=False)`` returns
    the tuple; with ``return_dict=True`` the reference falls off the end of ``forward`` and returns ``None``
    (:832-868, the dataclass return is commented out) — reproduced.  ``config`` mirrors ``register_to_config``."""

    def __init__(self, in_channels: int = 3,
                 down_block_types: Tuple[str, ...] = ("SimpleDownEncoderBlock2D",) * 4,
                 block_out_channels: Tuple[int, ...] = (32, 64, 128, 256), layers_per_block: int = 1, act_fn: str = "silu",
                 norm_num_groups: int = 32,
                 lora_pre_down_block_types: Tuple[Optional[str], ...] = (None,) + ("SimpleDownEncoderBlock2D",) * 3,
                 lora_pre_down_layers_per_block: int = 1, lora_pre_conv_skipped: bool = False,
                 lora_pre_conv_types: Tuple[str, ...] = ("SimpleDownEncoderBlock2D",) * 4,
                 lora_pre_conv_layers_per_block: int = 1, lora_pre_conv_layers_kernel_size: int = 1,
                 lora_block_in_channels: Tuple[int, ...] = (256, 256, 256, 256),
                 lora_block_out_channels: Tuple[int, ...] = (320, 640, 1280, 1280),
                 lora_cross_attention_dims: Tuple[List[Optional[int]], ...] = ([None, 768] * 5, [None, 768] * 5,
                                                                              [None, 768] * 5, [None, 768]),
                 lora_rank: int = 4, lora_control_rank: Optional[int] = None, lora_post_add: bool = False,
                 lora_concat_hidden: bool = False, lora_control_channels: Tuple[Optional[int], ...] = (None,) * 4,
                 lora_control_self_add: bool = True, lora_key_states_skipped: bool = False,
                 lora_value_states_skipped: bool = False, lora_output_states_skipped: bool = False,
                 lora_control_version: int = 1):
        super().__init__()
        self.config = {k: v for k, v in locals().items() if k not in ("self", "__class__")}
        cls = ControlLoRACrossAttnProcessorV2 if lora_control_version == 2 else ControlLoRACrossAttnProcessor
        assert lora_block_in_channels[0] == block_out_channels[-1]
        if lora_pre_conv_skipped:
            lora_control_channels = lora_block_in_channels
            lora_control_self_add = False
        self.layers_per_block = layers_per_block
        self.lora_pre_down_layers_per_block = lora_pre_down_layers_per_block
        self.lora_pre_conv_layers_per_block = lora_pre_conv_layers_per_block
        self.conv_in = nn.Conv2d(in_channels, block_out_channels[0], kernel_size=3, stride=1, padding=1)
        self.down_blocks = nn.ModuleList([])
        self.pre_lora_layers = nn.ModuleList([])
        self.lora_layers = nn.ModuleList([])
        common = dict(resnet_eps=1e-6, downsample_padding=0, resnet_act_fn=act_fn, resnet_groups=norm_num_groups,
                      attn_num_head_channels=None, temb_channels=None)

        def pre_conv(i, cin):
            if lora_pre_conv_skipped:
                return nn.Identity()
            cout = lora_block_out_channels[i] if lora_control_channels[i] is None else lora_control_channels[i]
            return get_down_block(lora_pre_conv_types[i], num_layers=lora_pre_conv_layers_per_block, in_channels=cin,
                                  out_channels=cout, add_downsample=False,
                                  resnet_kernel_size=lora_pre_conv_layers_kernel_size, **common)

        def processors(i):
            return nn.ModuleList([cls(lora_block_out_channels[i], cross_attention_dim=cad, rank=lora_rank,
                                      control_rank=lora_control_rank, post_add=lora_post_add,
                                      concat_hidden=lora_concat_hidden, control_channels=lora_control_channels[i],
                                      control_self_add=lora_control_self_add, key_states_skipped=lora_key_states_skipped,
                                      value_states_skipped=lora_value_states_skipped,
                                      output_states_skipped=lora_output_states_skipped)
                                  for cad in lora_cross_attention_dims[i]])

        pre, cout = [], block_out_channels[0]
        for i, t in enumerate(down_block_types):
            cin, cout = cout, block_out_channels[i]
            pre.append(get_down_block(t, num_layers=layers_per_block, in_channels=cin, out_channels=cout,
                                      add_downsample=i != len(block_out_channels) - 1, **common))
        self.down_blocks.append(nn.Sequential(*pre))
        self.pre_lora_layers.append(pre_conv(0, lora_block_in_channels[0]))
        self.lora_layers.append(processors(0))
        cout = lora_block_in_channels[0]
        for i, t in enumerate(lora_pre_down_block_types):
            if i == 0:
                continue
            cin, cout = cout, lora_block_in_channels[i]
            self.down_blocks.append(get_down_block(t, num_layers=lora_pre_down_layers_per_block, in_channels=cin,
                                                   out_channels=cout, add_downsample=True, **common))
            self.pre_lora_layers.append(pre_conv(i, cout))
            self.lora_layers.append(processors(i))

    @staticmethod
    def _run_block(block, h: torch.Tensor) -> torch.Tensor:
        if isinstance(block, nn.Identity):
            return h
        if isinstance(block, nn.Sequential):
            for b in block:
                h = b._run(h)
            return h
        return block._run(h)

    @torch.no_grad()
    def forward(self, x: torch.Tensor, return_dict: bool = True) -> Union[None, Tuple[torch.Tensor, ...]]:
        ops._cuda(x, "ControlLoRA.forward")
        orig_dtype = x.dtype
        h = ops.nchw_to_nhwc(x.float())
        b = None if self.conv_in.bias is None else self.conv_in.bias.float().contiguous()
        h = ops.conv_direct(h, ops.conv_direct_weight(self.conv_in.weight), b, self.conv_in.in_channels,
                            self.conv_in.out_channels, 3, 1)
        control_states_list = []
        for down_block, pre_lora_layer, lora_layer_list in zip(self.down_blocks, self.pre_lora_layers, self.lora_layers):
            h = self._run_block(down_block, h)
            cs = ops.nhwc_to_nchw(self._run_block(pre_lora_layer, h)).to(orig_dtype)   # [B, C, H, W], caller's dtype
            for lora_layer in lora_layer_list:
                lora_layer.inject_control_states(cs)
            control_states_list.append(cs)
        if not return_dict:
            return tuple(control_states_list)
        return None   # sic (see the class docstring)
