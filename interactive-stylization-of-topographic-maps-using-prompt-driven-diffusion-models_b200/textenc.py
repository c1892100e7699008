"""Text conditioning producer: frozen HuggingFace BERT-base + the trainable last-state
adapter MLP, on the sm_100a kernels.  Mirrors ``build_text_encoder`` / ``encode_on_device`` of
``LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:121-192`` (2 hidden adapter layers) and
``…_addlrSD.py:121-192`` (5 hidden layers, ``:138-170``):

    tokenizer(prompts, padding="max_length", truncation=True, max_length=77) -> input_ids, attention_mask
    tokens = hf_bert(input_ids, attention_mask).last_hidden_state            # [B, 77, 768]
    tokens = adapter(tokens)   # (Linear -> ReLU -> LayerNorm) x k, Linear, final LayerNorm   -> c_crossattn

The HF ``BertModel`` and the ``LastStateAdapter`` are used as PARAMETER CONTAINERS (same attribute / state-dict names
as the reference: ``te.hf_bert``, ``te.tokenizer``, ``te.max_len``, ``te.adapter``, ``te.encode``); the arithmetic is
packed once and executed as C-ABI calls: embedding sum + LayerNorm, per layer one fused q/k/v GEMM, flash attention
with the per-prompt key length (the attention mask of a right-padded prompt), out-projection with the residual in the
epilogue, LayerNorm, GELU GEMM, GEMM + residual, LayerNorm; then the adapter's ReLU GEMMs and LayerNorms.
Runs once per tile batch (not per DDIM step).  No CPU path: the fp32 checker is HF's own ``BertModel.forward``.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from . import capi, ops
from .capi import ACT_GELU, ACT_NONE, ACT_RELU
from .engine import Lin, Norm, _dev, gemm, gemm_multi, ln
from .ops import BF16

__all__ = ["LastStateAdapter", "BERTTextEncoder", "build_text_encoder", "BertAlign", "build_cond_uc_safe_pad"]


class LastStateAdapter(nn.Module):
    """``ff = (Linear, ReLU, LayerNorm) x n_hidden + Linear`` then ``final_ln`` ("mimics CLIP's final LayerNorm").
    ``n_hidden = 2``: ``…_ctrlora.py:139-158``; ``n_hidden = 5``: ``…_addlrSD.py:138-170``.  Same ``ff.<i>`` indices."""

    def __init__(self, dim: int = 768, hidden: int = 768, n_hidden: int = 2):
        super().__init__()
        layers: List[nn.Module] = []
        cin = dim
        for _ in range(n_hidden):
            layers += [nn.Linear(cin, hidden), nn.ReLU(), nn.LayerNorm(hidden)]
            cin = hidden
        layers.append(nn.Linear(cin, dim, bias=True))
        self.ff = nn.Sequential(*layers)
        self.final_ln = nn.LayerNorm(dim)

    def forward(self, x):
        raise capi.TopoError("LastStateAdapter is a parameter container here: call BERTTextEncoder.encode / "
                             ".encode_ids (sm_100a kernels); the fp32 checker lives in tests/")


def _lin(m: nn.Linear, device) -> Lin:
    return Lin(m.weight.detach().to(device=device, dtype=BF16).contiguous(), _dev(m.bias, device, torch.float32))


def _norm(m: nn.LayerNorm, device) -> Norm:
    return Norm(_dev(m.weight, device, torch.float32), _dev(m.bias, device, torch.float32), float(m.eps))


class _PackedLayer:
    def __init__(self, layer, device):
        a = layer.attention
        self.q, self.k, self.v = _lin(a.self.query, device), _lin(a.self.key, device), _lin(a.self.value, device)
        self.o, self.ln1 = _lin(a.output.dense, device), _norm(a.output.LayerNorm, device)
        self.i, self.f, self.ln2 = _lin(layer.intermediate.dense, device), _lin(layer.output.dense, device), _norm(layer.output.LayerNorm, device)


class BERTTextEncoder(nn.Module):
    """The object ``build_text_encoder`` returns: ``hf_bert`` (frozen), ``tokenizer``, ``max_len``, ``adapter``,
    ``encode(prompts) -> [B, max_len, n_embed]`` fp32 on the device."""

    def __init__(self, hf_bert: nn.Module, tokenizer=None, adapter: Optional[LastStateAdapter] = None, max_len: int = 77):
        super().__init__()
        cfg = hf_bert.config
        if getattr(cfg, "hidden_act", "gelu") != "gelu":
            raise capi.TopoError(f"BERT activation {cfg.hidden_act!r}: only the erf GELU of bert-base-uncased is implemented")
        if getattr(cfg, "position_embedding_type", "absolute") != "absolute":
            raise capi.TopoError("only absolute position embeddings (bert-base-uncased) are implemented")
        self.hf_bert, self.tokenizer, self.max_len = hf_bert, tokenizer, max_len
        self.adapter = adapter if adapter is not None else LastStateAdapter(cfg.hidden_size)
        for p in self.hf_bert.parameters():
            p.requires_grad = False
        self._packed = None
        self.encode = self.encode_on_device   # the reference assigns the closure to te.encode (…_ctrlora.py:190)

    # ---- packing (once per weight load) -------------------------------------------------------------------
    def _fingerprint(self):
        ps = list(self.hf_bert.parameters()) + list(self.adapter.parameters())
        return (len(ps), sum(p._version for p in ps), str(ps[0].device))

    def _pack(self):
        fp = self._fingerprint()
        if self._packed is not None and self._packed["fp"] == fp:
            return self._packed
        bert = self.hf_bert
        dev = next(bert.parameters()).device
        if dev.type != "cuda":
            raise capi.TopoError("BERTTextEncoder: move the encoder to a B200 (`te.to('cuda')`); there is no CPU forward")
        capi.handle(dev.index if dev.index is not None else torch.cuda.current_device())
        e = bert.embeddings
        P = dict(fp=fp, dev=dev, heads=bert.config.num_attention_heads, hid=bert.config.hidden_size,
                 word=_dev(e.word_embeddings.weight, dev, torch.float32), pos=_dev(e.position_embeddings.weight, dev, torch.float32),
                 type0=_dev(e.token_type_embeddings.weight[0], dev, torch.float32), emb_ln=_norm(e.LayerNorm, dev),
                 layers=[_PackedLayer(l, dev) for l in bert.encoder.layer])
        ff = list(self.adapter.ff)
        steps = []
        for i, m in enumerate(ff):
            if isinstance(m, nn.Linear):
                relu = i + 1 < len(ff) and isinstance(ff[i + 1], nn.ReLU)
                steps.append(("lin", _lin(m, dev), relu))
            elif isinstance(m, nn.LayerNorm):
                steps.append(("ln", _norm(m, dev), False))
            elif not isinstance(m, nn.ReLU):
                raise capi.TopoError(f"adapter layer {type(m).__name__} is not part of the reference adapter")
        P["adapter"] = steps
        P["final_ln"] = _norm(self.adapter.final_ln, dev)
        self._packed = P
        return P

    # ---- forward ---------------------------------------------------------------------------------------------
    @torch.no_grad()
    def bert_last_hidden_state(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``hf_bert(input_ids, attention_mask).last_hidden_state`` as bf16 rows ``[B*T, hidden]``."""
        P = self._pack()
        dev = P["dev"]
        ids = input_ids.to(dev)
        B, T = ids.shape
        if T > P["pos"].shape[0]:
            raise capi.TopoError(f"sequence length {T} exceeds the position table ({P['pos'].shape[0]})")
        if attention_mask is None:
            kv_len = torch.full((B,), T, dtype=torch.int32, device=dev)
        else:
            m = attention_mask.to(dev).to(torch.int32)
            kv_len = m.sum(1).to(torch.int32)
            # the flash kernel masks a SUFFIX of the keys: that is what padding="max_length" produces
            prefix = (torch.arange(T, device=dev)[None, :] < kv_len[:, None]).to(torch.int32)
            if not torch.equal(prefix, m):
                raise capi.TopoError("attention_mask must be a right-padded prefix mask (tokenizer padding='max_length')")
        heads, hid = P["heads"], P["hid"]
        d = hid // heads
        dp = ops.head_pad(d)
        if dp != d:
            raise capi.TopoError(f"BERT head dim {d} needs padding to {dp}: not implemented (bert-base has d = 64)")
        x = ln(ops.embed_sum(ids, P["word"], P["pos"], P["type0"]).view(B * T, hid), P["emb_ln"])
        for L in P["layers"]:
            qkv = torch.empty(B * T, 3 * hid, dtype=BF16, device=dev)
            gemm_multi(x, [L.q, L.k, L.v], qkv, [0, hid, 2 * hid])
            t = qkv.view(B, T, 3 * hid)
            o = ops.attention(t[:, :, :hid], t[:, :, hid:2 * hid], t[:, :, 2 * hid:], heads, dp, d ** -0.5, kv_len=kv_len)
            x = ln(gemm(o.view(B * T, hid), L.o, res=x), L.ln1)
            x = ln(gemm(gemm(x, L.i, act=ACT_GELU), L.f, res=x), L.ln2)
        return x

    @torch.no_grad()
    def encode_ids(self, input_ids: torch.Tensor, attention_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
        """``adapter(hf_bert(...).last_hidden_state)`` -> fp32 ``[B, T, n_embed]`` (the ``c_crossattn`` context)."""
        P = self._pack()
        B, T = input_ids.shape
        x = self.bert_last_hidden_state(input_ids, attention_mask)
        for kind, w, relu in P["adapter"]:
            x = gemm(x, w, act=ACT_RELU if relu else ACT_NONE) if kind == "lin" else ln(x, w)
        x = ln(x, P["final_ln"])
        return x.view(B, T, -1).float()

    def encode_on_device(self, prompts: Sequence[str]) -> torch.Tensor:
        if self.tokenizer is None:
            raise capi.TopoError("BERTTextEncoder.encode needs a tokenizer (BertTokenizer of bert-base-uncased); "
                                 "use encode_ids(input_ids, attention_mask) with pre-tokenised prompts otherwise")
        enc = self.tokenizer(list(prompts), return_tensors="pt", padding="max_length", truncation=True, max_length=self.max_len)
        return self.encode_ids(enc["input_ids"], enc["attention_mask"])

    def forward(self, text):
        return self.encode_on_device(text)


def build_text_encoder(n_embed: int = 768, n_layer: int = 12, max_len: int = 77, device="cuda", use_tokenizer: bool = True,
                       hf_bert: Optional[nn.Module] = None, tokenizer=None, adapter_hidden_layers: int = 2) -> BERTTextEncoder:
    """``build_text_encoder`` of the training scripts.  The reference downloads ``bert-base-uncased``; offline callers
    pass a loaded ``hf_bert`` (any ``transformers.BertModel``) and ``tokenizer``.  ``adapter_hidden_layers``: 2 for
    the onlyCNLora / ctrlora scripts, 5 for addlrSD."""
    if hf_bert is None:
        from transformers import BertModel, BertTokenizer
        hf_bert = BertModel.from_pretrained("bert-base-uncased")
        tokenizer = tokenizer or (BertTokenizer.from_pretrained("bert-base-uncased") if use_tokenizer else None)
    from .lora import disable_bert_dropout
    disable_bert_dropout(hf_bert)
    hf_bert.eval()
    te = BERTTextEncoder(hf_bert, tokenizer, LastStateAdapter(n_embed, n_embed, adapter_hidden_layers), max_len)
    return te.to(device)


# ----------------------------------------------------------------------------------------------------------------
# CLIP (+) BERT context fusion (TextEncoder_Finetuning/token_utils.py:95-158)
# ----------------------------------------------------------------------------------------------------------------
def _plain_ln(x2: torch.Tensor) -> torch.Tensor:
    """``F.layer_norm(x, (C,))`` without affine on bf16 rows ``[M, C]`` (the LayerNorm kernel with gamma = 1, beta = 0)."""
    C = x2.shape[1]
    one = torch.ones(C, dtype=torch.float32, device=x2.device)
    return ln(x2, Norm(one, torch.zeros_like(one), 1e-5))


class BertAlign(nn.Module):
    """``token_utils.py:95-112``: ``sigmoid(beta) * proj(layer_norm(x))`` with a zero-initialised ``proj`` and a scalar
    gate (``gate_init = -2``) — a light alignment of the BERT tokens only.  Same parameter names (``proj``, ``beta``)."""

    def __init__(self, dim: int = 768, gate_init: float = -2.0):
        super().__init__()
        self.proj = nn.Linear(dim, dim, bias=True)
        nn.init.zeros_(self.proj.weight)
        nn.init.zeros_(self.proj.bias)
        self.beta = nn.Parameter(torch.tensor(gate_init))

    @torch.no_grad()
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        if not x.is_cuda:
            raise capi.TopoError("BertAlign: x must be a CUDA tensor (no CPU forward)")
        B, T, C = x.shape
        lin = _lin(self.proj, x.device)
        y = gemm(_plain_ln(x.reshape(B * T, C).to(BF16).contiguous()), lin, alpha=float(torch.sigmoid(self.beta)))
        return y.view(B, T, C).float()


@torch.no_grad()
def build_cond_uc_safe_pad(seg, prompts, model, clip_textenc, bert_textenc, device, clip_style_text="map in swisstopo style",
                           gamma=1.0, w_clip=0.4, bert_align=None, start_idx=8):
    """``token_utils.py:115-158``: ``cond = w_clip * clip_ctx + gamma * tail(bert)``, ``uc = w_clip * clip_ctx`` — the CFG
    difference comes from BERT only.  ``clip_textenc`` is any object with ``get_learned_conditioning(list[str]) ->
    [B,77,768]`` (CLIP weights are outside this package); the BERT tokens are (optionally aligned,) layer-normed without
    affine and written into the 77-token window from ``start_idx``.  Returns ``(cond, uc)`` with the same hint list."""
    seg = seg.to(device=device, dtype=torch.float32)
    B = seg.size(0)
    clip_ctx = clip_textenc.get_learned_conditioning([clip_style_text] * B).to(device=device, dtype=torch.float32)
    bert_tok = bert_textenc.encode(prompts).to(device=device, dtype=torch.float32)
    if bert_align is not None:
        bert_tok = bert_align(bert_tok)
    Bb, Lb, C = bert_tok.shape
    bert_tok = _plain_ln(bert_tok.reshape(Bb * Lb, C).to(BF16).contiguous()).view(Bb, Lb, C).float()
    max_len = clip_ctx.size(1)
    tail = torch.zeros(B, max_len, clip_ctx.size(-1), device=device, dtype=torch.float32)
    end = min(max_len, start_idx + Lb)
    if end > start_idx:
        tail[:, start_idx:end, :] = bert_tok[:, :end - start_idx, :]
    cond_ctx = w_clip * clip_ctx + gamma * tail
    uc_ctx = w_clip * clip_ctx
    hints = [seg]
    return {"c_concat": hints, "c_crossattn": [cond_ctx]}, {"c_concat": hints, "c_crossattn": [uc_ctx]}
