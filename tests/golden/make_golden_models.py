"""Generates tests/golden/models_ref.pt by running the reference's own ``LoRA_modified/models.py`` (processor algebra
:70-431 and the ControlLoRA pyramid :434-833) in the build container.

The reference module imports ``diffusers`` (0.12-0.14 API, ``models.py:7-12``), which is not installed here.  The handful
of diffusers primitives it touches are provided as STAND-INS below — ``LoRALinearLayer``, the ``CrossAttention`` host
module (``to_q/to_k/to_v/to_out``, ``head_to_batch_dim``, ``get_attention_scores`` …), ``Downsample2D``, ``ModelMixin`` /
``ConfigMixin`` / ``register_to_config`` / ``BaseOutput`` — restating diffusers 0.12's published definitions.  What the
fixture pins is everything the REFERENCE wrote on top of them: which tensor feeds which LoRA (``post_add``), the skip
flags, where the control states enter (v1: the q-LoRA input; V2: the hidden states and the attention output), the
pre / post LoRA stacks and their missing ``scale`` on the value path, ``concat_hidden`` / ``control_self_add`` / batch
repeat, and the pyramid's block wiring.  oracle/models_oracle.py (CPU) and topo_b200/models.py (CUDA) are both checked
against these outputs (tests/test_oracle_golden.py, tests/test_gpu_models.py).

Run:  python tests/golden/make_golden_models.py
"""
import importlib.util
import os
import sys
import types

import torch
import torch.nn as nn
import torch.nn.functional as F

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
C, HEADS, D, CTX, HW, B = 128, 2, 64, 192, 8, 2


# ---------------------------------------------------------------------------------------------------------------------
# diffusers 0.12 stand-ins
# ---------------------------------------------------------------------------------------------------------------------
class LoRALinearLayer(nn.Module):
    def __init__(self, in_features, out_features, rank=4):
        super().__init__()
        if rank > min(in_features, out_features):
            raise ValueError(f"LoRA rank {rank} must be less or equal than {min(in_features, out_features)}")
        self.down = nn.Linear(in_features, rank, bias=False)
        self.up = nn.Linear(rank, out_features, bias=False)
        nn.init.normal_(self.down.weight, std=1 / rank)
        nn.init.zeros_(self.up.weight)

    def forward(self, hidden_states):
        orig = hidden_states.dtype
        return self.up(self.down(hidden_states.to(self.down.weight.dtype))).to(orig)


class CrossAttention(nn.Module):
    def __init__(self, query_dim, cross_attention_dim=None, heads=8, dim_head=64, dropout=0.0, bias=False):
        super().__init__()
        inner = dim_head * heads
        cross_attention_dim = cross_attention_dim if cross_attention_dim is not None else query_dim
        self.scale, self.heads = dim_head ** -0.5, heads
        self.to_q = nn.Linear(query_dim, inner, bias=bias)
        self.to_k = nn.Linear(cross_attention_dim, inner, bias=bias)
        self.to_v = nn.Linear(cross_attention_dim, inner, bias=bias)
        self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(dropout)])

    def batch_to_head_dim(self, t):
        h = self.heads
        b, s, d = t.shape
        return t.reshape(b // h, h, s, d).permute(0, 2, 1, 3).reshape(b // h, s, d * h)

    def head_to_batch_dim(self, t):
        h = self.heads
        b, s, d = t.shape
        return t.reshape(b, s, h, d // h).permute(0, 2, 1, 3).reshape(b * h, s, d // h)

    def get_attention_scores(self, query, key, attention_mask=None):
        empty = torch.empty(query.shape[0], query.shape[1], key.shape[1], dtype=query.dtype, device=query.device)
        scores = torch.baddbmm(empty, query, key.transpose(-1, -2), beta=0, alpha=self.scale)
        return scores.softmax(dim=-1).to(query.dtype)

    def prepare_attention_mask(self, attention_mask, target_length, batch_size=None):
        assert attention_mask is None
        return None


class Downsample2D(nn.Module):
    def __init__(self, channels, use_conv=False, out_channels=None, padding=1, name="conv"):
        super().__init__()
        self.channels, self.out_channels = channels, out_channels or channels
        self.use_conv, self.padding, self.name = use_conv, padding, name
        conv = nn.Conv2d(channels, self.out_channels, 3, stride=2, padding=padding) if use_conv else nn.AvgPool2d(2, 2)
        if name == "conv":
            self.Conv2d_0 = conv
        self.conv = conv

    def forward(self, hidden_states):
        if self.use_conv and self.padding == 0:
            hidden_states = F.pad(hidden_states, (0, 1, 0, 1), mode="constant", value=0)
        return self.conv(hidden_states)


def install_stubs():
    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    class BaseOutput(dict):
        pass

    class ConfigMixin:
        pass

    class ModelMixin(nn.Module):
        pass

    def unavailable(*a, **k):
        raise NotImplementedError("not needed by the fixture")

    mod("diffusers")
    mod("diffusers.utils")
    mod("diffusers.utils.outputs", BaseOutput=BaseOutput)
    mod("diffusers.configuration_utils", ConfigMixin=ConfigMixin, register_to_config=lambda f: f)
    mod("diffusers.models")
    mod("diffusers.models.modeling_utils", ModelMixin=ModelMixin)
    mod("diffusers.models.unet_2d_blocks", get_down_block=unavailable)
    import functools
    mod("diffusers.models.resnet", Mish=nn.Mish, Upsample2D=unavailable, Downsample2D=Downsample2D, upsample_2d=unavailable,
        downsample_2d=unavailable, partial=functools.partial)
    mod("diffusers.models.cross_attention", CrossAttention=CrossAttention, LoRALinearLayer=LoRALinearLayer)


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def randomize(mod_, seed):
    g = torch.Generator().manual_seed(seed)
    with torch.no_grad():
        for name, p in mod_.named_parameters():
            if name.endswith("up.weight"):
                p.copy_(torch.randn(p.shape, generator=g) * 0.05)
            elif name.endswith("down.weight"):
                p.copy_(torch.randn(p.shape, generator=g) * p.shape[1] ** -0.5)
    return mod_


# every case: (name, class name, constructor kwargs, cross?, call scale, control channels or None, extras)
CASES = [
    ("lora_self", "LoRACrossAttnProcessor", dict(rank=4), False, 1.0, None, {}),
    ("lora_cross_postadd", "LoRACrossAttnProcessor", dict(rank=8, post_add=True), True, 0.5, None, {}),
    ("lora_cross_skips", "LoRACrossAttnProcessor", dict(rank=4, key_states_skipped=True, value_states_skipped=True,
                                                       output_states_skipped=True), True, 1.0, None, {}),
    ("ctrl_v1_self", "ControlLoRACrossAttnProcessor", dict(rank=4, control_rank=6), False, 1.0, C, {}),
    ("ctrl_v1_cross_concat", "ControlLoRACrossAttnProcessor", dict(rank=4, concat_hidden=True, control_channels=64), True, 0.7, 64, {}),
    ("ctrl_v1_stacks", "ControlLoRACrossAttnProcessor", dict(rank=4), True, 0.8, C, {"stacks": True}),
    ("ctrl_v1_batch_repeat", "ControlLoRACrossAttnProcessor", dict(rank=4, concat_hidden=True, control_channels=64), False, 1.0, 64,
     {"control_batch": 1}),
    ("ctrl_v2_self", "ControlLoRACrossAttnProcessorV2", dict(rank=4, control_rank=6), False, 1.0, C, {}),
    ("ctrl_v2_cross_concat", "ControlLoRACrossAttnProcessorV2", dict(rank=4, concat_hidden=True, control_channels=64), True, 0.6, 64, {}),
]
PYRAMID = dict(block_out_channels=(32, 32, 64, 64), lora_block_in_channels=(64, 64, 64, 64),
               lora_block_out_channels=(64, 128, 128, 128),
               lora_cross_attention_dims=([None, CTX], [None, CTX], [None, CTX], [None, CTX]), lora_rank=4)


def main():
    install_stubs()
    ref = _load(os.path.join(REF, "LoRA_modified", "models.py"), "ref_models")
    out = {"dims": dict(C=C, heads=HEADS, d=D, ctx=CTX, hw=HW, B=B), "cases": {}}
    attns = {}
    for cross in (False, True):
        torch.manual_seed(10 + cross)
        attns[cross] = CrossAttention(C, CTX if cross else None, HEADS, D).eval()
        out["attn_cross" if cross else "attn_self"] = {k: v.clone() for k, v in attns[cross].state_dict().items()}
    g = torch.Generator().manual_seed(1)
    hidden = torch.randn(B, HW * HW, C, generator=g)
    enc = torch.randn(B, 77, CTX, generator=g)
    out["hidden"], out["enc"] = hidden, enc
    for i, (name, cls, kw, cross, scale, cch, extra) in enumerate(CASES):
        torch.manual_seed(100 + i)
        proc = randomize(getattr(ref, cls)(C, cross_attention_dim=CTX if cross else None, **kw), 200 + i)
        rec = {"cls": cls, "kw": kw, "cross": cross, "scale": scale, "sd": {k: v.clone() for k, v in proc.state_dict().items()}}
        if cch is not None:
            cs = torch.randn(extra.get("control_batch", B), cch, HW, HW, generator=g)
            proc.inject_control_states(cs.clone())
            rec["control"] = cs
        if extra.get("stacks"):
            pre = randomize(ref.LoRACrossAttnProcessor(C, CTX, rank=4), 300 + i)
            post = randomize(ref.LoRACrossAttnProcessor(C, CTX, rank=8, post_add=True), 400 + i)
            proc.inject_pre_lora(pre)
            proc.inject_post_lora(post)
            rec["pre_sd"] = {k: v.clone() for k, v in pre.state_dict().items()}
            rec["post_sd"] = {k: v.clone() for k, v in post.state_dict().items()}
        with torch.no_grad():
            rec["out"] = proc(attns[cross], hidden, enc if cross else None, None, scale)
        out["cases"][name] = rec
    torch.manual_seed(7)
    pyr = randomize(ref.ControlLoRA(**PYRAMID), 8).eval()
    x = torch.rand(B, 3, 64, 64, generator=g)
    with torch.no_grad():
        maps = pyr(x, return_dict=False)
        none = pyr(x)                     # the reference's forward falls off its end with return_dict=True (models.py:832)
    out["pyramid"] = {"kw": PYRAMID, "sd": {k: v.clone() for k, v in pyr.state_dict().items()}, "x": x,
                      "maps": [m.clone() for m in maps], "return_dict_true_is_none": none is None,
                      "injected_shapes": [tuple(l[0].control_states.shape) for l in pyr.lora_layers]}
    torch.save(out, os.path.join(HERE, "models_ref.pt"))
    size = os.path.getsize(os.path.join(HERE, "models_ref.pt"))
    print(f"models_ref.pt: {len(out['cases'])} processor cases + pyramid {[tuple(m.shape) for m in maps]}, {size / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
