"""Generates tests/golden/*.pt|json by IMPORTING the reference's own code (only possible in the build container,
where /root/reference exists).  The fixtures pin the oracle (and through it the CUDA path) to the reference:

  lora_linear.pt   — LoRA_modified/LoRA_utils.py:20-42 LoRALinear.forward on seeded inputs (fp32), incl. the qkvo
                     variant's scaling (Training_sdfusion_bert_lora_qkvo.py:149-171)
  lora_qkv.json    — LoRA_utils.lora_qkv applied to this repo's module tree: wrap counts and resulting state-dict keys
  stitch.pt        — BaseModel/stitchTiles.py:13-41 hstack/vstack order on a small tile stack (restated with numpy
                     because the script is top-level code that reads D:\\ paths)

  control_lora.pt  — LoRA_modified/attention_lora.py: the reference's own ``ControlLoRA`` three-conv encoder (:87-101) and the
                     control branch of ``CrossAttention_LoRA_Control`` (``process_control_states`` :61-71 and the
                     add-iff-shapes-match rule of ``forward`` :73-81) on seeded inputs.  The file imports upstream
                     ``ControlNet.ldm.modules.attention.CrossAttention`` (empty submodule here), so a STAND-IN base class
                     restating upstream's attention is injected for the import; what the fixture pins is the reference's
                     own code on top of it (encoder, to_control algebra, flatten order, scale, shape rule)
  drivers.pt       — TextEncoder_Finetuning/vis_metrics.py: the reference's own ``validate_step`` (:73-116) and
                     ``sample_preview`` (:407-452) executed against a deterministic mock model / recording sampler:
                     pins the driver logic (cond / uc layout, zeros uc, seeds and generator use, argument values,
                     decode + clamp, the eps-MSE average) — not the network arithmetic

Run:  python tests/golden/make_golden.py
"""
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference"


def _load(path, name):
    spec = importlib.util.spec_from_file_location(name, path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def main():
    ref = _load(os.path.join(REF, "LoRA_modified", "LoRA_utils.py"), "ref_lora_utils")
    ref_qkvo = None
    cases = []
    for i, (fin, fout, r, bias, alpha) in enumerate([(128, 64, 4, False, None), (128, 96, 16, False, None),
                                                      (192, 192, 8, True, None), (64, 96, 12, True, 24.0),
                                                      (128, 128, 32, False, None)]):
        torch.manual_seed(100 + i)
        base = torch.nn.Linear(fin, fout, bias=bias)
        m = ref.LoRALinear(base, r=r, alpha=alpha)
        torch.nn.init.normal_(m.B.weight, std=0.05)  # reference inits B to zero; non-zero makes the check meaningful
        x = torch.randn(3, 7, fin)
        with torch.no_grad():
            y = m(x)
        cases.append({"x": x, "y": y, "scaling": float(m.scaling), "r": r,
                      "sd": {k: v.clone() for k, v in m.state_dict().items()}})
    torch.save(cases, os.path.join(HERE, "lora_linear.pt"))

    from topo_b200 import ldm
    torch.manual_seed(0)
    unet = ldm.ControlledUnetModel(ldm.ModelConfig.tiny())
    counts = ref.lora_qkv(unet, r_q=8, r_kv=8)
    keys = sorted(k for k in unet.state_dict().keys() if ".attn" in k and ("to_q" in k or "to_k" in k or "to_v" in k))
    json.dump({"counts": counts, "n_keys": len(keys), "sample_keys": keys[:12],
               "lora_filter_keys": sum(1 for k in unet.state_dict() if ".A." in k or ".B." in k)},
              open(os.path.join(HERE, "lora_qkv.json"), "w"), indent=1)

    g = np.random.default_rng(5)
    tiles = g.integers(0, 256, size=(12, 8, 8, 3), dtype=np.uint8)
    rows = []
    for rI in range(3):   # stitchTiles.py: hstack 'cols' tiles into a row, vstack the rows
        rows.append(np.hstack([tiles[rI * 4 + c] for c in range(4)]))
    torch.save({"tiles": torch.from_numpy(tiles), "cols": 4, "mosaic": torch.from_numpy(np.vstack(rows))},
               os.path.join(HERE, "stitch.pt"))
    print("golden fixtures written to", HERE)


class MockLDM(torch.nn.Module):
    """Deterministic stand-in for ControlLDM used by BOTH the generator (under the reference's drivers) and the tests
    (under this repo's drivers): every method is a cheap closed form, so equal outputs mean equal driver logic."""
    num_timesteps = 1000

    def __init__(self):
        super().__init__()
        self.calls = []
        ac = torch.cumprod(1 - torch.linspace(0.00085 ** 0.5, 0.012 ** 0.5, 1000, dtype=torch.float64) ** 2, 0).float()
        self.register_buffer("alphas_cumprod", ac)
        self.control_scales, self.only_mid_control = [1.0] * 13, False

    def encode_first_stage(self, gt):
        return torch.nn.functional.avg_pool2d(gt, 8).repeat(1, 2, 1, 1)[:, :4]

    def get_first_stage_encoding(self, z):
        return 0.18215 * z

    def q_sample(self, z, t, noise=None):
        a = self.alphas_cumprod[t].view(-1, 1, 1, 1)
        return a.sqrt() * z + (1 - a).sqrt() * noise

    def apply_model(self, x, t, cond):
        self.calls.append(("apply_model", tuple(x.shape), t.tolist(), sorted(cond)))
        c, h = torch.cat(cond["c_crossattn"], 1), torch.cat(cond["c_concat"], 1)
        return 0.9 * x + 1e-3 * t.view(-1, 1, 1, 1).float() + c.mean(dim=(1, 2)).view(-1, 1, 1, 1) + h.mean(dim=(1, 2, 3)).view(-1, 1, 1, 1)

    def decode_first_stage(self, z):
        return 3.0 * torch.nn.functional.interpolate(z[:, :3], scale_factor=8, mode="nearest")


class MockEnc:
    def encode(self, prompts):
        g = torch.Generator().manual_seed(len(prompts) + sum(len(p) for p in prompts))
        return torch.randn(len(prompts), 77, 768, generator=g)


class RecordingSampler:
    """Stands in for ldm's DDIMSampler under the reference's ``sample_preview``: records every argument and returns the
    first RNG draw after the driver's ``manual_seed`` (the x_T ldm would draw), scaled."""
    last = None

    def __init__(self, model, **kw):
        self.model = model

    def sample(self, **kw):
        x = torch.randn((kw["batch_size"], *kw["shape"]))
        RecordingSampler.last = {k: (v if not isinstance(v, dict) else {n: [t.clone() for t in lst] for n, lst in v.items()})
                                 for k, v in kw.items()}
        RecordingSampler.last["same_hint_object"] = kw["conditioning"]["c_concat"][0] is kw["unconditional_conditioning"]["c_concat"][0]
        return 0.5 * x, {}


def driver_batch():
    g = torch.Generator().manual_seed(77)
    return {"seg": torch.rand(2, 3, 64, 64, generator=g), "gt": torch.rand(2, 3, 64, 64, generator=g) * 2 - 1,
            "prompt": ["Set Forest color to dark green.", "Hide roads."]}


def make_control_lora():
    import types
    import torch.nn as nn

    class UpstreamCrossAttention(nn.Module):   # stand-in for ControlNet/ldm/modules/attention.py::CrossAttention
        def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0.0):
            super().__init__()
            inner = heads * dim_head
            context_dim = context_dim or query_dim
            self.heads, self.scale = heads, dim_head ** -0.5
            self.to_q = nn.Linear(query_dim, inner, bias=False)
            self.to_k = nn.Linear(context_dim, inner, bias=False)
            self.to_v = nn.Linear(context_dim, inner, bias=False)
            # upstream: nn.Sequential(Linear, Dropout); the reference's wrapper only handles an indexable ModuleList
            # (attention_lora.py:42-45), so the stand-in exposes that form
            self.to_out = nn.ModuleList([nn.Linear(inner, query_dim), nn.Dropout(dropout)])

        def forward(self, x, context=None, mask=None):
            ctx = x if context is None else context
            q, k, v = self.to_q(x), self.to_k(ctx), self.to_v(ctx)
            b, n, c = q.shape
            h = self.heads
            q, k, v = (t.reshape(b, -1, h, c // h).permute(0, 2, 1, 3) for t in (q, k, v))
            o = torch.softmax(q.float() @ k.float().transpose(-1, -2) * self.scale, -1) @ v
            return self.to_out[1](self.to_out[0](o.permute(0, 2, 1, 3).reshape(b, n, c)))

    for name in ("ControlNet", "ControlNet.ldm", "ControlNet.ldm.modules"):
        sys.modules.setdefault(name, types.ModuleType(name))
    att = types.ModuleType("ControlNet.ldm.modules.attention")
    att.CrossAttention = UpstreamCrossAttention
    sys.modules["ControlNet.ldm.modules.attention"] = att
    ref = _load(os.path.join(REF, "LoRA_modified", "attention_lora.py"), "ref_attention_lora")

    torch.manual_seed(31)
    enc = ref.ControlLoRA()
    g = torch.Generator().manual_seed(32)
    hint = torch.rand(2, 3, 64, 64, generator=g)
    with torch.no_grad():
        enc_y = enc(hint)

    torch.manual_seed(33)
    attn = ref.CrossAttention_LoRA_Control(query_dim=128, context_dim=192, heads=2, dim_head=64, lora_rank=4, control_rank=6)
    for m in attn.modules():
        if isinstance(m, ref.LoRALinear):
            torch.nn.init.normal_(m.B.weight, std=0.05)
    torch.nn.init.normal_(attn.to_control[1].weight, std=0.2)
    B, H, W = 2, 4, 4
    x = torch.randn(B, H * W, 128, generator=g)
    ctx = torch.randn(B, 7, 192, generator=g)
    cs = torch.randn(B, 128, H, W, generator=g)
    with torch.no_grad():
        y_none = attn(x, ctx)
        attn.inject_control_states(cs)
        ctrl = attn.process_control_states(x, scale=0.7)
        y_ctrl = attn(x, ctx, scale=0.7)
        attn.inject_control_states(cs[:, :, :2])          # spatial mismatch: silently ignored (:77-80)
        y_bad = attn(x, ctx, scale=0.7)
    torch.save({"enc_sd": {k: v.clone() for k, v in enc.state_dict().items()}, "enc_x": hint, "enc_y": enc_y,
                "attn_sd": {k: v.clone() for k, v in attn.state_dict().items()}, "x": x, "ctx": ctx, "cs": cs, "scale": 0.7,
                "ctrl": ctrl, "y_none": y_none, "y_ctrl": y_ctrl, "y_bad_equals_none": bool(torch.equal(y_bad, y_none)),
                "scaling": float(attn.to_q.scaling)}, os.path.join(HERE, "control_lora.pt"))


def make_drivers():
    import types
    sys.path.insert(0, REF)
    for name in ("ControlNet", "ControlNet.ldm", "ControlNet.ldm.models", "ControlNet.ldm.models.diffusion"):
        sys.modules.setdefault(name, types.ModuleType(name))
    ddim = types.ModuleType("ControlNet.ldm.models.diffusion.ddim")
    ddim.DDIMSampler = RecordingSampler
    sys.modules["ControlNet.ldm.models.diffusion.ddim"] = ddim
    import TextEncoder_Finetuning.vis_metrics as vm
    batch = driver_batch()
    model = MockLDM()
    torch.manual_seed(5)
    val = vm.validate_step(model, MockEnc(), batch, "cpu", iters=3)
    calls = list(model.calls)
    t_fix = torch.tensor([700, 120])
    n_fix = torch.randn(2, 4, 8, 8, generator=torch.Generator().manual_seed(9))
    val_fixed = vm.validate_step(model, MockEnc(), batch, "cpu", iters=2, fixed_t=t_fix, fixed_noise=n_fix)
    pred = vm.sample_preview(model, MockEnc(), batch, "cpu", steps=12, scale=6.5, eta=0.0, seed=4321)
    torch.save({"validate": val, "validate_calls": calls, "validate_fixed": val_fixed, "fixed_t": t_fix, "fixed_noise": n_fix,
                "preview": pred, "sampler_kwargs": RecordingSampler.last}, os.path.join(HERE, "drivers.pt"))


if __name__ == "__main__":
    main()
    make_control_lora()
    make_drivers()
