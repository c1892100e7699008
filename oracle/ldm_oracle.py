"""ORACLE — test infrastructure, not product code.

CPU/fp32 PyTorch restatement of the reference's hot path: ``ControlLDM.apply_model`` (ControlNet +
ControlledUnetModel forward), the ldm ``DDIMSampler`` with batched classifier-free guidance and the
``AutoencoderKL`` decoder, with the LoRA forms the reference injects.  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs may import
this module; the product package must never do so.

PARITY UNPINNED for the UNet / ControlNet / DDIM / VAE arithmetic: it lives in two git submodules that are empty in
/root/reference (``ControlNet`` @ ed85cd1e25a5ed592f7d8178495b4483de0331bf and ``ctrlora`` @
8940c941a00cc45da1154fc25ace5a7368a3b3a0, ``.SUBMODULES.json:8-21``) and the reference ships no tests, golden vectors
or fixtures; those functions restate the published upstream algorithm and are
anchored on the reference's call sites, cited per function below.  PINNED against the reference's own code, by fixtures
generated from it in the build container (``tests/golden/make_golden.py``, checked by ``tests/test_oracle_golden.py``):

* ``lora_linear``                — ``LoRA_modified/LoRA_utils.py:20-42`` outputs (``lora_linear.pt``)
* ``lora_control_attention``     — the control branch, flatten order, scale and add-iff-shapes-match rule of
  ``LoRA_modified/attention_lora.py:57-81`` (``control_lora.pt``; the upstream base class it subclasses is a stand-in)
* ``control_lora_encoder``       — ``attention_lora.py:87-101`` outputs (``control_lora.pt``)
* ``stitch_tiles`` / ``make_schedule`` — ``BaseModel/stitchTiles.py:13-41`` order (``stitch.pt``), schedule known answers
* the drivers ``sample_preview`` / ``validate_step`` of the PRODUCT (``topo_b200.sampling``) are held to the reference's own
  ``TextEncoder_Finetuning/vis_metrics.py:73-116,407-452`` run on a mock model (``drivers.pt``)

The oracle is functional: it consumes a ``state_dict`` whose keys are the upstream key names (the
product's module tree in ``topo_b200.ldm`` produces exactly those), so the same weights can be fed
to both sides.  LoRA forms recognised on any linear ``p``:

* ``p.base.weight / p.A.weight / p.B.weight``  — ``LoRALinear`` (``LoRA_utils.py:20-42``),
  scale from ``scales[p]`` (default 2.0 = ``(alpha or 2r)/r``, ``LoRA_utils.py:38``)
* ``p.weight + p.lora_layer.down.weight / p.lora_layer.up.weight`` — ctrlora ``lora_layer``
  (``…_ctrlora.py:251``), scale from ``scales[p]`` (default 1.0)
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


# ----------------------------------------------------------------------------- primitives
def lora_linear(sd: SD, p: str, x: torch.Tensor, scales: Optional[Dict[str, float]] = None) -> torch.Tensor:
    """y = base(x) + B(A(x))*scaling — LoRA_modified/LoRA_utils.py:40-42 (a1); ctrlora lora_layer (u11)."""
    scales = scales or {}
    if p + ".base.weight" in sd:
        y = F.linear(x, sd[p + ".base.weight"], sd.get(p + ".base.bias"))
        s = scales.get(p, 2.0)
        return y + F.linear(F.linear(x, sd[p + ".A.weight"]), sd[p + ".B.weight"]) * s
    y = F.linear(x, sd[p + ".weight"], sd.get(p + ".bias"))
    if p + ".lora_layer.down.weight" in sd:
        s = scales.get(p, 1.0)
        y = y + s * F.linear(F.linear(x, sd[p + ".lora_layer.down.weight"]), sd[p + ".lora_layer.up.weight"])
    return y


def conv(sd: SD, p: str, x, stride=1, padding=0):
    return F.conv2d(x, sd[p + ".weight"], sd.get(p + ".bias"), stride=stride, padding=padding)


def group_norm(sd: SD, p: str, x, eps, groups=None):
    """GroupNorm32: nn.GroupNorm(32, C) evaluated on x.float() (upstream util.py)."""
    c = x.shape[1]
    g = groups or (32 if c % 32 == 0 else c // 4)
    return F.group_norm(x.float(), g, sd[p + ".weight"], sd[p + ".bias"], eps).to(x.dtype)


def timestep_embedding(t: torch.Tensor, dim: int) -> torch.Tensor:
    """[cos(t f), sin(t f)], f = exp(-ln(1e4) i / half) — upstream util.py (u8, K11)."""
    half = dim // 2
    freqs = torch.exp(-math.log(10000.0) * torch.arange(half, dtype=torch.float32) / half).to(t.device)
    args = t[:, None].float() * freqs[None]
    return torch.cat([torch.cos(args), torch.sin(args)], dim=-1)


# ----------------------------------------------------------------------------- attention
def attention_core(q, k, v, heads: int):
    """softmax(q k^T d^-1/2) v, fp32, explicit (non-xformers) branch — upstream attention.py (u6, K2/K3)."""
    b, n, c = q.shape
    d = c // heads
    q, k, v = (t.reshape(b, -1, heads, d).permute(0, 2, 1, 3).reshape(b * heads, -1, d) for t in (q, k, v))
    sim = torch.einsum("bid,bjd->bij", q.float(), k.float()) * (d ** -0.5)
    sim = sim.softmax(dim=-1)
    out = torch.einsum("bij,bjd->bid", sim, v)
    return out.reshape(b, heads, n, d).permute(0, 2, 1, 3).reshape(b, n, c)


def cross_attention(sd: SD, p: str, x, context, heads: int, scales=None):
    """CrossAttention.forward — to_q/to_k/to_v/to_out[0] called as modules (LoRA_utils.py:50-58)."""
    ctx = x if context is None else context
    q = lora_linear(sd, p + ".to_q", x, scales)
    k = lora_linear(sd, p + ".to_k", ctx, scales)
    v = lora_linear(sd, p + ".to_v", ctx, scales)
    o = attention_core(q, k, v, heads)
    return lora_linear(sd, p + ".to_out.0", o, scales)


def lora_control_attention(sd: SD, p: str, x, context, heads, control_states=None, scale=1.0, scales=None):
    """CrossAttention_LoRA_Control.forward — LoRA_modified/attention_lora.py:73-81 (a5).

    out = attention(x, ctx) with LoRA q/k/v/out; if control_states is set, add
    scale * to_control(flatten(control)) when shapes match (attention_lora.py:61-80).
    """
    out = cross_attention(sd, p, x, context, heads, scales)
    if control_states is not None:
        c = control_states
        if c.ndim == 4 and x.ndim == 3:
            B, C, H, W = c.shape
            c = c.permute(0, 2, 3, 1).reshape(B, H * W, C)
        ctrl = scale * F.linear(F.linear(c.to(x.dtype), sd[p + ".to_control.0.weight"]), sd[p + ".to_control.1.weight"])
        if ctrl.shape == out.shape:
            out = out + ctrl
    return out


def control_lora_encoder(sd: SD, p: str, x):
    """ControlLoRA.forward — LoRA_modified/attention_lora.py:87-101: conv3(3->64,s1) SiLU conv3(64->128,s2) SiLU
    conv3(128->256,s2) SiLU on the NCHW hint."""
    for i, s_ in ((0, 1), (2, 2), (4, 2)):
        x = F.silu(conv(sd, f"{p}.encoder.{i}", x, stride=s_, padding=1))
    return x


def basic_transformer_block(sd, p, x, ctx, heads, scales=None, ctrl=None):
    """x += attn1(LN x); x += attn2(LN x, ctx); x += ff(LN x) — upstream attention.py (u5).  ``ctrl`` maps an attention
    prefix to the control states injected into that ``CrossAttention_LoRA_Control`` (attention_lora.py:57-59)."""
    c = x.shape[-1]
    ln = lambda n, t: F.layer_norm(t, (c,), sd[f"{p}.{n}.weight"], sd[f"{p}.{n}.bias"], 1e-5)
    ctrl = ctrl or {}
    x = lora_control_attention(sd, p + ".attn1", ln("norm1", x), None, heads, ctrl.get(p + ".attn1"), 1.0, scales) + x
    x = lora_control_attention(sd, p + ".attn2", ln("norm2", x), ctx, heads, ctrl.get(p + ".attn2"), 1.0, scales) + x
    h = lora_linear(sd, p + ".ff.net.0.proj", ln("norm3", x), scales)
    a, g = h.chunk(2, dim=-1)
    x = lora_linear(sd, p + ".ff.net.2", a * F.gelu(g), scales) + x
    return x


def spatial_transformer(sd, p, x, ctx, heads, scales=None, ctrl=None):
    """GN(eps 1e-6) → 1x1 → tokens → block(s) → 1x1 → + x_in — upstream attention.py (u5)."""
    b, c, h, w = x.shape
    x_in = x
    x = group_norm(sd, p + ".norm", x, 1e-6)
    x = conv(sd, p + ".proj_in", x)
    x = x.permute(0, 2, 3, 1).reshape(b, h * w, -1)
    i = 0
    while f"{p}.transformer_blocks.{i}.norm1.weight" in sd:
        x = basic_transformer_block(sd, f"{p}.transformer_blocks.{i}", x, ctx, heads, scales, ctrl)
        i += 1
    x = x.reshape(b, h, w, -1).permute(0, 3, 1, 2)
    return conv(sd, p + ".proj_out", x) + x_in


def res_block(sd, p, x, emb):
    """ResBlock — upstream openaimodel.py (u4): conv3(SiLU(GN x)) + Linear(SiLU emb); conv3(SiLU(GN h)); + skip."""
    h = conv(sd, p + ".in_layers.2", F.silu(group_norm(sd, p + ".in_layers.0", x, 1e-5)), padding=1)
    e = F.linear(F.silu(emb), sd[p + ".emb_layers.1.weight"], sd[p + ".emb_layers.1.bias"])
    h = h + e[:, :, None, None]
    h = conv(sd, p + ".out_layers.3", F.silu(group_norm(sd, p + ".out_layers.0", h, 1e-5)), padding=1)
    if p + ".skip_connection.weight" in sd:
        x = conv(sd, p + ".skip_connection", x)
    return x + h


def _run_block(sd, p, h, emb, ctx, heads, scales, ctrl=None):
    """One TimestepEmbedSequential: dispatch on which keys exist under ``p.<i>``."""
    i = 0
    while True:
        q = f"{p}.{i}"
        if q + ".in_layers.0.weight" in sd:
            h = res_block(sd, q, h, emb)
        elif q + ".proj_in.weight" in sd:
            h = spatial_transformer(sd, q, h, ctx, heads, scales, ctrl)
        elif q + ".op.weight" in sd:
            h = conv(sd, q + ".op", h, stride=2, padding=1)
        elif q + ".conv.weight" in sd:
            h = F.interpolate(h, scale_factor=2, mode="nearest")
            h = conv(sd, q + ".conv", h, padding=1)
        elif q + ".weight" in sd and sd[q + ".weight"].ndim == 4:
            k = sd[q + ".weight"].shape[-1]
            h = conv(sd, q, h, padding=k // 2)
        else:
            break
        i += 1
    return h


def _time_embed(sd, p, t, model_channels):
    e = timestep_embedding(t, model_channels)
    e = F.linear(e, sd[p + ".time_embed.0.weight"], sd[p + ".time_embed.0.bias"])
    return F.linear(F.silu(e), sd[p + ".time_embed.2.weight"], sd[p + ".time_embed.2.bias"])


def _count(sd, prefix):
    n = 0
    while any(k.startswith(f"{prefix}.{n}.") for k in sd):
        n += 1
    return n


def hint_block(sd, p, hint):
    """input_hint_block: 8 convs, SiLU between, strides 1,1,2,1,2,1,2,1 (vis_metrics.py:368-369)."""
    strides = [1, 1, 2, 1, 2, 1, 2, 1]
    g = hint
    for j, s in enumerate(strides):
        g = conv(sd, f"{p}.input_hint_block.{2 * j}", g, stride=s, padding=1)
        if j != len(strides) - 1:
            g = F.silu(g)
    return g


def controlnet_forward(sd, x, hint, t, ctx, heads, scales=None, p="control_model", ctrl=None) -> List[torch.Tensor]:
    """ControlNet.forward — upstream cldm/cldm.py (u2): 13 zero-conv outputs."""
    mc = sd[p + ".time_embed.0.weight"].shape[1]
    emb = _time_embed(sd, p, t, mc)
    g = hint_block(sd, p, hint)
    outs = []
    h = x
    for i in range(_count(sd, p + ".input_blocks")):
        h = _run_block(sd, f"{p}.input_blocks.{i}", h, emb, ctx, heads, scales, ctrl)
        if g is not None:
            h = h + g
            g = None
        outs.append(conv(sd, f"{p}.zero_convs.{i}.0", h))
    h = _run_block(sd, p + ".middle_block", h, emb, ctx, heads, scales, ctrl)
    outs.append(conv(sd, p + ".middle_block_out.0", h))
    return outs


def unet_forward(sd, x, t, ctx, control, heads, only_mid_control=False, scales=None,
                 p="model.diffusion_model", ctrl=None):
    """ControlledUnetModel.forward — upstream cldm/cldm.py (u3)."""
    mc = sd[p + ".time_embed.0.weight"].shape[1]
    emb = _time_embed(sd, p, t, mc)
    control = list(control) if control is not None else None
    hs = []
    h = x
    for i in range(_count(sd, p + ".input_blocks")):
        h = _run_block(sd, f"{p}.input_blocks.{i}", h, emb, ctx, heads, scales, ctrl)
        hs.append(h)
    h = _run_block(sd, p + ".middle_block", h, emb, ctx, heads, scales, ctrl)
    if control is not None:
        h = h + control.pop()
    for i in range(_count(sd, p + ".output_blocks")):
        if only_mid_control or control is None:
            h = torch.cat([h, hs.pop()], dim=1)
        else:
            h = torch.cat([h, hs.pop() + control.pop()], dim=1)
        h = _run_block(sd, f"{p}.output_blocks.{i}", h, emb, ctx, heads, scales, ctrl)
    h = F.silu(group_norm(sd, p + ".out.0", h, 1e-5))
    return conv(sd, p + ".out.2", h, padding=1)


def apply_model(sd, x_noisy, t, cond, heads=8, control_scales=None, only_mid_control=False, scales=None, ctrl=None):
    """ControlLDM.apply_model — upstream cldm/cldm.py (u1); call sites …_ctrlora.py:388, vis_metrics.py:112.
    ``ctrl``: {attention prefix: control_states} for modules replaced by inject_control_lora_into_ldm (a5/a6)."""
    ctx = torch.cat(cond["c_crossattn"], 1)
    if cond.get("c_concat") is None:
        control = None
    else:
        hint = torch.cat(cond["c_concat"], 1)
        control = controlnet_forward(sd, x_noisy, hint, t, ctx, heads, scales, ctrl=ctrl)
        cs = control_scales or [1.0] * len(control)
        control = [c * s for c, s in zip(control, cs)]
    return unet_forward(sd, x_noisy, t, ctx, control, heads, only_mid_control, scales, ctrl=ctrl)


# ----------------------------------------------------------------------------- DDIM
def make_schedule(S: int, T: int = 1000, linear_start=0.00085, linear_end=0.0120, eta=0.0):
    """ldm make_ddim_timesteps('uniform') + make_ddim_sampling_parameters.

    c = T // S; ts = arange(0, T, c) + 1  (S=30 → 31 entries); a_prev = [ᾱ[0]] + ᾱ[ts[:-1]].
    """
    betas = torch.linspace(linear_start ** 0.5, linear_end ** 0.5, T, dtype=torch.float64) ** 2
    ac = torch.cumprod(1.0 - betas, 0).float()
    c = T // S
    ts = torch.arange(0, T, c) + 1
    a = ac[ts]
    a_prev = torch.cat([ac[:1], ac[ts[:-1]]])
    sig = eta * torch.sqrt((1 - a_prev) / (1 - a) * (1 - a / a_prev))
    return ts, a, a_prev, sig


def ddim_sample(sd, S, cond, shape, scale=1.0, uc=None, eta=0.0, x_T=None, heads=8, seed=None,
                scales=None, hacked=False, return_intermediates=True, control_scales=None, ctrl=None):
    """ldm DDIMSampler.sample / p_sample_ddim as driven by vis_metrics.py:430-447 (u9, K10).

    CFG batches [uncond ‖ cond] in ONE apply_model call (uncond first); ``hacked=True`` is the
    cldm/ddim_hacked.py variant used by BaseModel/evaluateCN.py:93-96 (two calls).
    """
    ts, a, a_prev, sig = make_schedule(S, eta=eta)
    if seed is not None:
        torch.manual_seed(seed)
    x = torch.randn(shape) if x_T is None else x_T.clone()
    B = x.shape[0]
    a, a_prev, sig = a.to(x.device), a_prev.to(x.device), sig.to(x.device)
    inter = []
    for i in reversed(range(len(ts))):
        t = torch.full((B,), int(ts[i]), dtype=torch.long, device=x.device)
        if uc is None or scale == 1.0:
            e = apply_model(sd, x, t, cond, heads, control_scales, scales=scales, ctrl=ctrl)
        elif hacked:
            e_c = apply_model(sd, x, t, cond, heads, control_scales, scales=scales, ctrl=ctrl)
            e_u = apply_model(sd, x, t, uc, heads, control_scales, scales=scales, ctrl=ctrl)
            e = e_u + scale * (e_c - e_u)
        else:
            c_in = {k: [torch.cat([uc[k][j], cond[k][j]]) for j in range(len(cond[k]))] for k in cond}
            e_u, e_c = apply_model(sd, torch.cat([x] * 2), torch.cat([t] * 2), c_in, heads, control_scales,
                                   scales=scales, ctrl=ctrl).chunk(2)
            e = e_u + scale * (e_c - e_u)
        pred_x0 = (x - (1 - a[i]).sqrt() * e) / a[i].sqrt()
        dir_xt = (1.0 - a_prev[i] - sig[i] ** 2).sqrt() * e
        noise = sig[i] * torch.randn_like(x)
        x = a_prev[i].sqrt() * pred_x0 + dir_xt + noise
        if return_intermediates:
            inter.append(x.clone())
    return x, inter


# ----------------------------------------------------------------------------- VAE decoder
def _vae_res(sd, p, x):
    h = conv(sd, p + ".conv1", F.silu(group_norm(sd, p + ".norm1", x, 1e-6)), padding=1)
    h = conv(sd, p + ".conv2", F.silu(group_norm(sd, p + ".norm2", h, 1e-6)), padding=1)
    if p + ".nin_shortcut.weight" in sd:
        x = conv(sd, p + ".nin_shortcut", x)
    return x + h


def _vae_attn(sd, p, x):
    b, c, hh, ww = x.shape
    h = group_norm(sd, p + ".norm", x, 1e-6)
    q = conv(sd, p + ".q", h).reshape(b, c, -1).permute(0, 2, 1)
    k = conv(sd, p + ".k", h).reshape(b, c, -1)
    v = conv(sd, p + ".v", h).reshape(b, c, -1)
    w = torch.bmm(q, k) * (c ** -0.5)
    w = w.softmax(dim=2)
    h = torch.bmm(v, w.permute(0, 2, 1)).reshape(b, c, hh, ww)
    return x + conv(sd, p + ".proj_out", h)


def decode_first_stage(sd, z, scale_factor=0.18215, p="first_stage_model"):
    """AutoencoderKL.decode(z / scale_factor) — upstream (u10); call site vis_metrics.py:450."""
    z = z / scale_factor
    z = conv(sd, p + ".post_quant_conv", z)
    d = p + ".decoder"
    h = conv(sd, d + ".conv_in", z, padding=1)
    h = _vae_res(sd, d + ".mid.block_1", h)
    h = _vae_attn(sd, d + ".mid.attn_1", h)
    h = _vae_res(sd, d + ".mid.block_2", h)
    nlev = _count(sd, d + ".up")
    for lvl in reversed(range(nlev)):
        j = 0
        while f"{d}.up.{lvl}.block.{j}.norm1.weight" in sd:
            h = _vae_res(sd, f"{d}.up.{lvl}.block.{j}", h)
            j += 1
        if f"{d}.up.{lvl}.upsample.conv.weight" in sd:
            h = F.interpolate(h, scale_factor=2.0, mode="nearest")
            h = conv(sd, f"{d}.up.{lvl}.upsample.conv", h, padding=1)
    h = F.silu(group_norm(sd, d + ".norm_out", h, 1e-6))
    return conv(sd, d + ".conv_out", h, padding=1)


def encode_first_stage(sd, x, p="first_stage_model"):
    """AutoencoderKL.encode(x).parameters — upstream ldm Encoder (ch 128, mult 1/2/4/4, 2 res blocks, Downsample with
    pad (0,1,0,1) + stride-2 conv, mid res/attn/res, GN+SiLU, conv_out -> 2z) then quant_conv; call site
    TextEncoder_Finetuning/vis_metrics.py:86.  Returns the moments [B, 2z, H/8, W/8] (mean | logvar)."""
    e = p + ".encoder"
    h = conv(sd, e + ".conv_in", x, padding=1)
    nlev = _count(sd, e + ".down")
    for lvl in range(nlev):
        j = 0
        while f"{e}.down.{lvl}.block.{j}.norm1.weight" in sd:
            h = _vae_res(sd, f"{e}.down.{lvl}.block.{j}", h)
            j += 1
        if f"{e}.down.{lvl}.downsample.conv.weight" in sd:
            h = F.pad(h, (0, 1, 0, 1), mode="constant", value=0)
            h = conv(sd, f"{e}.down.{lvl}.downsample.conv", h, stride=2)
    h = _vae_res(sd, e + ".mid.block_1", h)
    h = _vae_attn(sd, e + ".mid.attn_1", h)
    h = _vae_res(sd, e + ".mid.block_2", h)
    h = F.silu(group_norm(sd, e + ".norm_out", h, 1e-6))
    h = conv(sd, e + ".conv_out", h, padding=1)
    return conv(sd, p + ".quant_conv", h)


def get_first_stage_encoding(moments, noise=None, scale_factor=0.18215):
    """scale_factor * DiagonalGaussianDistribution(moments).sample() (logvar clamped to [-30, 20]); ``noise=None`` -> mode."""
    mean, logvar = torch.chunk(moments, 2, dim=1)
    std = torch.exp(0.5 * torch.clamp(logvar, -30.0, 20.0))
    return scale_factor * (mean if noise is None else mean + std * noise)


def validate_step(sd, z, c_cross, seg, ts, noises, heads=8, scales=None, T=1000, linear_start=0.00085, linear_end=0.0120):
    """vis_metrics.py:73-116 with the encoders factored out: mean over the iterations of
    mse(apply_model(q_sample(z, t, noise), t, cond), noise); q_sample = sqrt(ac[t]) z + sqrt(1 - ac[t]) noise."""
    betas = torch.linspace(linear_start ** 0.5, linear_end ** 0.5, T, dtype=torch.float64) ** 2
    ac = torch.cumprod(1.0 - betas, 0).float().to(z.device)
    cond = {"c_crossattn": [c_cross], "c_concat": [seg]}
    total = 0.0
    for t, n in zip(ts, noises):
        a = ac[t].view(-1, 1, 1, 1)
        zt = a.sqrt() * z + (1 - a).sqrt() * n
        eps = apply_model(sd, zt, t, cond, heads=heads, scales=scales)
        total += F.mse_loss(eps.float(), n.float()).item()
    return total / len(ts)


def sample_preview(sd, c_cross, seg, steps=30, scale=7.5, eta=0.0, seed=1234, x_T=None, heads=8, scales=None):
    """TextEncoder_Finetuning/vis_metrics.py:407-452 (a12) with the text encoder factored out:
    uc = zeros_like(c) (:425), same hint for both branches (:427-428), seed before x_T (:431),
    decode + clamp (:450-451)."""
    cond = {"c_crossattn": [c_cross], "c_concat": [seg]}
    uc = {"c_crossattn": [torch.zeros_like(c_cross)], "c_concat": [seg]}
    B, _, H, W = seg.shape
    x, inter = ddim_sample(sd, steps, cond, (B, 4, H // 8, W // 8), scale, uc, eta, x_T, heads, seed, scales)
    img = decode_first_stage(sd, x.float()).clamp(-1.0, 1.0)
    return img, x, inter


def tiles_to_uint8(x_samples: torch.Tensor):
    """BaseModel/evaluateCN.py:101-103 (a16): [-1,1] BCHW → uint8 BHWC via *127.5+127.5, clip, truncate."""
    return (x_samples.permute(0, 2, 3, 1) * 127.5 + 127.5).cpu().numpy().clip(0, 255).astype("uint8")


def stitch_tiles(tiles, cols: int):
    """BaseModel/stitchTiles.py:6-41 (a16): row-major hstack/vstack mosaic of [T,H,W,3] uint8."""
    import numpy as np
    rows = [np.hstack(list(tiles[r * cols:(r + 1) * cols])) for r in range(len(tiles) // cols)]
    return np.vstack(rows)
