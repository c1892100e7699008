"""Module tree of the ControlLDM (SD-1.5 UNet + ControlNet + VAE decoder) the hot path runs.

This file is the *Python face* of the drop-in boundary: the attribute and
parameter names the reference's scripts touch by duck-typing are kept name-for-name
(``model.model.diffusion_model``, ``model.control_model``, ``input_blocks / middle_block /
output_blocks[i][1].transformer_blocks[j].{attn1,attn2,ff,norm1..3}``, ``to_q/to_k/to_v`` as
bias-free ``nn.Linear``, ``to_out`` as ``[Linear, Dropout]``, ``zero_convs``,
``middle_block_out``, ``lora_layer``) so that

* the reference's surgery functions (``LoRA_modified/LoRA_utils.py:45-59`` ``lora_qkv``,
  ``TextEncoder_Finetuning/Training_sdfusion_bert_lora_qkvo.py:175-187``) mutate it the same way,
* state-dict keys equal the upstream lllyasviel/ControlNet @ ed85cd1 keys, so trained
  checkpoints load with ``strict=False`` as in ``LoRA_modified/quantitative_metrics.py:44-55``.

The modules here are parameter containers.  They hold no arithmetic: the forward pass of the
whole model is executed by ``engine.TopoEngine`` (hand-written sm_100a kernels behind the C ABI in
``include/topo_b200.h``).  There is deliberately no PyTorch/CPU forward in this package — the
fp32 restatement used for checking lives in ``oracle/`` and is never imported from here.

Architecture constants follow upstream ``models/cldm_v15.yaml``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import torch
import torch.nn as nn


# --------------------------------------------------------------------------------------
# configuration (the reference's three script-level variants become data)
# --------------------------------------------------------------------------------------
@dataclass
class ModelConfig:
    in_channels: int = 4
    out_channels: int = 4
    hint_channels: int = 3
    model_channels: int = 320
    channel_mult: Sequence[int] = (1, 2, 4, 4)
    num_res_blocks: int = 2
    attention_resolutions: Sequence[int] = (4, 2, 1)
    num_heads: int = 8
    context_dim: int = 768
    transformer_depth: int = 1
    # ctrlora-style low rank branches on the ControlNet linears (0 = stock ControlNet)
    control_lora_rank: int = 0
    control_lora_scale: float = 1.0
    control_lora_targets: Sequence[str] = ("to_q", "to_k", "to_v", "to_out.0")
    # DDPM schedule (cldm_v15.yaml)
    timesteps: int = 1000
    linear_start: float = 0.00085
    linear_end: float = 0.0120
    scale_factor: float = 0.18215
    # VAE decoder
    vae_ch: int = 128
    vae_ch_mult: Sequence[int] = (1, 2, 4, 4)
    vae_num_res_blocks: int = 2
    vae_z_channels: int = 4
    with_vae: bool = True

    @staticmethod
    def tiny() -> "ModelConfig":
        """Structurally identical, 10x narrower model for fast CPU tests."""
        return ModelConfig(model_channels=32, num_heads=2, context_dim=64,
                           vae_ch=32, vae_ch_mult=(1, 2, 2, 2))


VARIANTS = {
    # name: (control_lora_rank, control_lora_scale, unet surgery)
    # LoRA_modified/multiseg_19prompts_batch1_shuffle_onlyCNLora.py:210,244-248 (BASELINE cfg 1: r=4)
    "onlyCNLora": dict(control_lora_rank=4, unet=None),
    # LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:216,250-258 (BASELINE cfg 2: r=16)
    "ctrlora": dict(control_lora_rank=16, unet=None),
    # LoRA_modified/multiseg_19prompts_batch1_shuffle_addlrSD.py:281,316 (CN r=32 + lora_qkv(8,8))
    "addlrSD": dict(control_lora_rank=32, unet=("qkv", 8, 8)),
    # TextEncoder_Finetuning/Training_sdfusion_bert_lora_qkvo.py:175-187 (stock CN + qkvo r=8, s=4)
    "qkvo": dict(control_lora_rank=0, unet=("qkvo", 8, 8)),
    # BaseModel/evaluateCN.py:24 (stock cldm_v15)
    "stock": dict(control_lora_rank=0, unet=None),
}


class _NoForward(nn.Module):
    """Base for container modules: forward is executed by the engine, not by PyTorch."""

    def forward(self, *a, **k):  # pragma: no cover - loud by design
        raise RuntimeError(
            f"{type(self).__name__}.forward: this package has no PyTorch/CPU forward. "
            "Run the model through ControlLDM.apply_model on a CUDA (sm_100a) device; "
            "the fp32 checker lives in oracle/ and is test infrastructure only.")


def zero_module(m: nn.Module) -> nn.Module:
    with torch.no_grad():
        for p in m.parameters():
            p.zero_()
    return m


# --------------------------------------------------------------------------------------
# LoRA-capable linear of the ctrlora ControlNet
# --------------------------------------------------------------------------------------
class LoRALinearLayer(nn.Module):
    """``up(down(x))`` low-rank branch; parameter path ``<linear>.lora_layer.{down,up}.weight``
    (the substring ``lora_layer`` is what the reference's trainable-parameter filter keys on,
    ``LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:251``)."""

    def __init__(self, in_features: int, out_features: int, rank: int):
        super().__init__()
        self.down = nn.Linear(in_features, rank, bias=False)
        self.up = nn.Linear(rank, out_features, bias=False)
        nn.init.normal_(self.down.weight, std=1.0 / rank)
        nn.init.zeros_(self.up.weight)
        self.rank = rank


class LoRACompatibleLinear(nn.Linear):
    """``y = W x (+b) + scale * up(down(x))`` when ``lora_layer`` is set, else a plain Linear."""

    def __init__(self, *args, lora_rank: int = 0, lora_scale: float = 1.0, **kwargs):
        super().__init__(*args, **kwargs)
        self.lora_layer: Optional[LoRALinearLayer] = None
        self.scale = lora_scale
        if lora_rank > 0:
            self.lora_layer = LoRALinearLayer(self.in_features, self.out_features, lora_rank)

    def forward(self, x):
        from .ops import lora_linear_module
        return lora_linear_module(self, x)


# --------------------------------------------------------------------------------------
# attention / transformer containers (upstream ldm/modules/attention.py names)
# --------------------------------------------------------------------------------------
class CrossAttention(nn.Module):
    """Carries ``query_dim, heads, dim_head, dropout, context_dim`` because the reference's
    injector reads them (``LoRA_modified/attention_lora.py:112-116``)."""

    def __init__(self, query_dim, context_dim=None, heads=8, dim_head=64, dropout=0.0, bias=False,
                 lora_rank=0, lora_scale=1.0, lora_targets=()):
        super().__init__()
        inner = heads * dim_head
        self.query_dim, self.heads, self.dim_head, self.dropout = query_dim, heads, dim_head, dropout
        self.context_dim = context_dim
        ctx = context_dim if context_dim is not None else query_dim
        self.scale = dim_head ** -0.5

        def lin(i, o, b, name):
            r = lora_rank if name in lora_targets else 0
            if lora_rank > 0:
                return LoRACompatibleLinear(i, o, bias=b, lora_rank=r, lora_scale=lora_scale)
            return nn.Linear(i, o, bias=b)

        self.to_q = lin(query_dim, inner, bias, "to_q")
        self.to_k = lin(ctx, inner, bias, "to_k")
        self.to_v = lin(ctx, inner, bias, "to_v")
        self.to_out = nn.Sequential(lin(inner, query_dim, True, "to_out.0"), nn.Dropout(dropout))

    def forward(self, x, context=None, mask=None):
        """Stand-alone attention on CUDA tensors ``x [B,N,C]`` (row a5/u6 of SURVEY.md §8a):
        calls ``self.to_q(x)`` etc. as modules, so any replacement module changes the result."""
        from .ops import cross_attention_module
        return cross_attention_module(self, x, context, mask)


class GEGLU(_NoForward):
    def __init__(self, dim_in, dim_out):
        super().__init__()
        self.proj = nn.Linear(dim_in, dim_out * 2)


class FeedForward(_NoForward):
    def __init__(self, dim, mult=4, dropout=0.0):
        super().__init__()
        inner = dim * mult
        self.net = nn.Sequential(GEGLU(dim, inner), nn.Dropout(dropout), nn.Linear(inner, dim))


class BasicTransformerBlock(_NoForward):
    def __init__(self, dim, n_heads, d_head, context_dim, **lora):
        super().__init__()
        self.attn1 = CrossAttention(dim, None, n_heads, d_head, **lora)
        self.ff = FeedForward(dim)
        self.attn2 = CrossAttention(dim, context_dim, n_heads, d_head, **lora)
        self.norm1 = nn.LayerNorm(dim)
        self.norm2 = nn.LayerNorm(dim)
        self.norm3 = nn.LayerNorm(dim)


class SpatialTransformer(_NoForward):
    def __init__(self, in_channels, n_heads, d_head, depth=1, context_dim=None, **lora):
        super().__init__()
        inner = n_heads * d_head
        self.in_channels = in_channels
        self.norm = nn.GroupNorm(32 if in_channels % 32 == 0 else in_channels // 4,
                                 in_channels, eps=1e-6, affine=True)
        self.proj_in = nn.Conv2d(in_channels, inner, 1)
        self.transformer_blocks = nn.ModuleList(
            [BasicTransformerBlock(inner, n_heads, d_head, context_dim, **lora) for _ in range(depth)])
        self.proj_out = zero_module(nn.Conv2d(inner, in_channels, 1))


# --------------------------------------------------------------------------------------
# UNet building blocks (upstream ldm/modules/diffusionmodules/openaimodel.py names)
# --------------------------------------------------------------------------------------
def _gn(c):  # GroupNorm32: nn.GroupNorm(32, c) evaluated in fp32 upstream
    return nn.GroupNorm(32 if c % 32 == 0 else c // 4, c)


class ResBlock(_NoForward):
    def __init__(self, channels, emb_channels, out_channels=None):
        super().__init__()
        out_channels = out_channels or channels
        self.channels, self.out_channels = channels, out_channels
        self.in_layers = nn.Sequential(_gn(channels), nn.SiLU(), nn.Conv2d(channels, out_channels, 3, padding=1))
        self.emb_layers = nn.Sequential(nn.SiLU(), nn.Linear(emb_channels, out_channels))
        self.out_layers = nn.Sequential(_gn(out_channels), nn.SiLU(), nn.Dropout(0.0),
                                        zero_module(nn.Conv2d(out_channels, out_channels, 3, padding=1)))
        self.skip_connection = nn.Identity() if out_channels == channels else nn.Conv2d(channels, out_channels, 1)


class Downsample(_NoForward):
    def __init__(self, channels):
        super().__init__()
        self.op = nn.Conv2d(channels, channels, 3, stride=2, padding=1)


class Upsample(_NoForward):
    def __init__(self, channels):
        super().__init__()
        self.conv = nn.Conv2d(channels, channels, 3, padding=1)


class TimestepEmbedSequential(nn.Sequential):
    def forward(self, *a, **k):  # pragma: no cover
        return _NoForward.forward(self, *a, **k)


def _encoder_blocks(cfg: ModelConfig, time_dim: int, lora: dict):
    """input_blocks + the channel bookkeeping shared by UNet and ControlNet (Appendix A.2)."""
    mc = cfg.model_channels
    blocks = nn.ModuleList([TimestepEmbedSequential(nn.Conv2d(cfg.in_channels, mc, 3, padding=1))])
    chans = [mc]
    ch, ds = mc, 1
    for level, mult in enumerate(cfg.channel_mult):
        for _ in range(cfg.num_res_blocks):
            layers: List[nn.Module] = [ResBlock(ch, time_dim, mult * mc)]
            ch = mult * mc
            if ds in cfg.attention_resolutions:
                layers.append(SpatialTransformer(ch, cfg.num_heads, ch // cfg.num_heads,
                                                 cfg.transformer_depth, cfg.context_dim, **lora))
            blocks.append(TimestepEmbedSequential(*layers))
            chans.append(ch)
        if level != len(cfg.channel_mult) - 1:
            blocks.append(TimestepEmbedSequential(Downsample(ch)))
            chans.append(ch)
            ds *= 2
    return blocks, chans, ch, ds


def _middle(cfg, ch, time_dim, lora):
    return TimestepEmbedSequential(
        ResBlock(ch, time_dim),
        SpatialTransformer(ch, cfg.num_heads, ch // cfg.num_heads, cfg.transformer_depth, cfg.context_dim, **lora),
        ResBlock(ch, time_dim))


class ControlledUnetModel(_NoForward):
    """SD-1.5 UNet with ControlNet residual inputs (upstream cldm/cldm.py)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        mc = cfg.model_channels
        time_dim = mc * 4
        self.model_channels = mc
        self.time_embed = nn.Sequential(nn.Linear(mc, time_dim), nn.SiLU(), nn.Linear(time_dim, time_dim))
        self.input_blocks, chans, ch, ds = _encoder_blocks(cfg, time_dim, {})
        self.middle_block = _middle(cfg, ch, time_dim, {})
        self.output_blocks = nn.ModuleList()
        for level, mult in list(enumerate(cfg.channel_mult))[::-1]:
            for i in range(cfg.num_res_blocks + 1):
                ich = chans.pop()
                layers = [ResBlock(ch + ich, time_dim, mc * mult)]
                ch = mc * mult
                if ds in cfg.attention_resolutions:
                    layers.append(SpatialTransformer(ch, cfg.num_heads, ch // cfg.num_heads,
                                                     cfg.transformer_depth, cfg.context_dim))
                if level and i == cfg.num_res_blocks:
                    layers.append(Upsample(ch))
                    ds //= 2
                self.output_blocks.append(TimestepEmbedSequential(*layers))
        self.out = nn.Sequential(_gn(ch), nn.SiLU(), zero_module(nn.Conv2d(mc, cfg.out_channels, 3, padding=1)))


class ControlNet(_NoForward):
    """Stock-architecture ControlNet; ``control_lora_rank>0`` gives the ctrlora variant whose
    attention projections carry ``lora_layer`` branches (placement is data)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        mc = cfg.model_channels
        time_dim = mc * 4
        self.model_channels = mc
        lora = {}
        if cfg.control_lora_rank > 0:
            lora = dict(lora_rank=cfg.control_lora_rank, lora_scale=cfg.control_lora_scale,
                        lora_targets=tuple(cfg.control_lora_targets))
        self.time_embed = nn.Sequential(nn.Linear(mc, time_dim), nn.SiLU(), nn.Linear(time_dim, time_dim))
        self.input_blocks, chans, ch, ds = _encoder_blocks(cfg, time_dim, lora)
        self.zero_convs = nn.ModuleList(
            [TimestepEmbedSequential(zero_module(nn.Conv2d(c, c, 1))) for c in chans])
        hc = [16, 16, 32, 32, 96, 96, 256]
        strides = [1, 1, 2, 1, 2, 1, 2]
        layers: List[nn.Module] = []
        cin = cfg.hint_channels
        for c, s in zip(hc, strides):
            layers += [nn.Conv2d(cin, c, 3, padding=1, stride=s), nn.SiLU()]
            cin = c
        layers.append(zero_module(nn.Conv2d(cin, mc, 3, padding=1)))
        self.input_hint_block = TimestepEmbedSequential(*layers)
        self.middle_block = _middle(cfg, ch, time_dim, lora)
        self.middle_block_out = TimestepEmbedSequential(zero_module(nn.Conv2d(ch, ch, 1)))


# --------------------------------------------------------------------------------------
# VAE decoder (upstream ldm/modules/diffusionmodules/model.py names)  — SURVEY.md A.5
# --------------------------------------------------------------------------------------
def _vgn(c):
    return nn.GroupNorm(32 if c % 32 == 0 else c // 4, c, eps=1e-6, affine=True)


class ResnetBlock(_NoForward):
    def __init__(self, in_channels, out_channels):
        super().__init__()
        self.in_channels, self.out_channels = in_channels, out_channels
        self.norm1 = _vgn(in_channels)
        self.conv1 = nn.Conv2d(in_channels, out_channels, 3, padding=1)
        self.norm2 = _vgn(out_channels)
        self.conv2 = nn.Conv2d(out_channels, out_channels, 3, padding=1)
        if in_channels != out_channels:
            self.nin_shortcut = nn.Conv2d(in_channels, out_channels, 1)


class AttnBlock(_NoForward):
    def __init__(self, c):
        super().__init__()
        self.norm = _vgn(c)
        self.q = nn.Conv2d(c, c, 1)
        self.k = nn.Conv2d(c, c, 1)
        self.v = nn.Conv2d(c, c, 1)
        self.proj_out = nn.Conv2d(c, c, 1)


class _VaeUpsample(_NoForward):
    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, padding=1)


class _Mid(_NoForward):
    def __init__(self, c):
        super().__init__()
        self.block_1 = ResnetBlock(c, c)
        self.attn_1 = AttnBlock(c)
        self.block_2 = ResnetBlock(c, c)


class _UpLevel(_NoForward):
    pass


class Decoder(_NoForward):
    def __init__(self, cfg: ModelConfig):
        super().__init__()
        ch, mult = cfg.vae_ch, tuple(cfg.vae_ch_mult)
        block_in = ch * mult[-1]
        self.conv_in = nn.Conv2d(cfg.vae_z_channels, block_in, 3, padding=1)
        self.mid = _Mid(block_in)
        self.up = nn.ModuleList()
        ups = []
        for i_level in reversed(range(len(mult))):
            lvl = _UpLevel()
            lvl.block = nn.ModuleList()
            block_out = ch * mult[i_level]
            for _ in range(cfg.vae_num_res_blocks + 1):
                lvl.block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
            if i_level != 0:
                lvl.upsample = _VaeUpsample(block_in)
            ups.insert(0, lvl)
        for lvl in ups:
            self.up.append(lvl)
        self.norm_out = _vgn(block_in)
        self.conv_out = nn.Conv2d(block_in, 3, 3, padding=1)


class _VaeDownsample(_NoForward):
    """ldm ``Downsample(with_conv=True)``: zero-pad (0,1,0,1) then ``Conv2d(c, c, 3, stride=2, padding=0)``."""

    def __init__(self, c):
        super().__init__()
        self.conv = nn.Conv2d(c, c, 3, stride=2, padding=0)


class _DownLevel(_NoForward):
    pass


class Encoder(_NoForward):
    """ldm ``Encoder(ch=128, ch_mult=(1,2,4,4), num_res_blocks=2, attn_resolutions=[], double_z=True)`` — upstream
    ``ldm/modules/diffusionmodules/model.py``; state-dict keys ``down.{i}.block.{j}.*``, ``down.{i}.downsample.conv.*``,
    ``mid.*``, ``norm_out``, ``conv_out``.  Called through ``encode_first_stage`` (``vis_metrics.py:86``)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        ch, mult = cfg.vae_ch, tuple(cfg.vae_ch_mult)
        self.conv_in = nn.Conv2d(3, ch, 3, padding=1)
        self.down = nn.ModuleList()
        block_in = ch
        for i_level in range(len(mult)):
            lvl = _DownLevel()
            lvl.block = nn.ModuleList()
            block_out = ch * mult[i_level]
            for _ in range(cfg.vae_num_res_blocks):
                lvl.block.append(ResnetBlock(block_in, block_out))
                block_in = block_out
            if i_level != len(mult) - 1:
                lvl.downsample = _VaeDownsample(block_in)
            self.down.append(lvl)
        self.mid = _Mid(block_in)
        self.norm_out = _vgn(block_in)
        self.conv_out = nn.Conv2d(block_in, 2 * cfg.vae_z_channels, 3, padding=1)


class DiagonalGaussianDistribution:
    """ldm ``DiagonalGaussianDistribution`` (upstream ``ldm/modules/distributions/distributions.py``): ``parameters`` =
    ``[B, 2z, h, w]`` moments; logvar clamped to [-30, 20]; ``sample() = mean + std * randn``; ``mode() = mean``."""

    def __init__(self, parameters: torch.Tensor, deterministic: bool = False):
        self.parameters = parameters
        self.mean, self.logvar = torch.chunk(parameters, 2, dim=1)
        self.logvar = torch.clamp(self.logvar, -30.0, 20.0)
        self.deterministic = deterministic
        self.std = torch.exp(0.5 * self.logvar)
        self.var = torch.exp(self.logvar)
        if deterministic:
            self.var = self.std = torch.zeros_like(self.mean)

    def sample(self):
        return self.mean + self.std * torch.randn(self.mean.shape, device=self.parameters.device)

    def mode(self):
        return self.mean


class AutoencoderKL(_NoForward):
    """``encode`` (Encoder + ``quant_conv`` -> posterior; ``validate_step`` / training targets) and ``decode``
    (``post_quant_conv`` + Decoder; every sampled tile).  Both halves run on the kernels (``engine.PVae``)."""

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.decoder = Decoder(cfg)
        self.post_quant_conv = nn.Conv2d(cfg.vae_z_channels, cfg.vae_z_channels, 1)
        self.encoder = Encoder(cfg)
        self.quant_conv = nn.Conv2d(2 * cfg.vae_z_channels, 2 * cfg.vae_z_channels, 1)


class DiffusionWrapper(nn.Module):
    def __init__(self, unet):
        super().__init__()
        self.diffusion_model = unet


# --------------------------------------------------------------------------------------
# ControlLDM
# --------------------------------------------------------------------------------------
class ControlLDM(nn.Module):
    """Drop-in for upstream ``cldm.cldm.ControlLDM`` restricted to the inference path.

    Call sites kept: ``apply_model`` (``LoRA_modified/multiseg_19prompts_batch1_shuffle_ctrlora.py:388``,
    ``TextEncoder_Finetuning/vis_metrics.py:112``), ``decode_first_stage`` (``vis_metrics.py:450``),
    ``q_sample`` / ``predict_start_from_noise``, ``sample_log`` (``LoRA_modified/gen_test.py:114``),
    ``control_scales`` (``BaseModel/evaluateCN.py:91``).
    """

    def __init__(self, cfg: ModelConfig):
        super().__init__()
        self.cfg = cfg
        self.model = DiffusionWrapper(ControlledUnetModel(cfg))
        self.control_model = ControlNet(cfg)
        self.first_stage_model = AutoencoderKL(cfg) if cfg.with_vae else None
        self.cond_stage_model = None
        self.control_scales = [1.0] * 13
        self.only_mid_control = False
        self.num_timesteps = cfg.timesteps
        self.scale_factor = cfg.scale_factor
        self.channels = cfg.in_channels
        betas = torch.linspace(cfg.linear_start ** 0.5, cfg.linear_end ** 0.5, cfg.timesteps,
                               dtype=torch.float64) ** 2
        ac = torch.cumprod(1.0 - betas, dim=0)
        self.register_buffer("betas", betas.float(), persistent=False)
        self.register_buffer("alphas_cumprod", ac.float(), persistent=False)
        self.register_buffer("alphas_cumprod_prev", torch.cat([torch.ones(1, dtype=torch.float64), ac[:-1]]).float(),
                             persistent=False)
        self.register_buffer("sqrt_alphas_cumprod", ac.sqrt().float(), persistent=False)
        self.register_buffer("sqrt_one_minus_alphas_cumprod", (1 - ac).sqrt().float(), persistent=False)
        self.register_buffer("sqrt_recip_alphas_cumprod", (1 / ac).sqrt().float(), persistent=False)
        self.register_buffer("sqrt_recipm1_alphas_cumprod", (1 / ac - 1).sqrt().float(), persistent=False)
        self._engine = None

    # ---- engine plumbing --------------------------------------------------------------
    def engine(self, refresh: bool = False):
        """The compiled sm_100a execution plan for the current weights / LoRA surgery."""
        from .engine import TopoEngine
        if self._engine is None or refresh or self._engine.stale(self):
            self._engine = TopoEngine(self)
        return self._engine

    def invalidate_engine(self):
        self._engine = None

    # ---- the hot path -----------------------------------------------------------------
    @torch.no_grad()
    def apply_model(self, x_noisy, t, cond, *args, **kwargs):
        """eps = UNet(x, t, ctx, control=ControlNet(x, hint, t, ctx) * control_scales)  (u1)."""
        assert isinstance(cond, dict)
        ctx = torch.cat(cond["c_crossattn"], 1)
        hint = None if cond.get("c_concat") is None else torch.cat(cond["c_concat"], 1)
        return self.engine().apply_model(x_noisy, t, ctx, hint, self.control_scales, self.only_mid_control)

    @torch.no_grad()
    def decode_first_stage(self, z):
        return self.engine().decode_first_stage(z)

    @torch.no_grad()
    def encode_first_stage(self, x):
        """``first_stage_model.encode(x)`` -> posterior (``vis_metrics.py:86``); x ``[B,3,H,W]`` in [-1,1]."""
        return DiagonalGaussianDistribution(self.engine().encode_moments(x))

    def get_first_stage_encoding(self, encoder_posterior):
        """ldm ``LatentDiffusion.get_first_stage_encoding``: ``scale_factor * posterior.sample()`` (a tensor passes through)."""
        if isinstance(encoder_posterior, DiagonalGaussianDistribution):
            z = encoder_posterior.sample()
        elif isinstance(encoder_posterior, torch.Tensor):
            z = encoder_posterior
        else:
            raise NotImplementedError(f"encoder_posterior of type '{type(encoder_posterior)}' not yet implemented")
        return self.scale_factor * z

    def q_sample(self, x_start, t, noise=None):
        noise = torch.randn_like(x_start) if noise is None else noise
        a = self.sqrt_alphas_cumprod.to(x_start.device)[t].view(-1, 1, 1, 1)
        b = self.sqrt_one_minus_alphas_cumprod.to(x_start.device)[t].view(-1, 1, 1, 1)
        return a * x_start + b * noise

    def predict_start_from_noise(self, x_t, t, noise):
        a = self.sqrt_recip_alphas_cumprod.to(x_t.device)[t].view(-1, 1, 1, 1)
        b = self.sqrt_recipm1_alphas_cumprod.to(x_t.device)[t].view(-1, 1, 1, 1)
        return a * x_t - b * noise

    def get_learned_conditioning(self, c):
        if self.cond_stage_model is None:
            raise RuntimeError("no cond_stage_model attached (text encoder is row N3, not on the hot path)")
        return self.cond_stage_model.encode(c)

    @torch.no_grad()
    def sample_log(self, cond, batch_size, ddim=True, ddim_steps=30, **kwargs):
        from .sampling import DDIMSampler
        _, _, h, w = cond["c_concat"][0].shape
        return DDIMSampler(self).sample(ddim_steps, batch_size, (self.channels, h // 8, w // 8), cond,
                                        verbose=False, **kwargs)


def create_model(config="stock", **overrides) -> ControlLDM:
    """``create_model(yaml).cpu()`` of the reference (``…_ctrlora.py:219``) with the yaml replaced
    by a variant name or a ``ModelConfig``.  Returns the model on CPU."""
    if isinstance(config, ModelConfig):
        cfg = config
        unet_surgery = None
    else:
        name = str(config)
        import re
        m = re.search(r"rank(\d+)", name)
        if m:  # the reference's yaml names: ctrlora_finetune_sd15_rank12.yaml / …rank32.yaml (…_ctrlora.py:216, …_addlrSD.py:281)
            v = dict(control_lora_rank=int(m.group(1)), unet=None)
        elif "cldm_v15" in name:
            v = VARIANTS["stock"]
        else:
            hit = [key for key in VARIANTS if key.lower() == name.lower()] or \
                  [key for key in VARIANTS if key.lower() in name.lower()]
            if not hit:
                raise KeyError(f"unknown model config {config!r}; variants: {sorted(VARIANTS)}")
            v = VARIANTS[hit[0]]
        cfg = ModelConfig(control_lora_rank=v["control_lora_rank"])
        unet_surgery = v["unet"]
    for k, val in overrides.items():
        setattr(cfg, k, val)
    model = ControlLDM(cfg)
    if unet_surgery is not None:
        from . import lora as _lora
        kind, r_q, r_kv = unet_surgery
        if kind == "qkv":
            _lora.lora_qkv(model.model.diffusion_model, r_q=r_q, r_kv=r_kv)
        else:
            _lora.lora_attn_qkvo(model.model.diffusion_model, r=r_q)
    return model.cpu()


def load_state_dict(path, location="cpu"):
    sd = torch.load(path, map_location=location)
    return sd.get("state_dict", sd)


def randomize_zero_modules(model: nn.Module, seed: int = 1, std: float = 0.02) -> int:
    """Re-draw every all-zero weight tensor from N(0, std^2): upstream zero-inits
    every ResBlock out-conv, proj_out, zero-conv, the UNet out conv and every LoRA up/B matrix, so
    with default init eps == 0 and any parity check would be vacuous."""
    g = torch.Generator().manual_seed(seed)
    n = 0
    with torch.no_grad():
        for name, p in sorted(model.named_parameters(), key=lambda kv: kv[0]):
            if p.ndim >= 2 and not bool(p.any()):
                p.copy_((torch.randn(p.shape, generator=g) * std).to(p.device))   # bumps p._version: engines go stale
                n += 1
    if hasattr(model, "invalidate_engine"):
        model.invalidate_engine()
    return n
