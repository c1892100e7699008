"""Sampling drivers of the hot path: ldm ``DDIMSampler`` (uniform schedule, batched classifier-free guidance),
the ``cldm/ddim_hacked`` two-call variant, and the reference's ``sample_preview`` / ``validate_step`` /
``MeterEMA`` (``TextEncoder_Finetuning/vis_metrics.py:63-116,407-452``).

The loop body is: ONE engine.apply_model over the [uncond ; cond] batch (uncond first, as ldm does) + ONE
``topo_ddim_step`` kernel doing the CFG combine and the DDIM update in fp32.  Step-invariant work (cross-attention
K/V of the prompt, ControlNet hint block) is hoisted through ``engine.prepare``.  SURVEY.md Appendix A.4.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from . import capi, ops


def make_ddim_timesteps(S: int, T: int = 1000) -> np.ndarray:
    """ldm 'uniform' discretisation: c = T // S; ts = range(0, T, c) + 1.  S=30 → 31 steps [1, 34, …, 991]."""
    c = T // S
    return np.asarray(list(range(0, T, c))) + 1


def make_ddim_sampling_parameters(alphacums: np.ndarray, ts: np.ndarray, eta: float):
    a = alphacums[ts]
    a_prev = np.asarray([alphacums[0]] + alphacums[ts[:-1]].tolist())
    sig = eta * np.sqrt((1 - a_prev) / (1 - a) * (1 - a / a_prev))
    return a, a_prev, sig


class DDIMSampler:
    """``DDIMSampler(model).sample(S=, conditioning=, batch_size=, shape=, unconditional_guidance_scale=,
    unconditional_conditioning=, eta=, x_T=)`` → ``(samples, intermediates)`` — the call made at
    ``vis_metrics.py:430-447``.  ``hacked=True`` gives ``cldm.ddim_hacked.DDIMSampler`` (``BaseModel/evaluateCN.py:93-96``):
    cond and uncond as two separate apply_model calls."""

    def __init__(self, model, schedule: str = "linear", hacked: bool = False, **kwargs):
        self.model = model
        self.ddpm_num_timesteps = model.num_timesteps
        self.hacked = hacked

    def make_schedule(self, ddim_num_steps: int, ddim_discretize: str = "uniform", ddim_eta: float = 0.0, verbose=False):
        assert ddim_discretize == "uniform", "only the uniform schedule is on the reference's path"
        ac = self.model.alphas_cumprod.detach().cpu().numpy()  # stored float32, as upstream
        self.ddim_timesteps = make_ddim_timesteps(ddim_num_steps, self.ddpm_num_timesteps)
        self.ddim_alphas, self.ddim_alphas_prev, self.ddim_sigmas = make_ddim_sampling_parameters(
            ac, self.ddim_timesteps, ddim_eta)

    @staticmethod
    def _cat_cond(uc: Dict, c: Dict) -> Dict:
        return {k: [torch.cat([uc[k][i], c[k][i]]) for i in range(len(c[k]))] for k in c}

    # keyword arguments of ldm's ``DDIMSampler.sample`` that the reference never passes; a value other than the
    # upstream default is refused instead of being silently dropped
    _UNSUPPORTED = {"mask": None, "x0": None, "temperature": 1.0, "noise_dropout": 0.0, "score_corrector": None,
                    "corrector_kwargs": None, "quantize_x0": False, "dynamic_threshold": None, "ucg_schedule": None,
                    "normals_sequence": None, "timesteps": None}

    @torch.no_grad()
    def sample(self, S, batch_size, shape, conditioning=None, callback=None, img_callback=None, eta=0.0, x_T=None,
               verbose=False, log_every_t=100, unconditional_guidance_scale=1.0, unconditional_conditioning=None,
               **kwargs) -> Tuple[torch.Tensor, Dict[str, List[torch.Tensor]]]:
        for k, v in kwargs.items():
            if k not in self._UNSUPPORTED:
                raise TypeError(f"DDIMSampler.sample: unexpected keyword argument {k!r}")
            default = self._UNSUPPORTED[k]
            if v is not None and not (isinstance(v, (bool, int, float)) and v == default):
                raise capi.TopoError(f"DDIMSampler.sample: {k}={v!r} is not on the reference's path and is not implemented")
        model = self.model
        eng = model.engine()
        dev = eng.device
        self.make_schedule(S, ddim_eta=eta)
        C, H, W = shape
        # first RNG draw after the caller's manual_seed, on the sampling device (as ldm does)
        img = torch.randn((batch_size, C, H, W), device=dev) if x_T is None else x_T.to(dev, torch.float32).clone()
        cfg = unconditional_conditioning is not None and unconditional_guidance_scale != 1.0
        w = float(unconditional_guidance_scale)

        def _pack(cond):
            ctx = torch.cat(cond["c_crossattn"], 1).to(dev, torch.float32)
            hint = None if cond.get("c_concat") is None else torch.cat(cond["c_concat"], 1).to(dev, torch.float32)
            return ctx, hint

        two_calls = self.hacked
        if cfg and not two_calls:
            uh, ch = unconditional_conditioning.get("c_concat"), conditioning.get("c_concat")
            if (uh is None) != (ch is None):
                two_calls = True   # guess mode (evaluateCN.py:86): the unconditional branch has no control -> cannot be batched
        if cfg and not two_calls:
            shared = (uh is not None and ch is not None and len(uh) == len(ch)
                      and all(a is b for a, b in zip(uh, ch)))  # same control images for both CFG halves
            if shared:
                ctx_u, _ = _pack({"c_crossattn": unconditional_conditioning["c_crossattn"]})
                ctx_c, hint = _pack(conditioning)
                conds = [(torch.cat([ctx_u, ctx_c]), hint)]
            else:
                conds = [_pack(self._cat_cond(unconditional_conditioning, conditioning))]
        elif cfg:
            conds = [_pack(unconditional_conditioning), _pack(conditioning)]  # eps buffer layout is [uncond ; cond]
        else:
            conds = [_pack(conditioning)]
        scales, omc = model.control_scales, model.only_mid_control
        rep = conds[0][0].shape[0] // batch_size      # 2 when the CFG pair is batched into one call
        use_graph = bool(getattr(model, "use_cuda_graphs", True))
        if use_graph:
            runs = []
            for br, (ctx, hint) in enumerate(conds):
                # the two-call variant needs two sets of static buffers: one cached graph per branch
                g = eng.graphed(ctx.shape[0], (H, W), tuple(ctx.shape[1:]), None if hint is None else tuple(hint.shape[2:]),
                                scales, omc, None if hint is None else hint.shape[0], rep, branch=br)
                g.set_cond(ctx, hint)
                runs.append(g)
        else:
            plans = [(ctx, hint, eng.prepare(ctx, hint)) for ctx, hint in conds]

        inter = {"x_inter": [img], "pred_x0": [img]}
        ts = self.ddim_timesteps
        n = len(ts)
        for i in reversed(range(n)):           # i = ldm's ``index`` (descending)
            if use_graph:
                outs = []
                for g in runs:
                    for r in range(rep):
                        g.x[r * batch_size:(r + 1) * batch_size].copy_(img)
                    g.t.fill_(float(ts[i]))
                    outs.append(g.step())
                eps = outs[0] if len(outs) == 1 else torch.cat(outs)
            else:
                t = torch.full((batch_size * rep,), int(ts[i]), device=dev, dtype=torch.long)
                xin = img if rep == 1 else torch.cat([img] * rep)
                outs = [eng.apply_model(xin, t, c_, h_, scales, omc, prepared=P, x_rep=rep) for c_, h_, P in plans]
                eps = outs[0] if len(outs) == 1 else torch.cat(outs)
            sig = float(self.ddim_sigmas[i])
            noise = torch.randn_like(img) if sig != 0.0 else None  # sigma = 0 (eta = 0) on every reference call
            log = (i % log_every_t == 0) or (i == n - 1)            # ldm: ``index % log_every_t == 0 or index == total_steps - 1``
            px0 = torch.empty_like(img) if (log or img_callback) else None
            img = ops.ddim_step(img, eps, cfg, w, float(self.ddim_alphas[i]), float(self.ddim_alphas_prev[i]), sig, noise,
                                pred_x0=px0)
            if callback:
                callback(i)
            if img_callback:
                img_callback(px0, i)
            if log:
                inter["x_inter"].append(img)
                inter["pred_x0"].append(px0)
        return img, inter


# ------------------------------------------------------------------------------------------------
# vis_metrics.py drivers
# ------------------------------------------------------------------------------------------------
class MeterEMA:
    """EMA smoother for displayed losses (``vis_metrics.py:63-71``)."""

    def __init__(self, beta: float = 0.98):
        self.beta, self.avg = beta, None

    def update(self, x: float) -> float:
        self.avg = x if self.avg is None else self.beta * self.avg + (1 - self.beta) * x
        return self.avg


@torch.no_grad()
def sample_preview(model, textenc, batch, device, steps=30, scale=7.5, eta=0.0, seed=1234, x_T=None):
    """Prompts → context, ``uc = zeros_like(c)`` (not the empty-prompt encoding), same hint on both branches, seed,
    DDIM with CFG, VAE decode, clamp → ``[B,3,H,W]`` in [-1,1] fp32 (``vis_metrics.py:407-452``).  ``x_T`` (not in the
    reference signature) lets a caller pin the starting latent for parity runs."""
    model.eval()
    seg = batch["seg"].to(device=device, dtype=torch.float32)
    prompts = batch["prompt"]
    B, _, H, W = seg.shape
    c_cross = textenc.encode(prompts).to(device=device, dtype=torch.float32)
    uc_cross = torch.zeros_like(c_cross)
    cond = {"c_crossattn": [c_cross], "c_concat": [seg]}
    uc = {"c_crossattn": [uc_cross], "c_concat": [seg]}
    sampler = DDIMSampler(model)
    torch.manual_seed(seed)
    samples, _ = sampler.sample(S=steps, conditioning=cond, batch_size=B, shape=(4, H // 8, W // 8), verbose=False,
                                unconditional_guidance_scale=scale, unconditional_conditioning=uc, eta=eta, x_T=x_T)
    pred = model.decode_first_stage(samples.to(torch.float32))
    return torch.clamp(pred, -1.0, 1.0)


@torch.no_grad()
def validate_step(model, textenc, batch, device, max_length=77, iters=64, fixed_t=None, fixed_noise=None, z=None):
    """ε-MSE over ``iters`` random timesteps (``vis_metrics.py:73-116``): ``z = get_first_stage_encoding(
    encode_first_stage(gt))`` (VAE encoder on the kernels), then per iteration ``q_sample`` → ``apply_model`` → mse.
    ``z`` (or ``batch['z']``) may be supplied to skip the encoder (pre-encoded validation sets)."""
    seg = batch["seg"].to(device=device, dtype=torch.float32)
    z = z if z is not None else batch.get("z")
    if z is None:
        gt = batch["gt"].to(device=device, dtype=torch.float32)
        z = model.get_first_stage_encoding(model.encode_first_stage(gt))
    z = z.to(device=device, dtype=torch.float32)
    B = z.size(0)
    c_cross = textenc.encode(batch["prompt"]).to(device=device, dtype=torch.float32)
    cond = {"c_crossattn": [c_cross], "c_concat": [seg]}
    g = None
    if fixed_t is None or fixed_noise is None:
        g = torch.Generator(device=device).manual_seed(12345)
    total = 0.0
    for _ in range(iters):
        t = fixed_t if fixed_t is not None else torch.randint(0, model.num_timesteps, (B,), generator=g, device=device).long()
        noise = fixed_noise if fixed_noise is not None else torch.randn_like(z)
        eps_hat = model.apply_model(model.q_sample(z, t, noise=noise), t, cond)
        total += torch.mean((eps_hat.float() - noise.float()) ** 2).item()
    return total / float(iters)
