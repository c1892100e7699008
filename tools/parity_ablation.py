"""Which rounding owns the per-step latent error?  ("Next round" 1b.)

Runs the fp32 oracle on BASELINE configs[0] (onlyCNLora r=4, one 512x512 tile, CFG 7.5, first of the 4 coarse DDIM steps,
t=751: the step with the largest error in round 1) with bf16 rounding injected at one class of places at a time and
prints the error of eps and of the teacher-forced DDIM step against the un-rounded fp32 oracle, next to the error of
the CUDA kernels on the same inputs:

  W  weights (every >=2-D tensor of the state dict) rounded to bf16
  A  operands: the input of every linear / conv and q, k, v, P of every attention rounded to bf16
  R  residual stream: every residual SUM (ResBlock output, the three adds of a transformer block, SpatialTransformer
     output, control adds) rounded to bf16
  O  every linear / conv / norm OUTPUT rounded to bf16 (what a bf16 activation store does)

The oracle is test infrastructure; this tool is a checker, not a product path.
Usage: python tools/parity_ablation.py [lat=64] [variant=onlyCNLora]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import torch.nn.functional as F

from oracle import bf16_emulation as EMU
from oracle import ldm_oracle as O
from topo_b200 import ops, synth

torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
lat = int(sys.argv[1]) if len(sys.argv) > 1 else 64
variant = sys.argv[2] if len(sys.argv) > 2 else "onlyCNLora"
dev = os.environ.get("TOPO_ABL_DEV", "cuda")     # "cpu": oracle rows only (dry run without a GPU)


model = synth.build_model(variant).to(dev)
sd32 = {k: v.float() for k, v in model.state_dict().items()}
sd16 = EMU.round_weights(sd32)
scales = synth.lora_scales(model)
hint, ctx, xT = synth.make_hint(1, lat * 8).to(dev), synth.make_ctx(1).to(dev), synth.make_xT(1, lat).to(dev)
c2, h2, x2 = torch.cat([torch.zeros_like(ctx), ctx]), torch.cat([hint, hint]), torch.cat([xT, xT])
cond = {"c_crossattn": [c2], "c_concat": [h2]}
ts, a, a_prev, _ = O.make_schedule(4)
i = 3
t2 = torch.full((2,), int(ts[i]), device=dev)
w = 7.5


def step(eps):
    e_u, e_c = eps.chunk(2)
    e = e_u + w * (e_c - e_u)
    x0 = (xT - (1 - a[i]).sqrt() * e) / a[i].sqrt()
    return a_prev[i].sqrt() * x0 + (1 - a_prev[i]).sqrt() * e


def run(sd, **flags):
    with torch.no_grad(), EMU.rounding(**flags):
        return O.apply_model(sd, x2, t2, cond, heads=8, scales=scales)


ref = run(sd32)
xref = step(ref)
print(f"# {variant} lat{lat} t={int(ts[i])} CFG {w}: max|eps|={ref.abs().max().item():.3f} max|x_next|={xref.abs().max().item():.3f} "
      f"max|e_c-e_u|={(ref[1] - ref[0]).abs().max().item():.3f}")
print(f"{'rounding':<28}{'eps rel-max':>12}{'eps cos':>11}{'step max-abs':>14}{'step rel':>11}")


def report(name, eps):
    x = step(eps)
    er = (eps - ref).abs().max().item() / ref.abs().max().item()
    cs = F.cosine_similarity(eps.flatten(), ref.flatten(), dim=0).item()
    sa = (x - xref).abs().max().item()
    print(f"{name:<28}{er:>12.3e}{cs:>11.6f}{sa:>14.3e}{sa / xref.abs().max().item():>11.3e}")


report("W  weights", run(sd16))
report("A  operands", run(sd32, a=True))
report("R  residual sums", run(sd32, r=True))
report("O  primitive outputs", run(sd32, o=True))
report("W+A", run(sd16, a=True))
report("W+A+O", run(sd16, a=True, o=True))
report("W+A+R+O (all)", run(sd16, a=True, r=True, o=True))
if dev == "cpu":
    sys.exit(0)
eng = model.engine()
with torch.no_grad():
    P = eng.prepare(c2, h2)
    eps = eng.apply_model(x2, t2.float(), c2, h2, model.control_scales, False, prepared=P)
report("CUDA kernels", eps)
