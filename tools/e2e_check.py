"""GPU debugging aid: CUDA path vs the oracle (run in fp32 on the same GPU) for apply_model / DDIM / VAE."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
torch.backends.cuda.matmul.allow_tf32 = False
torch.backends.cudnn.allow_tf32 = False
from topo_b200 import synth
from oracle import ldm_oracle as O
from topo_b200.sampling import DDIMSampler

variant = sys.argv[1] if len(sys.argv) > 1 else "ctrlora"
lat = int(sys.argv[2]) if len(sys.argv) > 2 else 32
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = "cuda"
t0 = time.time()
model = synth.build_model(variant).to(dev)
print(f"built {variant} in {time.time()-t0:.1f}s")
sd = {k: v.float() for k, v in model.state_dict().items()}
scales = synth.lora_scales(model)
B = 1
hint = synth.make_hint(B, lat * 8).to(dev)
ctx = synth.make_ctx(B).to(dev)
xT = synth.make_xT(B, lat).to(dev)


def stat(name, got, ref):
    err = (got.float() - ref.float()).abs().max().item()
    cos = torch.nn.functional.cosine_similarity(got.float().flatten(), ref.float().flatten(), dim=0).item()
    print(f"[e2e] {name}: max_abs={err:.3e} ref_max={ref.abs().max().item():.3e} ref_std={ref.std().item():.3e} cos={cos:.6f}", flush=True)


with torch.no_grad():
    cond = {"c_crossattn": [ctx], "c_concat": [hint]}
    uc = {"c_crossattn": [torch.zeros_like(ctx)], "c_concat": [hint]}
    t = torch.full((B,), 501, device=dev, dtype=torch.long)
    eps = model.apply_model(xT, t, cond)
    torch.cuda.synchronize()
    ref = O.apply_model(sd, xT, t, cond, heads=8, scales=scales)
    stat("apply_model", eps, ref)
    # DDIM with CFG, per-step latents
    x, inter = O.ddim_sample(sd, steps, cond, xT.shape, 7.5, uc, 0.0, xT, 8, scales=scales)
    s = DDIMSampler(model)
    got, gi = s.sample(S=steps, conditioning=cond, batch_size=B, shape=xT.shape[1:], unconditional_guidance_scale=7.5,
                       unconditional_conditioning=uc, eta=0.0, x_T=xT, log_every_t=1)
    for i, (a, b) in enumerate(zip(gi["x_inter"][1:], inter)):
        stat(f"latent step {i}", a, b)
    img_ref = O.decode_first_stage(sd, x).clamp(-1, 1)
    img_same = model.decode_first_stage(x).clamp(-1, 1)
    stat("vae decode (same latent)", img_same, img_ref)
    img = model.decode_first_stage(got).clamp(-1, 1)
    stat("final tile", img, img_ref)
    from topo_b200 import engine
    print("launches", engine.LAUNCHES[0])
    # timing of one CFG step
    eng = model.engine()
    c2 = torch.cat([torch.zeros_like(ctx), ctx]); h2 = torch.cat([hint, hint])
    P = eng.prepare(c2, h2)
    x2 = torch.cat([xT, xT]); t2 = torch.cat([t, t])
    for _ in range(3):
        eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(5):
        eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P)
    e1.record(); torch.cuda.synchronize()
    print(f"apply_model Be=2 lat={lat}: {e0.elapsed_time(e1)/5:.2f} ms/call (eager, incl. host launch overhead)")
