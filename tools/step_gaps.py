"""Where one bench step (19 tiles) goes besides the 31 graph replays: replay-only time vs DDIMSampler.sample vs
stylize_tiles (CUDA events, Be=38)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import synth
from topo_b200 import tiles
from topo_b200.sampling import DDIMSampler

T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
model = synth.build_model("ctrlora", control_lora_rank=16).cuda(); eng = model.engine()
ctx = synth.make_ctx(T).cuda(); hint = synth.make_hint(T).cuda(); xT = synth.make_xT(T).cuda()


def ev(fn, n=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


with torch.no_grad():
    full = ev(lambda: tiles.stylize_tiles(model, ctx, hint, xT, steps=30, scale=7.5, batch=T))
    cond = {"c_crossattn": [ctx], "c_concat": [hint]}; uc = {"c_crossattn": [torch.zeros_like(ctx)], "c_concat": [hint]}
    smp = DDIMSampler(model)
    samp = ev(lambda: smp.sample(S=30, conditioning=cond, batch_size=T, shape=(4, 64, 64), unconditional_guidance_scale=7.5,
                                 unconditional_conditioning=uc, eta=0.0, x_T=xT))
    g = next(iter(eng._graphs.values()))
    rep = ev(lambda: [g.step() for _ in range(31)])
    prep = ev(lambda: g.g_prep.replay())
    z = torch.randn(T, 4, 64, 64, device="cuda")
    dec = ev(lambda: eng.decode_to_u8(z))
print(f"T={T}: stylize_tiles {full:.1f} ms | sample {samp:.1f} ms | 31 replays {rep:.1f} ms ({rep/31:.2f}/step) | "
      f"prepare graph {prep:.2f} ms | decode_to_u8 {dec:.1f} ms | sample-replays-prepare = {samp-rep-prep:.1f} ms")
