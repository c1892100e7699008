"""Where one bench step (19 tiles, 31 model calls, CFG, VAE decode, uint8) spends its time, stage by stage
(CUDA events on the launching stream), plus a per-kernel-class breakdown of the VAE decode.
Usage: python tools/stage_breakdown.py [tiles]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import synth
from topo_b200 import engine as E

T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
dev = torch.device("cuda", 0)
model = synth.build_model("ctrlora", control_lora_rank=16).to(dev)
eng = model.engine()
hints, ctx, xT = synth.make_hint(T, 512).to(dev), synth.make_ctx(T).to(dev), synth.make_xT(T, 64).to(dev)
c2 = torch.cat([torch.zeros_like(ctx), ctx]); h2 = torch.cat([hints, hints]); x2 = torch.cat([xT, xT])
XREP = int(os.environ.get("XREP", "2"))   # 2: the sampler's batched CFG call (context-free prefix shared)
g = eng.graphed(2 * T, (64, 64), tuple(ctx.shape[1:]), (512, 512), model.control_scales, False, T, XREP)


def ev():
    return torch.cuda.Event(enable_timing=True)


def timed(fn, n=3):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        a, b = ev(), ev()
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


g.x.copy_(x2); g.t.fill_(501.0)
t_prep = timed(lambda: g.set_cond(c2, hints))
t_step = timed(lambda: g.step(), 5)
t_31 = timed(lambda: [g.step() for _ in range(31)], 2)
z = torch.randn(T, 4, 64, 64, device=dev)
t_vae = timed(lambda: eng.decode_to_u8(z))
print(f"tiles={T}  prepare {t_prep:.2f} ms | apply_model (graph) {t_step:.2f} ms | 31 calls {t_31:.1f} ms | "
      f"VAE decode+u8 {t_vae:.2f} ms | sum {t_prep + t_31 + t_vae:.1f} ms -> {T / (t_prep + t_31 + t_vae) * 1e3:.2f} tiles/s")

E.PROF.recs, E.PROF.on = [], True
eng.decode_to_u8(z)
torch.cuda.synchronize()
E.PROF.on = False
tot = 0.0
rows = {}
for kind, fl, nb, e0, e1, tag, kern in E.PROF.recs:
    ms = e0.elapsed_time(e1)
    d = rows.setdefault((kind, tag), [0, 0.0, 0.0, 0.0])
    d[0] += 1; d[1] += ms; d[2] += fl; d[3] += nb
    tot += ms
print(f"VAE decode, instrumented launches: {tot:.2f} ms")
for (kind, tag), d in sorted(rows.items(), key=lambda kv: -kv[1][1])[:25]:
    print(f"  {d[1]:8.3f} ms n={d[0]:3d} {kind:15s} {tag:40s} {d[2] / 1e9 / max(d[1], 1e-9):7.1f} TF/s {d[3] / 1e6 / max(d[1], 1e-9):7.1f} GB/s")

# the same per-shape table for one apply_model
P = eng.prepare(c2, hints)
t2 = torch.full((2 * T,), 501.0, device=dev)
eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=XREP)
torch.cuda.synchronize()
E.PROF.recs, E.PROF.on = [], True
eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=XREP)
torch.cuda.synchronize()
E.PROF.on = False
rows, tot = {}, 0.0
for kind, fl, nb, e0, e1, tag, kern in E.PROF.recs:
    ms = e0.elapsed_time(e1)
    d = rows.setdefault((kind, tag), [0, 0.0, 0.0, 0.0])
    d[0] += 1; d[1] += ms; d[2] += fl; d[3] += nb
    tot += ms
print(f"apply_model, instrumented launches: {tot:.2f} ms")
for (kind, tag), d in sorted(rows.items(), key=lambda kv: -kv[1][1])[:45]:
    print(f"  {d[1]:8.3f} ms n={d[0]:3d} {kind:15s} {tag:40s} {d[2] / 1e9 / max(d[1], 1e-9):7.1f} TF/s {d[3] / 1e6 / max(d[1], 1e-9):7.1f} GB/s")
