"""Per-shape time breakdown of one eager apply_model (Be=38) from CUDA events."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import synth
from topo_b200 import engine as E
Be = int(sys.argv[1]) if len(sys.argv) > 1 else 38
variant = sys.argv[2] if len(sys.argv) > 2 else "ctrlora"
model = synth.build_model(variant).cuda(); eng = model.engine(); B = Be // 2
ctx = synth.make_ctx(B).cuda(); hint = synth.make_hint(B).cuda(); xT = synth.make_xT(B).cuda()
c2 = torch.cat([torch.zeros_like(ctx), ctx]); h2 = torch.cat([hint, hint]); x2 = torch.cat([xT, xT]); t2 = torch.full((Be,), 501.0, device="cuda")
with torch.no_grad():
    P = eng.prepare(c2, hint)   # as the sampler: one hint per tile, context-free prefix shared
    for _ in range(2): eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    torch.cuda.synchronize(); E.PROF.recs, E.PROF.on = [], True
    eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    torch.cuda.synchronize(); E.PROF.on = False
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for kind, fl, nb, e0, e1, tag, kern in E.PROF.recs:
    a = agg[(kind, tag + "  " + kern)]; a[0] += 1; a[1] += e0.elapsed_time(e1); a[2] += fl
tot = sum(a[1] for a in agg.values())
print(f"total {tot:.2f} ms")
for (kind, tag), (n, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:60]:
    print(f"{ms:7.3f} ms {100*ms/tot:5.1f}% n={n:3d} {fl/1e9/max(ms,1e-9):7.1f} TF/s  {kind:12s} {tag}")
