"""One eager apply_model at the bench batch (Be=38, 64x64 latents, ctrlora r=16) inside a cudaProfilerStart/Stop
range, for `ncu --profile-from-start off`.  Usage: python tools/profile_step.py [Be]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import synth

Be = int(sys.argv[1]) if len(sys.argv) > 1 else 38
dev = "cuda"
variant = sys.argv[2] if len(sys.argv) > 2 else "ctrlora"
model = synth.build_model(variant).to(dev)
eng = model.engine()
B = Be // 2
ctx = synth.make_ctx(B).to(dev); hint = synth.make_hint(B).to(dev); xT = synth.make_xT(B).to(dev)
c2 = torch.cat([torch.zeros_like(ctx), ctx]); h2 = torch.cat([hint, hint]); x2 = torch.cat([xT, xT])
t2 = torch.full((Be,), 501.0, device=dev)
with torch.no_grad():
    P = eng.prepare(c2, hint)   # one hint per tile (the CFG halves share it), context-free prefix shared (x_rep=2)
    eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStart()
    eps = eng.apply_model(x2, t2, c2, h2, model.control_scales, False, prepared=P, x_rep=2)
    torch.cuda.synchronize()
    torch.cuda.cudart().cudaProfilerStop()
print("ok", float(eps.abs().mean()))
