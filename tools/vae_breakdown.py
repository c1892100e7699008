"""Per-kernel-class time of one VAE decode + uint8 conversion at the bench batch (T=19), CUDA events per launch."""
import os, sys, collections
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import synth
from topo_b200 import engine as E
T = int(sys.argv[1]) if len(sys.argv) > 1 else 19
model = synth.build_model("ctrlora", control_lora_rank=16).cuda(); eng = model.engine()
z = torch.randn(T, 4, 64, 64, device="cuda")
with torch.no_grad():
    for _ in range(2): eng.decode_to_u8(z)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record(); eng.decode_to_u8(z); e1.record(); torch.cuda.synchronize()
    print(f"decode_to_u8 T={T}: {e0.elapsed_time(e1):.2f} ms")
    E.PROF.recs, E.PROF.on = [], True
    eng.decode_to_u8(z); torch.cuda.synchronize(); E.PROF.on = False
agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
for kind, fl, nb, a, b, tag, kern in E.PROF.recs:
    r = agg[(kind, tag)]; r[0] += 1; r[1] += a.elapsed_time(b); r[2] += fl; r[3] += nb
tot = sum(r[1] for r in agg.values())
print(f"instrumented total {tot:.2f} ms over {sum(r[0] for r in agg.values())} launches")
for (kind, tag), (n, ms, fl, nb) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:24]:
    print(f"{ms:7.3f} ms {100*ms/tot:5.1f}% n={n:3d} {fl/1e9/max(ms,1e-9):7.1f} TF/s {nb/1e6/max(ms,1e-9):7.1f} GB/s  {kind:14s} {tag}")
