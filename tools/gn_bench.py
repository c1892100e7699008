"""GroupNorm(32)+SiLU micro-benchmark at the shapes of one Be=38 apply_model: stats+apply pair, CUDA events,
inputs rotated through > L2 worth of buffers.  GB/s on the 6 B/element convention (stats read + apply read + write)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 38
SHAPES = [(4096, 320), (4096, 640), (4096, 960), (1024, 320), (1024, 640), (1024, 960), (1024, 1280), (1024, 1920),
          (256, 640), (256, 1280), (256, 1920), (256, 2560), (64, 1280), (64, 2560)]
if len(sys.argv) > 2:
    SHAPES = [tuple(int(v) for v in t.split("x")) for t in sys.argv[2].split(",")]
REP = int(sys.argv[3]) if len(sys.argv) > 3 else 4
tot = 0.0
for HW, C in SHAPES:
    nbuf = max(2, int(400e6 // (B * HW * C * 2)) + 1)
    xs = [torch.randn(B, HW, C, device="cuda").to(torch.bfloat16) for _ in range(min(nbuf, 64))]
    y = torch.empty_like(xs[0])
    g = torch.rand(C, device="cuda") + 0.5; b = torch.randn(C, device="cuda")
    for x in xs[:2]: ops.groupnorm_silu(x, g, b, 32, 1e-5, True, out=y)
    torch.cuda.synchronize()
    n = REP * len(xs)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(n): ops.groupnorm_silu(xs[i % len(xs)], g, b, 32, 1e-5, True, out=y)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    tot += us
    print(f"B{B} HW{HW:5d} C{C:5d}: {us:7.1f} us  {B*HW*C*6/us/1e3:7.1f} GB/s (6 B/el)  {B*HW*C*2/1e6:6.1f} MB")
print(f"sum {tot:.1f} us")
