"""LayerNorm micro-benchmark at the token-matrix shapes of one Be=38 apply_model (CUDA events, inputs rotated through
> L2 worth of buffers).  GB/s on the 4 B/element convention (read + write)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 38
tot = 0.0
for N, C, cnt in [(4096, 320, 21), (1024, 640, 21), (256, 1280, 27)]:
    M = B * N
    xs = [torch.randn(M, C, device="cuda").to(torch.bfloat16) for _ in range(max(2, int(400e6 // (M * C * 2)) + 1))]
    g = torch.rand(C, device="cuda") + 0.5; b = torch.randn(C, device="cuda")
    for x in xs[:2]: ops.layernorm(x, g, b)
    torch.cuda.synchronize()
    n = 6 * len(xs)
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for i in range(n): ops.layernorm(xs[i % len(xs)], g, b)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / n * 1e3
    tot += us * cnt
    print(f"M{M:7d} C{C:5d}: {us:7.1f} us  {M*C*4/us/1e3:7.1f} GB/s (4 B/el)  x{cnt}")
print(f"weighted sum {tot/1e3:.3f} ms per apply_model")
