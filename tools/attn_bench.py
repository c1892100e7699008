"""Self-attention core in isolation at the bench shapes (Be = 38): us per launch, TFLOP/s (true head dim), with the row sum
accumulated by the softmax warps (l_col off) and taken from the P V product (ones column, l_col on).  argv[1] = batch."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
from topo_b200 import capi, ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 38
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for N, d, dp in ((4096, 40, 48), (1024, 80, 80), (256, 160, 160)):
    h = 8
    g = torch.Generator(device="cuda").manual_seed(0)
    q, k, v = (torch.randn(B, N, h, dp, device="cuda", generator=g) for _ in range(3))
    for t in (q, k, v):
        t[..., d:] = 0
    v1 = v.clone()
    if dp > d:
        v1[..., d] = 1.0
    q, k, v, v1 = (t.reshape(B, N, h * dp).bfloat16() for t in (q, k, v, v1))
    for lcol in ([-1, d] if dp == 48 else [-1]):
        vv = v1 if lcol >= 0 else v
        for _ in range(2):
            ops.attention(q, k, vv, h, dp, d ** -0.5, l_col=lcol)
        ts = []
        for _ in range(7):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ops.attention(q, k, vv, h, dp, d ** -0.5, l_col=lcol); e1.record()
            torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
        us = sorted(ts)[len(ts) // 2]
        print(f"B{B} N{N} d{d} l_col={lcol:3d}: {us:8.1f} us  {4.0 * B * h * N * N * d / us / 1e6:7.1f} TFLOP/s  {capi.last_kernel()}", flush=True)
