"""Cross-attention core in isolation at the bench shapes (Be = 38, 77 keys): us per launch, algorithmic GB/s (q + o + k + v
bytes) against the measured HBM peak.  TOPO_XATTN=0 selects the flash kernel (round-1 path) for an A/B run."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT]
import torch
from topo_b200 import capi, ops

pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6650.0
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for B, N, d, dp in ((38, 4096, 40, 48), (19, 4096, 40, 48), (38, 1024, 80, 80), (38, 256, 160, 160)):
    h = 8
    g = torch.Generator(device="cuda").manual_seed(0)
    q = torch.randn(B, N, h * dp, device="cuda", generator=g).bfloat16()
    k = torch.randn(B, 77, h * dp, device="cuda", generator=g).bfloat16()
    v = torch.randn(B, 77, h * dp, device="cuda", generator=g).bfloat16()
    o = torch.empty_like(q)
    for _ in range(3):
        ops.attention(q, k, v, h, dp, d ** -0.5, out=o)
    ts = []
    for _ in range(10):
        flush.zero_()                       # L2 flush between timed launches (256 MB > 126 MB L2)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ops.attention(q, k, v, h, dp, d ** -0.5, out=o); e1.record()
        torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort(); us = ts[len(ts) // 2]
    nbytes = 2.0 * (2 * B * N * h * dp + 2 * B * 77 * h * dp)
    print(f"B{B} N{N} d{d}: {us:7.1f} us  {nbytes / us / 1e3:7.1f} GB/s = {nbytes / us / 1e3 / pk:.2f} of measured HBM peak "
          f"({4.0 * B * h * N * 77 * d / us / 1e6:6.1f} TFLOP/s)  {capi.last_kernel()}")
