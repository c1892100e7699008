"""Correctness of the CTA-pair (cta_group::2) GEMM / conv kernels at sizes that select them (odd m-tile count, N tails)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
import torch.nn.functional as F
from topo_b200 import ops
from gpu_util import assert_close
BF = torch.bfloat16
torch.manual_seed(0)
for (M, K, N) in [(20000, 320, 640), (155648, 320, 320), (38912, 640, 2560), (20096, 1280, 1000)]:
    x = torch.randn(M, K, device="cuda").to(BF)
    W = (torch.randn(N, K, device="cuda") * K ** -0.5).to(BF)
    b = torch.randn(N, device="cuda")
    res = torch.randn(M, N, device="cuda").to(BF)
    y = ops.linear(x, W, b, res=res, alpha=0.5)
    ref = 0.5 * (x.float() @ W.float().t() + b) + res.float()
    assert_close(y, ref, 1e-2, f"pair gemm {M}x{K}x{N}")
M, K, inner = 40000, 320, 1280
x = torch.randn(M, K, device="cuda").to(BF); W = (torch.randn(2 * inner, K, device="cuda") * K ** -0.5).to(BF); b = torch.randn(2 * inner, device="cuda")
y = ops.linear(x, W, b, act=ops.ACT_GEGLU)
h = x.float() @ W.float().t() + b
a, g = h.chunk(2, -1)
assert_close(y, a * F.gelu(g), 1e-2, "pair gemm GEGLU")
for (B, H, C, N) in [(38, 32, 128, 320), (19, 64, 64, 256), (5, 128, 64, 96)]:
    x = torch.randn(B, H, H, C, device="cuda").to(BF)
    w = (torch.randn(N, C, 3, 3, device="cuda") * (9 * C) ** -0.5).to(BF)
    b = torch.randn(N, device="cuda")
    y = ops.conv3x3(x, w.float(), b)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, padding=1).permute(0, 2, 3, 1)
    assert_close(y, ref, 1e-2, f"pair conv B{B} H{H} C{C} N{N}")
print("pair_check OK")
