python - <<'PY'
import sys, os
sys.path[:0] = [os.getcwd(), os.path.join(os.getcwd(), "tests")]
import torch
from topo_b200 import ops
BF = torch.bfloat16
def t(B, N, Nk, d):
    dp = ops.head_pad(d); h = 8
    q = torch.randn(B, N, h * dp, device="cuda").to(BF); k = torch.randn(B, Nk, h * dp, device="cuda").to(BF); v = torch.randn(B, Nk, h * dp, device="cuda").to(BF)
    for _ in range(3): ops.attention(q, k, v, h, dp, d ** -0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(20): ops.attention(q, k, v, h, dp, d ** -0.5)
    e1.record(); torch.cuda.synchronize()
    print(f"B{B} N{N} Nk{Nk} d{d}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
for Nk in (64, 77, 128):
    t(38, 4096, Nk, 40)
for Nk in (64, 77, 128):
    t(38, 1024, Nk, 80)
PY
