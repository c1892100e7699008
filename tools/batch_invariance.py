"""Is every kernel batch-invariant (op(cat([x, x]))[:half] == op(x) bit for bit)?  Needed by the CFG shared-prefix path."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import torch
from topo_b200 import ops
from topo_b200 import engine as E
BF = torch.bfloat16
torch.manual_seed(0)


def chk(name, a, b):
    d = (a.float() - b.float()).abs().max().item()
    print(f"{name:44s} max|diff| = {d:.3e}  {'OK' if d == 0 else 'DIFFERS'}")


for (B, HW, C, N) in [(2, 256, 320, 320), (19, 4096, 320, 320), (2, 256, 320, 1280)]:
    x = torch.randn(B * HW, C, device="cuda").to(BF)
    x2 = torch.cat([x, x])
    W = (torch.randn(N, C, device="cuda") * C ** -0.5).to(BF)
    b = torch.randn(N, device="cuda")
    lin = E.Lin(W, b)
    chk(f"gemm M{B*HW} N{N} K{C}", E.gemm(x2, lin)[:B * HW], E.gemm(x, lin))
    r, r2 = torch.randn(B * HW, N, device="cuda").to(BF), None
    r2 = torch.cat([r, r])
    chk(f"gemm+res M{B*HW} N{N} K{C}", E.gemm(x2, lin, res=r2)[:B * HW], E.gemm(x, lin, res=r))
    H = int(HW ** 0.5)
    Wc = (torch.randn(N, 9 * C, device="cuda") * (9 * C) ** -0.5).to(BF)
    linc = E.Lin(Wc, b)
    chk(f"conv3 B{B} H{H} C{C} N{N}", E.conv3(x2, 2 * B, H, H, linc)[:B * HW], E.conv3(x, B, H, H, linc))
    g, be = torch.randn(C, device="cuda"), torch.randn(C, device="cuda")
    chk(f"groupnorm B{B} HW{HW} C{C}", ops.groupnorm_silu(x2.view(2 * B, HW, C), g, be, 32, 1e-5, True)[:B],
        ops.groupnorm_silu(x.view(B, HW, C), g, be, 32, 1e-5, True))
    chk(f"layernorm M{B*HW} C{C}", ops.layernorm(x2, g, be, 1e-5)[:B * HW], ops.layernorm(x, g, be, 1e-5))
    heads, d = 8, C // 8
    dp = ops.head_pad(d)
    q = torch.zeros(B, HW, heads, dp, device="cuda"); q[..., :d] = torch.randn(B, HW, heads, d, device="cuda")
    q = q.view(B, HW, heads * dp).to(BF)
    q2 = torch.cat([q, q])
    chk(f"attention B{B} N{HW} d{d}", ops.attention(q2, q2, q2, heads, dp, d ** -0.5)[:B], ops.attention(q, q, q, heads, dp, d ** -0.5))
