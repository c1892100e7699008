"""smoke(): one small invocation of the hot path on cuda:0 — a LoRA-wrapped self-attention + cross-attention block
(fused q/k/v LoRA projection GEMM → flash attention → LoRA out projection) and a GroupNorm+SiLU → 3x3 conv, each
checked against the oracle (``oracle/ldm_oracle.py``, imported here only as the checker)."""
from __future__ import annotations

import torch


def run() -> None:
    from oracle import ldm_oracle as O
    from . import ldm, lora, ops
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    attn = ldm.CrossAttention(320, 768, heads=8, dim_head=40)
    lora.lora_attn_qkvo(attn, r=8)
    for m in attn.modules():
        if isinstance(m, lora.LoRALinear):
            torch.nn.init.normal_(m.B.weight, std=0.05)
    attn = attn.to(dev)
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 1024, 320, generator=g).to(dev)
    ctx = torch.randn(2, 77, 768, generator=g).to(dev)
    y = attn(x.to(torch.bfloat16), ctx.to(torch.bfloat16)).float()
    sd = {"a." + k: v.float() for k, v in attn.state_dict().items()}
    scales = {"a.to_q": 4.0, "a.to_k": 4.0, "a.to_v": 4.0, "a.to_out.0": 4.0}
    ref = O.cross_attention(sd, "a", x.to(torch.bfloat16).float(), ctx.to(torch.bfloat16).float(), 8, scales)
    err = (y - ref).abs().max().item() / ref.abs().max().item()
    assert err < 2e-2, f"smoke: LoRA cross-attention rel err {err:.3e}"

    gn = torch.nn.GroupNorm(32, 320).to(dev)
    conv = torch.nn.Conv2d(320, 320, 3, padding=1).to(dev)
    xi = torch.randn(2, 16, 16, 320, generator=g).to(dev).to(torch.bfloat16)
    h = ops.groupnorm_silu(xi, gn.weight.float(), gn.bias.float(), 32, 1e-5, True)
    yo = ops.conv3x3(h, conv.weight, conv.bias).float().permute(0, 3, 1, 2)
    sd2 = {"n.weight": gn.weight, "n.bias": gn.bias, "c.weight": conv.weight, "c.bias": conv.bias}
    r2 = O.conv(sd2, "c", torch.nn.functional.silu(O.group_norm(sd2, "n", xi.float().permute(0, 3, 1, 2), 1e-5)), padding=1)
    err2 = (yo - r2).abs().max().item() / r2.abs().max().item()
    assert err2 < 2e-2, f"smoke: GroupNorm+SiLU+conv3x3 rel err {err2:.3e}"
    torch.cuda.synchronize()
    print(f"[smoke] LoRA attention rel err {err:.2e}; GN+SiLU+conv3x3 rel err {err2:.2e} — OK")
