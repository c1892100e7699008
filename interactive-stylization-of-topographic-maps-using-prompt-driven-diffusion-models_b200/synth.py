"""Seeded synthetic inputs for the BASELINE.json configs: weights, hint tiles, contexts, x_T.
Everything is drawn from CPU generators so the oracle and the CUDA path see bit-identical inputs."""
import torch

PALETTE = [(255, 255, 255), (34, 139, 34), (0, 0, 255), (128, 128, 128), (255, 0, 0), (255, 255, 0), (139, 69, 19),
           (0, 255, 255)]


def make_hint(B, size=512, seed=7):
    """Colour-coded 'seg' tile in [0,1]: 8-16 axis-aligned rectangles from a map palette on white, NCHW fp32."""
    out = torch.ones(B, 3, size, size)
    for b in range(B):
        g = torch.Generator().manual_seed(seed + b)
        n = int(torch.randint(8, 17, (1,), generator=g))
        for _ in range(n):
            x0, y0 = (int(v) for v in torch.randint(0, size - 8, (2,), generator=g))
            w, h = (int(v) for v in torch.randint(8, size // 2, (2,), generator=g))
            col = PALETTE[int(torch.randint(0, len(PALETTE), (1,), generator=g))]
            for c in range(3):
                out[b, c, y0:y0 + h, x0:x0 + w] = col[c] / 255.0
    return out


def make_ctx(B, dim=768, seed=11, tokens=77):
    """Adapter-like context: LayerNorm of Gaussian rows (the adapter MLP ends in LayerNorm, …_ctrlora.py:153-158)."""
    rows = []
    for b in range(B):
        g = torch.Generator().manual_seed(seed + b)
        rows.append(torch.nn.functional.layer_norm(torch.randn(tokens, dim, generator=g), (dim,)))
    return torch.stack(rows)


def make_xT(B, size=64, seed=1234):
    return torch.randn(B, 4, size, size, generator=torch.Generator().manual_seed(seed))


def build_model(variant="ctrlora", seed=0, lora_seed=2, **overrides):
    """Random-init model of a reference variant with every all-zero tensor (zero-convs, proj_out, ResBlock out convs,
    LoRA up/B) re-drawn from N(0, 0.02^2) so parity is not vacuous."""
    from topo_b200 import ldm
    torch.manual_seed(seed)
    model = ldm.create_model(variant, **overrides)
    ldm.randomize_zero_modules(model, seed=lora_seed)
    return model.eval()


def lora_scales(model):
    """prefix -> scaling for the oracle (LoRALinear.scaling / LoRACompatibleLinear.scale)."""
    out = {}
    for name, m in model.named_modules():
        if hasattr(m, "base") and hasattr(m, "A") and hasattr(m, "B"):
            out[name] = float(m.scaling)
        elif getattr(m, "lora_layer", None) is not None:
            out[name] = float(m.scale)
    return out
