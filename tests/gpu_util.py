"""Shared helpers for the -m gpu parity tests (CUDA path vs fp32 torch math on the same bf16-rounded inputs)."""
import torch

BF16 = torch.bfloat16


def rnd(*shape, seed=0, scale=1.0, device="cuda"):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(device)


def bf(x):
    """bf16-round but keep fp32 (what the kernel actually sees)."""
    return x.to(BF16).float()


def assert_close(got, ref, tol, what=""):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    mag = ref.abs().max().item() + 1e-12
    rel = err / mag
    print(f"[parity] {what}: max_abs={err:.3e} ref_max={mag:.3e} rel={rel:.3e} tol={tol:.1e}")
    assert torch.isfinite(got).all(), f"{what}: non-finite output"
    assert rel <= tol, f"{what}: rel err {rel:.3e} > {tol:.1e} (max_abs {err:.3e})"
