"""-m gpu: HBM-bound kernels (K4 GroupNorm+SiLU, K8 LayerNorm, K10 DDIM/CFG, K11 timestep embedding) and helpers."""
import math

import pytest
import torch
import torch.nn.functional as F

from gpu_util import BF16, assert_close, rnd

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,HW,C,silu,eps", [(2, 4096, 320, True, 1e-5), (2, 1024, 1920, True, 1e-5),
                                             (3, 64, 2560, True, 1e-5), (2, 256, 1280, False, 1e-6),
                                             (1, 16384, 128, True, 1e-6), (2, 1024, 960, True, 1e-5),
                                             # the fused one-kernel path of the 16x16 / 8x8 levels (register-resident slices)
                                             (3, 64, 1280, True, 1e-5), (2, 256, 640, True, 1e-5),
                                             (2, 256, 1920, True, 1e-5), (2, 256, 2560, True, 1e-5),
                                             (5, 256, 320, False, 1e-5), (2, 100, 1280, True, 1e-5)])
def test_groupnorm_silu(B, HW, C, silu, eps):
    from topo_b200 import ops
    x = (rnd(B, HW, C, seed=1) * 1.5 + 0.3).to(BF16)
    g, b = rnd(C, seed=2) * 0.2 + 1, rnd(C, seed=3) * 0.1
    y = ops.groupnorm_silu(x, g, b, 32, eps, silu)
    ref = F.group_norm(x.float().permute(0, 2, 1), 32, g, b, eps)
    ref = (F.silu(ref) if silu else ref).permute(0, 2, 1)
    assert_close(y, ref, 8e-3, f"groupnorm B{B} HW{HW} C{C} silu={silu}")


@pytest.mark.parametrize("M,C", [(8192, 320), (2048, 640), (77, 1280), (513, 2048), (8190, 320), (1001, 640), (77, 1280), (231, 768)])
def test_layernorm(M, C):
    from topo_b200 import ops
    x = (rnd(M, C, seed=1) * 2 + 0.5).to(BF16)
    g, b = rnd(C, seed=2) * 0.2 + 1, rnd(C, seed=3) * 0.1
    y = ops.layernorm(x, g, b, 1e-5)
    assert_close(y, F.layer_norm(x.float(), (C,), g, b, 1e-5), 8e-3, f"layernorm {M}x{C}")


def test_timestep_embedding():
    from topo_b200 import ops
    t = torch.tensor([1.0, 34.0, 501.0, 991.0], device="cuda")
    e = ops.timestep_embedding(t, 320)
    f = torch.exp(-math.log(10000.0) * torch.arange(160, device="cuda").float() / 160)
    a = t[:, None] * f[None]
    assert_close(e, torch.cat([a.cos(), a.sin()], -1), 8e-3, "timestep embedding")


def test_ddim_step_cfg():
    from topo_b200 import ops
    n = 2 * 4 * 64 * 64
    x, eps = rnd(n, seed=1), rnd(2 * n, seed=2)
    a_t, a_prev, w = 0.31, 0.42, 7.5
    y = ops.ddim_step(x, eps, True, w, a_t, a_prev)
    e = eps[:n] + w * (eps[n:] - eps[:n])
    x0 = (x - math.sqrt(1 - a_t) * e) / math.sqrt(a_t)
    assert_close(y, math.sqrt(a_prev) * x0 + math.sqrt(1 - a_prev) * e, 1e-5, "ddim step (cfg)")
    y1 = ops.ddim_step(x, eps[:n], False, 1.0, a_t, a_prev)
    x0 = (x - math.sqrt(1 - a_t) * eps[:n]) / math.sqrt(a_t)
    assert_close(y1, math.sqrt(a_prev) * x0 + math.sqrt(1 - a_prev) * eps[:n], 1e-5, "ddim step (no cfg)")


@pytest.mark.parametrize("Cin,Cout,k,s,H", [(3, 16, 3, 1, 64), (16, 32, 3, 2, 64), (96, 256, 3, 2, 32), (4, 320, 3, 1, 64),
                                            (4, 4, 1, 1, 64), (128, 3, 3, 1, 32)])
def test_conv_direct(Cin, Cout, k, s, H):
    from topo_b200 import ops
    x = rnd(2, H, H, Cin, seed=1).to(BF16)
    w = rnd(Cout, Cin, k, k, seed=2, scale=(k * k * Cin) ** -0.5).to(BF16)
    b = rnd(Cout, seed=3)
    y = ops.conv_direct(x, ops.conv_direct_weight(w), b, Cin, Cout, k, s, act=ops.ACT_SILU)
    ref = F.silu(F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, stride=s, padding=k // 2))
    assert_close(y.float().permute(0, 3, 1, 2), ref, 1e-2, f"conv_direct {Cin}->{Cout} k{k} s{s}")


def test_layout_upsample_im2col_add_u8_softmax():
    from topo_b200 import ops
    x = rnd(2, 5, 16, 24, seed=1)
    nhwc = ops.nchw_to_nhwc(x)
    assert_close(nhwc, x.permute(0, 2, 3, 1), 8e-3, "nchw->nhwc")
    assert_close(ops.nhwc_to_nchw(nhwc), nhwc.float().permute(0, 3, 1, 2), 0, "nhwc->nchw")
    a = rnd(2, 8, 8, 64, seed=2).to(BF16)
    up = ops.upsample2x(a)
    assert_close(up.float().permute(0, 3, 1, 2), F.interpolate(a.float().permute(0, 3, 1, 2), scale_factor=2), 0, "upsample2x")
    cols = ops.im2col3x3_s2(a)
    w = rnd(32, 64, 3, 3, seed=3).to(BF16)
    ref = F.conv2d(a.float().permute(0, 3, 1, 2), w.float(), stride=2, padding=1).permute(0, 2, 3, 1).reshape(-1, 32)
    assert_close(cols.float() @ ops.conv3x3_weight(w).float().t(), ref, 1e-5, "im2col s2")
    b2 = rnd(2, 8, 8, 64, seed=4).to(BF16)
    assert_close(ops.add(a, b2), (a.float() + b2.float()), 8e-3, "add")
    img = rnd(1, 16, 16, 3, seed=5) * 0.8
    u8 = ops.tile_to_u8(img)
    ref8 = (img * 127.5 + 127.5).clamp(0, 255).to(torch.uint8)
    assert (u8.int() - ref8.int()).abs().max() <= 1
    s = rnd(300, 1000, seed=6).to(BF16)
    assert_close(ops.softmax_rows(s, 0.25), torch.softmax(s.float() * 0.25, -1), 1e-2, "softmax rows")
