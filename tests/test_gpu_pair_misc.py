"""-m gpu: the CTA-pair (cta_group::2) GEMM / conv kernels at the sizes whose dispatch selects them (odd m-tile counts,
N tails, GEGLU pairs), and the host-visible ops added with the text encoder / ControlLoRA pyramid (key-length masked
attention, BERT embedding sum, the one-K-block im2col of the narrow input convs, Downsample2D(padding=0))."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import BF16, assert_close, rnd

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _ops():
    from topo_b200 import ops
    return ops


@pytest.mark.parametrize("M,K,N", [(20000, 320, 640), (155648, 320, 320), (38912, 640, 2560), (20096, 1280, 1000),
                                   (9728, 5120, 1280), (40000, 1280, 320), (155648, 1280, 320)])
def test_pair_gemm(M, K, N):
    ops = _ops()
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    b, res = rnd(N, seed=3), rnd(M, N, seed=4).to(BF16)
    y = ops.linear(x, W, b, res=res, alpha=0.5)
    ref = 0.5 * (x.float() @ W.float().t() + b) + res.float()
    assert_close(y, ref, TOL, f"pair gemm {M}x{K}x{N}")


def test_pair_gemm_geglu():
    ops = _ops()
    M, K, inner = 40000, 1280, 1280
    x, W, b = rnd(M, K, seed=1).to(BF16), rnd(2 * inner, K, seed=2, scale=K ** -0.5).to(BF16), rnd(2 * inner, seed=3)
    y = ops.linear(x, W, b, act=ops.ACT_GEGLU)
    a, g = (x.float() @ W.float().t() + b).chunk(2, -1)
    assert_close(y, a * F.gelu(g), TOL, "pair gemm GEGLU")


@pytest.mark.parametrize("B,H,C,N", [(38, 32, 128, 320), (19, 64, 64, 256), (5, 128, 64, 96), (38, 8, 320, 1280),
                                     (38, 16, 192, 640), (19, 64, 320, 320), (38, 64, 64, 320)])
def test_pair_conv3x3(B, H, C, N):
    ops = _ops()
    x = rnd(B, H, H, C, seed=1).to(BF16)
    w, b = rnd(N, C, 3, 3, seed=2, scale=(9 * C) ** -0.5).to(BF16), rnd(N, seed=3)
    res = rnd(B, H, H, N, seed=4).to(BF16)
    y = ops.conv3x3(x, w.float(), b, res=res)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, padding=1).permute(0, 2, 3, 1) + res.float()
    assert_close(y, ref, TOL, f"pair conv B{B} H{H} C{C} N{N}")


@pytest.mark.parametrize("heads,d,Nq,Nk", [(12, 64, 77, 77), (8, 40, 300, 200), (4, 64, 128, 257)])
def test_attention_key_length(heads, d, Nq, Nk):
    """softmax over the first kv_len[b] keys only (HF BERT's additive -inf mask of a right-padded prompt)."""
    ops = _ops()
    B = 5
    dp = ops.head_pad(d)
    kv = torch.tensor([Nk, 1, Nk // 2, 64, min(Nk, 65)], dtype=torch.int32, device="cuda")
    def mk(n, seed):
        t = torch.zeros(B, n, heads, dp, device="cuda")
        t[..., :d] = rnd(B, n, heads, d, seed=seed)
        return t.view(B, n, heads * dp).to(BF16)
    q, k, v = mk(Nq, 1), mk(Nk, 2), mk(Nk, 3)
    o = ops.attention(q, k, v, heads, dp, d ** -0.5, kv_len=kv).view(B, Nq, heads, dp)[..., :d]
    qf, kf, vf = (t.float().view(B, -1, heads, dp)[..., :d].transpose(1, 2) for t in (q, k, v))
    s = qf @ kf.transpose(-1, -2) * d ** -0.5
    mask = torch.arange(Nk, device="cuda")[None, None, None, :] >= kv[:, None, None, None]
    ref = (torch.softmax(s.masked_fill(mask, float("-inf")), -1) @ vf).transpose(1, 2)
    assert_close(o, ref, 2e-2, f"attention kv_len h{heads} d{d} {Nq}x{Nk}")


def test_embed_sum():
    ops = _ops()
    B, T, C, V = 7, 77, 768, 1000
    ids = torch.randint(0, V, (B, T), generator=torch.Generator().manual_seed(0)).cuda()
    word, pos, typ = rnd(V, C, seed=1), rnd(512, C, seed=2), rnd(C, seed=3)
    y = ops.embed_sum(ids, word, pos, typ)
    ref = word[ids] + pos[:T][None] + typ
    assert_close(y, ref, 5e-3, "embed_sum")


@pytest.mark.parametrize("C", [3, 4, 7])
def test_im2col_k64_gemm_is_conv(C):
    ops = _ops()
    B, H, N = 3, 40, 320
    x = rnd(B, H, H, C, seed=1).to(BF16)
    w = rnd(N, C, 3, 3, seed=2, scale=0.2).to(BF16)
    cols = ops.im2col3x3_s1_k64(x)
    assert cols.shape == (B * H * H, 64) and float(cols[:, 9 * C:].abs().max()) == 0.0
    Wm = torch.zeros(N, 64, device="cuda", dtype=BF16)
    Wm[:, :9 * C] = w.permute(0, 2, 3, 1).reshape(N, 9 * C)
    y = ops.linear(cols, Wm).view(B, H, H, N)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), padding=1).permute(0, 2, 3, 1)
    assert_close(y, ref, TOL, f"im2col_k64 C{C}")


def test_conv_direct_pad_after_only():
    """diffusers Downsample2D(padding=0): F.pad(x, (0,1,0,1)) then a stride-2 3x3 conv without padding."""
    ops = _ops()
    B, H, C, N = 2, 32, 32, 32
    x = rnd(B, H, H, C, seed=1).to(BF16)
    w, b = rnd(N, C, 3, 3, seed=2, scale=(9 * C) ** -0.5).to(BF16), rnd(N, seed=3)
    y = ops.conv_direct(x, ops.conv_direct_weight(w), b, C, N, 3, 2, pad_lo=0)
    ref = F.conv2d(F.pad(x.float().permute(0, 3, 1, 2), (0, 1, 0, 1)), w.float(), b, stride=2).permute(0, 2, 3, 1)
    assert y.shape == ref.shape
    assert_close(y, ref, TOL, "conv_direct pad_lo=0 stride 2")


def test_im2col_s2_pad_after_gemm_is_vae_downsample():
    """ldm VAE-encoder Downsample: F.pad(x, (0,1,0,1)) + Conv2d(C, C, 3, stride=2, padding=0) as im2col(pad_lo=0) + GEMM."""
    ops = _ops()
    B, H, C, N = 2, 64, 128, 128
    x = rnd(B, H, H, C, seed=1).to(BF16)
    w, b = rnd(N, C, 3, 3, seed=2, scale=(9 * C) ** -0.5).to(BF16), rnd(N, seed=3)
    y = ops.linear(ops.im2col3x3_s2(x, pad_lo=0), ops.conv3x3_weight(w), b).view(B, H // 2, H // 2, N)
    ref = F.conv2d(F.pad(x.float().permute(0, 3, 1, 2), (0, 1, 0, 1)), w.float(), b, stride=2).permute(0, 2, 3, 1)
    assert_close(y, ref, TOL, "im2col_s2 pad_lo=0 + gemm")
    y1 = ops.linear(ops.im2col3x3_s2(x), ops.conv3x3_weight(w), b).view(B, H // 2, H // 2, N)
    ref1 = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, stride=2, padding=1).permute(0, 2, 3, 1)
    assert_close(y1, ref1, TOL, "im2col_s2 pad_lo=1 + gemm")
