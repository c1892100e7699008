"""-m gpu: the tcgen05 projection GEMM family (K1/K5/K6/K7/K9) through the C ABI vs fp32 torch math.
Tolerance: outputs are bf16 (8 mantissa bits) of fp32-accumulated bf16 products -> relative 1e-2 of the
tensor's max magnitude (the LoRA branch adds one extra bf16 rounding of s*T)."""
import pytest
import torch
import torch.nn.functional as F

from gpu_util import BF16, assert_close, bf, rnd

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _ops():
    from topo_b200 import ops
    return ops


@pytest.mark.parametrize("M,K,N", [(128, 64, 128), (300, 320, 320), (8192, 320, 960), (154, 768, 384), (64, 1280, 1280)])
def test_gemm_plain(M, K, N):
    ops = _ops()
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    y = ops.linear(x, W)
    assert_close(y, x.float() @ W.float().t(), TOL, f"gemm {M}x{K}x{N}")


def test_gemm_bias_res_silu_fp32out():
    ops = _ops()
    M, K, N = 520, 640, 328
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    b, res = rnd(N, seed=3), rnd(M, N, seed=4).to(BF16)
    y = ops.linear(x, W, b, res=res, act=ops.ACT_SILU, alpha=0.7, out_fp32=True)
    ref = 0.7 * F.silu(x.float() @ W.float().t() + b) + res.float()
    assert y.dtype == torch.float32
    assert_close(y, ref, TOL, "gemm bias+silu+alpha+res fp32")


def test_gemm_geglu():
    ops = _ops()
    M, K, inner = 1000, 320, 1280
    x, W, b = rnd(M, K, seed=1).to(BF16), rnd(2 * inner, K, seed=2, scale=K ** -0.5).to(BF16), rnd(2 * inner, seed=3)
    y = ops.linear(x, W, b, act=ops.ACT_GEGLU)
    h = x.float() @ W.float().t() + b
    a, g = h.chunk(2, -1)
    assert y.shape == (M, inner)
    assert_close(y, a * F.gelu(g), TOL, "gemm GEGLU")


@pytest.mark.parametrize("r", [4, 12, 16, 32])
def test_lora_linear(r):
    """y = base(x) + B(A(x)) * s  — LoRA_modified/LoRA_utils.py:40-42."""
    ops = _ops()
    M, K, N, s = 1024 + 40, 320, 320, 2.0
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    A, Bm = rnd(r, K, seed=3, scale=K ** -0.5).to(BF16), rnd(N, r, seed=4, scale=0.3).to(BF16)
    b = rnd(N, seed=5)
    y = ops.linear(x, W, b, A, Bm, s)
    xf = x.float()
    ref = xf @ W.float().t() + b + s * ((xf @ A.float().t()) @ Bm.float().t())
    base = xf @ W.float().t() + b
    assert (ref - base).abs().max() > 0.1, "LoRA delta must be non-trivial for the test to mean anything"
    assert_close(y, ref, TOL, f"lora linear r={r}")


def test_fused_qkv_three_segments():
    """n_seg=3: one read of x, three (W, A, B, out) segments written into one [M, 3N] buffer."""
    from topo_b200 import capi
    ops = _ops()
    M, K, N, r = 2048, 320, 384, 8
    x = rnd(M, K, seed=1).to(BF16)
    out = torch.zeros(M, 3 * N, dtype=BF16, device="cuda")
    segs, refs, keep = [], [], []
    for i in range(3):
        W = rnd(N, K, seed=10 + i, scale=K ** -0.5).to(BF16)
        A, Bm = rnd(r, K, seed=20 + i, scale=K ** -0.5), rnd(N, r, seed=30 + i, scale=0.3)
        Ap, Bp, rp = ops.prep_lora(A, Bm)
        keep += [W, Ap, Bp]
        segs.append(capi.make_linear(W, out, N, A=Ap, B=Bp, r_pad=rp, s=2.0, out_ld=3 * N, out_col=i * N))
        xf = x.float()
        refs.append(xf @ W.float().t() + 2.0 * ((xf @ bf(A).t()) @ bf(Bm).t()))
    ops.lora_linear_raw(x, M, K, K, segs)
    assert_close(out, torch.cat(refs, 1), TOL, "fused qkv (3 segments)")


def test_zeroconv_residual_into_concat_buffer():
    """(e) out[:, off:off+C] = skip + scale*(h Wz^T + bz), other columns untouched."""
    ops = _ops()
    M, Cc, off, ld = 2048, 320, 640, 960
    h, skip = rnd(M, Cc, seed=1).to(BF16), rnd(M, Cc, seed=2).to(BF16)
    Wz, bz = rnd(Cc, Cc, seed=3, scale=0.05).to(BF16), rnd(Cc, seed=4)
    out = torch.full((M, ld), 7.0, dtype=BF16, device="cuda")
    ops.zeroconv_residual(h, Wz, bz, 0.5, skip, out, off)
    ref = skip.float() + 0.5 * (h.float() @ Wz.float().t() + bz)
    assert_close(out[:, off:off + Cc], ref, TOL, "zero-conv residual")
    assert (out[:, :off] == 7.0).all(), "columns outside the slice were touched"


@pytest.mark.parametrize("B,H,W,Cin,Cout", [(2, 64, 64, 320, 320), (2, 8, 8, 1280, 1280), (3, 16, 16, 640, 1280),
                                            (1, 32, 32, 1920, 640), (1, 128, 128, 128, 128), (1, 64, 64, 320, 4)])
def test_conv3x3_implicit_gemm(B, H, W, Cin, Cout):
    ops = _ops()
    x = rnd(B, H, W, Cin, seed=1).to(BF16)
    w = rnd(Cout, Cin, 3, 3, seed=2, scale=(9 * Cin) ** -0.5).to(BF16)
    b = rnd(Cout, seed=3)
    emb = rnd(B, Cout, seed=4)
    y = ops.conv3x3(x, w, b, rowbias=emb)
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, padding=1) + emb[:, :, None, None]
    assert_close(y.float().permute(0, 3, 1, 2), ref, TOL, f"conv3x3 {B}x{H}x{W} {Cin}->{Cout}")


@pytest.mark.parametrize("M,K,N,r,nseg,res", [
    (148 * 128 + 70, 320, 320, 16, 1, True),     # 160-wide LoRA tiles, contiguous chunks that cross (row tile) groups, ragged M
    (148 * 128 * 2, 320, 384, 8, 3, False),      # fused q/k/v, 192-wide LoRA tiles: groups change with the SEGMENT inside a row tile
    (9728, 1280, 1280, 32, 1, True),             # r_pad 32, K = 20 blocks (pre-pass longer than the ring), 8 n tiles per group
    (38912, 640, 640, 4, 3, False),              # 160-wide, three segments
    (640, 384, 200, 12, 1, True),                # 128-wide fallback (N not a multiple of 160/192), ragged N tile, strided walk
    (128, 64, 16, 16, 1, False),                 # one tile, one K block
])
def test_lora_prepass_walks(M, K, N, r, nseg, res):
    """K1 round 2: T = x A^T comes from a pre-pass issued one tile ahead of each (row tile, segment) group and every tile
    starts from (sT) B^T.  Shapes chosen so that a CTA's tile chunk begins / ends in the middle of a group, groups change by
    segment and by row tile, the pre-pass is longer than the stage ring, and the strided (one tile per CTA) walk is hit."""
    from topo_b200 import capi
    ops = _ops()
    x = rnd(M, K, seed=1).to(BF16)
    out = torch.zeros(M, nseg * N, dtype=BF16, device="cuda")
    rs = rnd(M, N, seed=2).to(BF16) if res else None
    segs, refs, keep = [], [], []
    xf = x.float()
    for i in range(nseg):
        W = rnd(N, K, seed=10 + i, scale=K ** -0.5).to(BF16)
        A, Bm = rnd(r, K, seed=20 + i, scale=K ** -0.5), rnd(N, r, seed=30 + i, scale=0.3)
        Ap, Bp, rp = ops.prep_lora(A, Bm)
        b = rnd(N, seed=40 + i)
        keep += [W, Ap, Bp, b]
        segs.append(capi.make_linear(W, out, N, bias=b, A=Ap, B=Bp, r_pad=rp, s=1.5, out_ld=nseg * N, out_col=i * N,
                                     res=rs if nseg == 1 else None, res_ld=0 if rs is None else rs.stride(0)))
        ref = xf @ W.float().t() + b + 1.5 * ((xf @ bf(A).t()) @ bf(Bm).t())
        delta = 1.5 * ((xf @ bf(A).t()) @ bf(Bm).t())
        assert delta.abs().max() > 0.05 * ref.abs().max()
        refs.append(ref + (rs.float() if (res and nseg == 1) else 0))
    ops.lora_linear_raw(x, M, K, K, segs)
    print(f"[parity] LoRA pre-pass M{M} K{K} N{nseg}x{N} r{r}: {capi.last_kernel()}")
    assert_close(out, torch.cat(refs, 1), TOL, f"LoRA pre-pass M{M} K{K} N{nseg}x{N} r{r}")


def test_fused_qkv_plain_192_wide_tiles():
    """Three bias-free segments of 384 columns at large M: the 192-wide tile path of the level-0 q/k/v projection."""
    from topo_b200 import capi
    ops = _ops()
    M, K, N = 148 * 128 + 5, 320, 384
    x = rnd(M, K, seed=1).to(BF16)
    out = torch.zeros(M, 3 * N, dtype=BF16, device="cuda")
    Ws = [rnd(N, K, seed=10 + i, scale=K ** -0.5).to(BF16) for i in range(3)]
    segs = [capi.make_linear(W, out, N, out_ld=3 * N, out_col=i * N) for i, W in enumerate(Ws)]
    ops.lora_linear_raw(x, M, K, K, segs)
    assert capi.last_kernel() in ("gemm_resident_kernel<192, 4, -5>", "gemm_kernel<192, 0, 4, 0, 0>"), capi.last_kernel()
    assert_close(out, torch.cat([x.float() @ W.float().t() for W in Ws], 1), TOL, "fused qkv, 192-wide tiles")


# ---- short-K projections with one operand resident in shared memory (gemm_resident_kernel) -------------------------------
@pytest.mark.parametrize("M,K,N,res,kernel", [
    (148 * 128 * 2 + 77, 320, 320, True, "gemm_resident_kernel<160, 5, 5>"),     # attention out-projection / proj_out, 64x64 level
    (148 * 128 + 1, 384, 320, True, "gemm_resident_kernel<160, 4, 6>"),          # out-projection from padded heads (8 x 48)
    (148 * 128 + 130, 256, 320, False, "gemm_resident_kernel<160, 5, 5>"),       # fewer k-blocks than slots
    (148 * 128 + 64, 320, 384, False, "gemm_resident_kernel<192, 4, 5>"),        # padded-head q projection
    (148 * 128 * 2, 320, 160, True, "gemm_resident_kernel<160, 5, 5>"),          # one n tile: W fully resident
])
def test_resident_w_projection(M, K, N, res, kernel):
    """W tile loaded once per CTA, x streamed; every CTA keeps one n tile over a strided row-tile walk (ragged M tail,
    bias, residual).  Also run twice: the second launch must not depend on anything the first left in shared memory."""
    from topo_b200 import capi
    ops = _ops()
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    b = rnd(N, seed=3)
    rs = rnd(M, N, seed=4).to(BF16) if res else None
    ref = x.float() @ W.float().t() + b + (rs.float() if res else 0)
    for _ in range(2):
        y = ops.linear(x, W, b, res=rs)
        assert capi.last_kernel() == kernel, capi.last_kernel()
        assert_close(y, ref, TOL, f"resident-W M{M} K{K} N{N}")


def test_resident_x_geglu():
    """GEGLU of the 64x64 level (N = 2560, K = 320): the x tile of a row tile stays in shared memory while the CTA walks
    that row tile's ten n tiles; chunk boundaries fall inside row tiles (1480 + 3 row tiles over 148 CTAs)."""
    from topo_b200 import capi
    ops = _ops()
    M, K, inner = 148 * 128 + 300, 320, 1280
    x, W, b = rnd(M, K, seed=1).to(BF16), rnd(2 * inner, K, seed=2, scale=K ** -0.5).to(BF16), rnd(2 * inner, seed=3)
    h = x.float() @ W.float().t() + b
    a, g = h.chunk(2, -1)
    for _ in range(2):
        y = ops.linear(x, W, b, act=ops.ACT_GEGLU)
        assert capi.last_kernel() == "gemm_resident_kernel<256, 3, -5>", capi.last_kernel()
        assert_close(y, a * F.gelu(g), TOL, "resident-x GEGLU")


def test_resident_x_fused_qkv_with_bias_and_ragged_tail():
    from topo_b200 import capi
    ops = _ops()
    M, K, N = 148 * 128 + 129, 320, 384
    x = rnd(M, K, seed=1).to(BF16)
    out = torch.zeros(M, 3 * N, dtype=BF16, device="cuda")
    Ws = [rnd(N, K, seed=10 + i, scale=K ** -0.5).to(BF16) for i in range(3)]
    bs = [rnd(N, seed=20 + i) for i in range(3)]
    segs = [capi.make_linear(W, out, N, bias=b, out_ld=3 * N, out_col=i * N) for i, (W, b) in enumerate(zip(Ws, bs))]
    ops.lora_linear_raw(x, M, K, K, segs)
    assert capi.last_kernel() == "gemm_resident_kernel<192, 4, -5>", capi.last_kernel()
    assert_close(out, torch.cat([x.float() @ W.float().t() + b for W, b in zip(Ws, bs)], 1), TOL, "resident-x fused qkv")


@pytest.mark.parametrize("M,K,N,kernel", [
    (9728, 1280, 1280, "gemm_kernel<256, 0, 4, 0, 0>"),      # 16x16 level: 256-wide tiles over three segments
    (38912, 640, 640, "gemm_kernel<160, 0, 5, 0, 0>"),       # 32x32 level: four 160-wide tiles per segment
])
def test_fused_qkv_wide_tiles(M, K, N, kernel):
    from topo_b200 import capi
    ops = _ops()
    x = rnd(M, K, seed=1).to(BF16)
    out = torch.zeros(M, 3 * N, dtype=BF16, device="cuda")
    Ws = [rnd(N, K, seed=10 + i, scale=K ** -0.5).to(BF16) for i in range(3)]
    segs = [capi.make_linear(W, out, N, out_ld=3 * N, out_col=i * N) for i, W in enumerate(Ws)]
    ops.lora_linear_raw(x, M, K, K, segs)
    assert capi.last_kernel() == kernel, capi.last_kernel()
    assert_close(out, torch.cat([x.float() @ W.float().t() for W in Ws], 1), TOL, f"fused qkv {M}x{K}x3x{N}")


def test_conv3x3_single_wave_192_wide_pair_tiles():
    """8x8 level at the bench batch (38 x 64 = 19 row tiles): 70 pair tiles of 192 columns instead of 50 of 256."""
    from topo_b200 import capi
    ops = _ops()
    B, H, W, Cin, Cout = 38, 8, 8, 1280, 1280
    x = rnd(B, H, W, Cin, seed=1).to(BF16)
    w = rnd(Cout, Cin, 3, 3, seed=2, scale=(9 * Cin) ** -0.5).to(BF16)
    b = rnd(Cout, seed=3)
    y = ops.conv3x3(x, w, b)
    assert capi.last_kernel() == "gemm_kernel<192, 0, 6, 1, 1>", capi.last_kernel()
    ref = F.conv2d(x.float().permute(0, 3, 1, 2), w.float(), b, padding=1).permute(0, 2, 3, 1)
    assert_close(y, ref, TOL, "conv3x3 8x8 192-wide pair tiles")


@pytest.mark.parametrize("M,K,N,res", [
    (2432, 1280, 1280, True),        # 8x8 level: 10 x 7 pair tiles in one wave (tail n tile 128 wide)
    (2432 - 70, 5120, 1280, False),  # ragged M, long K
    (148 * 128 + 33, 1280, 320, True),   # feed-forward output of the 64x64 level: 192 + 128 columns
])
def test_pair_tiles_192_wide_projection(M, K, N, res):
    from topo_b200 import capi
    ops = _ops()
    x, W = rnd(M, K, seed=1).to(BF16), rnd(N, K, seed=2, scale=K ** -0.5).to(BF16)
    b = rnd(N, seed=3)
    rs = rnd(M, N, seed=4).to(BF16) if res else None
    y = ops.linear(x, W, b, res=rs)
    assert capi.last_kernel() == "gemm_kernel<192, 0, 6, 0, 1>", capi.last_kernel()
    assert_close(y, x.float() @ W.float().t() + b + (rs.float() if res else 0), TOL, f"192-wide pair tiles M{M} K{K} N{N}")
