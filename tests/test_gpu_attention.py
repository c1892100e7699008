"""-m gpu: flash-style tcgen05 attention core (K2 self, K3 cross) vs explicit fp32 softmax(QK^T d^-1/2)V."""
import pytest
import torch

from gpu_util import BF16, assert_close, rnd

pytestmark = pytest.mark.gpu
TOL = 1.5e-2  # P is rounded to bf16 before P.V (as every flash kernel does) and O is stored bf16


def _ref(q, k, v, heads, d, dp, scale):
    B, Nq, _ = q.shape
    qf, kf, vf = (t.float().reshape(B, -1, heads, dp)[..., :d].permute(0, 2, 1, 3) for t in (q, k, v))
    p = torch.softmax(qf @ kf.transpose(-1, -2) * scale, -1)
    return (p @ vf).permute(0, 2, 1, 3)  # B, N, h, d


def _mk(B, N, heads, d, dp, seed):
    t = torch.zeros(B, N, heads, dp, device="cuda")
    t[..., :d] = rnd(B, N, heads, d, seed=seed)
    return t.reshape(B, N, heads * dp).to(BF16)


@pytest.mark.parametrize("B,heads,Nq,Nk,d", [
    (1, 2, 128, 128, 40), (1, 2, 256, 384, 40), (2, 8, 4096, 4096, 40), (2, 8, 1024, 1024, 80),
    (2, 8, 256, 256, 160), (2, 8, 64, 64, 160),
    (4, 8, 4096, 77, 40), (4, 8, 1024, 77, 80), (4, 8, 256, 77, 160), (2, 8, 64, 77, 160),
])
def test_attention(B, heads, Nq, Nk, d):
    from topo_b200 import ops
    dp = ops.head_pad(d)
    q, k, v = _mk(B, Nq, heads, d, dp, 1), _mk(B, Nk, heads, d, dp, 2), _mk(B, Nk, heads, d, dp, 3)
    scale = d ** -0.5
    o = ops.attention(q, k, v, heads, dp, scale)
    ref = _ref(q, k, v, heads, d, dp, scale)
    got = o.float().reshape(B, Nq, heads, dp)
    assert_close(got[..., :d], ref, TOL, f"attention B{B} h{heads} Nq{Nq} Nk{Nk} d{d}")
    if dp > d:
        assert got[..., d:].abs().max() == 0, "padded head columns must stay zero"


@pytest.mark.parametrize("gain", [6.0, 60.0, 400.0])
def test_attention_peaked_softmax(gain):
    """Large logits: online max/rescale path must stay exact (no overflow, correct renormalisation); gain 60 / 400 make
    the row maximum jump by hundreds of nats between key blocks."""
    from topo_b200 import ops
    B, heads, N, d, dp = 1, 2, 512, 40, 48
    q, k, v = _mk(B, N, heads, d, dp, 1), _mk(B, N, heads, d, dp, 2), _mk(B, N, heads, d, dp, 3)
    q = (q.float() * gain).to(BF16)
    o = ops.attention(q, k, v, heads, dp, d ** -0.5)
    ref = _ref(q, k, v, heads, d, dp, d ** -0.5)
    assert_close(o.float().reshape(B, N, heads, dp)[..., :d], ref, 3e-2, "attention peaked")


@pytest.mark.parametrize("B,heads,Nq,Nk,d", [
    (1, 1, 128, 77, 40),      # one tile, one CTA
    (1, 2, 320, 77, 40),      # ragged last query tile (Nq % 128 != 0)
    (2, 8, 4096, 80, 40),     # every padded key column is a real key
    (3, 8, 2048, 1, 40),      # a single key: softmax == 1, o == v[0]
    (2, 8, 1024, 16, 80), (2, 8, 512, 64, 160), (5, 8, 256, 77, 160), (3, 8, 1152, 77, 80),
    (38, 8, 4096, 77, 40),    # the bench shape: 9728 tiles over 296 persistent CTAs, ranges crossing (batch, head) boundaries
    (19, 8, 1024, 77, 80),
])
def test_cross_attention_resident_kv(B, heads, Nq, Nk, d):
    """K3, the K/V-resident single-pass kernel (Nk <= 80, Nq >= 128): persistent CTAs walk contiguous ranges of the flat
    (batch, head, tile) list, so ranges that start / end in the middle of a (batch, head) and K/V double buffering across
    segment boundaries are all exercised; key columns >= Nk must not leak into the softmax."""
    from topo_b200 import capi, ops
    dp = ops.head_pad(d)
    q, k, v = _mk(B, Nq, heads, d, dp, 11), _mk(B, Nk, heads, d, dp, 12), _mk(B, Nk, heads, d, dp, 13)
    scale = d ** -0.5
    o = ops.attention(q, k, v, heads, dp, scale)
    assert capi.last_kernel().startswith("cross_attention_kernel"), capi.last_kernel()
    ref = _ref(q, k, v, heads, d, dp, scale)
    got = o.float().reshape(B, Nq, heads, dp)
    assert_close(got[..., :d], ref, TOL, f"cross-attention B{B} h{heads} Nq{Nq} Nk{Nk} d{d}")
    if dp > d:
        assert got[..., d:].abs().max() == 0, "padded head columns must stay zero"
    if Nk == 1:
        assert_close(got[..., :d], v.float().reshape(B, 1, heads, dp)[..., :d].expand(B, Nq, heads, d), 1e-2, "single key")


@pytest.mark.parametrize("gain", [30.0, 400.0])
def test_cross_attention_peaked(gain):
    """Large logits on the single-pass kernel: the exact row maximum keeps exp2 in range."""
    from topo_b200 import ops
    B, heads, N, Nk, d, dp = 2, 8, 512, 77, 40, 48
    q, k, v = _mk(B, N, heads, d, dp, 1), _mk(B, Nk, heads, d, dp, 2), _mk(B, Nk, heads, d, dp, 3)
    q = (q.float() * gain).to(BF16)
    o = ops.attention(q, k, v, heads, dp, d ** -0.5)
    ref = _ref(q, k, v, heads, d, dp, d ** -0.5)
    assert_close(o.float().reshape(B, N, heads, dp)[..., :d], ref, 3e-2, f"cross-attention peaked x{gain}")


def test_cross_attention_strided_views_and_out():
    """The engine's call forms: k / v are column slices of one [B, Nk, 2*h*dp] buffer, q a slice of a shared tensor, and
    the result goes to a caller-provided slice of a larger output (run_first_block_shared)."""
    from topo_b200 import ops
    B, heads, N, Nk, d, dp = 4, 8, 1024, 77, 40, 48
    hdp = heads * dp
    kv = torch.cat([_mk(B, Nk, heads, d, dp, 21), _mk(B, Nk, heads, d, dp, 22)], dim=-1)
    k, v = kv[:, :, :hdp], kv[:, :, hdp:]
    q = _mk(B // 2, N, heads, d, dp, 23)
    out = torch.zeros(B, N, hdp, dtype=BF16, device="cuda")
    for i in range(2):
        sl = slice(i * (B // 2), (i + 1) * (B // 2))
        ops.attention(q, k[sl], v[sl], heads, dp, d ** -0.5, out=out[sl])
        ref = _ref(q, k[sl].contiguous(), v[sl].contiguous(), heads, d, dp, d ** -0.5)
        assert_close(out[sl].float().reshape(B // 2, N, heads, dp)[..., :d], ref, TOL, f"cross-attention slice {i}")


@pytest.mark.parametrize("B,N", [(2, 4096), (3, 320), (1, 128), (2, 1024), (3, 640), (1, 448)])
def test_attention_row_sum_from_ones_column(B, N):
    """topo_attention_lcol: V carries 1.0 in padded column 40 of every d = 40 head, the kernel takes the softmax row sum
    from the P V product (O[:, 40]) instead of accumulating it in the softmax warps; same result as the explicit sum, and
    the padded output columns (incl. column 40) stay zero.  N >= 384 selects attention_split_kernel (two threads per query
    row); N = 640 / 448 leave the last CTA with inactive streams and a ragged last key block."""
    from topo_b200 import ops
    heads, d, dp = 8, 40, 48
    q, k, v = _mk(B, N, heads, d, dp, 31), _mk(B, N, heads, d, dp, 32), _mk(B, N, heads, d, dp, 33)
    v1 = v.clone().reshape(B, N, heads, dp)
    v1[..., d] = 1.0
    v1 = v1.reshape(B, N, heads * dp)
    scale = d ** -0.5
    o = ops.attention(q, k, v1, heads, dp, scale, l_col=d)
    plain = ops.attention(q, k, v, heads, dp, scale)
    ref = _ref(q, k, v, heads, d, dp, scale)
    got = o.float().reshape(B, N, heads, dp)
    assert_close(got[..., :d], ref, TOL, f"attention l_col B{B} N{N}")
    assert got[..., d:].abs().max() == 0, "padded head columns (incl. the ones column) must stay zero"
    assert_close(o, plain, 1e-2, "ones-column sum vs explicit sum")


def test_attention_row_sum_from_ones_column_peaked():
    from topo_b200 import ops
    B, heads, N, d, dp = 1, 2, 512, 40, 48
    q, k, v = _mk(B, N, heads, d, dp, 1), _mk(B, N, heads, d, dp, 2), _mk(B, N, heads, d, dp, 3)
    q = (q.float() * 60.0).to(BF16)
    v1 = v.clone().reshape(B, N, heads, dp)
    v1[..., d] = 1.0
    o = ops.attention(q, k, v1.reshape(B, N, heads * dp), heads, dp, d ** -0.5, l_col=d)
    ref = _ref(q, k, v, heads, d, dp, d ** -0.5)
    assert_close(o.float().reshape(B, N, heads, dp)[..., :d], ref, 3e-2, "attention l_col peaked (rescale path)")
