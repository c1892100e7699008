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
