"""Tile metrics (row N4): host-side binning table vs numpy on CPU; the CUDA reduction vs the numpy oracle on the GPU
(bit-exact integer counts, float64 metrics to 1e-12)."""
import numpy as np
import pytest
import torch


def test_histogram_lut_is_numpy_binning():
    from topo_b200.metrics import histogram_lut
    for bins in (32, 16, 7, 64):
        lut = histogram_lut(bins)
        for v in range(256):
            h, _ = np.histogram(np.array([float(v)]), bins=bins, range=(0, 255))
            assert int(np.argmax(h)) == int(lut[v]), (bins, v)
    assert histogram_lut(32)[255] == 31 and histogram_lut(32)[0] == 0


def test_metrics_refuse_cpu_tensors():
    from topo_b200 import metrics, capi
    a = torch.zeros(3, 8, 8)
    with pytest.raises(capi.TopoError):
        metrics.mse_rgb(a, a)


def _pairs(T, H, W, seed):
    g = torch.Generator().manual_seed(seed)
    gt = torch.randint(0, 256, (T, H, W, 3), generator=g, dtype=torch.uint8)
    noise = torch.randint(-20, 21, (T, H, W, 3), generator=g)
    pred = (gt.int() + noise).clamp(0, 255).to(torch.uint8)
    pred[0, : H // 2] = 255                  # a saturated flat region: every thread of a warp hits one bin
    gt[-1] = pred[-1]                        # identical pair: mse 0, psnr inf, hist distance 0
    return pred, gt


@pytest.mark.gpu
@pytest.mark.parametrize("T,H,W", [(3, 64, 64), (2, 512, 512), (5, 40, 24)])
def test_tile_metrics_vs_numpy_oracle(T, H, W):
    from oracle import metrics_oracle as O
    from topo_b200 import metrics
    pred, gt = _pairs(T, H, W, seed=T * 1000 + H)
    out = metrics.tile_metrics(pred.cuda(), gt.cuda())
    # the reference's layout: [3,H,W] float 0..255 — must give the very same counts
    outf = metrics.tile_metrics(pred.permute(0, 3, 1, 2).float().cuda(), gt.permute(0, 3, 1, 2).float().cuda())
    assert torch.equal(out["sse"], outf["sse"]) and torch.equal(out["hist"], outf["hist"])
    for t in range(T):
        p, g = pred[t].permute(2, 0, 1).float().numpy(), gt[t].permute(2, 0, 1).float().numpy()
        sse = int(((p.astype(np.int64) - g.astype(np.int64)) ** 2).sum())
        assert int(out["sse"][t]) == sse
        for i, img in enumerate((p, g)):
            for c in range(3):
                h, _ = np.histogram(img[c], bins=32, range=(0, 255))
                assert np.array_equal(out["hist"][t, i, c].cpu().numpy(), h), (t, i, c)
        mse, psnr, hl1 = O.mse_rgb(p, g), O.psnr_rgb(p, g), O.color_histogram_l1_255(p, g)
        assert abs(float(out["mse"][t]) - mse) <= 1e-12 * max(mse, 1)
        assert float(out["psnr"][t]) == psnr or abs(float(out["psnr"][t]) - psnr) <= 1e-12 * abs(psnr)
        assert abs(float(out["hist_l1"][t]) - hl1) <= 1e-12
    print(f"[parity] tile_metrics T{T} {H}x{W}: sse/hist exact; mse[0]={float(out['mse'][0]):.4f} "
          f"psnr[0]={float(out['psnr'][0]):.4f} hist_l1[0]={float(out['hist_l1'][0]):.6f}")
    p0, g0 = pred[0].permute(2, 0, 1).float().cuda(), gt[0].permute(2, 0, 1).float().cuda()
    assert metrics.mse_rgb(p0, g0) == float(out["mse"][0])
    assert metrics.psnr_rgb(p0, g0) == float(out["psnr"][0])
    assert metrics.color_histogram_l1_255(p0, g0) == float(out["hist_l1"][0])


@pytest.mark.gpu
def test_to_rgb255_rounding():
    from oracle import metrics_oracle as O
    from topo_b200 import metrics, ops
    g = torch.Generator().manual_seed(3)
    x = (torch.rand(2, 3, 64, 64, generator=g) * 2.4 - 1.2)
    k = torch.arange(0, 256).float()
    x[0, 0, 0, :256 // 4] = ((k[::4] + 0.5) / 127.5 - 1.0)       # values that land on .5 before rounding
    got = metrics.to_rgb255(x.cuda()).cpu()
    want = torch.from_numpy(O.to_rgb255(x.numpy())).permute(0, 2, 3, 1).to(torch.uint8)
    assert torch.equal(got, want)
    tr = ops.tile_to_u8(x.permute(0, 2, 3, 1).contiguous().cuda()).cpu()     # evaluateCN.py:101-103: truncate
    want_tr = torch.from_numpy(np.clip(x.permute(0, 2, 3, 1).numpy() * np.float32(127.5) + np.float32(127.5), 0, 255)
                               .astype(np.uint8))
    assert torch.equal(tr, want_tr)
