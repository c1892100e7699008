"""TEST INFRASTRUCTURE — numpy restatement of the per-image metrics of ``LoRA_modified/quantitative_metrics.py``
(the checker for ``topo_b200/metrics.py`` / ``csrc/metrics.cu``).  Never imported by the product package.

Pinned: ``color_histogram_l1_255`` is the reference function body itself minus the torch->numpy conversion (:179-195,
pure numpy); ``mse_rgb`` (:164-165) is one torch expression; ``psnr_rgb`` (:167-170) calls
``skimage.metrics.peak_signal_noise_ratio(gt, pred, data_range=255)`` — skimage is not installed here, so its published
definition is restated: ``10 * log10(data_range**2 / mean((gt - pred)**2, dtype=float64))``.
"""
import numpy as np


def to_rgb255(x):
    """:206-207  ((x + 1) * 127.5).clamp(0, 255).round() — x float32 array in [-1, 1]; numpy rint = torch.round."""
    return np.rint(np.clip((x.astype(np.float32) + np.float32(1.0)) * np.float32(127.5), 0, 255)).astype(np.float32)


def mse_rgb(pred, gt):
    """:164-165  torch.mean((pred - gt) ** 2).item() on float32 [3,H,W] arrays of 0..255 values."""
    return float(np.mean((pred.astype(np.float32) - gt.astype(np.float32)) ** 2, dtype=np.float64))


def psnr_rgb(pred, gt):
    """:167-170  skimage PSNR with data_range=255."""
    err = np.mean((gt.astype(np.float64) - pred.astype(np.float64)) ** 2, dtype=np.float64)
    return float("inf") if err == 0 else float(10 * np.log10(255.0 ** 2 / err))


def color_histogram_l1_255(pred, gt, bins=32):
    """:179-195  pred, gt: [3,H,W] arrays of 0..255 values."""
    pred = np.transpose(pred, (1, 2, 0))
    gt = np.transpose(gt, (1, 2, 0))
    total = 0.0
    for ch in range(3):
        hist_p, _ = np.histogram(pred[:, :, ch], bins=bins, range=(0, 255), density=True)
        hist_g, _ = np.histogram(gt[:, :, ch], bins=bins, range=(0, 255), density=True)
        total += np.abs(hist_p - hist_g).sum()
    return total / 3.0
