"""Tile metrics on the device — the per-image numbers of
``LoRA_modified/quantitative_metrics.py:164-195`` as ONE CUDA reduction pass per tile batch:

* ``mse_rgb``   (:164-165)  ``mean((pred - gt)**2)`` over the 0..255 images,
* ``psnr_rgb``  (:167-170)  skimage ``peak_signal_noise_ratio(gt, pred, data_range=255)`` = ``10 log10(255² / mse)``,
* ``color_histogram_l1_255`` (:179-195)  mean over R,G,B of the L1 distance between the ``density=True`` 32-bin
  ``np.histogram(range=(0, 255))`` of prediction and ground truth.

LPIPS (:172-177) needs pretrained VGG weights and stays out.  The kernel (``csrc/metrics.cu``) returns exact
integers — the squared-error sum and the bin counts — so the only floating-point work left is a handful of float64
divisions, done here on the device without a host sync.  Inputs are what the script feeds its metric functions:
``[3,H,W]`` float tensors of rounded 0..255 values (``pred255``, ``gt255``), or uint8 HWC tiles from the tile loop.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch

from . import capi, ops

__all__ = ["histogram_lut", "to_rgb255", "tile_metrics", "mse_rgb", "psnr_rgb", "color_histogram_l1_255"]

_LUTS: Dict = {}


def histogram_lut(bins: int = 32, lo: float = 0.0, hi: float = 255.0) -> np.ndarray:
    """value (0..255) -> bin index under ``np.histogram(x, bins, range=(lo, hi))``'s uniform-bin rule: index =
    ``int((v - lo) * bins / (hi - lo))``, the right edge belongs to the last bin, then the edge-rounding corrections
    numpy applies against ``linspace(lo, hi, bins + 1)``.  256 integer evaluations of that rule, no image data."""
    edges = np.linspace(lo, hi, bins + 1, endpoint=True, dtype=np.float64)
    v = np.arange(256, dtype=np.float64)
    norm = bins / (hi - lo)
    idx = ((v - lo) * norm).astype(np.intp)
    idx[idx == bins] -= 1
    idx[v < edges[idx]] -= 1
    inc = (v >= edges[idx + 1]) & (idx != bins - 1)
    idx[inc] += 1
    if (v < lo).any() or (v > hi).any():
        raise capi.TopoError("histogram range must cover 0..255")
    return idx.astype(np.uint8)


def _lut(bins: int, device) -> torch.Tensor:
    key = (bins, str(device))
    if key not in _LUTS:
        _LUTS[key] = torch.from_numpy(histogram_lut(bins)).to(device)
    return _LUTS[key]


def to_rgb255(x: torch.Tensor) -> torch.Tensor:
    """``((x + 1) * 127.5).clamp(0, 255).round()`` of an NCHW [-1,1] image batch (``quantitative_metrics.py:206-207``)
    as uint8 HWC tiles ``[B, H, W, 3]`` on the device."""
    if x.dim() != 4:
        raise capi.TopoError("to_rgb255 expects [B, 3, H, W]")
    return ops.tile_to_u8(x.float().permute(0, 2, 3, 1).contiguous(), round_half_even=True)


def _as_tiles(t: torch.Tensor):
    """-> (tensor [T, ...], planar) for uint8 HWC ``[T,H,W,3]``/``[H,W,3]`` or float CHW ``[T,3,H,W]``/``[3,H,W]``."""
    if t.dtype == torch.uint8:
        t = t if t.dim() == 4 else t[None]
        if t.shape[-1] != 3:
            raise capi.TopoError("uint8 tiles must be HWC with 3 channels")
        return t, False
    t = t.float()
    t = t if t.dim() == 4 else t[None]
    if t.shape[1] != 3:
        raise capi.TopoError("float images must be [3,H,W] / [T,3,H,W] with values 0..255")
    return t, True


@torch.no_grad()
def tile_metrics(pred: torch.Tensor, gt: torch.Tensor, bins: int = 32) -> Dict[str, torch.Tensor]:
    """Per-tile ``mse``, ``psnr``, ``hist_l1`` (float64 device tensors ``[T]``) plus the raw ``sse`` / ``hist`` counts."""
    p, planar = _as_tiles(pred)
    g, planar_g = _as_tiles(gt)
    if planar != planar_g:
        raise capi.TopoError("pred and gt must use the same layout / dtype")
    sse, hist = ops.tile_metrics(p, g, _lut(bins, p.device), bins, planar)
    n = p[0].numel()
    mse = sse.double() / n
    psnr = 10.0 * torch.log10(255.0 ** 2 / mse)        # inf for identical images, as skimage
    db = torch.from_numpy(np.diff(np.linspace(0.0, 255.0, bins + 1))).to(p.device)
    dens = hist.double() / db / hist.sum(-1, keepdim=True).double()      # np.histogram(density=True): n / db / n.sum()
    hist_l1 = (dens[:, 0] - dens[:, 1]).abs().sum(-1).sum(-1) / 3.0
    return {"mse": mse, "psnr": psnr, "hist_l1": hist_l1, "sse": sse, "hist": hist}


def mse_rgb(pred, gt) -> float:
    return float(tile_metrics(pred, gt)["mse"][0])


def psnr_rgb(pred, gt) -> float:
    return float(tile_metrics(pred, gt)["psnr"][0])


def color_histogram_l1_255(pred, gt, bins: int = 32) -> float:
    return float(tile_metrics(pred, gt, bins)["hist_l1"][0])
