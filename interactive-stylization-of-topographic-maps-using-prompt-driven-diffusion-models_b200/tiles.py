"""Full-map stylization: sheet → 512² tiles → per-rank shards → DDIM+CFG on each shard → uint8 tiles → ONE gather →
row-major mosaic.  Restates the tile loop of ``BaseModel/evaluateCN.py:60-120`` (one tile at a time, strictly
serial) and the mosaic of ``BaseModel/stitchTiles.py:6-46`` as a data-parallel job:

* tiles are independent, so rank ``g`` of ``G`` owns the contiguous block ``[g*T/G, (g+1)*T/G)`` — a contiguous
  slab of the row-major mosaic; weights are replicated; there is NO collective inside the denoising loop;
* the decoded tile is converted to uint8 HWC on the device (``*127.5+127.5``, clip, truncate —
  ``evaluateCN.py:101-103``) so only 786 432 B per tile cross NVLink;
* one ``all_gather`` (NCCL on GPUs, gloo in the CPU tests) brings the slabs together for ``stitch_tiles``.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import ops
from .sampling import DDIMSampler


def shard_range(n_tiles: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block partition; the first ``n_tiles % world`` ranks take one extra tile."""
    base, extra = divmod(n_tiles, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def split_sheet(sheet: torch.Tensor, tile: int = 512) -> torch.Tensor:
    """``[H, W, C]`` sheet → ``[T, tile, tile, C]`` row-major tiles (``BaseModel/datasetCreation00.py:32-59``:
    ``np.split`` along rows then columns).  H and W must be multiples of ``tile``."""
    H, W, C = sheet.shape
    if H % tile or W % tile:
        raise ValueError(f"sheet {H}x{W} is not a multiple of the {tile}px tile")
    r, c = H // tile, W // tile
    return sheet.reshape(r, tile, c, tile, C).permute(0, 2, 1, 3, 4).reshape(r * c, tile, tile, C).contiguous()


def stitch_tiles(tiles: torch.Tensor, cols: int) -> torch.Tensor:
    """``[T, h, w, C]`` → ``[T/cols * h, cols * w, C]``: hstack rows of ``cols`` tiles, vstack the rows
    (``BaseModel/stitchTiles.py:13-41``, which hard-codes 10x10).  Pure data movement; works on any device."""
    T, h, w, C = tiles.shape
    if T % cols:
        raise ValueError(f"{T} tiles do not fill rows of {cols}")
    rows = T // cols
    return tiles.reshape(rows, cols, h, w, C).permute(0, 2, 1, 3, 4).reshape(rows * h, cols * w, C).contiguous()


def gather_tiles(local: torch.Tensor, n_tiles: int, group=None) -> torch.Tensor:
    """All ranks contribute their uint8 slab ``[t_g, h, w, C]``; returns the full ``[T, h, w, C]`` on every rank.
    Ragged shards are padded to the largest shard for the collective and trimmed afterwards."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = [shard_range(n_tiles, r, world) for r in range(world)]
    mx = max(hi - lo for lo, hi in sizes)
    pad = local
    if local.shape[0] < mx:
        pad = torch.zeros((mx, *local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
    out = torch.empty((world * mx, *local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad.contiguous(), group=group)
    parts = [out[r * mx:r * mx + (hi - lo)] for r, (lo, hi) in enumerate(sizes)]
    return torch.cat(parts) if any(p.shape[0] != mx for p in parts) else out


@torch.no_grad()
def stylize_tiles(model, ctx: torch.Tensor, hints: torch.Tensor, x_T: Optional[torch.Tensor] = None, steps: int = 30,
                  scale: float = 7.5, eta: float = 0.0, batch: int = 8, seed: int = 1234, hacked: bool = False,
                  uc: Optional[torch.Tensor] = None, out_host: Optional[torch.Tensor] = None) -> torch.Tensor:
    """Stylize ``T`` tiles: ``ctx [T,77,768]``, ``hints [T,3,512,512]`` in [0,1], optional ``x_T [T,4,64,64]`` (any of
    them may live in (pinned) host memory — they are copied batch by batch).  Returns uint8 ``[T,512,512,3]`` on the
    model's device, or fills and returns ``out_host`` when given.  ``uc`` defaults to the all-zero context of
    ``vis_metrics.py:425``."""
    dev = next(model.parameters()).device
    T = hints.shape[0]
    outs: List[torch.Tensor] = []
    sampler = DDIMSampler(model, hacked=hacked)
    for lo in range(0, T, batch):
        hi = min(T, lo + batch)
        c = ctx[lo:hi].to(dev, torch.float32, non_blocking=True)
        h = hints[lo:hi].to(dev, torch.float32, non_blocking=True)
        u = torch.zeros_like(c) if uc is None else uc[lo:hi].to(dev, torch.float32, non_blocking=True)
        if x_T is None:
            torch.manual_seed(seed + lo)
            xt = None
        else:
            xt = x_T[lo:hi].to(dev, torch.float32, non_blocking=True)
        cond = {"c_crossattn": [c], "c_concat": [h]}
        ucd = {"c_crossattn": [u], "c_concat": [h]}
        z, _ = sampler.sample(S=steps, conditioning=cond, batch_size=hi - lo, shape=(4, h.shape[2] // 8, h.shape[3] // 8),
                              unconditional_guidance_scale=scale, unconditional_conditioning=ucd, eta=eta, x_T=xt)
        u8 = model.engine().decode_to_u8(z)
        if out_host is not None:
            out_host[lo:hi].copy_(u8, non_blocking=True)
        else:
            outs.append(u8)
    if out_host is not None:
        torch.cuda.current_stream(dev).synchronize()
        return out_host
    return torch.cat(outs)


@torch.no_grad()
def stylize_map(model, ctx: torch.Tensor, hints: torch.Tensor, cols: int, x_T: Optional[torch.Tensor] = None,
                rank: int = 0, world: int = 1, group=None, **kw) -> torch.Tensor:
    """Data-parallel full map: this rank stylizes its contiguous shard, one gather, stitch.  Returns the mosaic."""
    T = hints.shape[0]
    lo, hi = shard_range(T, rank, world)
    local = stylize_tiles(model, ctx[lo:hi], hints[lo:hi], None if x_T is None else x_T[lo:hi], **kw)
    return stitch_tiles(gather_tiles(local, T, group), cols)
