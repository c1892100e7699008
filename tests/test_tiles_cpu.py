"""not-gpu: tile sharding / stitching host logic and the N>1 gather over gloo (world_size 2)."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from topo_b200 import tiles


def test_shard_range_partitions_contiguously():
    for T in (1, 7, 19, 100, 256):
        for G in (1, 2, 4, 8):
            spans = [tiles.shard_range(T, r, G) for r in range(G)]
            assert spans[0][0] == 0 and spans[-1][1] == T
            assert all(spans[i][1] == spans[i + 1][0] for i in range(G - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
    assert tiles.shard_range(256, 3, 8) == (96, 128)


def test_split_stitch_roundtrip_and_errors():
    g = torch.Generator().manual_seed(0)
    sheet = torch.randint(0, 256, (64, 96, 3), generator=g, dtype=torch.uint8)
    t = tiles.split_sheet(sheet, 16)
    assert t.shape == (24, 16, 16, 3)
    assert torch.equal(t[7], sheet[16:32, 16:32])          # row-major: tile 7 = row 1, col 1 of a 6-wide grid
    assert torch.equal(tiles.stitch_tiles(t, 6), sheet)
    with pytest.raises(ValueError):
        tiles.split_sheet(sheet, 20)
    with pytest.raises(ValueError):
        tiles.stitch_tiles(t, 5)
    one = tiles.stitch_tiles(t[:4], 1)                      # 1-column grids work (stitchTiles.py's j==1 seed breaks there)
    assert one.shape == (64, 16, 3)


def _worker(rank, world, port, T, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(T * 4 * 4 * 3, dtype=torch.int64).remainder(251).to(torch.uint8).reshape(T, 4, 4, 3)
    lo, hi = tiles.shard_range(T, rank, world)
    got = tiles.gather_tiles(full[lo:hi].clone(), T)
    ok = torch.equal(got, full)
    mosaic = tiles.stitch_tiles(got, cols=T // 1 if T < 4 else (T if T % 2 else 2))
    q.put((rank, ok, tuple(mosaic.shape)))
    dist.destroy_process_group()


@pytest.mark.parametrize("T", [8, 7])
def test_gather_tiles_gloo_world2(T):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, T, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in res), res


@pytest.mark.parametrize("noise", ["shared", "per_tile"])
def test_start_latents_depend_on_seed_and_global_index_only(noise):
    """the starting noise of a tile must not change with the sharding (world) or the batch size."""
    T, shape = 8, (4, 8, 8)

    def assemble(world, batch):
        parts = []
        for r in range(world):
            lo, hi = tiles.shard_range(T, r, world)
            for b0 in range(lo, hi, batch):           # stylize_map passes tile_offset=lo, stylize_tiles adds its batch offset
                n = min(hi, b0 + batch) - b0
                parts.append(tiles.start_latents(n, shape, 1234, "cpu", noise, tile_offset=b0))
        return torch.cat(parts)

    ref = assemble(1, 8)
    for world, batch in ((1, 4), (4, 8), (4, 1), (2, 3)):
        assert torch.equal(assemble(world, batch), ref), (world, batch)
    if noise == "shared":     # evaluateCN.py:77-79: seed_everything(seed) before every tile, batch 1 -> one latent for all tiles
        g = torch.Generator().manual_seed(1234)
        assert torch.equal(ref[5], torch.randn((1, *shape), generator=g)[0]) and torch.equal(ref[0], ref[7])
    else:
        assert not torch.equal(ref[0], ref[1])
    assert not torch.equal(tiles.start_latents(1, shape, 1, "cpu", noise), tiles.start_latents(1, shape, 2, "cpu", noise))
    with pytest.raises(ValueError):
        tiles.start_latents(1, shape, 1, "cpu", "nope")


def test_gather_refuses_a_sharded_map_without_process_group():
    local = torch.zeros(2, 4, 4, 3, dtype=torch.uint8)
    assert tiles.gather_tiles(local, 2) is local and tiles.gather_tiles(local, 2, world=1) is local
    with pytest.raises(RuntimeError):
        tiles.gather_tiles(local, 4, world=2)
