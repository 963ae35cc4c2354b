"""N>1 host logic on CPU: two gloo ranks own disjoint interleaved tiles (spt_owned_tiles through the C ABI) and the
framebuffer reduce reconstructs the full image exactly — the same plumbing renderer.py runs over NCCL."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, w, h, out):
    sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
    from spt import core
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    full = torch.arange(w * h * 3, dtype=torch.float32).reshape(h, w, 3) + 1.0   # stands in for the rendered image
    mine = torch.zeros_like(full)
    tx = (w + 31) // 32
    for t in core.owned_tiles(w, h, rank, world):
        x0, y0 = (int(t) % tx) * 32, (int(t) // tx) * 32
        mine[y0:y0 + 32, x0:x0 + 32] = full[y0:y0 + 32, x0:x0 + 32]
    counts = torch.tensor([float((mine.sum(-1) > 0).sum())])
    dist.reduce(mine, dst=0, op=dist.ReduceOp.SUM)
    dist.all_reduce(counts)
    if rank == 0:
        out.put((bool(torch.equal(mine, full)), float(counts.item())))
    dist.destroy_process_group()


def test_two_rank_tile_partition_and_reduce(built):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    w, h = 200, 120
    ps = [ctx.Process(target=_worker, args=(r, 2, port, w, h, q)) for r in range(2)]
    for p in ps:
        p.start()
    ok, n = q.get(timeout=120)
    for p in ps:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert ok and n == w * h


def _gather_worker(rank, world, port, w, h, tile, out):
    """The exchange spt_render_gather runs over NCCL, replayed over gloo on the CPU: every rank packs the film of the tiles it owns
    (tile-major, 3 x int64 per pixel, padding pixels of edge tiles zero), rank 0 receives one equal-sized segment per rank and unpacks
    them by every rank's tile list — the same segment size rule (ceil(n_tiles / world) tiles) and padding marker as csrc/spt_api.cu."""
    sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
    from spt import core
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tw, th = tile
    tx, ty = (w + tw - 1) // tw, (h + th - 1) // th
    per_rank = (tx * ty + world - 1) // world
    full = (torch.arange(w * h * 3, dtype=torch.int64).reshape(h, w, 3) * 7919 + 13)          # stands in for the integer film
    mine = core.owned_tiles(w, h, rank, world, tile=tile)
    assert len(mine) <= per_rank
    seg = torch.zeros((per_rank, th, tw, 3), dtype=torch.int64)
    for k, t in enumerate(mine):
        x0, y0 = (int(t) % tx) * tw, (int(t) // tx) * th
        blk = full[y0:y0 + th, x0:x0 + tw]
        seg[k, :blk.shape[0], :blk.shape[1]] = blk
    segs = [torch.zeros_like(seg) for _ in range(world)] if rank == 0 else None
    dist.gather(seg, segs, dst=0)
    if rank == 0:
        film = torch.zeros_like(full)
        seen = torch.zeros((h, w), dtype=torch.int32)
        for r in range(world):
            tiles = list(core.owned_tiles(w, h, r, world, tile=tile)) + [0xFFFFFFFF] * per_rank
            for k in range(per_rank):
                t = int(tiles[k])
                if t == 0xFFFFFFFF:
                    continue
                x0, y0 = (t % tx) * tw, (t // tx) * th
                hh, ww = min(th, h - y0), min(tw, w - x0)
                film[y0:y0 + hh, x0:x0 + ww] = segs[r][k, :hh, :ww]
                seen[y0:y0 + hh, x0:x0 + ww] += 1
        out.put((bool(torch.equal(film, full)), int(seen.min()), int(seen.max())))
    dist.destroy_process_group()


def test_two_rank_tile_packed_film_gather(built):
    ctx = mp.get_context("spawn")
    for (w, h, tile) in ((200, 120, (8, 4)), (97, 61, (16, 8))):   # the second: ragged edge tiles, odd tile count
        q = ctx.Queue()
        port = _free_port()
        ps = [ctx.Process(target=_gather_worker, args=(r, 2, port, w, h, tile, q)) for r in range(2)]
        for p in ps:
            p.start()
        ok, lo, hi = q.get(timeout=120)
        for p in ps:
            p.join(timeout=60)
            assert p.exitcode == 0
        assert ok and lo == 1 and hi == 1          # every pixel arrives exactly once
