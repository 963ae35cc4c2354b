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
