"""N > 1 on real GPUs (skipped on a one-GPU box): two ranks, one process per GPU, joined through the library's own communicator
(spt_comm_create; torch.distributed only broadcasts the id), spt_render_gather.  The gathered frame must equal the one-GPU frame
bit for bit — for the plain box filter and for a radius that needs neighbour pixels owned by the other rank."""
import os
import socket
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    for p in (ROOT, os.path.join(ROOT, "simple-path-tracer_b200")):
        sys.path.insert(0, p)
    from spt import core, scenes
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    ctx = core.Context(rank)
    sc = core.Scene(ctx, scenes.load("c2"))
    comm = core.Comm.from_torch(ctx)
    w, h = 200, 120
    res = {}
    for radius in (0.5, 1.5):
        fb = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
        st = sc.render_gather(comm, fb.data_ptr(), w, h, root=0, spp=8, seed=5, filter_radius=radius)
        torch.cuda.synchronize()
        if rank == 0:
            one = torch.zeros_like(fb)
            sc.render_device(one.data_ptr(), w, h, spp=8, seed=5, filter_radius=radius)
            torch.cuda.synchronize()
            res[radius] = (bool(torch.equal(fb, one)), float(fb.mean().item()), st["paths"])
        else:
            res[radius] = (True, 0.0, st["paths"])
    paths = torch.tensor([res[0.5][2]], dtype=torch.float64, device="cuda")
    dist.all_reduce(paths)
    if rank == 0:
        out.put((res[0.5][0], res[1.5][0], res[0.5][1], float(paths.item())))
    comm.close(); sc.close(); ctx.close()
    dist.destroy_process_group()


def test_two_gpus_gather_is_bit_identical_to_one_gpu(built):
    import torch
    import torch.multiprocessing as mp
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    ps = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in ps:
        p.start()
    same, same_r15, mean, paths = q.get(timeout=300)
    for p in ps:
        p.join(timeout=120)
        assert p.exitcode == 0
    assert same and same_r15 and mean > 0 and paths == 200 * 120 * 8
