"""Every entry point and kernel once, on tiny workloads: a smoke run of the whole ABI, and the workload for `compute-sanitizer`
where that tool is available (it is closed on this pool: "runs under it have left GPUs needing a reset"):
    compute-sanitizer --tool memcheck --error-exitcode 3 python tools/sanitize_smoke.py
Small on purpose: the sanitizer slows kernels down 10-100x.  torch is imported AFTER the first spt_comm_* call on purpose: the
binding must have loaded it first (one libnccl.so.2 per process)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("simple-path-tracer_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import rays as R  # noqa: E402
from spt import abi, core, scenes  # noqa: E402

ctx = core.Context(0)
w, h = 64, 40
for name in ("c1", "c2", "c3", "c4", "c5"):
    fs = scenes.load(name)
    sc = core.Scene(ctx, fs)
    img, st = sc.render(w, h, spp=2, seed=1, pool_size=2048)                      # many iterations incl. the drain (ext_queue)
    assert st["paths"] == w * h * 2 and np.isfinite(img).all()
    sc.render(w, h, spp=2, seed=1, flags=abi.RENDER_NO_SORT)                        # the all-materials kernel
    sc.render(w, h, spp=2, seed=1, flags=abi.RENDER_COUNT_TRAVERSAL)                # counting instantiations
    sc.render(w, h, spp=3, seed=1, filter_radius=1.5, sampler="jittered", div=(3, 1))  # k_resolve_box
    sc.render(w, h, spp=1, seed=1, flags=abi.RENDER_DEBUG_NORMAL)                   # k_aov_normal
    os.environ["SPT_SHADE_LEAN"] = "0"
    sc.render(w, h, spp=1, seed=1)                                                  # general (non-lean) shade kernels
    del os.environ["SPT_SHADE_LEAN"]
    sc.render(w, h, spp=2, seed=1, tile=(64, 64), pool_size=1024, max_depth=1)     # padding-only iterations
    pr = R.primary_rays(fs, 48, 32)
    hits = sc.trace(pr)
    assert np.array_equal(sc.trace_noalloc(pr).view(np.uint8), hits.view(np.uint8))
    sc.occluded(R.incoherent_rays(fs, 3001, finite_tmax=True))
    comm = core.Comm(ctx, core.Comm.unique_id(ctx.lib), 0, 1)
    import torch
    fb = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
    sc.render_gather(comm, fb.data_ptr(), w, h, spp=1, seed=1)
    comm.close()
    film = sc.film_save(w, h)
    sc.film_load(film, 1)
    sc.close()
    print(name, "ok", flush=True)
# device-built BLAS
os.environ["SPT_BVH_BUILD"] = "device"
fs = scenes.load("c2")
sc = core.Scene(ctx, fs)
sc.trace(R.primary_rays(fs, 32, 24))
sc.close()
print("sanitize smoke done")
