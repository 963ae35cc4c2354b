"""ad-hoc: how much would sorting incoherent extension rays buy?  D2 = second-bounce rays (from the hits of D), traced in
their natural order, in random order, and sorted by (direction octant, Morton code of the origin)."""
import os, sys, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "simple-path-tracer_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)
import torch
import rays as R
from spt import core, scenes

def morton(o, lo, hi):
    q = np.clip(((o - lo) / (hi - lo) * 1023).astype(np.uint64), 0, 1023)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F; v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2)

ctx = core.Context(0)
S = torch.cuda.Stream()
for name in sys.argv[1:]:
    fs = scenes.load(name); sc = core.Scene(ctx, fs)
    prim = R.primary_rays(fs, 2048, 1024)
    h1 = sc.trace(prim); d1 = R.bounce_rays(prim, h1)
    h2 = sc.trace(d1); d2 = R.bounce_rays(d1, h2, seed=77)
    h3 = sc.trace(d2); d3 = R.bounce_rays(d2, h3, seed=78)
    for label, batch in (("D1", d1), ("D2", d2), ("D3", d3)):
        lo, hi = batch["origin"].min(0), batch["origin"].max(0) + 1e-6
        oct_ = ((batch["direction"][:, 0] < 0).astype(np.uint64) << 32) | ((batch["direction"][:, 1] < 0).astype(np.uint64) << 31) | ((batch["direction"][:, 2] < 0).astype(np.uint64) << 30)
        key = oct_ | morton(batch["origin"].astype(np.float64), lo, hi)
        orders = {"natural": np.arange(len(batch)), "random": np.random.RandomState(1).permutation(len(batch)), "sorted": np.argsort(key, kind="stable"),
                  "morton_only": np.argsort(morton(batch["origin"].astype(np.float64), lo, hi), kind="stable")}
        for oname, order in orders.items():
            b = np.ascontiguousarray(batch[order])
            reps = int(np.ceil((1 << 22) / len(b))); b = np.tile(b, reps)
            d_r = torch.from_numpy(b.view(np.float32).reshape(-1, 8)).cuda()
            d_h = torch.zeros(len(b) * 5, dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            best = 1e9
            for _ in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(S); sc.trace_device(d_r.data_ptr(), len(b), d_h.data_ptr(), stream=S.cuda_stream); e1.record(S); torch.cuda.synchronize()
                best = min(best, e0.elapsed_time(e1))
            print(name, label, oname, len(b), round(len(b) / best / 1e6, 2), "Grays/s", flush=True)
    sc.close()
