"""Seam-B2 throughput (SURVEY.md §8d "deterministic ray batches"): closest-hit and any-hit traversal of 2^24-ray batches
resident in HBM, per scene and per batch kind, with the measured nodes / triangles / instances per ray that turn
rays/s into algorithmic bytes/s (B_ray = Q + 80 N_node + 48 N_tri + 64 N_inst).

    python tools/ray_batch_bench.py c2 c5 c5b [--log2n 24] [--out gpurun_out/ray_batches.json]

Batches: P = primary rays of the scene camera on a sqrt(n) x sqrt(n) grid; D = cosine-hemisphere bounce rays from the
P hit points; U = incoherent rays (uniform origins in 1.5x the scene box, uniform directions), any-hit with finite t_max.
c5b = the 2000 bunny instances of c5 flattened into ONE mesh of 10 M unique triangles (HBM-resident BVH)."""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "simple-path-tracer_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)

import torch  # noqa: E402
import rays as R  # noqa: E402
from spt import abi, core, scenes  # noqa: E402


flatten_instances = scenes.flatten_instances


def timed(fn, reps=5):
    """best of `reps`, CUDA events on STREAM — the stream the kernels are launched on (handed to the library)"""
    fn()
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(STREAM); fn(); e1.record(STREAM)
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) * 1e-3)
    return best


ap = argparse.ArgumentParser()
ap.add_argument("names", nargs="+")
ap.add_argument("--log2n", type=int, default=24)
ap.add_argument("--out", default="")
args = ap.parse_args()
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {}
hbm = float(peaks.get("hbm_gbs", 6543.1))
n = 1 << args.log2n
side = 1 << (args.log2n // 2)
ctx = core.Context(0)
STREAM = torch.cuda.Stream()
results = []
for name in args.names:
    t0 = time.time()
    fs = scenes.load("c5" if name == "c5b" else name)
    if name == "c5b":
        fs = flatten_instances(fs)
    t1 = time.time()
    sc = core.Scene(ctx, fs)
    t2 = time.time()
    info = sc.info()
    prim = R.primary_rays(fs, side, n // side)
    batches = {"P": (prim, False)}
    d_hits = torch.zeros(n * 5, dtype=torch.int32, device="cuda")
    d_occ = torch.zeros(n, dtype=torch.uint8, device="cuda")
    for kind in ("P", "D", "U"):
        if kind == "D":
            hp = np.frombuffer(d_hits_host, core.HIT_DT)
            b = R.bounce_rays(prim, hp)
            reps_ = int(np.ceil(n / max(1, len(b))))
            b = np.tile(b, reps_)[:n] if len(b) else prim
            batches["D"] = (b, False)
        if kind == "U":
            batches["U"] = (R.incoherent_rays(fs, n, finite_tmax=True), True)
        rays_np, anyhit = batches[kind]
        d_rays = torch.from_numpy(rays_np.view(np.float32).reshape(-1, 8)).cuda()
        for mode in (("closest", "any") if kind != "P" else ("closest",)):
            if mode == "closest":
                run = lambda fl=0, ws=False: sc.trace_device(d_rays.data_ptr(), len(rays_np), d_hits.data_ptr(), flags=fl, want_stats=ws, stream=STREAM.cuda_stream)
            else:
                run = lambda fl=0, ws=False: sc.occluded_device(d_rays.data_ptr(), len(rays_np), d_occ.data_ptr(), flags=fl, want_stats=ws, stream=STREAM.cuda_stream)
            torch.cuda.synchronize()
            sec = timed(run)
            st = run(abi.RENDER_COUNT_TRAVERSAL, True)
            torch.cuda.synchronize()
            nr = float(len(rays_np))
            nn, nt, ni = st["nodes_visited"] / nr, st["tris_tested"] / nr, st["insts_visited"] / nr
            q = 52 if mode == "closest" else 36
            b_ray = q + 80 * nn + 48 * nt + 64 * ni
            if mode == "closest":
                h = d_hits.cpu().numpy()
                if kind == "P":
                    d_hits_host = h.tobytes()
                hit_rate = float((np.frombuffer(h.tobytes(), core.HIT_DT)["inst_id"] >= 0).mean())
            else:
                hit_rate = float(d_occ.float().mean().item())
            r = dict(scene=name, batch=kind, mode=mode, rays=int(nr), seconds=sec, grays_per_s=nr / sec / 1e9, hit_rate=hit_rate,
                     nodes_per_ray=nn, tris_per_ray=nt, insts_per_ray=ni, bytes_per_ray=b_ray, algorithmic_gbs=nr * b_ray / sec / 1e9,
                     frac_of_hbm_peak=nr * b_ray / sec / 1e9 / hbm, compulsory_gbs=nr * q / sec / 1e9,
                     scene_bytes=info["bytes_total"], triangles=info["triangles"], instances=info["instances"],
                     load_s=t1 - t0, build_upload_s=t2 - t1)
            results.append(r)
            print(json.dumps(r), flush=True)
    sc.close()
    del d_hits, d_occ
if args.out:
    json.dump(results, open(args.out, "w"), indent=1)
