"""Development sweep on the GPU box: stage split and throughput per scene / pool size.
    python tools/sweep.py c2:32 c3:16 c5:4 --pools 262144,1048576 [--lib path/to/libspt_cuda.so]
Each NAME:SPP renders the named config at its full resolution with SPP samples per pixel (3 repeats, best kept)."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))

ap = argparse.ArgumentParser()
ap.add_argument("items", nargs="+")
ap.add_argument("--pools", default="0")
ap.add_argument("--lib", default="")
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--tag", default="")
ap.add_argument("--serial", action="store_true", help="one stream (SPT_RENDER_SERIAL)")
args = ap.parse_args()
if args.lib:
    os.environ["SPT_LIB"] = args.lib
import torch  # noqa: E402
from spt import abi, core, scenes  # noqa: E402

ctx = core.Context(0)
for item in args.items:
    name, spp = item.split(":")
    spp = int(spp)
    cfg = scenes.CONFIGS[name]
    w, h = cfg["width"], cfg["height"]
    sc = core.Scene(ctx, scenes.load(name))
    fb = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
    for pool in [int(p) for p in args.pools.split(",")]:
        best = None
        for rep in range(args.reps + 1):
            st = sc.render_device(fb.data_ptr(), w, h, spp=spp, seed=rep + 1, pool_size=pool, flags=abi.RENDER_TIME_STAGES if rep else 0)
            if rep and (best is None or st["seconds_total"] < best["seconds_total"]):
                best = st
        plain = min((sc.render_device(fb.data_ptr(), w, h, spp=spp, seed=9 + i, pool_size=pool, flags=abi.RENDER_SERIAL if args.serial else 0) for i in range(3)),
                    key=lambda st: st["seconds_total"])
        rays = best["ext_rays"] + best["shadow_rays"] + best["probe_rays"]
        print(json.dumps({"tag": args.tag, "scene": name, "spp": spp, "pool": pool, "mpaths_s": best["paths"] / best["seconds_total"] / 1e6,
                          "mpaths_s_untimed": plain["paths"] / plain["seconds_total"] / 1e6,
                          "mrays_s": rays / best["seconds_total"] / 1e6, "iters": best["iterations"], "launches": best["kernel_launches"],
                          "sec": best["seconds_total"], "stage": [round(x, 5) for x in best["seconds_stage"]],
                          "ext": best["ext_rays"], "shadow": best["shadow_rays"], "probe": best["probe_rays"], "culled": best["shadow_culled"],
                          "mean": float(fb.mean().item())}), flush=True)
    sc.close()
