"""Strong-scaling anatomy on ONE GPU: render the tiles of rank r of W (no NCCL involved) and compare the device time with
1/W of the full frame.  Separates the per-render fixed cost (pool ramp-up / drain) and the tile imbalance from the collective.
    python tools/rank_share.py c2 256 [--worlds 1,2,4,8]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
import torch  # noqa: E402
from spt import core, scenes  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("name")
ap.add_argument("spp", type=int)
ap.add_argument("--worlds", default="1,2,4,8")
ap.add_argument("--pool", type=int, default=0)
ap.add_argument("--tile", default="0x0")
args = ap.parse_args()
tile = tuple(int(x) for x in args.tile.split("x"))
cfg = scenes.CONFIGS[args.name]
w, h = cfg["width"], cfg["height"]
ctx = core.Context(0)
sc = core.Scene(ctx, scenes.load(args.name))
fb = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
full = None
for world in [int(x) for x in args.worlds.split(",")]:
    per_rank = []
    for rank in range(world):
        best = None
        for rep in range(3):
            st = sc.render_device(fb.data_ptr(), w, h, spp=args.spp, seed=rep + 1, rank=rank, world_size=world, pool_size=args.pool, tile=tile)
            if rep and (best is None or st["seconds_total"] < best["seconds_total"]):
                best = st
        per_rank.append(best)
    t = [s["seconds_total"] * 1e3 for s in per_rank]
    if world == 1:
        full = t[0]
    print(json.dumps({"scene": args.name, "spp": args.spp, "world": world, "tile": args.tile, "ms_max": max(t), "ms_min": min(t), "ideal_ms": full / world,
                      "efficiency": full / world / max(t), "iters": [s["iterations"] for s in per_rank][:2],
                      "mpaths_s_rank0": per_rank[0]["paths"] / per_rank[0]["seconds_total"] / 1e6,
                      "ms_per_rank": [round(x, 2) for x in t]}), flush=True)
