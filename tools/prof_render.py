"""Small fixed render for profiling under ncu: NAME scene at its full resolution, SPP samples, POOL slots."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
from spt import core, scenes
name = sys.argv[1] if len(sys.argv) > 1 else "c2"
spp = int(sys.argv[2]) if len(sys.argv) > 2 else 4
pool = int(sys.argv[3]) if len(sys.argv) > 3 else 0
cfg = scenes.CONFIGS[name]
ctx = core.Context(0)
sc = core.Scene(ctx, scenes.load(name))
img, st = sc.render(cfg["width"], cfg["height"], spp=spp, seed=1, pool_size=pool)
print(name, spp, "paths", st["paths"], "iters", st["iterations"], "launches", st["kernel_launches"], "dev s", st["seconds_total"],
      "Mpaths/s", st["paths"] / st["seconds_total"] / 1e6)
