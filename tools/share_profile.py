"""Per-iteration profile of one rank's share of a render (fill / steady state / drain), on ONE GPU:
    python tools/share_profile.py c2 256 8      # C2 at 256 spp, the tiles rank 0 of 8 owns
Prints the untimed device time of the share, the share's ideal (whole frame / world), and the per-iteration table
(live paths, shadow rays, ms of logic..generate / connect / extend) from the SPT_RENDER_TIME_STAGES run."""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
import torch  # noqa: E402
from spt import abi, core, scenes  # noqa: E402

name, spp, world = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
cfg = scenes.CONFIGS[name]
w, h = cfg["width"], cfg["height"]
ctx = core.Context(0)
sc = core.Scene(ctx, scenes.load(name))
fb = torch.zeros((h, w, 3), dtype=torch.float32, device="cuda")
full = min(sc.render_device(fb.data_ptr(), w, h, spp=spp, seed=s + 1)["seconds_total"] for s in range(4))
share = [sc.render_device(fb.data_ptr(), w, h, spp=spp, seed=s + 1, rank=0, world_size=world) for s in range(5)]
best = min(share, key=lambda st: st["seconds_total"])
print(f"{name} {spp} spp: whole frame {1e3 * full:.2f} ms; rank 0 of {world}: {1e3 * best['seconds_total']:.2f} ms = {full / world / best['seconds_total']:.3f} of ideal, "
      f"{best['iterations']} iterations, {best['paths']} paths")
path = tempfile.mktemp()
os.environ["SPT_TRACE_ITERS"] = path
st = sc.render_device(fb.data_ptr(), w, h, spp=spp, seed=1, rank=0, world_size=world, flags=abi.RENDER_TIME_STAGES)
del os.environ["SPT_TRACE_ITERS"]
print(f"timed (serial) run: {1e3 * st['seconds_total']:.2f} ms; stage sums ext {1e3 * st['seconds_stage'][1]:.2f} shade {1e3 * st['seconds_stage'][2]:.2f} con {1e3 * st['seconds_stage'][3]:.2f}")
print("iter pool live cursor shadow ms_shade ms_con ms_ext   Mrays/ms(ext)")
prev_live = 0
for line in open(path):
    it, pool, live, cursor, sq, a, b, c = line.split()
    print(f"{int(it):4d} {int(pool):9d} {int(live):9d} {int(cursor):11d} {int(sq):9d} {float(a):7.3f} {float(b):7.3f} {float(c):7.3f}   {int(live) / 1e6 / max(float(c), 1e-6):6.2f}")
