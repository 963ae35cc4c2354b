"""GPU-vs-oracle paired-render diagnostics for one scene: where and by how much the two images differ.
    python tools/paired_diag.py c5 [W H SPP]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("simple-path-tracer_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import orc  # noqa: E402
from spt import abi, core, scenes  # noqa: E402

name = sys.argv[1]
w, h, spp = (int(x) for x in sys.argv[2:5]) if len(sys.argv) >= 5 else (128, 72, 4)
fs = scenes.load(name)
ctx = core.Context(0)
sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs, fix_q1=True)
for depth in (1, 2, 3, 8):
    for flags, tag in ((0, "sorted"), (abi.RENDER_NO_SORT, "nosort")):
        gi, gst = sc.render(w, h, spp=spp, seed=11, max_depth=depth, flags=flags)
        oi, ost = osc.render(w, h, spp=spp, seed=11, max_depth=depth, rng_mode=1)
        fin = np.isfinite(oi).all(-1)
        rel = (np.abs(gi - oi) / (np.abs(oi) + 1e-3)).max(-1)
        bad = (rel >= 1e-3) & fin
        print(f"{name} depth {depth} {tag}: within 1e-3 {float((rel[fin] < 1e-3).mean()):.4f} within 1e-2 {float((rel[fin] < 1e-2).mean()):.4f} "
              f"mean ratio {float(gi[fin].mean() / oi[fin].mean()):.5f} median ratio of bad px {float(np.median((gi[bad].sum(-1) + 1e-6) / (oi[bad].sum(-1) + 1e-6))) if bad.any() else 1:.4f} "
              f"ext {gst['ext_rays']} vs {ost['ext_rays']} shadow+culled {gst['shadow_rays'] + gst['shadow_culled']} vs {ost['shadow_rays']} nonfinite {gst['nonfinite']} vs {ost['nonfinite']}")
        if depth == 2 and tag == "sorted" and bad.any():
            ys, xs = np.nonzero(bad)
            for y, x in list(zip(ys, xs))[:8]:
                print("   px", x, y, "gpu", gi[y, x], "oracle", oi[y, x])
# single-sample renders: which samples differ, and is the difference a different path (integer-like) or a small factor
gi, _ = sc.render(w, h, spp=1, seed=11, max_depth=2)
oi, _ = osc.render(w, h, spp=1, seed=11, max_depth=2, rng_mode=1)
rel = (np.abs(gi - oi) / (np.abs(oi) + 1e-3)).max(-1)
bad = rel >= 1e-3
print("1 spp depth 2: bad fraction", float(bad.mean()))
ratios = (gi[bad].sum(-1) + 1e-9) / (oi[bad].sum(-1) + 1e-9)
print("   ratio quantiles of bad samples", np.quantile(ratios, [0.05, 0.25, 0.5, 0.75, 0.95]) if bad.any() else None)
import rays as R  # noqa: E402
pr = R.primary_rays(fs, w, h)
hg, ho = sc.trace(pr), osc.trace(pr)[0]
inst_bad = ho["inst_id"].reshape(h, w)[bad]
print("   instance ids under bad pixels (first 20):", inst_bad[:20], " materials:", [fs.materials[fs.surfaces[fs.instances[i]["surface"]]["material"]]["tag"] if i >= 0 else -1 for i in inst_bad[:20]])
