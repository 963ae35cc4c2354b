"""Development check to run on the GPU box: seam-B2 parity and paired-RNG image parity against the CPU oracle."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("oracle", "simple-path-tracer_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np
import orc, rays as R
from spt import core, scenes
from PIL import Image

out = os.path.join(ROOT, "gpurun_out"); os.makedirs(out, exist_ok=True)
names = sys.argv[1:] or ["c1", "c1b", "c2", "c3", "c4"]
ctx = core.Context(0)
for name in names:
    fs = scenes.load(name)
    t = time.time(); sc = core.Scene(ctx, fs); tb = time.time() - t
    osc = orc.OracleScene(fs, fix_q1=True)
    print(f"== {name}: build {tb:.3f}s info {sc.info()}", flush=True)
    w, h = 320, 180
    pr = R.primary_rays(fs, w, h)
    ir = R.incoherent_rays(fs, 200000)
    ref, _ = osc.trace(pr); got = sc.trace(pr)
    print("  primary   ", R.compare_hits(ref, got), flush=True)
    br = R.bounce_rays(pr, ref)
    ref, _ = osc.trace(br); got = sc.trace(br)
    print("  bounce    ", R.compare_hits(ref, got), flush=True)
    ref, _ = osc.trace(ir); got = sc.trace(ir)
    print("  incoherent", R.compare_hits(ref, got), flush=True)
    fr = R.incoherent_rays(fs, 200000, finite_tmax=True)
    a = osc.occluded(fr); b = sc.occluded(fr)
    print("  occluded agree", float((a == b).mean()), "rate", float(a.mean()), flush=True)
    spp = 4
    t = time.time(); gi, gst = sc.render(w, h, spp=spp, seed=7); tg = time.time() - t
    oi, ost = osc.render(w, h, spp=spp, seed=7, rng_mode=1)
    d = np.abs(gi - oi); den = np.abs(oi) + 1e-3
    finite = np.isfinite(oi).all(-1)
    relerr = (d / den)[finite]
    print(f"  render gpu {tg:.3f}s dev {gst['seconds_total']:.4f}s iters {gst['iterations']} paths {gst['paths']} ext {gst['ext_rays']} sh {gst['shadow_rays']} probe {gst['probe_rays']} nonfinite {gst['nonfinite']}")
    print(f"  oracle paths {ost['paths']} ext {ost['ext_rays']} sh {ost['shadow_rays']} probe {ost['probe_rays']} nonfinite {ost['nonfinite']}")
    print(f"  paired: mean gpu {gi[finite].mean(0)} oracle {oi[finite].mean(0)}; pixels within 1e-3 rel: {float((relerr.max(-1) < 1e-3).mean()):.5f}, within 1e-2: {float((relerr.max(-1) < 1e-2).mean()):.5f}", flush=True)
    Image.fromarray((np.clip(np.nan_to_num(gi), 0, 1) ** (1 / 2.2) * 255).astype(np.uint8)).save(os.path.join(out, f"{name}_gpu.png"))
    Image.fromarray((np.clip(np.nan_to_num(oi), 0, 1) ** (1 / 2.2) * 255).astype(np.uint8)).save(os.path.join(out, f"{name}_orc.png"))
    sc.close()
