"""Render a named config with the CPU oracle at low resolution (debug helper)."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle")); sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
import numpy as np
import orc
from spt import scenes
from PIL import Image
name = sys.argv[1]; spp = int(sys.argv[2]) if len(sys.argv) > 2 else 16
w, h = (320, 180) if scenes.CONFIGS[name]["width"] != 512 else (256, 256)
fs = scenes.load(name)
t = time.time(); sc = orc.OracleScene(fs); tb = time.time() - t
t = time.time(); img, st = sc.render(w, h, spp=spp, rng_mode=0); dt = time.time() - t
print(name, "build %.2fs render %.2fs" % (tb, dt), "mean", img.mean(axis=(0, 1)), "%.2f Mpaths/s" % (st["paths"] / dt / 1e6),
      "ext/path %.2f shadow/path %.2f probe/path %.3f nonfinite %d" % (st["ext_rays"] / st["paths"], st["shadow_rays"] / st["paths"], st["probe_rays"] / st["paths"], st["nonfinite"]),
      "nodes/ray %.1f tris/ray %.1f" % (st["nodes_visited"] / (st["ext_rays"] + st["shadow_rays"]), st["tris_tested"] / (st["ext_rays"] + st["shadow_rays"])))
Image.fromarray((np.clip(np.nan_to_num(img), 0, 1) ** (1 / 2.2) * 255).astype(np.uint8)).save(f"/tmp/{name}.png")
