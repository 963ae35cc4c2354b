"""Generates tests/golden/*.npz from the CPU oracle (run in the build container).  The reference has no golden vectors
of its own, so these fixtures pin the ORACLE against regressions and give the CUDA core fixed vectors to
hit on the GPU box; they are not outputs of the Rust binary (which cannot be built here)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("oracle", "simple-path-tracer_b200", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
import numpy as np
import orc, rays as R
from spt import scenes
from test_oracle_kat import MATS, material_scene, _queries

G = os.path.join(ROOT, "tests", "golden")
os.makedirs(G, exist_ok=True)
NAMES = [a for a in sys.argv[1:] if not a.startswith("-")] or ["c1", "c1b", "c2", "c3", "c4", "c5"]  # `make_golden.py c5` writes only that file
for name in [n for n in NAMES if n != "bxdf"]:
    fs = scenes.load(name)
    sc = orc.OracleScene(fs)          # the reference-faithful BVH (Q1/Q2/Q3 tie order)
    pr = R.primary_rays(fs, 32, 24)
    hp, _ = sc.trace(pr)
    ir = R.incoherent_rays(fs, 2048, seed=11)
    hi, _ = sc.trace(ir)
    fr = R.incoherent_rays(fs, 2048, seed=12, finite_tmax=True)
    oc = sc.occluded(fr)
    img, st = sc.render(24, 16, spp=4, seed=5, rng_mode=1)
    np.savez_compressed(os.path.join(G, f"{name}.npz"), primary_hits=hp, incoherent_hits=hi, occluded=oc, image=img,
                        paths=st["paths"], shadow_rays=st["shadow_rays"])
    print(name, float((hp["inst_id"] >= 0).mean()), img.mean())
if NAMES != ["c1", "c1b", "c2", "c3", "c4", "c5"] and "bxdf" not in NAMES:
    sys.exit(0)
# BxDF vectors
g = np.random.default_rng(99)
out = {}
for mname, (mat, delta) in MATS.items():
    sc = orc.OracleScene(material_scene(mat))
    wo = g.normal(size=(64, 3)); wo /= np.linalg.norm(wo, axis=1, keepdims=True)
    wi = g.normal(size=(64, 3)); wi /= np.linalg.norm(wi, axis=1, keepdims=True)
    rand = g.random((64, 8))
    res = [sc.bxdf(_queries(wo[i], [wi[i]], rand[i]))[0] for i in range(64)]
    out[mname + "_wo"] = wo.astype(np.float32); out[mname + "_wi"] = wi.astype(np.float32); out[mname + "_rand"] = rand.astype(np.float32)
    out[mname + "_res"] = np.array([[*r.sample_wi, *r.sample_f, r.sample_pdf, r.sample_lobe, r.sample_dir, r.n_rand_used, *r.eval_f, r.eval_pdf, r.is_delta, r.bxdf_tag] for r in res], np.float64)
np.savez_compressed(os.path.join(G, "bxdf.npz"), **out)
np.savez_compressed(os.path.join(G, "misc.npz"), rng=orc.rng_stream(7, 123, 45, 64),
                    alias_u=orc.alias_table(np.arange(1, 17) / 136.0)[0], alias_k=orc.alias_table(np.arange(1, 17) / 136.0)[1])
print("golden written")
