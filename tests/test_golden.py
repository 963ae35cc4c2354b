"""The oracle against the committed golden fixtures (tests/golden, made by tools/make_golden.py).  Geometry is exact
(+,-,*,/,sqrt only); shading goes through libm, so images compare with a small tolerance across host CPUs."""
import os

import numpy as np
import pytest

import orc
import rays as R
from spt import scenes
from test_oracle_kat import MATS, material_scene, _queries

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", ["c1", "c1b", "c2", "c3", "c4"])
def test_oracle_reproduces_golden(name):
    z = np.load(os.path.join(G, f"{name}.npz"))
    fs = scenes.load(name)
    sc = orc.OracleScene(fs)
    hp, _ = sc.trace(R.primary_rays(fs, 32, 24))
    assert np.array_equal(hp, z["primary_hits"])
    hi, _ = sc.trace(R.incoherent_rays(fs, 2048, seed=11))
    assert np.array_equal(hi["prim_id"], z["incoherent_hits"]["prim_id"]) and np.array_equal(hi["t"], z["incoherent_hits"]["t"])
    assert np.array_equal(sc.occluded(R.incoherent_rays(fs, 2048, seed=12, finite_tmax=True)), z["occluded"])
    img, st = sc.render(24, 16, spp=4, seed=5, rng_mode=1)
    assert st["paths"] == int(z["paths"])
    rel = np.abs(img - z["image"]) / (np.abs(z["image"]) + 1e-3)
    assert (rel.max(-1) < 1e-3).mean() > 0.99
    # the fast checker BVH (fix_q1) returns the same hits as the reference-faithful tree
    fast = orc.OracleScene(fs, fix_q1=True)
    hf, _ = fast.trace(R.primary_rays(fs, 32, 24))
    assert R.compare_hits(hp, hf)["match_rate"] == 1.0 and np.array_equal(hp["t"], hf["t"])


def test_oracle_bxdf_golden():
    z = np.load(os.path.join(G, "bxdf.npz"))
    for mname, (mat, _) in MATS.items():
        sc = orc.OracleScene(material_scene(mat))
        wo, wi, rand, ref = z[mname + "_wo"], z[mname + "_wi"], z[mname + "_rand"], z[mname + "_res"]
        for i in range(len(wo)):
            r = sc.bxdf(_queries(wo[i], [wi[i]], rand[i]))[0]
            got = np.array([*r.sample_wi, *r.sample_f, r.sample_pdf, r.sample_lobe, r.sample_dir, r.n_rand_used, *r.eval_f, r.eval_pdf, r.is_delta, r.bxdf_tag])
            ok = np.isclose(got, ref[i], rtol=1e-4, atol=1e-6) | (~np.isfinite(ref[i]) & ~np.isfinite(got))
            assert ok.all(), (mname, i, got, ref[i])


def test_misc_golden():
    z = np.load(os.path.join(G, "misc.npz"))
    assert np.array_equal(orc.rng_stream(7, 123, 45, 64), z["rng"])
    u, k = orc.alias_table(np.arange(1, 17) / 136.0)
    assert np.allclose(u, z["alias_u"], atol=1e-6) and np.array_equal(k, z["alias_k"])
