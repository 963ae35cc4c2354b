"""The host acceleration-structure build checked WITHOUT a GPU: flatten_scene() (csrc/flatten.cpp + bvh_build.cpp: SAH build, collapse to
8-wide, 8-bit quantised child boxes, TLAS with the tightened instance boxes) is traversed on the CPU by tools/trav_sim.cpp — the same node
decoding, slab test and stage order as csrc/traverse.cuh — and compared with the oracle's BvhAccel::intersect / intersect_test
(bvh.rs:238-283).  Boxes only cull: if the quantisation or a tightened TLAS box ever cut a true hit away, it shows here as a missed hit."""
import ctypes as C
import importlib.util
import os

import numpy as np
import pytest

import orc
import rays as R
from spt import scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("trav_sim", os.path.join(ROOT, "tools", "trav_sim.py"))
ts = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ts)


@pytest.fixture(scope="module")
def sim():
    return ts.build()


def _trace(L, h, rays, any_hit):
    rays = np.ascontiguousarray(rays, R.RAY_DT)
    t = np.zeros(len(rays), np.float32)
    hit = np.zeros(len(rays), np.uint8)
    assert L.sim_trace(h, rays.ctypes.data, len(rays), 1 if any_hit else 0, t.ctypes.data, hit.ctypes.data) == 0
    return t, hit


@pytest.mark.parametrize("name", ["c1b", "c2", "c4", "c5"])
def test_host_built_wide_bvh_never_culls_a_true_hit(sim, name):
    fs = scenes.load(name)
    desc = fs.to_desc()
    h = sim.sim_scene_create(C.byref(desc))
    assert h
    osc = orc.OracleScene(fs, fix_q1=True)
    pr = R.primary_rays(fs, 160, 90)
    ref, _ = osc.trace(pr)
    batches = [pr, R.bounce_rays(pr, ref), R.incoherent_rays(fs, 30000), R.incoherent_rays(fs, 30000, seed=5, finite_tmax=True)]
    for rays in batches:
        ref, _ = osc.trace(rays)
        t, hit = _trace(sim, h, rays, False)
        ref_hit = ref["inst_id"] >= 0
        assert (hit.astype(bool) == ref_hit).mean() >= 0.9999, name
        both = hit.astype(bool) & ref_hit
        rel = np.abs(t[both].astype(np.float64) - ref["t"][both]) / np.maximum(np.abs(ref["t"][both]), 1e-6)
        assert (rel <= 1e-5).mean() >= 0.9999 and both.sum() > 100, (name, float(rel.max()))
        occ = osc.occluded(rays)
        _, ahit = _trace(sim, h, rays, True)
        assert (ahit == occ).mean() >= 0.9999, name
    sim.sim_scene_destroy(h)
