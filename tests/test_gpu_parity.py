"""Parity tests proper: the CUDA core (through the C ABI) against the CPU oracle on the same seeded inputs and
against the committed golden fixtures.  Bars:
  * ray batches: primitive id equal (or equal-t tie) on >= 99.99 % of rays, |dt|/t <= 1e-5 on every id match,
    occlusion bit equal on >= 99.99 %
  * paired-RNG images (same random numbers on both sides): >= 99.5 % of pixels within 1e-3 relative
  * statistical image parity against the reference-faithful oracle mode: relMSE(GPU, A) <= 2 x relMSE(A, B)
"""
import os

import numpy as np
import pytest

import rays as R
from test_oracle_kat import MATS, material_scene, _queries

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SCENES = ["c1", "c1b", "c2", "c3", "c4", "c5"]  # c5 = the north star's multi-GPU config: 2000 instances, deep TLAS


def _assert_paired_images(name, gi, oi, frac_1e3=0.995):
    """Paired renders trace the same paths; pixels agree to float noise.  C5 (and C5b) is the exception that needs a looser count:
    its environment is an HDR map whose sun texels are 200x their neighbours, and EnvLight's bilinear lookup (environment.rs:52-84)
    takes its weights from acos / atan2 of the direction — a 1-ulp difference between the device and the host libm in those weights
    is multiplied by that contrast, so a few per cent of the pixels differ by 1e-3 .. 1e-2 (tools/paired_diag.py: same ray counts,
    per-sample ratios 0.99 .. 1.006).  Those pixels must still agree to 1e-2, and the converged-image test covers C5 statistically."""
    fin = np.isfinite(oi).all(-1)
    rel = (np.abs(gi - oi) / (np.abs(oi) + 1e-3))[fin].max(-1)
    if name.startswith("c5"):
        assert float((rel < 1e-3).mean()) >= 0.95 and float((rel < 1e-2).mean()) >= 0.98, (name, float((rel < 1e-3).mean()), float((rel < 1e-2).mean()))
    else:
        assert float((rel < 1e-3).mean()) >= frac_1e3, (name, float((rel < 1e-3).mean()))


def _assert_hits(cmp_):
    assert cmp_["match_rate"] >= 0.9999, cmp_
    assert cmp_["t_ok_rate"] == 1.0 and cmp_["max_dt_same_id"] <= 1e-5, cmp_


@pytest.mark.parametrize("name", SCENES)
def test_ray_batches_match_oracle(scene_pair, name):
    fs, sc, osc = scene_pair(name)
    pr = R.primary_rays(fs, 512, 288)
    ref, _ = osc.trace(pr)
    _assert_hits(R.compare_hits(ref, sc.trace(pr)))
    br = R.bounce_rays(pr, ref)
    _assert_hits(R.compare_hits(osc.trace(br)[0], sc.trace(br)))
    ir = R.incoherent_rays(fs, 1 << 19)
    c = R.compare_hits(osc.trace(ir)[0], sc.trace(ir))
    _assert_hits(c)
    assert c["bit_exact_t"] >= 0.9999  # the leaf tests replay the reference's f32 operations without FMA
    fr = R.incoherent_rays(fs, 1 << 19, finite_tmax=True)
    assert (osc.occluded(fr) == sc.occluded(fr)).mean() >= 0.9999
    sr = R.incoherent_rays(fs, 1 << 17, seed=77)
    sr["t_max"] = R.FMAX
    assert (osc.occluded(sr) == sc.occluded(sr)).mean() >= 0.9999


def _batches_vs_oracle(fs, sc, osc, pw=512, ph=288, n_inc=1 << 19):
    """P / D / U closest-hit batches + finite and unbounded any-hit batches of one CUDA scene against the oracle"""
    pr = R.primary_rays(fs, pw, ph)
    ref, _ = osc.trace(pr)
    _assert_hits(R.compare_hits(ref, sc.trace(pr)))
    br = R.bounce_rays(pr, ref)
    _assert_hits(R.compare_hits(osc.trace(br)[0], sc.trace(br)))
    ir = R.incoherent_rays(fs, n_inc)
    c = R.compare_hits(osc.trace(ir)[0], sc.trace(ir))
    _assert_hits(c)
    assert c["bit_exact_t"] >= 0.9999
    fr = R.incoherent_rays(fs, n_inc, finite_tmax=True)
    assert (osc.occluded(fr) == sc.occluded(fr)).mean() >= 0.9999
    sb = br.copy(); sb["t_max"] = (np.float32(0.5) + np.arange(len(sb), dtype=np.float32) % np.float32(7.0))  # shadow-ray-like: from surfaces, finite
    assert (osc.occluded(sb) == sc.occluded(sb)).mean() >= 0.9999


@pytest.mark.parametrize("name", ["c2", "c5", "c4"])
@pytest.mark.parametrize("builder", ["device", "device_lbvh"])
def test_ray_batches_match_oracle_with_gpu_built_blas(ctx, scene_pair, monkeypatch, name, builder):
    """The same oracle comparison with every BLAS built by bvh_build_gpu.cu (PLOC / LBVH) instead of the host SAH builder —
    the builder that is the default for meshes of >= 2^18 triangles, forced here on the benchmark scenes."""
    from spt import core
    fs, _, osc = scene_pair(name)
    monkeypatch.setenv("SPT_BVH_BUILD", builder)
    sc = core.Scene(ctx, fs)
    monkeypatch.delenv("SPT_BVH_BUILD")
    assert sc.info()["blas_on_device"] == len(fs.meshes)
    _batches_vs_oracle(fs, sc, osc, 384, 216, 1 << 18)
    gi, gst = sc.render(128, 72, spp=4, seed=11)
    oi, ost = osc.render(128, 72, spp=4, seed=11, rng_mode=1)
    _assert_paired_images(name, gi, oi, 0.99)
    assert gst["paths"] == ost["paths"]
    sc.close()


def test_c5b_flattened_mesh_with_default_gpu_built_blas_matches_oracle(ctx):
    """C5b at 1/16 scale: 120 bunny instances (+ the floor and light quads) baked into ONE mesh of 600 004 unique triangles (>= 2^18, so flatten() builds its BLAS
    with the GPU PLOC builder by default), against the oracle's own BVH over the same mesh (BvhAccel::intersect, bvh.rs:262-283)."""
    import orc
    from spt import core, scenes
    fs = scenes.flatten_instances(scenes.load("c5", nx=12, nz=10))
    assert len(fs.meshes) == 1 and len(fs.meshes[0]["indices"]) == 600004
    sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs, fix_q1=True)
    assert sc.info()["blas_on_device"] == 1
    _batches_vs_oracle(fs, sc, osc, 512, 288, 1 << 18)
    gi, gst = sc.render(128, 72, spp=4, seed=3)
    oi, ost = osc.render(128, 72, spp=4, seed=3, rng_mode=1)
    _assert_paired_images("c5b", gi, oi)
    assert gst["paths"] == ost["paths"]
    sc.close()


@pytest.mark.parametrize("name", SCENES)
def test_golden_vectors(scene_pair, name):
    fs, sc, _ = scene_pair(name)
    z = np.load(os.path.join(G, f"{name}.npz"))
    _assert_hits(R.compare_hits(z["primary_hits"], sc.trace(R.primary_rays(fs, 32, 24))))
    _assert_hits(R.compare_hits(z["incoherent_hits"], sc.trace(R.incoherent_rays(fs, 2048, seed=11))))
    assert (sc.occluded(R.incoherent_rays(fs, 2048, seed=12, finite_tmax=True)) == z["occluded"]).mean() >= 0.999
    img, st = sc.render(24, 16, spp=4, seed=5)
    assert st["paths"] == int(z["paths"])
    _assert_paired_images(name, img, z["image"], 0.98)


def test_edge_cases(scene_pair):
    fs, sc, osc = scene_pair("c1b")
    assert len(sc.trace(np.zeros(0, R.RAY_DT))) == 0 and len(sc.occluded(np.zeros(0, R.RAY_DT))) == 0   # empty batch
    one = R.primary_rays(fs, 1, 1)                                                                          # ragged: 1 ray
    assert R.compare_hits(osc.trace(one)[0], sc.trace(one))["match_rate"] == 1.0
    r = R.incoherent_rays(fs, 4097, seed=3)                                                                 # not a multiple of the block
    r["direction"][::3, 0] = 0.0                                                                            # axis-parallel components
    r["direction"][1::3, 1] = 0.0
    r["direction"][::5] *= 7.5                                                                              # un-normalised directions
    r["t_min"][::4] = 0.25
    r["t_max"][::2] = 3.0
    c = R.compare_hits(osc.trace(r)[0], sc.trace(r))
    assert c["match_rate"] >= 0.999 and c["t_ok_rate"] == 1.0, c
    assert (osc.occluded(r) == sc.occluded(r)).mean() >= 0.999
    z = r.copy(); z["t_max"] = 0.0                                                                          # empty interval: nothing can be hit
    assert (sc.trace(z)["inst_id"] == -1).all() and (sc.occluded(z) == 0).all()


@pytest.mark.parametrize("name,sampler,div", [("c1", "recurrence", (1, 1)), ("c1b", "random", (1, 1)), ("c2", "jittered", (3, 2)),
                                              ("c3", "recurrence", (1, 1)), ("c4", "recurrence", (1, 1)), ("c2", "recurrence", (1, 1)),
                                              ("c5", "recurrence", (1, 1))])
def test_paired_render_matches_oracle(scene_pair, name, sampler, div):
    """Same random numbers on both sides: the two renderers trace the same paths, so images agree to float noise
    except at the rare pixels where a 1-ulp difference flips a discrete decision."""
    fs, sc, osc = scene_pair(name)
    w, h, spp = 160, 96, 6
    gi, gst = sc.render(w, h, spp=spp, seed=11, sampler=sampler, div=div)
    oi, ost = osc.render(w, h, spp=spp, seed=11, sampler=sampler, div=div, rng_mode=1)
    assert gst["paths"] == ost["paths"] == w * h * (div[0] * div[1] if sampler == "jittered" else spp)
    fin = np.isfinite(oi).all(-1)
    _assert_paired_images(name, gi, oi)
    # the reference calls intersect_test for every NEE sample; the CUDA core skips those with an exactly zero contribution
    assert abs((gst["shadow_rays"] + gst["shadow_culled"]) / ost["shadow_rays"] - 1) < 2e-3
    # whole-image mean; C5's is dominated by a few sun-lit firefly pixels (frame mean 5.6 for a median of 0.5), where ONE diverged path
    # moves it by per cents: compare it with the pixels clamped at 32 there
    cl = (lambda x: np.minimum(x, np.float32(32.0))) if name == "c5" else (lambda x: x)
    assert abs(float(cl(gi[fin]).mean()) / float(cl(oi[fin]).mean()) - 1) < 0.05


@pytest.mark.parametrize("name", ["c1b", "c2"])
def test_statistical_image_parity_vs_reference_faithful_oracle(scene_pair, name):
    """The reference's own RNG discipline (one entropy-seeded stream and sampler state per thread) cannot be paired,
    so here parity is statistical: relMSE(GPU, oracle seed A) <= 2 x relMSE(oracle seed A, oracle seed B)."""
    fs, sc, osc = scene_pair(name)
    w, h, spp = 96, 54, 256
    gi, _ = sc.render(w, h, spp=spp, seed=3)
    a, _ = osc.render(w, h, spp=spp, seed=1, rng_mode=0)
    b, _ = osc.render(w, h, spp=spp, seed=2, rng_mode=0)
    clamp = lambda x: np.minimum(np.nan_to_num(x), 8.0)  # fireflies (Fresnel > 1 on back-facing conductors) dominate any L2 metric
    relmse = lambda x, r: float((((clamp(x) - clamp(r)) ** 2) / (clamp(r) ** 2 + 1e-2)).mean())
    assert relmse(gi, a) <= 2.0 * relmse(b, a) + 1e-4, (relmse(gi, a), relmse(b, a))
    assert abs(clamp(gi).mean() / clamp(a).mean() - 1) < 0.02


def test_bxdf_hook_matches_oracle_and_golden(ctx):
    from spt import core
    z = np.load(os.path.join(G, "bxdf.npz"))
    for mname, (mat, _) in MATS.items():
        sc = core.Scene(ctx, material_scene(mat))
        wo, wi, rand, ref = z[mname + "_wo"], z[mname + "_wi"], z[mname + "_rand"], z[mname + "_res"]
        qs = [_queries(wo[i], [wi[i]], rand[i])[0] for i in range(len(wo))]
        res = sc.bxdf(qs)
        for i, r in enumerate(res):
            got = np.array([*r.sample_wi, *r.sample_f, r.sample_pdf, r.sample_lobe, r.sample_dir, r.n_rand_used, *r.eval_f, r.eval_pdf, r.is_delta, r.bxdf_tag])
            ok = np.isclose(got, ref[i], rtol=2e-3, atol=1e-5) | (~np.isfinite(ref[i]) & ~np.isfinite(got))
            assert ok.all(), (mname, i, got, ref[i])
        sc.close()


def test_image_is_independent_of_pool_size_and_rank_split(scene_pair):
    """Streams are keyed by (seed, pixel, sample) and the film is integer: the image is bit-identical for any pool
    size, and the sum of the per-rank images of a 2- or 3-way tile split equals the 1-rank image exactly (this is the
    N-GPU invariance test, run as N sequential ranks on one GPU)."""
    fs, sc, _ = scene_pair("c2")
    w, h, spp = 200, 120, 8
    base, st = sc.render(w, h, spp=spp, seed=5)
    small, _ = sc.render(w, h, spp=spp, seed=5, pool_size=4096)
    assert np.array_equal(base, small)
    for world in (2, 3):
        parts = [sc.render(w, h, spp=spp, seed=5, rank=r, world_size=world) for r in range(world)]
        acc = np.zeros_like(base)
        for img, pst in parts:
            assert np.all((img == 0) | (acc == 0))           # tiles are disjoint
            acc += img
        assert np.array_equal(acc, base)
        assert sum(p[1]["paths"] for p in parts) == st["paths"]
    again, _ = sc.render(w, h, spp=spp, seed=5)
    assert np.array_equal(base, again)                          # run-to-run determinism
    other, _ = sc.render(w, h, spp=spp, seed=6)
    assert not np.array_equal(base, other)


def test_render_finishes_every_sample_when_an_iteration_is_all_tile_padding(scene_pair):
    """Big tiles on an image whose height is not a multiple of the tile: the bottom tiles hold runs of padding ordinals longer than
    a small pool, so with max_depth 1 (every path ends at its first hit) a whole iteration can start nothing.  The loop must go on
    until the sample cursor is exhausted (it used to stop at live == 0 and leave black regions)."""
    fs, sc, _ = scene_pair("c1b")
    w, h, spp = 200, 100, 3
    base, st = sc.render(w, h, spp=spp, seed=8, max_depth=1)
    big, st2 = sc.render(w, h, spp=spp, seed=8, max_depth=1, tile=(64, 64), pool_size=1024)
    assert st["paths"] == st2["paths"] == w * h * spp
    assert np.array_equal(base, big)
    deep, st3 = sc.render(w, h, spp=spp, seed=8, tile=(64, 64), pool_size=1024)
    ref, _ = sc.render(w, h, spp=spp, seed=8)
    assert st3["paths"] == w * h * spp and np.array_equal(deep, ref)


@pytest.mark.parametrize("name", ["c2", "c4", "c3"])
def test_image_is_independent_of_kernel_scheduling(scene_pair, name):
    """Multi-stream overlap, stage timing and the counting build only change WHEN a path is advanced, never what it
    computes: with the integer film the images must be bit-identical.  Without material binning (NO_SORT) the same
    formulas run in the all-materials kernel, whose compiled arithmetic differs in the last ulp: equal to float noise."""
    from spt import abi
    fs, sc, _ = scene_pair(name)
    w, h, spp = 160, 90, 6
    base, st = sc.render(w, h, spp=spp, seed=4)
    for flags in (abi.RENDER_SERIAL, abi.RENDER_NO_SORT, abi.RENDER_TIME_STAGES, abi.RENDER_COUNT_TRAVERSAL, abi.RENDER_NO_SORT | abi.RENDER_SERIAL):
        img, st2 = sc.render(w, h, spp=spp, seed=4, flags=flags)
        if flags & abi.RENDER_NO_SORT:
            rel = np.abs(base - img) / (np.abs(base) + 1e-3)
            assert float((rel.max(-1) < 1e-3).mean()) >= 0.995 and st["paths"] == st2["paths"]
            assert abs(st2["ext_rays"] / st["ext_rays"] - 1) < 1e-3
            continue
        assert np.array_equal(base, img), flags
        for k in ("paths", "ext_rays", "shadow_rays", "probe_rays", "shadow_culled", "nonfinite"):
            assert st[k] == st2[k], (flags, k)
        if flags & abi.RENDER_COUNT_TRAVERSAL:
            rays = st2["ext_rays"] + st2["shadow_rays"] + st2["probe_rays"]
            busy, pop, node, inst, tri = st2["trav_profile"][:5]
            assert st2["nodes_visited"] >= rays > 0 and node <= st2["nodes_visited"] and inst <= st2["insts_visited"]
            if st2["probe_rays"] == 0:  # BSSRDF probes are traced inside the shade kernel, outside the staged driver
                assert node == st2["nodes_visited"] and inst == st2["insts_visited"]
            assert busy >= node and tri <= st2["tris_tested"]   # spheres count as primitive tests but have no triangle stage


def test_render_gather_with_a_one_rank_communicator_equals_render_device(ctx, scene_pair):
    """spt_comm_* / spt_render_gather on the degenerate one-rank communicator (NCCL is loaded and a communicator created, nothing
    travels): the film is left unresolved by the render and reconstructed by the gather path — same frame as spt_render_device."""
    import torch
    from spt import core
    fs, sc, _ = scene_pair("c2")
    comm = core.Comm(ctx, core.Comm.unique_id(ctx.lib), 0, 1)
    for radius in (0.5, 1.0):
        a = torch.zeros((90, 160, 3), dtype=torch.float32, device="cuda")
        b = torch.zeros_like(a)
        st = sc.render_gather(comm, a.data_ptr(), 160, 90, spp=4, seed=2, filter_radius=radius)
        sc.render_device(b.data_ptr(), 160, 90, spp=4, seed=2, filter_radius=radius)
        torch.cuda.synchronize()
        assert torch.equal(a, b) and st["paths"] == 160 * 90 * 4 and st["seconds_stage"][5] > 0
    comm.close()


@pytest.mark.parametrize("name", ["c2", "c4", "c5"])
def test_lean_and_general_shade_kernels_agree(scene_pair, monkeypatch, name):
    """Scenes whose textures are all plain scalars run the SPT_LEAN compilation of the shade kernels (no texture-expression evaluator,
    no TextureInput record); SPT_SHADE_LEAN=0 forces the general kernels on the same scene.  Same formulas, separately compiled:
    the images agree to float noise and every counter is equal."""
    fs, sc, _ = scene_pair(name)
    w, h, spp = 160, 90, 6
    lean, st = sc.render(w, h, spp=spp, seed=4)
    monkeypatch.setenv("SPT_SHADE_LEAN", "0")
    full, st2 = sc.render(w, h, spp=spp, seed=4)
    monkeypatch.delenv("SPT_SHADE_LEAN")
    _assert_paired_images(name, lean, full)
    assert st["paths"] == st2["paths"] and abs(st2["ext_rays"] / st["ext_rays"] - 1) < 1e-3


def test_rgb8_encode_and_device_output(scene_pair):
    import torch
    fs, sc, _ = scene_pair("c1b")
    img, rgb8, _ = sc.render(64, 48, spp=4, seed=2, want_rgb8=True)
    assert np.array_equal(rgb8, np.clip(img * np.float32(255.0), 0, 255).astype(np.uint8))   # film.rs:94-99
    buf = torch.zeros((48, 64, 3), dtype=torch.float32, device="cuda")
    sc.render_device(buf.data_ptr(), 64, 48, spp=4, seed=2)
    torch.cuda.synchronize()
    assert np.array_equal(buf.cpu().numpy(), img)


def test_textured_scene_matches_oracle(ctx):
    """Image texture with mip pyramid + sRGB + tiling/offset/wrap modifier + binary operators + normal map + emissive map."""
    import orc
    from spt import core, loader, scenes
    g = np.random.default_rng(4)
    tex = (g.random((37, 53, 4)) * 255).astype(np.uint8)
    tex[..., 3] = 255
    nm = np.zeros((16, 16, 4), np.uint8); nm[..., 0] = 128 + (g.random((16, 16)) * 60 - 30).astype(np.int32); nm[..., 1] = 128; nm[..., 2] = 255; nm[..., 3] = 255
    d = scenes.c1b()
    d["textures"] += [{"type": "image_array", "name": "img", "rgba": tex.tolist(), "is_srgb": True, "tiling": [3.0, 2.0], "offset": [0.25, 0.5], "wrap": "mirror_repeat"},
                      {"type": "image_array", "name": "nmap", "rgba": nm.tolist()},
                      {"type": "scalar", "name": "half", "value": [0.5, 0.5, 0.5]},
                      {"type": "mul", "name": "img_half", "t1": "img", "t2": "half"},
                      {"type": "add", "name": "img_plus", "t1": "img_half", "t2": "blue"},
                      {"type": "image_array", "name": "em", "rgba": tex[::-1].tolist(), "wrap": "clamp"}]
    d["materials"] += [{"type": "lambert", "name": "textured", "albedo": "img_plus"},
                       {"type": "pbr_metallic", "name": "pbr_tex", "base_color": "img", "metallic": "img", "roughness": "half"}]
    d["surfaces"] += [{"name": "bumpy", "material": "textured", "normal_map": "nmap"},
                      {"name": "glow", "material": "lambert_black", "emissive": [20.0, 20.0, 20.0], "emissive_map": "em"}]
    d["instances"][1] = dict(d["instances"][1], surface="bumpy"); d["instances"][1].pop("material")
    d["instances"][0]["material"] = "pbr_tex"
    d["instances"][3]["surface"] = "glow"
    fs = loader.load_scene(d, scenes.BASE_PATH)
    sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs)
    gi, gst = sc.render(160, 160, spp=4, seed=9)
    oi, ost = osc.render(160, 160, spp=4, seed=9, rng_mode=1)
    rel = np.abs(gi - oi) / (np.abs(oi) + 1e-3)
    assert float((rel.max(-1) < 2e-3).mean()) >= 0.99, float((rel.max(-1) < 2e-3).mean())
    assert gst["paths"] == ost["paths"]
    sc.close()


def test_empty_scene_and_degenerate_triangles(ctx, tmp_path):
    """Edge cases of the domain: (a) a scene without any instance (empty aggregate: every ray misses, bvh.rs:238-283 on an empty tree) under an
    environment — the frame is the environment seen by the camera; (b) a mesh holding zero-area triangles (two coincident vertices, three collinear
    ones) and a sliver next to regular ones: Triangle::intersect_ray rejects det == 0 only (triangle.rs:131), everything else must agree with the oracle."""
    import orc
    from spt import core, loader, scenes
    d = scenes.c3()
    d["instances"] = []
    fs = loader.load_scene(d, scenes.BASE_PATH)
    sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs)
    r = R.primary_rays(fs, 64, 36)
    h = sc.trace(r)
    assert (h["inst_id"] == -1).all() and np.array_equal(h["t"], r["t_max"]) and (sc.occluded(r) == 0).all()
    gi, gst = sc.render(64, 36, spp=4, seed=3)
    oi, ost = osc.render(64, 36, spp=4, seed=3, rng_mode=1)
    assert gst["paths"] == ost["paths"] and gst["shadow_rays"] == 0 and gst["ext_rays"] == gst["paths"]
    assert np.allclose(gi, oi, rtol=2e-3, atol=1e-5) and gi.mean() > 0
    sc.close()
    pos = np.array([[-1, 0, -1], [1, 0, -1], [1, 0, 1], [-1, 0, 1],            # a regular quad ...
                    [0, 0.5, 0], [0, 0.5, 0], [1, 0.5, 0],                        # ... a triangle with two coincident vertices
                    [-1, 0.7, 0], [0, 0.7, 0], [1, 0.7, 0],                       # ... three collinear vertices
                    [-1, 0.9, -1], [1, 0.9, -1], [1, 0.9 + 1e-7, -1 + 1e-6]], np.float32)   # ... and a sliver
    idx = np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6], [7, 8, 9], [10, 11, 12]], np.uint32)
    np.savez(tmp_path / "degenerate.npz", positions=pos, indices=idx)
    d = scenes.c1b()
    d["primitives"] = d["primitives"] + [{"type": "trimesh", "name": "deg", "obj_file": str(tmp_path / "degenerate.npz")}]
    d["instances"][1] = {"name": "floor", "primitive": "deg", "material": "lambert_blue", "translate": [0.0, -1.0, 0.0], "scale": [5.0, 1.0, 5.0]}
    fs = loader.load_scene(d, scenes.BASE_PATH)
    sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs, fix_q1=True)
    for rays in (R.primary_rays(fs, 256, 144), R.incoherent_rays(fs, 1 << 16)):
        _assert_hits(R.compare_hits(osc.trace(rays)[0], sc.trace(rays)))
    down = np.zeros(5, R.RAY_DT)                                                  # straight down through each degenerate triangle's "interior"
    down["origin"] = [[0.5 * 5, 3, 0], [0, 3, 0], [2.5, 3, 0], [0.1, 3, 0.2], [4.9, 3, -4.999]]
    down["direction"] = [0, -1, 0]; down["t_min"] = 1e-4; down["t_max"] = R.FMAX
    a, b = osc.trace(down)[0], sc.trace(down)
    assert np.array_equal(a["prim_id"], b["prim_id"]) and np.array_equal(a["t"], b["t"])
    fr = R.incoherent_rays(fs, 1 << 14, finite_tmax=True)
    assert (osc.occluded(fr) == sc.occluded(fr)).mean() >= 0.9999
    gi, _ = sc.render(96, 64, spp=4, seed=2)
    oi, _ = osc.render(96, 64, spp=4, seed=2, rng_mode=1)
    _assert_paired_images("c1b", gi, oi, 0.99)
    sc.close()


def test_delta_lights_and_spot(ctx):
    import orc
    from spt import core, loader, scenes
    d = scenes.c1b()
    d["lights"] = [{"type": "point", "name": "p", "position": [0.0, 3.0, 2.0], "strength": [30.0, 20.0, 10.0]},
                   {"type": "spot", "name": "s", "position": [2.0, 4.0, 3.0], "direction": [-0.4, -0.8, -0.5], "inner_angle": 10.0,
                    "outer_angle": 35.0, "strength": [80.0, 80.0, 80.0]},
                   {"type": "directional", "name": "d", "direction": [0.3, -1.0, -0.2], "strength": [1.0, 1.0, 2.0]}]
    for ls in ("uniform", "power_is"):
        d["light_sampler"] = ls
        fs = loader.load_scene(d, scenes.BASE_PATH)
        sc, osc = core.Scene(ctx, fs), orc.OracleScene(fs)
        gi, _ = sc.render(128, 96, spp=4, seed=1)
        oi, _ = osc.render(128, 96, spp=4, seed=1, rng_mode=1)
        rel = np.abs(gi - oi) / (np.abs(oi) + 1e-3)
        assert float((rel.max(-1) < 1e-3).mean()) >= 0.995
        sc.close()


def _r2_offsets(w, h, spp):
    """closed-form sub-pixel offsets of the recurrence sampler (csrc/rng.cuh r2_offset): arrays [h*w, spp]"""
    pix = np.arange(w * h, dtype=np.uint64)[:, None]
    n = ((pix * np.uint64(spp) + np.arange(spp, dtype=np.uint64)[None, :]) & np.uint64(0xFFFFFFFF)).astype(np.uint32) + np.uint32(1)
    X = (np.uint32(0x80000000) + n * np.uint32(3242174889)).astype(np.uint32)
    Y = (np.uint32(0x80000000) + n * np.uint32(2447445413)).astype(np.uint32)
    return (X >> np.uint32(8)).astype(np.float32) * np.float32(1 / 16777216.0), (Y >> np.uint32(8)).astype(np.float32) * np.float32(1 / 16777216.0)


@pytest.mark.parametrize("radius", [1.0, 1.5, 0.3, 2.5])
def test_box_filter_radius_reproduces_film_rs(scene_pair, radius):
    """Film::filter_pixel with a box radius other than 0.5 (film.rs:68-91, boxf.rs:10-34, Q14): unweighted colour sum over the
    neighbourhood / number of samples inside the box.  (a) against the oracle, which keeps the reference's per-sample film, with
    paired random numbers, for all three samplers; (b) exactly, against the same reconstruction done in numpy on the GPU's own
    radius-0.5 frame with the closed-form recurrence offsets (isolates k_resolve_box from path noise)."""
    fs, sc, osc = scene_pair("c1b")
    w, h = 72, 40
    for sampler, div, spp in (("recurrence", (1, 1), 6), ("random", (1, 1), 5), ("jittered", (3, 2), 6)):
        kw = dict(spp=spp, seed=4, sampler=sampler, div=div, filter_radius=radius)
        gi, g8, gst = sc.render(w, h, want_rgb8=True, **kw)
        oi, o8, ost = osc.render(w, h, want_rgb8=True, rng_mode=1, **kw)
        assert gst["paths"] == ost["paths"] == w * h * spp
        fin = np.isfinite(oi).all(-1)
        assert np.array_equal(fin, np.isfinite(gi).all(-1))          # pixels without a sample inside the box are inf / NaN on both sides
        rel = (np.abs(gi - oi) / (np.abs(oi) + 1e-3))[fin]
        assert float((rel.max(-1) < 2e-3).mean()) >= 0.97, (sampler, float((rel.max(-1) < 2e-3).mean()))
        assert (np.abs(g8.astype(int) - o8.astype(int)) <= 1).mean() >= 0.97
        assert np.array_equal(g8[fin], np.clip(gi[fin] * np.float32(255.0), 0, 255).astype(np.uint8))
    spp = 6
    base, _ = sc.render(w, h, spp=spp, seed=4)                       # own-pixel means: sums = mean * spp
    got, _ = sc.render(w, h, spp=spp, seed=4, filter_radius=radius)
    ox, oy = _r2_offsets(w, h, spp)
    R = int(np.ceil(np.float32(radius) - np.float32(0.5)))
    num, cnt = np.zeros((h, w, 3), np.float64), np.zeros((h, w), np.float64)
    sums = base.astype(np.float64) * spp
    fx, fy = ox.reshape(h, w, spp) - np.float32(0.5), oy.reshape(h, w, spp) - np.float32(0.5)
    for j in range(-R, R + 1):
        for i in range(-R, R + 1):
            ys, xs = slice(max(0, -j), min(h, h - j)), slice(max(0, -i), min(w, w - i))          # output pixels whose neighbour (x+i, y+j) exists
            yn, xn = slice(max(0, -j) + j, min(h, h - j) + j), slice(max(0, -i) + i, min(w, w - i) + i)
            num[ys, xs] += sums[yn, xn]
            inside = (np.abs(np.float32(i) + fx[yn, xn]) <= np.float32(radius)) & (np.abs(np.float32(j) + fy[yn, xn]) <= np.float32(radius))
            cnt[ys, xs] += inside.sum(-1)
    with np.errstate(divide="ignore", invalid="ignore"):
        want = num / cnt[..., None]
    ok = cnt > 0
    assert np.allclose(got[ok], want[ok], rtol=2e-6, atol=1e-7)
    assert not np.isfinite(got[~ok]).any()


@pytest.mark.parametrize("name", ["c1", "c2", "c4", "c5"])
def test_debug_normal_aov_matches_oracle(scene_pair, name):
    """The reference's `debug_normal` feature (pt.rs:113-118): first-hit interpolated normal as a colour.  Paired sub-pixel offsets:
    the two images are the same function of the same camera rays."""
    from spt import abi
    fs, sc, osc = scene_pair(name)
    w, h, spp = 192, 108, 4
    gi, gst = sc.render(w, h, spp=spp, seed=2, flags=abi.RENDER_DEBUG_NORMAL)
    oi, ost = osc.render(w, h, spp=spp, seed=2, rng_mode=1, debug_normal=True)
    assert gst["paths"] == ost["paths"] == w * h * spp and gst["shadow_rays"] == 0 and gst["ext_rays"] == gst["paths"]
    d = np.abs(gi - oi).max(-1)
    assert (d < 1e-4).mean() >= 0.999, float((d < 1e-4).mean())      # the rest: a silhouette sample that grazes another triangle
    assert 0.0 <= gi.min() and gi.max() <= 1.0 + 1e-6 or fs.env is not None
    again, _ = sc.render(w, h, spp=spp, seed=2, flags=abi.RENDER_DEBUG_NORMAL, pool_size=4096)
    assert np.array_equal(gi, again)


def test_errors_are_reported_not_swallowed(ctx, scene_pair):
    from spt import abi, core, loader, scenes
    fs, sc, _ = scene_pair("c1")
    with pytest.raises(abi.SptError, match="radius"):
        sc.render(32, 32, spp=1, filter_radius=4.0)
    with pytest.raises(abi.SptError, match="radius"):
        sc.render(32, 32, spp=1, filter_radius=0.0)
    with pytest.raises(abi.SptError, match="consecutive"):
        sc.render(32, 32, spp=8, spp_range=(0, 2))
        sc.render(32, 32, spp=8, spp_range=(4, 2), flags=abi.RENDER_ACCUMULATE, filter_radius=1.0)
    with pytest.raises(abi.SptError, match="camera"):
        sc.render(32, 32, spp=1, camera=3)
    with pytest.raises(abi.SptError):
        sc.render(0, 32, spp=1)
    d = scenes.c1(); d["lights"] = []
    nolight = core.Scene(ctx, loader.load_scene(d, scenes.BASE_PATH))
    with pytest.raises(abi.SptError, match="no light"):
        nolight.render(16, 16, spp=1)
    nolight.close()
    # malformed descriptors are refused by spt_scene_create instead of reading out of bounds on the device
    for field, value, msg in (("light_sampler", 7, "light_sampler"), ("env_light_index", 5, "env_light_index")):
        bad = loader.load_scene(scenes.c1(), scenes.BASE_PATH)
        setattr(bad, field, value)
        with pytest.raises(abi.SptError, match=msg):
            core.Scene(ctx, bad)
    bad = loader.load_scene(scenes.c2(), scenes.BASE_PATH)
    bad.materials[3]["rough_chan"] = 9
    with pytest.raises(abi.SptError, match="rough_chan"):
        core.Scene(ctx, bad)


@pytest.mark.parametrize("name", ["c2", "c4"])
def test_device_built_bvh_is_interchangeable_with_the_host_built_one(ctx, monkeypatch, name):
    """bvh_build_gpu.cu (Morton sort -> radix tree -> refit -> wide collapse on the GPU) against the host SAH builder: hits
    depend on the leaf tests, not on the tree, so ids agree (up to equal-t ties), t is bit-identical, occlusion agrees and
    the rendered image is the same."""
    from spt import core, scenes
    fs = scenes.load(name)
    monkeypatch.setenv("SPT_BVH_BUILD", "host")
    host = core.Scene(ctx, fs)
    monkeypatch.setenv("SPT_BVH_BUILD", "device")
    dev = core.Scene(ctx, fs)
    monkeypatch.delenv("SPT_BVH_BUILD")
    hi, di = host.info(), dev.info()
    assert hi["blas_on_device"] == 0 and di["blas_on_device"] == len(fs.meshes) and di["triangles"] == hi["triangles"]
    assert di["seconds_bvh"] > 0 and di["seconds_create"] >= di["seconds_bvh"]
    import orc
    osc = orc.OracleScene(fs, fix_q1=True)
    prim = R.primary_rays(fs, 256, 144)
    for batch in (prim, R.bounce_rays(prim, host.trace(prim)), R.incoherent_rays(fs, 1 << 15)):
        a, b = host.trace(batch), dev.trace(batch)
        cmp_ = R.compare_hits(a, b)
        assert cmp_["match_rate"] == 1.0 and cmp_["bit_exact_t"] == 1.0 and cmp_["tie_rate"] < 1e-3, cmp_
        _assert_hits(R.compare_hits(osc.trace(batch)[0], b))    # and the device-built scene against the ORACLE, not only against CUDA
    shadow = R.incoherent_rays(fs, 1 << 15, finite_tmax=True)
    assert np.array_equal(host.occluded(shadow), dev.occluded(shadow))
    ia, _ = host.render(160, 96, spp=8, seed=11)
    ib, _ = dev.render(160, 96, spp=8, seed=11)
    assert (np.abs(ia - ib).max(-1) <= 1e-6 * (np.abs(ia).max(-1) + 1e-3)).mean() > 0.999  # an equal-t tie may pick the other triangle
    host.close(); dev.close()


@pytest.mark.parametrize("name, sampler, div", [("c2", "recurrence", (1, 1)), ("c4", "random", (1, 1)), ("c1b", "jittered", (4, 3))])
def test_progressive_passes_and_film_resume_add_up_bit_for_bit(ctx, name, sampler, div):
    """SPT_RENDER_ACCUMULATE + spp_first/spp_count: a frame rendered as passes over consecutive sample ranges — also with the
    film saved, the scene destroyed and the film loaded into a fresh scene in between — equals the one-call frame exactly;
    every intermediate output is the mean over the samples so far."""
    from spt import abi, core, scenes
    fs = scenes.load(name)
    w, h = 176, 96
    spp = div[0] * div[1] if sampler == "jittered" else 12
    kw = dict(spp=spp, sampler=sampler, div=div, seed=21)
    sc = core.Scene(ctx, fs)
    full, st = sc.render(w, h, **kw)
    film_full = sc.film_save(w, h)
    a, sa = sc.render(w, h, spp_range=(0, 5), **kw)
    assert sa["paths"] == w * h * 5
    alone, _ = sc.render(w, h, spp=5, sampler=sampler, div=div, seed=21) if sampler == "random" else (None, None)
    if alone is not None:
        assert np.array_equal(a, alone)                       # the first 5 samples of 12 are the 5 samples of a 5-spp frame (random sampler)
    film5 = sc.film_save(w, h)
    b, sb = sc.render(w, h, spp_range=(5, 3), flags=abi.RENDER_ACCUMULATE, **kw)
    film8 = sc.film_save(w, h)
    assert np.array_equal(np.float32((film8.astype(np.float64) / 16777216.0).astype(np.float32) * np.float32(1.0 / 8)), b)
    sc.close()
    sc2 = core.Scene(ctx, fs)                                   # "another process": fresh scene, film from the checkpoint
    with pytest.raises(abi.SptError, match="needs the film of a previous pass"):
        sc2.render(w, h, spp_range=(8, spp - 8), flags=abi.RENDER_ACCUMULATE, **kw)
    sc2.film_load(film8, 8)
    c, sc_st = sc2.render(w, h, spp_range=(8, spp - 8), flags=abi.RENDER_ACCUMULATE, **kw)
    assert np.array_equal(c, full) and np.array_equal(sc2.film_save(w, h), film_full)
    assert sa["paths"] + sb["paths"] + sc_st["paths"] == st["paths"]
    assert not np.array_equal(film5, film8)
    with pytest.raises(abi.SptError, match="exceeds the frame's spp"):
        sc2.render(w, h, spp_range=(spp - 1, 2), **kw)
    sc2.close()
