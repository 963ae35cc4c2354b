"""ctypes binding of oracle/_build/liboracle.so — TEST INFRASTRUCTURE ONLY.

May be imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never
by the product package (simple-path-tracer_b200/).  It reuses only the *interface* structs of spt.abi.
"""
import ctypes as C
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "simple-path-tracer_b200"))
from spt import abi  # noqa: E402  (interface structs only)

LIB = os.path.join(HERE, "_build", "liboracle.so")
_lib = None


def build(force=False):
    if force or not os.path.exists(LIB):
        subprocess.check_call(["make", "-C", HERE] + (["-B"] if force else []), stdout=subprocess.DEVNULL)
    return LIB


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(LIB)
        vp = C.c_void_p
        L.orc_last_error.restype = C.c_char_p
        L.orc_scene_create.argtypes = [C.POINTER(abi.SceneDesc), C.POINTER(vp)]
        L.orc_scene_destroy.argtypes = [vp]
        L.orc_scene_destroy.restype = None
        L.orc_set_fix_q1.argtypes = [C.c_int32]
        L.orc_set_fix_q1.restype = None
        L.orc_render.argtypes = [vp, C.POINTER(abi.RenderDesc), C.c_int32, C.c_int32, vp, vp, C.POINTER(abi.Stats)]
        L.orc_render_ex.argtypes = [vp, C.POINTER(abi.RenderDesc), C.c_int32, C.c_int32, vp, vp, C.POINTER(abi.Stats), C.POINTER(RenderOpts)]
        L.orc_trace_batch.argtypes = [vp, vp, C.c_int64, vp, C.c_int32, C.POINTER(abi.Stats)]
        L.orc_occluded_batch.argtypes = [vp, vp, C.c_int64, vp, C.c_int32]
        L.orc_bxdf_batch.argtypes = [vp, vp, C.c_int64, vp]
        L.orc_alias_table.argtypes = [vp, C.c_int32, vp, vp]
        L.orc_alias_sample.argtypes = [vp, C.c_int32, vp, C.c_int64, vp]
        L.orc_light_probs.argtypes = [vp, vp]
        L.orc_env_lookup.argtypes = [vp, vp, C.c_int64, vp]
        L.orc_hg.argtypes = [C.c_float, vp, C.c_int64, vp, vp]
        L.orc_ss_table.argtypes = [vp, vp, C.c_int64, vp]
        L.orc_calc_differential.argtypes = [vp, vp]
        L.orc_rng_stream.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_int32, vp]
        _lib = L
    return _lib


def _chk(st):
    if st != 0:
        raise RuntimeError(f"oracle status {st}: {lib().orc_last_error().decode()}")


RAY_DT = np.dtype([("origin", "<f4", 3), ("t_min", "<f4"), ("direction", "<f4", 3), ("t_max", "<f4")])
HIT_DT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim_id", "<i4"), ("inst_id", "<i4")])


class RenderOpts(C.Structure):  # OrcRenderOpts, oracle.cpp
    _fields_ = [("debug_normal", C.c_int32), ("sample_clamp", C.c_float), ("out_clamped_mean", C.c_void_p)]


class OracleScene:
    def __init__(self, flat_scene, fix_q1=False):
        """fix_q1=True builds the checker's BVH with the intended bucket rule (fast; same hits up to exact ties)."""
        self.fs = flat_scene
        lib().orc_set_fix_q1(1 if fix_q1 else 0)
        self._desc = flat_scene.to_desc()
        self.h = C.c_void_p()
        _chk(lib().orc_scene_create(C.byref(self._desc), C.byref(self.h)))
        lib().orc_set_fix_q1(0)

    def close(self):
        if self.h:
            lib().orc_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, width, height, spp=16, max_depth=8, sampler="recurrence", div=(1, 1), seed=1, camera=0,
               rng_mode=1, n_threads=0, want_rgb8=False, filter_radius=0.5, debug_normal=False, sample_clamp=0.0):
        """rng_mode 0 = reference-faithful per-thread streams, 1 = keyed (paired with the CUDA core).
        filter_radius != 0.5 keeps the reference's per-sample film (film.rs); debug_normal = the cargo feature of pt.rs:113-118;
        sample_clamp > 0 also returns the per-pixel mean of min(sample, clamp) as st["clamped_mean"]."""
        rd = abi.RenderDesc(camera, width, height, max_depth, abi.SAMPLER[sampler], spp, div[0], div[1], filter_radius, seed,
                            0, 0, 0, 1, 0, 0)
        img = np.zeros((height, width, 3), np.float32)
        rgb8 = np.zeros((height, width, 3), np.uint8) if want_rgb8 else None
        clamped = np.zeros((height, width, 3), np.float32) if sample_clamp > 0 else None
        st = abi.Stats()
        opts = RenderOpts(1 if debug_normal else 0, sample_clamp, clamped.ctypes.data if clamped is not None else None)
        _chk(lib().orc_render_ex(self.h, C.byref(rd), rng_mode, n_threads, img.ctypes.data,
                                 rgb8.ctypes.data if want_rgb8 else None, C.byref(st), C.byref(opts)))
        d = st.as_dict()
        if clamped is not None:
            d["clamped_mean"] = clamped
        return (img, rgb8, d) if want_rgb8 else (img, d)

    def trace(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, RAY_DT)
        hits = np.zeros(len(rays), HIT_DT)
        st = abi.Stats()
        _chk(lib().orc_trace_batch(self.h, rays.ctypes.data, len(rays), hits.ctypes.data, n_threads, C.byref(st)))
        return hits, st.as_dict()

    def occluded(self, rays, n_threads=0):
        rays = np.ascontiguousarray(rays, RAY_DT)
        out = np.zeros(len(rays), np.uint8)
        _chk(lib().orc_occluded_batch(self.h, rays.ctypes.data, len(rays), out.ctypes.data, n_threads))
        return out

    def bxdf(self, queries):
        n = len(queries)
        q = (abi.BxdfQuery * n)(*queries)
        r = (abi.BxdfResult * n)()
        _chk(lib().orc_bxdf_batch(self.h, q, n, r))
        return list(r)

    def env_lookup(self, dirs):
        """EnvLight::strength_dist_pdf(wi) -> [n, 4] = (r, g, b, pdf)"""
        d = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
        out = np.zeros((len(d), 4), np.float32)
        _chk(lib().orc_env_lookup(self.h, d.ctypes.data, len(d), out.ctypes.data))
        return out

    def light_probs(self):
        out = np.zeros(len(self.fs.lights), np.float32)
        _chk(lib().orc_light_probs(self.h, out.ctypes.data))
        return out


def hg(g, x):
    """(henyey_greenstein(g, x), henyey_greenstein_cdf_inverse(g, x)) elementwise (medium/util.rs:1-19)"""
    x = np.ascontiguousarray(x, np.float32)
    a, b = np.zeros_like(x), np.zeros_like(x)
    _chk(lib().orc_hg(g, x.ctypes.data, len(x), a.ctypes.data, b.ctypes.data))
    return a, b


def ss_table(rand=()):
    """(SS_CDF_TABLE as [512, 2], Subsurface::sample_r(rand)) (substrate.rs:187-228)"""
    tab = np.zeros((512, 2), np.float32)
    r = np.ascontiguousarray(rand, np.float32)
    out = np.zeros_like(r)
    _chk(lib().orc_ss_table(tab.ctypes.data, r.ctypes.data, len(r), out.ctypes.data))
    return tab, out


def calc_differential(t, o, d, normal, tangent, bitangent, xo, xd, yo, yd):
    """Intersection::calc_differential (intersection.rs:28-83) -> (duvdx, duvdy)"""
    a = np.concatenate([[t], o, d, normal, tangent, bitangent, xo, xd, yo, yd]).astype(np.float32)
    out = np.zeros(4, np.float32)
    _chk(lib().orc_calc_differential(a.ctypes.data, out.ctypes.data))
    return out[:2], out[2:]


def alias_table(props):
    p = np.ascontiguousarray(props, np.float32)
    u = np.zeros(len(p), np.float32)
    k = np.zeros(len(p), np.int64)
    _chk(lib().orc_alias_table(p.ctypes.data, len(p), u.ctypes.data, k.ctypes.data))
    return u, k


def alias_sample(props, rand):
    p = np.ascontiguousarray(props, np.float32)
    r = np.ascontiguousarray(rand, np.float32)
    out = np.zeros(len(r), np.int64)
    _chk(lib().orc_alias_sample(p.ctypes.data, len(p), r.ctypes.data, len(r), out.ctypes.data))
    return out


def rng_stream(seed, pixel, sample, n):
    out = np.zeros(n, np.float32)
    _chk(lib().orc_rng_stream(seed, pixel, sample, n, out.ctypes.data))
    return out
