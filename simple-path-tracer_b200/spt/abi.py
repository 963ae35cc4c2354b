"""ctypes mirror of include/spt.h (the C ABI of the CUDA core) and the loader of the product library.

Only the interface lives here: struct layouts, enum values and the `load_library()` that opens
``simple-path-tracer_b200/lib/libspt_cuda.so``.  There is no CPU fallback — if the shared library is
missing, `load_library()` raises.
"""
import ctypes as C
import os

c_f = C.c_float
P = C.POINTER

# enums (include/spt.h)
PRIM_SPHERE, PRIM_TRIMESH = 0, 1
MAT = dict(lambert=0, pseudo=1, conductor=2, dielectric=3, plastic=4, pbr_metallic=5, pbr_specular=6, subsurface=7)
TEX = dict(scalar=0, image=1, add=2, sub=3, mul=4, div=5, srgb=6, input_modifier=7)
TEXMODE = dict(specified=0, texcoords=1, position=2, normal=3, tangent=4, bitangent=5)
TEXWRAP = dict(repeat=0, mirror_repeat=1, clamp=2, mirror_clamp=3)
LIGHT = dict(directional=0, point=1, spot=2, shape=3, env=4)
LS = dict(uniform=0, power_is=1)
SAMPLER = dict(random=0, jittered=1, recurrence=2)
AGG = dict(group=0, bvh=1)
RENDER_COUNT_TRAVERSAL = 1
RENDER_NO_SORT = 2
RENDER_TIME_STAGES = 4
RENDER_SERIAL = 8
RENDER_ACCUMULATE = 16
RENDER_DEBUG_NORMAL = 32


class MeshDesc(C.Structure):
    _fields_ = [("n_verts", C.c_int32), ("n_tris", C.c_int32), ("position", P(c_f)), ("normal", P(c_f)),
                ("texcoord", P(c_f)), ("tangent", P(c_f)), ("bitangent", P(c_f)), ("index", P(C.c_uint32))]


class SphereDesc(C.Structure):
    _fields_ = [("center", c_f * 3), ("radius", c_f)]


class InstanceDesc(C.Structure):
    _fields_ = [("prim_kind", C.c_int32), ("prim_index", C.c_int32), ("surface", C.c_int32), ("xform", c_f * 12)]


class SurfaceDesc(C.Structure):
    _fields_ = [("material", C.c_int32), ("normal_map", C.c_int32), ("emissive", c_f * 3),
                ("emissive_map", C.c_int32), ("double_sided", C.c_int32), ("inside_medium", C.c_int32)]


class MaterialDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("tex", C.c_int32 * 4), ("ior", c_f), ("rough_chan", C.c_int32),
                ("metal_chan", C.c_int32)]


class TextureDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("child", C.c_int32 * 2), ("value", c_f * 3), ("tiling", c_f * 3),
                ("offset", c_f * 3), ("mode", C.c_int32), ("wrap", C.c_int32), ("width", C.c_int32),
                ("height", C.c_int32), ("rgba", P(C.c_uint8))]


class MediumDesc(C.Structure):
    _fields_ = [("sigma_a", c_f * 3), ("sigma_s", c_f * 3), ("asymmetric", c_f)]


class LightDesc(C.Structure):
    _fields_ = [("tag", C.c_int32), ("position", c_f * 3), ("direction", c_f * 3), ("strength", c_f * 3),
                ("inner_angle", c_f), ("outer_angle", c_f), ("instance", C.c_int32)]


class EnvDesc(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("rgb", P(c_f)), ("scale", c_f * 3)]


class CameraDesc(C.Structure):
    _fields_ = [("eye", c_f * 3), ("forward", c_f * 3), ("up", c_f * 3), ("fov", c_f)]


class SceneDesc(C.Structure):
    _fields_ = [("n_meshes", C.c_int32), ("meshes", P(MeshDesc)),
                ("n_spheres", C.c_int32), ("spheres", P(SphereDesc)),
                ("n_instances", C.c_int32), ("instances", P(InstanceDesc)),
                ("n_surfaces", C.c_int32), ("surfaces", P(SurfaceDesc)),
                ("n_materials", C.c_int32), ("materials", P(MaterialDesc)),
                ("n_textures", C.c_int32), ("textures", P(TextureDesc)),
                ("n_mediums", C.c_int32), ("mediums", P(MediumDesc)),
                ("n_lights", C.c_int32), ("lights", P(LightDesc)),
                ("n_cameras", C.c_int32), ("cameras", P(CameraDesc)),
                ("env", P(EnvDesc)), ("env_light_index", C.c_int32), ("light_sampler", C.c_int32),
                ("aggregate", C.c_int32)]


class Ray(C.Structure):
    _fields_ = [("origin", c_f * 3), ("t_min", c_f), ("direction", c_f * 3), ("t_max", c_f)]


class Hit(C.Structure):
    _fields_ = [("t", c_f), ("u", c_f), ("v", c_f), ("prim_id", C.c_int32), ("inst_id", C.c_int32)]


class BxdfQuery(C.Structure):
    _fields_ = [("material", C.c_int32), ("texcoord", c_f * 2), ("wo", c_f * 3), ("wi", c_f * 3), ("rand", c_f * 8)]


class BxdfResult(C.Structure):
    _fields_ = [("sample_wi", c_f * 3), ("sample_f", c_f * 3), ("sample_pdf", c_f), ("sample_lobe", C.c_int32),
                ("sample_dir", C.c_int32), ("n_rand_used", C.c_int32), ("eval_f", c_f * 3), ("eval_pdf", c_f),
                ("is_delta", C.c_int32), ("bxdf_tag", C.c_int32)]


class RenderDesc(C.Structure):
    _fields_ = [("camera_index", C.c_int32), ("width", C.c_int32), ("height", C.c_int32), ("max_depth", C.c_int32),
                ("sampler_kind", C.c_int32), ("spp", C.c_int32), ("div_x", C.c_int32), ("div_y", C.c_int32),
                ("filter_radius", c_f), ("seed", C.c_uint64), ("tile_w", C.c_int32), ("tile_h", C.c_int32),
                ("rank", C.c_int32), ("world_size", C.c_int32), ("pool_size", C.c_int32), ("flags", C.c_uint32),
                ("spp_first", C.c_int32), ("spp_count", C.c_int32)]


class Stats(C.Structure):
    _fields_ = [("paths", C.c_uint64), ("ext_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("probe_rays", C.c_uint64), ("nodes_visited", C.c_uint64), ("tris_tested", C.c_uint64),
                ("insts_visited", C.c_uint64), ("nonfinite", C.c_uint64), ("iterations", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("seconds_total", C.c_double), ("seconds_stage", C.c_double * 6),
                ("shadow_culled", C.c_uint64), ("trav_profile", C.c_uint64 * 10)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("seconds_stage", "trav_profile")}
        d["trav_profile"] = list(self.trav_profile)
        d["seconds_stage"] = list(self.seconds_stage)
        return d


class SceneInfo(C.Structure):
    _fields_ = [("tlas_nodes", C.c_int64), ("blas_nodes", C.c_int64), ("triangles", C.c_int64),
                ("instances", C.c_int64), ("bytes_nodes", C.c_int64), ("bytes_tris", C.c_int64),
                ("bytes_total", C.c_int64), ("seconds_bvh", C.c_double), ("seconds_create", C.c_double),
                ("blas_on_device", C.c_int64)]


# every symbol include/spt.h declares (tests check the .so exports exactly these)
SYMBOLS = ["spt_last_error", "spt_version", "spt_context_create", "spt_context_destroy", "spt_scene_create",
           "spt_scene_upload", "spt_scene_info", "spt_scene_destroy", "spt_trace_batch", "spt_occluded_batch",
           "spt_trace_batch_device", "spt_occluded_batch_device", "spt_bxdf_batch", "spt_render",
           "spt_render_device", "spt_film_save", "spt_film_load", "spt_owned_tiles",
           "spt_comm_unique_id", "spt_comm_create", "spt_comm_destroy", "spt_render_gather",
           "spt_trace_batch_noalloc"]
NOALLOC_MAX_RAYS = 65536
COMM_ID_BYTES = 128

PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.path.join(PKG_ROOT, "lib", "libspt_cuda.so")
_lib = None


class SptError(RuntimeError):
    pass


def load_library(path=None):
    """Open the CUDA core.  Fails loudly when it has not been built (no CPU fallback exists)."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    path = path or os.environ.get("SPT_LIB") or LIB_PATH  # SPT_LIB: A/B builds of the same library while tuning
    if not os.path.exists(path):
        raise SptError(f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                       "(the product has no CPU fallback)")
    lib = C.CDLL(path, mode=C.RTLD_GLOBAL)
    vp = C.c_void_p
    lib.spt_last_error.restype = C.c_char_p
    lib.spt_version.restype = C.c_int32
    sig = {
        "spt_context_create": [C.c_int32, P(vp)],
        "spt_context_destroy": [vp],
        "spt_scene_create": [vp, P(SceneDesc), P(vp)],
        "spt_scene_upload": [vp, P(C.c_int64)],
        "spt_scene_info": [vp, P(SceneInfo)],
        "spt_scene_destroy": [vp],
        "spt_trace_batch": [vp, vp, C.c_int64, vp],
        "spt_occluded_batch": [vp, vp, C.c_int64, vp],
        "spt_trace_batch_noalloc": [vp, vp, C.c_int64, vp],
        "spt_trace_batch_device": [vp, vp, C.c_int64, vp, C.c_uint32, P(Stats), vp],
        "spt_occluded_batch_device": [vp, vp, C.c_int64, vp, C.c_uint32, P(Stats), vp],
        "spt_bxdf_batch": [vp, vp, C.c_int64, vp],
        "spt_render": [vp, P(RenderDesc), vp, vp, P(Stats)],
        "spt_render_device": [vp, P(RenderDesc), vp, P(Stats)],
        "spt_film_save": [vp, vp, C.c_int64],
        "spt_film_load": [vp, vp, C.c_int32, C.c_int32, C.c_int32],
        "spt_owned_tiles": [C.c_int32] * 6 + [vp, C.c_int64, P(C.c_int64)],
        "spt_comm_unique_id": [vp],
        "spt_comm_create": [vp, vp, C.c_int32, C.c_int32, P(vp)],
        "spt_comm_destroy": [vp],
        "spt_render_gather": [vp, vp, P(RenderDesc), C.c_int32, vp, P(Stats)],
    }
    for name, args in sig.items():
        fn = getattr(lib, name)
        fn.argtypes = args
        fn.restype = C.c_int32
    if path == LIB_PATH:
        _lib = lib
    return lib


def check(lib, status):
    if status != 0:
        msg = lib.spt_last_error()
        raise SptError(f"spt status {status}: {msg.decode() if msg else '?'}")
