"""Thin Python handles over the C ABI (include/spt.h): Context, Scene (device-resident flattened scene) and the
calls a host makes on them.  All compute happens in libspt_cuda.so; nothing here falls back to the CPU."""
import ctypes as C

import numpy as np

from . import abi

RAY_DT = np.dtype([("origin", "<f4", 3), ("t_min", "<f4"), ("direction", "<f4", 3), ("t_max", "<f4")])
HIT_DT = np.dtype([("t", "<f4"), ("u", "<f4"), ("v", "<f4"), ("prim_id", "<i4"), ("inst_id", "<i4")])


class Context:
    def __init__(self, device=0):
        self.lib = abi.load_library()
        self.h = C.c_void_p()
        abi.check(self.lib, self.lib.spt_context_create(device, C.byref(self.h)))
        self.device = device

    def close(self):
        if self.h:
            self.lib.spt_context_destroy(self.h)
            self.h = C.c_void_p()


class Comm:
    """spt_comm_create: this process' rank in the library's NCCL communicator (one process per GPU).  `id_bytes`: the 128 bytes rank 0
    got from Comm.unique_id(), handed to every rank by the host (see Comm.from_torch)."""

    @staticmethod
    def _load_other_nccl_consumers_first():
        """A process holds one libnccl.so.2.  PyTorch bundles its own (newer than the system's) and must bind to THAT one, so it has
        to be loaded before the library dlopen()s libnccl.so.2 — else a later `import torch` fails on a missing symbol."""
        try:
            import torch  # noqa: F401
        except ImportError:
            pass

    def __init__(self, ctx, id_bytes, rank, world_size):
        self._load_other_nccl_consumers_first()
        self.ctx, self.lib, self.rank, self.world_size = ctx, ctx.lib, rank, world_size
        buf = (C.c_uint8 * abi.COMM_ID_BYTES).from_buffer_copy(bytes(id_bytes))
        self.h = C.c_void_p()
        abi.check(self.lib, self.lib.spt_comm_create(ctx.h, buf, rank, world_size, C.byref(self.h)))

    @staticmethod
    def unique_id(lib=None):
        Comm._load_other_nccl_consumers_first()
        lib = lib or abi.load_library()
        buf = (C.c_uint8 * abi.COMM_ID_BYTES)()
        abi.check(lib, lib.spt_comm_unique_id(buf))
        return bytes(buf)

    @classmethod
    def from_torch(cls, ctx):
        """Join the ranks of an initialised torch.distributed process group: rank 0 creates the id, a broadcast hands it round."""
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        dev = torch.device("cuda", ctx.device) if dist.get_backend() == "nccl" else torch.device("cpu")
        t = torch.zeros(abi.COMM_ID_BYTES, dtype=torch.uint8, device=dev)
        if rank == 0:
            t.copy_(torch.frombuffer(bytearray(cls.unique_id(ctx.lib)), dtype=torch.uint8))
        dist.broadcast(t, src=0)
        return cls(ctx, bytes(t.cpu().numpy().tobytes()), rank, world)

    def close(self):
        if self.h:
            self.lib.spt_comm_destroy(self.h)
            self.h = C.c_void_p()


class Scene:
    """spt_scene_create on a FlatScene (spt.loader)."""

    def __init__(self, ctx, flat_scene):
        self.ctx, self.lib, self.fs = ctx, ctx.lib, flat_scene
        desc = flat_scene.to_desc()
        self.h = C.c_void_p()
        abi.check(self.lib, self.lib.spt_scene_create(ctx.h, C.byref(desc), C.byref(self.h)))

    def close(self):
        if self.h:
            self.lib.spt_scene_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        i = abi.SceneInfo()
        abi.check(self.lib, self.lib.spt_scene_info(self.h, C.byref(i)))
        return {k: getattr(i, k) for k, _ in i._fields_}

    def upload(self):
        n = C.c_int64()
        abi.check(self.lib, self.lib.spt_scene_upload(self.h, C.byref(n)))
        return n.value

    # ---- seam B2
    def trace(self, rays):
        rays = np.ascontiguousarray(rays, RAY_DT)
        hits = np.zeros(len(rays), HIT_DT)
        abi.check(self.lib, self.lib.spt_trace_batch(self.h, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def trace_noalloc(self, rays):
        """spt_trace_batch_noalloc: spt_trace_batch for at most abi.NOALLOC_MAX_RAYS rays, with its device scratch in a fixed
        file-scope array — no allocation on the call path."""
        rays = np.ascontiguousarray(rays, RAY_DT)
        hits = np.zeros(len(rays), HIT_DT)
        abi.check(self.lib, self.lib.spt_trace_batch_noalloc(self.h, rays.ctypes.data, len(rays), hits.ctypes.data))
        return hits

    def occluded(self, rays):
        rays = np.ascontiguousarray(rays, RAY_DT)
        out = np.zeros(len(rays), np.uint8)
        abi.check(self.lib, self.lib.spt_occluded_batch(self.h, rays.ctypes.data, len(rays), out.ctypes.data))
        return out

    def trace_device(self, d_rays_ptr, n, d_hits_ptr, flags=0, stream=None, want_stats=False):
        st = abi.Stats()
        abi.check(self.lib, self.lib.spt_trace_batch_device(self.h, d_rays_ptr, n, d_hits_ptr, flags,
                                                            C.byref(st) if want_stats else None, stream))
        return st.as_dict() if want_stats else None

    def occluded_device(self, d_rays_ptr, n, d_out_ptr, flags=0, stream=None, want_stats=False):
        st = abi.Stats()
        abi.check(self.lib, self.lib.spt_occluded_batch_device(self.h, d_rays_ptr, n, d_out_ptr, flags,
                                                               C.byref(st) if want_stats else None, stream))
        return st.as_dict() if want_stats else None

    def bxdf(self, queries):
        n = len(queries)
        q = (abi.BxdfQuery * n)(*queries)
        r = (abi.BxdfResult * n)()
        abi.check(self.lib, self.lib.spt_bxdf_batch(self.h, q, n, r))
        return list(r)

    # ---- seam B1
    def render_desc(self, width, height, spp=16, max_depth=8, sampler="recurrence", div=(1, 1), seed=1, camera=0,
                    rank=0, world_size=1, pool_size=0, flags=0, tile=(0, 0), filter_radius=0.5, spp_range=(0, 0)):
        return abi.RenderDesc(camera, width, height, max_depth, abi.SAMPLER[sampler], spp, div[0], div[1], filter_radius,
                              seed, tile[0], tile[1], rank, world_size, pool_size, flags, spp_range[0], spp_range[1])

    def render(self, width, height, want_rgb8=False, out=None, **kw):
        """spt_render: host buffers out.  Returns (mean image [H,W,3] f32, stats) or (mean, rgb8, stats).  `out`: an existing
        C-contiguous f32 [H,W,3] array to receive the image (e.g. page-locked memory, so the device-to-host copy is one DMA)."""
        rd = self.render_desc(width, height, **kw)
        if out is not None:
            if out.dtype != np.float32 or out.shape != (height, width, 3) or not out.flags["C_CONTIGUOUS"]:
                raise ValueError("out must be a C-contiguous float32 array of shape (height, width, 3)")
            img = out
        else:
            img = np.zeros((height, width, 3), np.float32)
        rgb8 = np.zeros((height, width, 3), np.uint8) if want_rgb8 else None
        st = abi.Stats()
        abi.check(self.lib, self.lib.spt_render(self.h, C.byref(rd), img.ctypes.data,
                                                rgb8.ctypes.data if want_rgb8 else None, C.byref(st)))
        return (img, rgb8, st.as_dict()) if want_rgb8 else (img, st.as_dict())

    # ---- progressive dump / resume
    def film_save(self, width, height):
        """spt_film_save: the int64 fixed-point film of the last render, [H,W,3]."""
        out = np.zeros((height, width, 3), np.int64)
        abi.check(self.lib, self.lib.spt_film_save(self.h, out.ctypes.data, out.size))
        return out

    def film_load(self, film, spp_done):
        """spt_film_load: resume from a saved film that holds spp_done samples per pixel."""
        film = np.ascontiguousarray(film, np.int64)
        abi.check(self.lib, self.lib.spt_film_load(self.h, film.ctypes.data, film.shape[1], film.shape[0], spp_done))

    def render_gather(self, comm, d_out_ptr, width, height, root=0, **kw):
        """spt_render_gather: every rank of `comm` renders its tiles; the root receives the tile-packed film of the others over NCCL and
        leaves the full frame at d_out_ptr (device memory, width*height*3 f32; ignored on the other ranks).  Returns this rank's stats."""
        kw.pop("rank", None); kw.pop("world_size", None)
        rd = self.render_desc(width, height, rank=comm.rank, world_size=comm.world_size, **kw)
        st = abi.Stats()
        abi.check(self.lib, self.lib.spt_render_gather(self.h, comm.h, C.byref(rd), root, d_out_ptr, C.byref(st)))
        return st.as_dict()

    def render_device(self, d_out_ptr, width, height, **kw):
        """spt_render_device: result stays in device memory at d_out_ptr (width*height*3 f32)."""
        rd = self.render_desc(width, height, **kw)
        st = abi.Stats()
        abi.check(self.lib, self.lib.spt_render_device(self.h, C.byref(rd), d_out_ptr, C.byref(st)))
        return st.as_dict()


def owned_tiles(width, height, rank, world_size, tile=(0, 0)):
    """spt_owned_tiles: host-only, works without a GPU."""
    lib = abi.load_library()
    n = C.c_int64()
    abi.check(lib, lib.spt_owned_tiles(width, height, tile[0], tile[1], rank, world_size, None, 0, C.byref(n)))
    out = np.zeros(max(1, n.value), np.uint32)
    abi.check(lib, lib.spt_owned_tiles(width, height, tile[0], tile[1], rank, world_size, out.ctypes.data, n.value, C.byref(n)))
    return out[:n.value]
