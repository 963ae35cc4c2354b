"""Host-side mirror of the reference's renderer interface (src/renderer/mod.rs:9-38), backed by the CUDA core.

    renderer = create_renderer("pt_wavefront", max_depth, pixel_sampler, filter)     # mod.rs:26-38
    renderer.render(scene, OutputConfig(width, height, output_filename, used_camera_name))   # mod.rs:16-19

`"pt_wavefront"` is the arm a Rust host would add next to `"pt"`; `"pt"` itself — the reference's
CPU integrator — is not part of this product and is rejected (there is no CPU fallback).

Multi-GPU (one process per GPU): when torch.distributed is initialised, each rank renders its interleaved tiles with
the scene replicated, and the owned tiles of every rank are gathered on rank 0 inside the library (spt_render_gather: integer film over NCCL).
"""
from dataclasses import dataclass
from typing import Optional

import numpy as np

from . import abi, core, loader


@dataclass
class PixelSampler:  # pixel_sampler/{random,jittered,recurrence}.rs
    kind: str = "recurrence"
    spp: int = 256
    division_x: int = 1
    division_y: int = 1

    def total_spp(self):
        return self.division_x * self.division_y if self.kind == "jittered" else self.spp


@dataclass
class BoxFilter:  # filter/boxf.rs
    radius: float = 0.5


@dataclass
class OutputConfig:  # renderer/mod.rs:9-14
    width: int = 512
    height: int = 512
    output_filename: Optional[str] = None
    used_camera_name: Optional[str] = None


class WavefrontPathTracer:
    """RendererT for the CUDA core (≙ PathTracer, src/renderer/pt.rs:24-37)."""

    def __init__(self, max_depth, pixel_sampler, filter, seed=1, pool_size=0, flags=0):
        self.max_depth, self.pixel_sampler, self.filter = max_depth, pixel_sampler, filter
        self.seed, self.pool_size, self.flags = seed, pool_size, flags
        self.last_stats = None
        self._comm = None

    def _kw(self, scene, config, rank=0, world=1):
        ps = self.pixel_sampler
        return dict(spp=ps.spp, max_depth=self.max_depth, sampler=ps.kind, div=(ps.division_x, ps.division_y),
                    seed=self.seed, camera=scene.fs.camera_index(config.used_camera_name), rank=rank, world_size=world,
                    pool_size=self.pool_size, flags=self.flags, filter_radius=self.filter.radius)

    def render(self, scene: "core.Scene", config: OutputConfig):
        """Returns the linear f32 mean image [H,W,3] (on rank 0 when distributed; None elsewhere) and writes the 8-bit
        image (film.rs:94-99: x255, clamp, no gamma) to config.output_filename if given."""
        try:
            import torch.distributed as dist
            distributed = dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1
        except ImportError:
            distributed = False
        if not distributed:
            img, rgb8, st = scene.render(config.width, config.height, want_rgb8=True, **self._kw(scene, config))
            self.last_stats = st
        else:
            import torch
            from . import core
            rank, world = dist.get_rank(), dist.get_world_size()
            dev = torch.device("cuda", scene.ctx.device)
            if self._comm is None:  # the library's own communicator: torch.distributed only hands the id round
                self._comm = core.Comm.from_torch(scene.ctx)
            buf = torch.empty((config.height, config.width, 3), dtype=torch.float32, device=dev)
            # every rank renders its tiles; only owned tiles travel (integer film, ncclSend/ncclRecv), rank 0 reconstructs the frame
            self.last_stats = scene.render_gather(self._comm, buf.data_ptr(), config.width, config.height, root=0, **self._kw(scene, config, rank, world))
            if rank != 0:
                return None
            img = buf.cpu().numpy()
            rgb8 = np.clip(np.nan_to_num(img * np.float32(255.0), nan=0.0), 0.0, 255.0).astype(np.uint8)
        if config.output_filename:
            from PIL import Image
            Image.fromarray(rgb8).save(config.output_filename)
        return img


def create_renderer(ty, max_depth, pixel_sampler, filter, **kw):
    """renderer::create_renderer, src/renderer/mod.rs:26-38."""
    if ty == "pt_wavefront":
        return WavefrontPathTracer(max_depth, pixel_sampler, filter, **kw)
    if ty == "pt":
        raise abi.SptError("renderer - type 'pt' is the reference's CPU integrator; this library only provides 'pt_wavefront'")
    raise loader.SceneError(f"renderer - unknown type '{ty}'")


def load_renderer(source, ty_override="pt_wavefront", **kw):
    """loader::load_renderer (src/loader/mod.rs:8-18): reads scenes/pt.json-style files; the `type` found in the file
    is replaced by `ty_override` unless that is None."""
    r = loader.load_renderer(source)
    s = r["sampler"]
    kind = {v: k for k, v in abi.SAMPLER.items()}[s["kind"]]
    ps = PixelSampler(kind, s["spp"], s["div_x"], s["div_y"])
    return create_renderer(ty_override or r["type"], r["max_depth"], ps, BoxFilter(r["filter_radius"]), **kw)
