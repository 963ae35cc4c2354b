"""One-off converter (run in the build container, where /root/reference exists): turns the reference's mesh
*assets* (not sources) into compact .npz files under assets/ so that tests and bench.py can run on the GPU box,
where /root/reference does not exist.  Output arrays are exactly what spt.meshio.load_obj returns."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
from spt import meshio  # noqa: E402

SRC = "/root/reference/scenes/models"
for name in ("bunny_5k", "cube", "plane"):
    pos, nrm, uv, idx = meshio.load_obj(os.path.join(SRC, name + ".obj"))
    out = os.path.join(ROOT, "assets", name + ".npz")
    np.savez_compressed(out, positions=pos, normals=nrm, texcoords=uv, indices=idx)
    print(name, pos.shape, idx.shape, os.path.getsize(out))
