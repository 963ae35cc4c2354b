"""Build time and size of the acceleration structure per builder: host SAH, GPU LBVH, GPU PLOC.
    python tools/bvh_build_times.py c2 c5 c5b"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (os.path.join(ROOT, "simple-path-tracer_b200"), os.path.join(ROOT, "tests"), os.path.join(ROOT, "tools")):
    sys.path.insert(0, p)
from spt import core, scenes  # noqa: E402

names = [a for a in sys.argv[1:] if not a.startswith("--")]
modes = [a[2:] for a in sys.argv[1:] if a.startswith("--")] or ["host", "device_lbvh", "device"]  # e.g. --device: only the GPU PLOC builder
ctx = core.Context(0)
for name in names:
    for mode in modes:
        os.environ["SPT_BVH_BUILD"] = mode
        fs = scenes.load("c5" if name == "c5b" else name)
        if name == "c5b":
            fs = scenes.flatten_instances(fs)
        t = time.time(); sc = core.Scene(ctx, fs); wall = time.time() - t
        i = sc.info()
        print(json.dumps(dict(scene=name, builder={"host": "host SAH", "device_lbvh": "GPU LBVH", "device": "GPU PLOC"}[mode], triangles=i["triangles"],
                              nodes=i["blas_nodes"] + i["tlas_nodes"], seconds_bvh=round(i["seconds_bvh"], 4), seconds_create=round(i["seconds_create"], 3),
                              blas_on_device=i["blas_on_device"])), flush=True)
        sc.close()
