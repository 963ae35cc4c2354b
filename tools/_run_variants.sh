timeout 800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/r1_pytest_v14.log
timeout 600 python tools/ray_batch_bench.py c5b --log2n 22 --out gpurun_out/ray_batches_c5b_dev.json > gpurun_out/ray_batches_c5b_dev.log 2>&1
python - <<'PY' > gpurun_out/bvh_build_times.log 2>&1
import os, sys, time, json
sys.path.insert(0, 'simple-path-tracer_b200'); sys.path.insert(0, 'tools')
from spt import core, scenes
import importlib.util
ctx = core.Context(0)
src = open('tools/ray_batch_bench.py').read().split('def timed')[0].split('import torch')[1]
import numpy as np
from spt import abi
exec('import rays as R\n' if False else '')
exec('def flatten_instances' + src.split('def flatten_instances')[1])
for name in ('c2', 'c5', 'c5b'):
    for mode in ('host', 'device'):
        os.environ['SPT_BVH_BUILD'] = mode
        fs = scenes.load('c5' if name == 'c5b' else name)
        if name == 'c5b': fs = flatten_instances(fs)
        t = time.time(); sc = core.Scene(ctx, fs); wall = time.time() - t
        i = sc.info()
        print(json.dumps(dict(scene=name, builder=mode, triangles=i['triangles'], nodes=i['blas_nodes'] + i['tlas_nodes'], seconds_bvh=round(i['seconds_bvh'], 4),
                              seconds_create=round(i['seconds_create'], 3), wall_incl_python=round(wall, 3), blas_on_device=i['blas_on_device'])), flush=True)
        sc.close()
PY
