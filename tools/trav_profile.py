"""Warp-round profile of the traversal kernels (COUNT build): rounds per ray and lanes active per stage.
    python tools/trav_profile.py c2:4 c5:1"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "simple-path-tracer_b200"))
from spt import abi, core, scenes
ctx = core.Context(0)
for item in sys.argv[1:]:
    name, spp = item.split(":")
    cfg = scenes.CONFIGS[name]
    sc = core.Scene(ctx, scenes.load(name))
    img, st = sc.render(cfg["width"], cfg["height"], spp=int(spp), seed=1, flags=abi.RENDER_COUNT_TRAVERSAL)
    rays = st["ext_rays"] + st["shadow_rays"] + st["probe_rays"]
    busy, pop, node, inst, tri, wr, wpop, wnode, winst, wtri = st["trav_profile"]
    print(f"{name}: rays {rays} nodes/ray {st['nodes_visited']/rays:.2f} tris/ray {st['tris_tested']/rays:.2f} insts/ray {st['insts_visited']/rays:.2f}")
    print(f"  rounds per ray {busy/rays:.2f}; lane occupancy (has ray) {busy/(32*wr):.2f}")
    for nm, l, w in (("pop", pop, wpop), ("node", node, wnode), ("inst", inst, winst), ("tri", tri, wtri)):
        print(f"  {nm:5s}: executed in {w/wr:.2f} of warp-rounds, {l/max(1,w):5.1f} lanes when executed, {l/rays:.2f} per ray")
    sc.close()
