"""CPU model of the persistent-warp traversal scheduler (tools/trav_sim.cpp): compares scheduling designs on the real CWBVH
without a GPU.   python tools/trav_sim.py c2 c5 [--rays 400000]
Ray mix: primary rays (8x4-pixel blocks per warp, like k_generate hands them out) + first- and second-bounce cosine rays from the
oracle-free hit points (the simulator's own closest hits are not exported, so bounce origins come from tests/rays.py + the oracle)."""
import argparse
import ctypes as C
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in ("simple-path-tracer_b200", "oracle", "tests"):
    sys.path.insert(0, os.path.join(ROOT, p))
LIB = os.path.join(ROOT, "tools", "_build", "libtravsim.so")


class Cfg(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("any", "n_warps", "chunk", "min_idle", "tri_per_round", "tri_dist", "node_twice", "inst_dist", "slots", "max_pass", "pop_end", "order", "gate_leaf", "gate_wait", "gate")] + \
               [(k, C.c_float) for k in ("c_round", "c_pop", "c_node", "c_inst", "c_tri", "c_tri_dist", "c_tri_setup", "c_slot")]


class Out(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("rounds_per_ray", "lanes_node", "lanes_inst", "lanes_tri", "lanes_pop", "busy_frac", "issue_per_ray",
                                          "lane_util", "nodes_per_ray", "tris_per_ray", "insts_per_ray", "hit_rate", "tri_passes_per_round")]


def build():
    src = [os.path.join(ROOT, "tools", "trav_sim.cpp"), os.path.join(ROOT, "simple-path-tracer_b200", "csrc", "flatten.cpp"),
           os.path.join(ROOT, "simple-path-tracer_b200", "csrc", "bvh_build.cpp")]
    if not os.path.exists(LIB) or any(os.path.getmtime(s) > os.path.getmtime(LIB) for s in src):
        os.makedirs(os.path.dirname(LIB), exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-shared", "-fPIC", "-pthread", "-ffp-contract=off", "-I/usr/local/cuda/include"] + src + ["-o", LIB])
    L = C.CDLL(LIB)
    L.sim_scene_create.restype = C.c_void_p
    L.sim_scene_create.argtypes = [C.c_void_p]
    L.sim_scene_destroy.argtypes = [C.c_void_p]
    L.sim_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.POINTER(Cfg), C.POINTER(Out)]
    L.sim_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p]
    return L


def ray_mix(fs, n, seed=1):
    """extension rays of a steady-state pool: primary : bounce-1 : bounce-2 about 1 : 1 : 0.6, interleaved in blocks of 32 primaries"""
    import orc
    import rays as R
    osc = orc.OracleScene(fs, fix_q1=True)
    side = int(np.sqrt(n / 2.6 * 16 / 9)) // 8 * 8
    w, h = side, max(4, int(side * 9 / 16) // 4 * 4)
    p = R.primary_rays(fs, w, h)
    # 8x4 block order
    idx = np.arange(w * h).reshape(h // 4, 4, w // 8, 8).transpose(0, 2, 1, 3).reshape(-1)
    p = p[idx]
    hp, _ = osc.trace(p)
    d1 = R.bounce_rays(p, hp, seed=seed)
    h1, _ = osc.trace(d1)
    d2 = R.bounce_rays(d1, h1, seed=seed + 1)
    g = np.random.default_rng(seed)
    d2 = d2[g.random(len(d2)) < 0.6]
    # pool slots: a slot holds a primary ray or a bounce ray depending on when its last path ended -> random interleave of runs of 32 primaries
    blocks = [p[i:i + 32] for i in range(0, len(p), 32)]
    rest = np.concatenate([d1, d2])
    g.shuffle(rest)
    blocks += [rest[i:i + 32] for i in range(0, len(rest), 32)]
    order = g.permutation(len(blocks))
    allr = np.concatenate([blocks[i] for i in order])
    # within a warp's 32 slots the pool mixes both kinds: shuffle locally in windows of 256
    out = allr.copy()
    for s in range(0, len(out) - 255, 256):
        out[s:s + 256] = out[s:s + 256][g.permutation(256)]
    return np.ascontiguousarray(out[:n])


def shadow_mix(fs, n, seed=3):
    import orc
    import rays as R
    osc = orc.OracleScene(fs, fix_q1=True)
    side = int(np.sqrt(n * 16 / 9)) // 8 * 8
    p = R.primary_rays(fs, side, max(4, int(side * 9 / 16)))
    hp, _ = osc.trace(p)
    d1 = R.bounce_rays(p, hp, seed=seed)
    lo, hi = R.scene_bounds(fs)
    d1["t_max"] = (np.random.default_rng(seed).random(len(d1)) * 0.6 * np.linalg.norm(hi - lo)).astype(np.float32)
    return np.ascontiguousarray(d1[:n])


BASE = dict(any=0, n_warps=48, chunk=64, min_idle=6, tri_per_round=2, tri_dist=0, node_twice=0, inst_dist=0, slots=32, max_pass=1, pop_end=0, order=0, gate_leaf=1, gate_wait=1, gate=1,
            c_round=45.0, c_pop=25.0, c_node=250.0, c_inst=110.0, c_tri=75.0, c_tri_dist=110.0, c_tri_setup=50.0, c_slot=30.0)
VARIANTS = {  # the design space explored in round 2 (profiles/README.md has the GPU measurements of the ones that were built)
    "round 1 (pop first)": {},
    "now (pop last)": dict(pop_end=1),
    "now, refill at 3 idle": dict(pop_end=1, min_idle=3),
    "helper triangle pass": dict(tri_dist=2, pop_end=1, tri_per_round=1, c_tri_dist=100.0),      # built: -10 % on the GPU
    "all pending triangles dealt to 32 lanes": dict(tri_dist=1, pop_end=1),
    "eager node step": dict(pop_end=1, inst_dist=1),                                              # identical: leaves drain within the round
    "leaf stages gated at 8 lanes": dict(pop_end=1, gate_leaf=8, gate_wait=1),
    "64 ray slots per warp, stages gated at 16": dict(slots=64, min_idle=8, pop_end=1, max_pass=2, gate=16),  # not built (state through shared memory)
}

if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("scenes", nargs="+")
    ap.add_argument("--rays", type=int, default=300000)
    ap.add_argument("--extend-only", action="store_true")
    args = ap.parse_args()
    from spt import scenes
    L = build()
    for name in args.scenes:
        fs = scenes.load(name)
        desc = fs.to_desc()
        h = L.sim_scene_create(C.byref(desc))
        for kind, rays in (("extend", ray_mix(fs, args.rays)),) + ((("connect", shadow_mix(fs, args.rays)),) if not args.extend_only else ()):
            for vn, kv in VARIANTS.items():
                cfg = Cfg(**dict(BASE, any=1 if kind == "connect" else 0, **kv))
                o = Out()
                L.sim_run(h, rays.ctypes.data, len(rays), C.byref(cfg), C.byref(o))
                print(f"{name} {kind:7s} {vn:16s} issue/ray {o.issue_per_ray:7.1f}  util {o.lane_util:5.2f}/32  rounds/ray {o.rounds_per_ray:5.2f} busy {o.busy_frac:.2f} "
                      f"lanes node {o.lanes_node:4.1f} inst {o.lanes_inst:4.1f} tri {o.lanes_tri:4.1f}  per ray: nodes {o.nodes_per_ray:.2f} tris {o.tris_per_ray:.2f} inst {o.insts_per_ray:.2f} "
                      f"hit {o.hit_rate:.2f} tri passes/round {o.tri_passes_per_round:.2f}", flush=True)
        L.sim_scene_destroy(h)
