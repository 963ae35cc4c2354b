"""The benchmark / parity scenes of BASELINE.json:configs, written in the reference's own scene-JSON schema
(src/loader/json.rs) as Python dicts so they can be generated anywhere (the GPU box has no /root/reference).
All numbers are float literals on purpose (Q23: the reference rejects integer literals where it expects floats).

C1  = scenes/test_scene_00.json of the reference, restated value for value.
C1b = emissive-quad variant (the geometry of test_scene_03.json with scalar textures inlined).
C2..C5 are the self-contained scenes SURVEY.md §8d specifies.
"""
import math
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
ASSETS = os.path.join(ROOT, "assets")
BASE_PATH = os.path.join(ASSETS, "scene.json")  # relative files resolve against assets/

PT_JSON = {"type": "pt", "max_depth": 8, "sampler": {"type": "recurrence", "spp": 256},
           "filter": {"type": "box", "radius": 0.5}}  # scenes/pt.json


def _cam(eye, fwd=(0.0, 0.0, -1.0), fov=45.0):
    return {"type": "perspective", "name": "default_camera", "eye": list(eye), "forward": list(fwd),
            "up": [0.0, 1.0, 0.0], "fov": fov}


def _scalar(name, v):
    return {"type": "scalar", "name": name, "value": [float(v[0]), float(v[1]), float(v[2])]}


PRIMS = [{"type": "sphere", "name": "sphere_1", "radius": 1.0},
         {"type": "trimesh", "name": "cube", "obj_file": "cube.npz"},
         {"type": "trimesh", "name": "plane", "obj_file": "plane.npz"},
         {"type": "trimesh", "name": "bunny", "obj_file": "bunny_5k.npz"}]


def c1():
    return {
        "cameras": _cam((0.0, 0.0, 5.0)),
        "textures": [_scalar("tex_blue", (0.5, 0.7, 1.0))],
        "materials": [{"type": "lambert", "name": "mat_sphere", "albedo": "tex_blue"}],
        "mediums": [],
        "primitives": [{"type": "sphere", "name": "sphere", "radius": 1.0}],
        "surfaces": [{"name": "surf_sphere", "material": "mat_sphere"}],
        "instances": [{"name": "sphere", "primitive": "sphere", "surface": "surf_sphere", "translate": [0.5, 0.0, 0.0]}],
        "lights": [{"type": "directional", "name": "light0", "direction": [-1.0, -1.0, -1.0], "strength": [5.0, 5.0, 5.0]}],
        "aggregate": "group",
    }


def c1b():
    return {
        "cameras": _cam((0.0, 0.0, 7.0)),
        "textures": [_scalar("magenta", (0.8, 0.2, 0.8)), _scalar("blue", (0.2, 0.3, 0.8)), _scalar("black", (0.0, 0.0, 0.0))],
        "materials": [{"type": "lambert", "name": "lambert_magenta", "albedo": "magenta"},
                      {"type": "lambert", "name": "lambert_blue", "albedo": "blue"},
                      {"type": "lambert", "name": "lambert_black", "albedo": "black"}],
        "mediums": [],
        "surfaces": [{"name": "light_white", "material": "lambert_black", "emissive": [45.0, 45.0, 45.0]}],
        "primitives": PRIMS,
        "instances": [
            {"name": "sphere", "primitive": "sphere_1", "material": "lambert_magenta", "translate": [1.0, 0.0, 0.0]},
            {"name": "floor", "primitive": "plane", "material": "lambert_blue", "translate": [0.0, -1.0, 0.0], "scale": [5.0, 1.0, 5.0]},
            {"name": "wall", "primitive": "plane", "material": "lambert_blue", "scale": [5.0, 1.0, 5.0],
             "rotate": [90.0, 0.0, 0.0], "translate": [0.0, 3.0, -5.0]},
            {"name": "light", "primitive": "plane", "surface": "light_white", "scale": [0.9, 1.0, 0.9],
             "rotate": [0.0, 0.0, -90.0], "translate": [-1.5, 0.0, 0.0]}],
        "lights": [],
        "aggregate": "bvh",
    }


def _room(half=6.0, height=6.0, floor_mat="lambert_grey", wall_mat="lambert_wall"):
    """floor + back/left/right walls (planes are 2x2 in xz, normal +y)"""
    return [
        {"name": "floor", "primitive": "plane", "material": floor_mat, "translate": [0.0, -1.0, 0.0], "scale": [half, 1.0, half]},
        {"name": "wall_back", "primitive": "plane", "material": wall_mat, "scale": [half, 1.0, height],
         "rotate": [90.0, 0.0, 0.0], "translate": [0.0, height - 1.0, -half]},
        {"name": "wall_left", "primitive": "plane", "material": wall_mat, "scale": [height, 1.0, half],
         "rotate": [0.0, 0.0, -90.0], "translate": [-half, height - 1.0, 0.0]},
        {"name": "wall_right", "primitive": "plane", "material": wall_mat, "scale": [height, 1.0, half],
         "rotate": [0.0, 0.0, 90.0], "translate": [half, height - 1.0, 0.0]},
    ]


def c2():
    """bunny_5k as conductor / pbr_metallic (isotropic + anisotropic) on a Lambert floor, emissive quad, power_is."""
    tex = [_scalar("grey", (0.6, 0.6, 0.6)), _scalar("wall", (0.7, 0.55, 0.4)), _scalar("black", (0.0, 0.0, 0.0)),
           _scalar("gold_ior", (0.143, 0.375, 1.442)), _scalar("gold_k", (3.983, 2.386, 1.603)),
           _scalar("r03", (0.3, 0.3, 0.3)), _scalar("r02", (0.2, 0.2, 0.2)), _scalar("r08", (0.8, 0.8, 0.8)),
           _scalar("one", (1.0, 1.0, 1.0)), _scalar("copper", (0.95, 0.64, 0.54)), _scalar("steel", (0.6, 0.65, 0.7))]
    mats = [{"type": "lambert", "name": "lambert_grey", "albedo": "grey"},
            {"type": "lambert", "name": "lambert_wall", "albedo": "wall"},
            {"type": "lambert", "name": "lambert_black", "albedo": "black"},
            {"type": "conductor", "name": "gold_rough", "ior": "gold_ior", "ior_k": "gold_k", "roughness": "r03"},
            {"type": "pbr_metallic", "name": "copper_pbr", "base_color": "copper", "metallic": "one", "roughness": "r02"},
            {"type": "pbr_metallic", "name": "steel_aniso", "base_color": "steel", "metallic": "one",
             "roughness_x": "r02", "roughness_y": "r08"}]
    inst = _room() + [
        {"name": "bunny_gold", "primitive": "bunny", "material": "gold_rough", "translate": [-2.8, -0.04, 0.0]},
        {"name": "bunny_copper", "primitive": "bunny", "material": "copper_pbr", "rotate": [0.0, 35.0, 0.0], "translate": [0.0, -0.04, -0.6]},
        {"name": "bunny_steel", "primitive": "bunny", "material": "steel_aniso", "rotate": [0.0, -40.0, 0.0], "translate": [2.8, -0.04, 0.0]},
        {"name": "light", "primitive": "plane", "surface": "light_white", "scale": [2.0, 1.0, 1.5],
         "rotate": [180.0, 0.0, 0.0], "translate": [0.0, 4.5, 1.0]}]
    return {"cameras": _cam((0.0, 1.2, 9.0), (0.0, -0.12, -1.0), 40.0), "textures": tex, "materials": mats, "mediums": [],
            "surfaces": [{"name": "light_white", "material": "lambert_black", "emissive": [18.0, 18.0, 18.0]}],
            "primitives": PRIMS, "instances": inst, "lights": [], "aggregate": "bvh", "light_sampler": "power_is"}


def c3():
    """glass (specular + rough dielectric) bunny/spheres under a procedural HDR lat-long env, power_is."""
    tex = [_scalar("grey", (0.5, 0.5, 0.5)), _scalar("white", (1.0, 1.0, 1.0)), _scalar("zero", (0.0, 0.0, 0.0)),
           _scalar("r02", (0.2, 0.2, 0.2))]
    mats = [{"type": "lambert", "name": "lambert_grey", "albedo": "grey"},
            {"type": "dielectric", "name": "glass", "int_ior": 1.5, "reflectance": "white", "transmittance": "white", "roughness": "zero"},
            {"type": "dielectric", "name": "glass_rough", "int_ior": 1.5, "reflectance": "white", "transmittance": "white", "roughness": "r02"}]
    inst = [{"name": "floor", "primitive": "plane", "material": "lambert_grey", "translate": [0.0, -1.0, 0.0], "scale": [20.0, 1.0, 20.0]},
            {"name": "bunny_glass", "primitive": "bunny", "material": "glass", "translate": [0.0, -0.04, 0.0]},
            {"name": "sphere_glass", "primitive": "sphere_1", "material": "glass", "translate": [-2.8, 0.0, 0.3]},
            {"name": "sphere_rough", "primitive": "sphere_1", "material": "glass_rough", "translate": [2.8, 0.0, 0.3]},
            {"name": "bunny_rough", "primitive": "bunny", "material": "glass_rough", "scale": [0.6, 0.6, 0.6],
             "rotate": [0.0, 60.0, 0.0], "translate": [1.4, -0.42, 2.2]}]
    return {"cameras": _cam((0.0, 1.0, 8.5), (0.0, -0.1, -1.0), 40.0), "textures": tex, "materials": mats, "mediums": [],
            "surfaces": [], "primitives": PRIMS, "instances": inst, "lights": [],
            "environment": {"type": "procedural_sky", "width": 1024, "height": 512, "sun_gain": 1000.0},
            "aggregate": "bvh", "light_sampler": "power_is"}


def c4(g=0.0):
    """homogeneous HG medium inside a pseudo-material sphere + subsurface bunny + emissive ceiling (test_scene_07-like)."""
    tex = [_scalar("green", (0.2, 0.5, 0.25)), _scalar("blue", (0.25, 0.3, 0.6)), _scalar("black", (0.0, 0.0, 0.0)),
           _scalar("skin", (0.8, 0.55, 0.45)), _scalar("ld", (3.0, 3.0, 3.0)), _scalar("r03", (0.3, 0.3, 0.3))]
    mats = [{"type": "lambert", "name": "lambert_green_dark", "albedo": "green"},
            {"type": "lambert", "name": "lambert_blue_dark", "albedo": "blue"},
            {"type": "lambert", "name": "lambert_black", "albedo": "black"},
            {"type": "pseudo", "name": "pseudo"},
            {"type": "subsurface", "name": "sss", "int_ior": 1.3, "albedo": "skin", "ld": "ld", "roughness": "r03"}]
    inst = [
        {"name": "medium", "primitive": "sphere_1", "surface": "medium", "translate": [-1.4, 0.2, 0.0]},
        {"name": "bunny_sss", "primitive": "bunny", "material": "sss", "rotate": [0.0, -20.0, 0.0], "translate": [1.4, -0.04, 0.0]},
        {"name": "floor", "primitive": "plane", "material": "lambert_green_dark", "translate": [0.0, -1.0, 0.0], "scale": [5.0, 1.0, 5.0]},
        {"name": "wall_back", "primitive": "plane", "material": "lambert_blue_dark", "scale": [5.0, 1.0, 3.0],
         "rotate": [90.0, 0.0, 0.0], "translate": [0.0, 2.0, -5.0]},
        {"name": "wall_front", "primitive": "plane", "material": "lambert_blue_dark", "scale": [5.0, 1.0, 3.0],
         "rotate": [-90.0, 0.0, 0.0], "translate": [0.0, 2.0, 5.0]},
        {"name": "wall_left", "primitive": "plane", "material": "lambert_blue_dark", "scale": [3.0, 1.0, 5.0],
         "rotate": [0.0, 0.0, -90.0], "translate": [-5.0, 2.0, 0.0]},
        {"name": "wall_right", "primitive": "plane", "material": "lambert_blue_dark", "scale": [3.0, 1.0, 5.0],
         "rotate": [0.0, 0.0, 90.0], "translate": [5.0, 2.0, 0.0]},
        {"name": "ceil", "primitive": "plane", "surface": "light", "scale": [5.0, 1.0, 5.0],
         "rotate": [180.0, 0.0, 0.0], "translate": [0.0, 5.0, 0.0]}]
    return {"cameras": _cam((0.0, 0.0, 4.0), fov=75.0), "textures": tex, "materials": mats,
            "mediums": [{"type": "homogeneous", "name": "med_homo", "sigma_a": [0.9, 1.1, 1.6], "sigma_s": [2.29, 2.39, 1.97],
                         "asymmetric": float(g)}],
            "surfaces": [{"name": "light", "material": "lambert_black", "emissive": [2.0, 2.0, 2.0]},
                         {"name": "medium", "material": "pseudo", "inside_medium": "med_homo"}],
            "primitives": PRIMS, "instances": inst, "lights": [], "aggregate": "bvh"}


def c5(nx=50, nz=40, seed=0xC5):
    """nx*nz instances of bunny_5k (2000 -> 10.0 M instanced triangles) on a jittered grid, cycling materials."""
    rng = np.random.RandomState(seed)
    tex = [_scalar("grey", (0.55, 0.55, 0.55)), _scalar("black", (0.0, 0.0, 0.0)), _scalar("white", (1.0, 1.0, 1.0)),
           _scalar("zero", (0.0, 0.0, 0.0)), _scalar("red", (0.75, 0.25, 0.2)), _scalar("one", (1.0, 1.0, 1.0)),
           _scalar("r03", (0.3, 0.3, 0.3)), _scalar("gold_ior", (0.143, 0.375, 1.442)), _scalar("gold_k", (3.983, 2.386, 1.603)),
           _scalar("copper", (0.95, 0.64, 0.54))]
    mats = [{"type": "lambert", "name": "lambert_grey", "albedo": "grey"},
            {"type": "lambert", "name": "lambert_black", "albedo": "black"},
            {"type": "lambert", "name": "m0", "albedo": "red"},
            {"type": "pbr_metallic", "name": "m1", "base_color": "copper", "metallic": "one", "roughness": "r03"},
            {"type": "conductor", "name": "m2", "ior": "gold_ior", "ior_k": "gold_k", "roughness": "r03"},
            {"type": "dielectric", "name": "m3", "int_ior": 1.5, "reflectance": "white", "transmittance": "white", "roughness": "zero"}]
    inst = [{"name": "floor", "primitive": "plane", "material": "lambert_grey", "translate": [0.0, -1.0, 0.0],
             "scale": [float(nx) * 2.0, 1.0, float(nz) * 2.0]},
            {"name": "light", "primitive": "plane", "surface": "light_white", "scale": [float(nx), 1.0, float(nz)],
             "rotate": [180.0, 0.0, 0.0], "translate": [0.0, 30.0, 0.0]}]
    k = 0
    for iz in range(nz):
        for ix in range(nx):
            jx, jz, ry, sc = rng.uniform(-0.4, 0.4), rng.uniform(-0.4, 0.4), rng.uniform(0, 360), rng.uniform(0.7, 1.1)
            inst.append({"name": f"b{k:05d}", "primitive": "bunny", "material": f"m{k % 4}",
                         "scale": [float(sc)] * 3, "rotate": [0.0, float(ry), 0.0],
                         "translate": [float((ix - nx / 2 + 0.5) * 3.0 + jx), float(-1.0 + 0.962671 * sc), float((iz - nz / 2 + 0.5) * 3.0 + jz)]})
            k += 1
    return {"cameras": _cam((0.0, 22.0, float(nz) * 2.2), (0.0, -0.45, -1.0), 50.0), "textures": tex, "materials": mats, "mediums": [],
            "surfaces": [{"name": "light_white", "material": "lambert_black", "emissive": [6.0, 6.0, 6.0]}],
            "primitives": PRIMS, "instances": inst, "lights": [],
            "environment": {"type": "procedural_sky", "width": 512, "height": 256, "sun_gain": 200.0},
            "aggregate": "bvh", "light_sampler": "power_is"}


CONFIGS = {
    "c1": dict(scene=c1, width=512, height=512, spp=256),
    "c1b": dict(scene=c1b, width=512, height=512, spp=256),
    "c2": dict(scene=c2, width=1920, height=1080, spp=256),
    "c3": dict(scene=c3, width=1920, height=1080, spp=1024),
    "c4": dict(scene=c4, width=1280, height=720, spp=1024),
    "c5": dict(scene=c5, width=3840, height=2160, spp=4096),
}


def load(name, **kw):
    """FlatScene for a named config."""
    from . import loader
    return loader.load_scene(CONFIGS[name]["scene"](**kw), BASE_PATH)


def _abi():
    from . import abi
    return abi


def flatten_instances(fs):
    """every trimesh instance of `fs` baked into one world-space mesh behind a single identity instance"""
    pos, nrm, uv, tan, bit, idx, base = [], [], [], [], [], [], 0
    keep = []
    for inst in fs.instances:
        if inst["prim_kind"] != _abi().PRIM_TRIMESH:
            keep.append(inst)
            continue
        m = fs.meshes[inst["prim_index"]]
        x = np.asarray(inst["xform"], np.float32)
        A, t = x[:9].reshape(3, 3).T, x[9:]
        nA = np.linalg.inv(A.astype(np.float64)).T.astype(np.float32)
        pos.append(m["positions"] @ A.T + t)
        n = m["normals"] @ nA.T
        nrm.append(n / np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20))
        uv.append(m["texcoords"]); tan.append(m["tangents"] @ A.T); bit.append(m["bitangents"] @ A.T)
        idx.append(m["indices"].reshape(-1) + base)
        base += len(m["positions"])
    mesh = dict(positions=np.concatenate(pos).astype(np.float32), normals=np.concatenate(nrm).astype(np.float32),
                texcoords=np.concatenate(uv).astype(np.float32), tangents=np.concatenate(tan).astype(np.float32),
                bitangents=np.concatenate(bit).astype(np.float32), indices=np.concatenate(idx).astype(np.uint32).reshape(-1, 3), path="<flattened>")
    surf = next(i["surface"] for i in fs.instances if i["prim_kind"] == _abi().PRIM_TRIMESH)
    fs.meshes = [mesh]
    fs.instances = keep + [dict(prim_kind=_abi().PRIM_TRIMESH, prim_index=0, surface=surf,
                                xform=[1.0, 0, 0, 0, 1.0, 0, 0, 0, 1.0, 0, 0, 0])]
    fs.lights = [l for l in fs.lights if l["tag"] != _abi().LIGHT["shape"]]
    return fs
