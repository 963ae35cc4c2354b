"""The compiled C++ host (simple-path-tracer_b200/host/) against the Python mirror of the reference loader: both read the
same scene JSON and must flatten it into the same SptSceneDesc (ints exact, floats to 1e-6: sin/cos of the instance
rotations come from different libm's), reject the same malformed inputs, and — on the GPU — produce the same image."""
import ctypes
import json
import os
import struct
import subprocess

import numpy as np
import pytest

from spt import abi, loader, scenes

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def host_bin(built):
    import __graft_entry__ as g
    return g.build_host()


def read_dump(path):
    out, buf, p = {}, open(path, "rb").read(), 0
    while p < len(buf):
        nl, = struct.unpack_from("<I", buf, p); p += 4
        name = buf[p:p + nl].decode(); p += nl
        nb, = struct.unpack_from("<Q", buf, p); p += 8
        out[name] = buf[p:p + nb]; p += nb
    return out


def struct_bytes(arr, n, size=None):
    size = size or ctypes.sizeof(arr._type_)
    return b"".join(ctypes.string_at(ctypes.addressof(arr[i]), size) for i in range(n))


def python_dump(fs):
    d = fs.to_desc()
    out = {}
    for i, m in enumerate(fs.meshes):
        for k in ("positions", "normals", "texcoords", "tangents", "bitangents"):
            out[f"mesh{i}.{k}"] = np.ascontiguousarray(m[k], np.float32).tobytes()
        out[f"mesh{i}.indices"] = np.ascontiguousarray(m["indices"], np.uint32).tobytes()
    out["spheres"] = struct_bytes(d.spheres, d.n_spheres)
    out["instances"] = struct_bytes(d.instances, d.n_instances)
    out["surfaces"] = struct_bytes(d.surfaces, d.n_surfaces)
    out["materials"] = struct_bytes(d.materials, d.n_materials)
    out["textures"] = struct_bytes(d.textures, d.n_textures, abi.TextureDesc.rgba.offset)
    for i, t in enumerate(fs.textures):
        if t["rgba"] is not None:
            out[f"texture{i}.rgba"] = np.ascontiguousarray(t["rgba"], np.uint8).tobytes()
    out["mediums"] = struct_bytes(d.mediums, d.n_mediums)
    out["lights"] = struct_bytes(d.lights, d.n_lights)
    out["cameras"] = struct_bytes(d.cameras, d.n_cameras)
    if fs.env is not None:
        out["env.size"] = struct.pack("<ii", fs.env["width"], fs.env["height"])
        out["env.scale"] = np.asarray(fs.env["scale"], np.float32).tobytes()
        out["env.rgb"] = np.ascontiguousarray(fs.env["rgb"], np.float32).tobytes()
    out["misc"] = struct.pack("<iii", fs.env_light_index, fs.light_sampler, fs.aggregate)
    out["warnings"] = "".join(w + "\n" for w in fs.warnings).encode()
    return out


def words_match(a, b, tol=2e-6):
    """every 4-byte word equal as an integer, or close as a float"""
    assert len(a) == len(b)
    ia, ib = np.frombuffer(a, np.int32), np.frombuffer(b, np.int32)
    fa, fb = np.frombuffer(a, np.float32), np.frombuffer(b, np.float32)
    with np.errstate(invalid="ignore", over="ignore"):
        close = np.abs(fa - fb) <= tol * np.maximum(1.0, np.abs(fa))
    return (ia == ib) | close


def run_host(host_bin, args, check=True):
    r = subprocess.run([host_bin] + args, capture_output=True, text=True)
    if check:
        assert r.returncode == 0, r.stderr
    return r


def dump_via_host(host_bin, tmp_path, scene_dict, name="scene"):
    sj = tmp_path / f"{name}.json"
    sj.write_text(json.dumps(scene_dict))
    out = tmp_path / f"{name}.dump"
    run_host(host_bin, ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--dump-desc", str(out)])
    return read_dump(out)


@pytest.mark.parametrize("name", ["c1", "c1b", "c2", "c3", "c4", "c5"])
def test_cpp_loader_flattens_like_the_python_loader(host_bin, tmp_path, name):
    sd = scenes.CONFIGS[name]["scene"]()
    got = dump_via_host(host_bin, tmp_path, sd)
    want = python_dump(loader.load_scene(sd, scenes.BASE_PATH))
    assert sorted(got) == sorted(want)
    for k in want:
        if k.endswith(".indices") or k.endswith(".rgba") or k in ("warnings", "misc", "env.size"):
            assert got[k] == want[k], k
        elif k == "env.rgb":  # float64 sin/cos/acos from two libm's: a texel on the sun disc's edge may flip
            ok = words_match(got[k], want[k], 1e-5)
            assert ok.mean() > 0.9999, k
        else:
            ok = words_match(got[k], want[k])
            assert ok.all(), (k, int((~ok).sum()))


TEXTURED = {
    "cameras": {"type": "perspective", "name": "cam_a", "eye": [0.0, 0.0, 5.0], "forward": [0.0, 0.0, -1.0], "up": [0.0, 1.0, 0.0], "fov": 45.0},
    "textures": [
        {"type": "scalar", "name": "a", "value": [0.5, 0.25, 1.0]},
        {"type": "scalar", "name": "b", "value": [0.2, 0.2, 0.2], "is_srgb": True},
        {"type": "image_array", "name": "img", "rgba": [[[255, 0, 0, 255], [0, 255, 0, 255]], [[0, 0, 255, 255], [255, 255, 255, 128]]],
         "tiling": [2.0, 3.0], "wrap": "mirror_repeat", "#comment": "ignored"},
        {"type": "mul", "name": "ab", "t1": "a", "t2": "b"},
        {"type": "div", "name": "pos", "t1": "ab", "t2": "a", "mode": "position", "offset": [0.5, 0.5, 0.25], "bogus": 1.0},
    ],
    "materials": [{"type": "lambert", "name": "m0", "albedo": "ab"},
                  {"type": "plastic", "name": "m1", "int_ior": 1.5, "ext_ior": 1.1, "albedo": "img", "roughness_x": "a", "roughness_y": "b"},
                  {"type": "pbr_specular", "name": "m2", "diffuse": "pos", "specular": "a", "roughness": "b"}],
    "mediums": [{"type": "homogeneous", "name": "fog", "sigma_a": [0.1, 0.2, 0.3], "sigma_s": [9.0, 9.0, 9.0], "asymmetric": 0.3}],
    "primitives": [{"type": "sphere", "name": "s", "radius": 0.5, "center": [0.0, 1.0, 0.0]},
                   {"type": "trimesh", "name": "cube", "obj_file": "cube.npz"},
                   {"type": "cubic_bezier", "name": "skipped", "whatever": 1}],
    "surfaces": [{"name": "sf", "material": "m1", "normal_map": "img", "emissive": [1.0, 2.0, 3.0], "emissive_map": "a",
                  "double_sided": True, "inside_medium": "fog"}],
    "instances": [{"name": "zz", "primitive": "s", "material": "m0", "translate": [1.0, 2.0, 3.0]},
                  {"name": "aa", "primitive": "cube", "surface": "sf", "scale": [1.0, 2.0, 0.5], "rotate": [10.0, 20.0, 30.0],
                   "matrix": [1.0, 0.0, 0.0, 0.0, 0.0, 1.0, 0.5, 0.0, 0.0, 0.0, 1.0, 0.0, 0.25, 0.5, 0.75, 1.0]},
                  {"name": "mm", "primitive": "cube", "material": "m2"}],
    "lights": [{"type": "spot", "name": "sp", "position": [0.0, 4.0, 0.0], "direction": [0.0, -1.0, 0.0], "inner_angle": 20.0,
                "strength": [10.0, 10.0, 10.0]},
               {"type": "point", "name": "pt", "position": [1.0, 1.0, 1.0], "strength": [1.0, 1.0, 1.0]},
               {"type": "directional", "name": "dl", "direction": [0.0, -1.0, 0.5], "strength": [2.0, 2.0, 2.0]}],
    "environment": {"type": "color", "color": [0.1, 0.2, 0.3], "scale": [2.0, 2.0, 2.0]},
    "light_sampler": "power_is",
}


def test_cpp_loader_covers_every_component_kind(host_bin, tmp_path):
    """textures with srgb / input-modifier wrappers, every light kind, medium (Q13), explicit surface, matrix+SRT
    instance transforms, name-sorted instances and lights (Q9), env light index (Q8), unused-key warnings."""
    got = dump_via_host(host_bin, tmp_path, TEXTURED)
    fs = loader.load_scene(TEXTURED, scenes.BASE_PATH)
    want = python_dump(fs)
    assert sorted(got) == sorted(want)
    for k in want:
        if k.endswith(".indices") or k.endswith(".rgba") or k in ("warnings", "misc", "env.size"):
            assert got[k] == want[k], k
        else:
            assert words_match(got[k], want[k]).all(), k
    assert b"bogus" in got["warnings"] and b"skipped" in got["warnings"]
    assert fs.env_light_index == 3 and len(fs.lights) == 5  # 3 explicit + env + 1 shape light


def _broken(**patch):
    sd = json.loads(json.dumps(scenes.c1b()))
    sd.update(patch)
    return sd


@pytest.mark.parametrize("mutate, message", [
    (lambda s: s["cameras"].update(fov=45), "'fov' should be float"),  # Q23
    (lambda s: s["materials"][0].update(albedo="nope"), "There is no texture named 'nope'"),
    (lambda s: s["textures"].append(dict(s["textures"][0])), "Duplicated texture name"),
    (lambda s: s.pop("lights"), "There is no 'lights' field"),
    (lambda s: s["instances"][0].update(primitive="ghost"), "There is no primitive named 'ghost'"),
    (lambda s: s["primitives"].append({"type": "torus", "name": "t"}), "unknown type 'torus'"),
    (lambda s: s.update(aggregate="octree"), "Unknown aggregate type 'octree'"),
    (lambda s: s.update(gltf="x.gltf"), "out of scope"),
    (lambda s: s.update(environment=[{"type": "color", "color": [1.0, 1.0, 1.0]}]), "Field should not be an array"),
])
def test_cpp_loader_rejects_what_the_python_loader_rejects(host_bin, tmp_path, mutate, message):
    sd = json.loads(json.dumps(scenes.c1b()))
    mutate(sd)
    with pytest.raises(loader.SceneError, match=message.replace("(", r"\(")):
        loader.load_scene(sd, scenes.BASE_PATH)
    sj = tmp_path / "bad.json"
    sj.write_text(json.dumps(sd))
    r = run_host(host_bin, ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--dump-desc", str(tmp_path / "x.dump")], check=False)
    assert r.returncode == 1 and message in r.stderr, r.stderr


def test_cpp_host_external_files_obj_and_png(host_bin, tmp_path):
    """external JSON sections (json.rs:212-242), an .obj mesh with shared corners + a quad (tobj triangulation, two objects:
    Q22 fixed), and a PNG image texture written by PIL and decoded by the host's own reader."""
    from PIL import Image
    rng = np.random.RandomState(5)
    img = rng.randint(0, 256, (5, 7, 3), dtype=np.uint8)
    Image.fromarray(img).save(tmp_path / "tex.png")
    (tmp_path / "quad.obj").write_text(
        "v -1 0 -1\nv 1 0 -1\nv 1 0 1\nv -1 0 1\nvt 0 0\nvt 1 0\nvt 1 1\nvt 0 1\nvn 0 1 0\n"
        "o first\nf 1/1/1 2/2/1 3/3/1\n"  # 3 vertices, 3 indices
        "o second\nv 0 1 0\nf 1/1/1 2/2/1 3/3/1 4/4/1\nf -1/3/1 2/2/1 3/3/1\n")
    (tmp_path / "mats.json").write_text(json.dumps([{"type": "lambert", "name": "m", "albedo": "t"}]))
    sd = {"cameras": {"type": "perspective", "name": "c", "eye": [0.0, 1.0, 4.0], "forward": [0.0, 0.0, -1.0], "up": [0.0, 1.0, 0.0], "fov": 50.0},
          "textures": [{"type": "image", "name": "t", "image_file": "tex.png", "is_srgb": True}],
          "materials": "mats.json", "mediums": [],
          "primitives": [{"type": "trimesh", "name": "q", "obj_file": "quad.obj"}],
          "surfaces": [], "instances": [{"name": "i", "primitive": "q", "material": "m"}],
          "lights": [{"type": "point", "name": "p", "position": [0.0, 3.0, 0.0], "strength": [9.0, 9.0, 9.0]}]}
    sj = tmp_path / "scene.json"
    sj.write_text(json.dumps(sd))
    out = tmp_path / "scene.dump"
    run_host(host_bin, ["--scene", str(sj), "--dump-desc", str(out)])
    got = read_dump(out)
    want = python_dump(loader.load_scene(str(sj)))
    assert sorted(got) == sorted(want)
    for k in want:
        if k.endswith(".indices") or k.endswith(".rgba") or k in ("warnings", "misc"):
            assert got[k] == want[k], k
        else:
            assert words_match(got[k], want[k]).all(), k
    assert len(got["mesh0.indices"]) == 4 * 3 * 4  # triangle + quad (2 triangles) + triangle
    # Q22 fixed (index offset = running VERTEX count): a two-object OBJ whose index count differs from its vertex count — the
    # reference's offset would overrun here — loads, and its triangles are the triangles of the same faces in one object
    (tmp_path / "quad.obj").write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\no a\nf 1 2 3 4\no b\nf 1 2 3\n")
    two = loader.load_scene(str(sj)).meshes[0]
    (tmp_path / "quad.obj").write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nf 1 2 3 4\nf 1 2 3\n")
    one = loader.load_scene(str(sj)).meshes[0]
    assert two["indices"].shape == one["indices"].shape == (3, 3) and len(two["positions"]) == 7 and len(one["positions"]) == 4
    assert np.array_equal(two["positions"][two["indices"]], one["positions"][one["indices"]])
    (tmp_path / "quad.obj").write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\no a\nf 1 2 3 4\no b\nf 1 2 3\n")
    run_host(host_bin, ["--scene", str(sj), "--dump-desc", str(out)])
    got2 = read_dump(out)
    assert got2["mesh0.indices"] == python_dump(loader.load_scene(str(sj)))["mesh0.indices"]


def test_cpp_renderer_config_and_cli_errors(host_bin, tmp_path):
    sj = tmp_path / "s.json"
    sj.write_text(json.dumps(scenes.c1()))
    rj = tmp_path / "r.json"
    rj.write_text(json.dumps({"type": "pt", "max_depth": 8.0, "sampler": {"type": "recurrence", "spp": 4}, "filter": {"type": "box", "radius": 0.5}}))
    r = run_host(host_bin, ["--scene", str(sj), "--renderer", str(rj), "--output", str(tmp_path / "o.png")], check=False)
    assert r.returncode == 1 and "'max_depth' shoule be integer" in r.stderr
    rj.write_text(json.dumps(scenes.PT_JSON))
    r = run_host(host_bin, ["--scene", str(sj), "--renderer", str(rj), "--renderer-type", "pt", "--output", str(tmp_path / "o.png")], check=False)
    assert r.returncode == 1 and "CPU integrator" in r.stderr
    r = run_host(host_bin, ["--scene", str(tmp_path / "s.txt"), "--dump-desc", "x"], check=False)
    assert r.returncode == 1 and "File extension 'txt' is not recognized" in r.stderr
    if not os.path.exists("/dev/nvidiactl"):  # no GPU here: the product must fail loudly, not fall back
        r = run_host(host_bin, ["--scene", str(sj), "--renderer", str(rj), "--output", str(tmp_path / "o.png")], check=False)
        assert r.returncode == 1 and "spt_context_create" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["c1b", "c4"])
def test_cpp_host_render_equals_python_host_render(host_bin, tmp_path, ctx, name):
    """same scene, same seed: the C++ host (spt_host CLI) and the Python host drive the same kernels -> same image."""
    from PIL import Image
    from spt import core
    sd = scenes.CONFIGS[name]["scene"]()
    sj, rj = tmp_path / "s.json", tmp_path / "pt.json"
    sj.write_text(json.dumps(sd))
    rj.write_text(json.dumps(dict(scenes.PT_JSON, sampler={"type": "recurrence", "spp": 8})))
    png, raw = tmp_path / "o.png", tmp_path / "o.f32"
    r = run_host(host_bin, ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--renderer", str(rj), "--width", "160", "--height", "96",
                            "--output", str(png), "--raw", str(raw), "--seed", "7"])
    stats = json.loads(r.stdout.strip().splitlines()[-1])
    assert stats["paths"] == 160 * 96 * 8 and stats["kernel_launches"] > 0
    got = np.fromfile(raw, np.float32).reshape(96, 160, 3)
    sc = core.Scene(ctx, loader.load_scene(sd, scenes.BASE_PATH))
    want, rgb8, _ = sc.render(160, 96, spp=8, seed=7, want_rgb8=True)
    sc.close()
    rel = np.abs(got - want) / (np.abs(want) + 1e-3)
    assert (rel.max(-1) < 1e-3).mean() > 0.995  # rotations differ in the last ulp between the two hosts' libm
    disk = np.asarray(Image.open(png).convert("RGB"))
    assert disk.shape == (96, 160, 3)
    assert (np.abs(disk.astype(int) - rgb8.astype(int)) <= 1).mean() > 0.995


def _c1b_files(tmp_path, spp=8):
    sj, rj = tmp_path / "s.json", tmp_path / "pt.json"
    sj.write_text(json.dumps(scenes.c1b()))
    rj.write_text(json.dumps(dict(scenes.PT_JSON, sampler={"type": "recurrence", "spp": spp})))
    return sj, rj


def test_cpp_host_rank_launcher_fails_loudly_without_gpus(host_bin, tmp_path):
    if os.path.exists("/dev/nvidiactl"):
        pytest.skip("this box has a GPU")
    sj, rj = _c1b_files(tmp_path)
    r = run_host(host_bin, ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--renderer", str(rj), "--output", str(tmp_path / "o.png"),
                            "--gpus", "2"], check=False)
    assert r.returncode == 1 and r.stderr.count("spt_context_create") == 2, r.stderr  # both ranks report, nothing is rendered


@pytest.mark.gpu
def test_cpp_host_two_ranks_over_nccl_equal_one_rank(host_bin, tmp_path):
    """spt_host --gpus 2: two processes, interleaved tiles, one ncclReduce -> the very same frame as one process."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    sj, rj = _c1b_files(tmp_path)
    common = ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--renderer", str(rj), "--width", "200", "--height", "120", "--seed", "5"]
    run_host(host_bin, common + ["--output", str(tmp_path / "one.png"), "--raw", str(tmp_path / "one.f32")])
    r = run_host(host_bin, common + ["--output", str(tmp_path / "two.png"), "--raw", str(tmp_path / "two.f32"), "--gpus", "2"])
    st = json.loads(r.stdout.strip().splitlines()[-1])
    assert st["world_size"] == 2 and 0 < st["paths"] < 200 * 120 * 8
    one, two = np.fromfile(tmp_path / "one.f32", np.float32), np.fromfile(tmp_path / "two.f32", np.float32)
    assert np.array_equal(one, two)
    assert open(tmp_path / "one.png", "rb").read() == open(tmp_path / "two.png", "rb").read()


@pytest.mark.parametrize("name", ["test_scene_00.json", "test_scene_01.json"])
def test_both_hosts_load_the_reference_scene_files_verbatim(host_bin, tmp_path, name):
    """The shipped scenes that need no image files (scenes/textures/ is not part of the reference checkout) go through both
    loaders unchanged — external common_*.json sections, OBJ models and all — and flatten identically."""
    path = os.path.join("/root/reference/scenes", name)
    if not os.path.exists(path):
        pytest.skip("reference checkout not present on this machine")
    out = tmp_path / "ref.dump"
    run_host(host_bin, ["--scene", path, "--dump-desc", str(out)])
    got, want = read_dump(out), python_dump(loader.load_scene(path))
    assert sorted(got) == sorted(want)
    for k in want:
        if k.endswith(".indices") or k in ("warnings", "misc"):
            assert got[k] == want[k], k
        else:
            assert words_match(got[k], want[k]).all(), k


EXR_FILES = [f"env_{c}_{t}.exr" for c in ("none", "rle", "zips", "zip") for t in ("float", "half")]


@pytest.mark.parametrize("fname", EXR_FILES)
def test_exr_environment_through_both_hosts(host_bin, tmp_path, fname):
    """environment type "exr" (environment.rs:95-98): fixtures written by OpenCV's OpenEXR codec (tools/make_golden_exr.py)
    decode to the committed source array bit for bit in spt/exrio.py and in host/exr.cpp."""
    import shutil
    golden = os.path.join(ROOT, "tests", "golden")
    src = np.load(os.path.join(golden, "env_golden.npy"))
    shutil.copy(os.path.join(golden, fname), tmp_path / fname)
    sd = json.loads(json.dumps(scenes.c1()))
    sd["environment"] = {"type": "exr", "exr_file": fname, "scale": [0.5, 1.0, 2.0]}
    sj = tmp_path / "scene.json"
    sj.write_text(json.dumps(sd))
    fs = loader.load_scene(str(sj))
    assert fs.env["width"] == 24 and fs.env["height"] == 12 and np.array_equal(fs.env["rgb"], src)
    assert fs.env_light_index == 1 and sum(l["tag"] == abi.LIGHT["env"] for l in fs.lights) == 1  # Q8
    out = tmp_path / "scene.dump"
    run_host(host_bin, ["--scene", str(sj), "--dump-desc", str(out)])
    got = read_dump(out)
    assert got["env.rgb"] == src.tobytes() and got["env.size"] == struct.pack("<ii", 24, 12)
    assert got["env.scale"] == np.asarray([0.5, 1.0, 2.0], np.float32).tobytes()


def test_exr_unsupported_compression_is_reported(host_bin, tmp_path):
    import shutil
    shutil.copy(os.path.join(ROOT, "tests", "golden", "env_piz_half.exr"), tmp_path / "e.exr")
    sd = json.loads(json.dumps(scenes.c1()))
    sd["environment"] = {"type": "exr", "exr_file": "e.exr"}
    sj = tmp_path / "scene.json"
    sj.write_text(json.dumps(sd))
    with pytest.raises(loader.SceneError, match="compression 'piz' is not supported"):
        loader.load_scene(str(sj))
    r = run_host(host_bin, ["--scene", str(sj), "--dump-desc", str(tmp_path / "x")], check=False)
    assert r.returncode == 1 and "compression 'piz' is not supported" in r.stderr
    sd["environment"]["exr_file"] = "missing.exr"
    sj.write_text(json.dumps(sd))
    with pytest.raises(loader.SceneError, match="can't read"):
        loader.load_scene(str(sj))
    r = run_host(host_bin, ["--scene", str(sj), "--dump-desc", str(tmp_path / "x")], check=False)
    assert r.returncode == 1 and "can't read" in r.stderr


@pytest.mark.gpu
def test_cpp_host_progressive_passes_checkpoint_and_resume(host_bin, tmp_path):
    """--passes / --checkpoint / --resume of the compiled host: 3 passes, and 1 pass + a resumed run from its checkpoint, both
    give the very image of the one-call render."""
    sj, rj = _c1b_files(tmp_path, spp=9)
    common = ["--scene", str(sj), "--base-path", scenes.BASE_PATH, "--renderer", str(rj), "--width", "120", "--height", "80", "--seed", "4"]
    run_host(host_bin, common + ["--output", str(tmp_path / "one.png"), "--raw", str(tmp_path / "one.f32")])
    run_host(host_bin, common + ["--output", str(tmp_path / "p3.png"), "--raw", str(tmp_path / "p3.f32"), "--passes", "3", "--checkpoint", str(tmp_path / "ck")])
    one = np.fromfile(tmp_path / "one.f32", np.float32)
    assert np.array_equal(one, np.fromfile(tmp_path / "p3.f32", np.float32))
    hdr = np.fromfile(tmp_path / "ck", np.int32, 4)
    assert hdr.tolist() == [0x46545053, 120, 80, 9]
    # a run that stops after 4 of 9 samples (spp override), then a second process that resumes from its checkpoint
    r4 = tmp_path / "pt4.json"
    r4.write_text(json.dumps(dict(scenes.PT_JSON, sampler={"type": "random", "spp": 9})))
    common_r = [x if x != str(rj) else str(r4) for x in common]
    run_host(host_bin, common_r + ["--output", str(tmp_path / "full.png"), "--raw", str(tmp_path / "full.f32")])
    # write a 4-sample checkpoint through the library, as a killed 9-spp run would have left it
    from spt import abi, core
    ctx = core.Context(0)
    sc = core.Scene(ctx, loader.load_scene(scenes.c1b(), scenes.BASE_PATH))
    sc.render(120, 80, spp=9, sampler="random", seed=4, spp_range=(0, 4))
    film = sc.film_save(120, 80)
    sc.close(); ctx.close()
    with open(tmp_path / "ck4", "wb") as fh:
        fh.write(np.asarray([0x46545053, 120, 80, 4], np.int32).tobytes()); fh.write(film.tobytes())
    run_host(host_bin, common_r + ["--output", str(tmp_path / "res.png"), "--raw", str(tmp_path / "res.f32"), "--resume", str(tmp_path / "ck4")])
    assert np.array_equal(np.fromfile(tmp_path / "full.f32", np.float32), np.fromfile(tmp_path / "res.f32", np.float32))
