"""Host logic: the scene front-end mirrors the reference loader's semantics (src/loader/json.rs, src/core/loader.rs)."""
import json
import math
import os

import numpy as np
import pytest

from spt import abi, loader, meshio, scenes, renderer


def test_c1_is_test_scene_00_value_for_value():
    fs = scenes.load("c1")
    assert len(fs.instances) == 1 and fs.instances[0]["prim_kind"] == abi.PRIM_SPHERE
    assert fs.instances[0]["xform"] == [1, 0, 0, 0, 1, 0, 0, 0, 1, 0.5, 0, 0]
    assert fs.lights[0]["tag"] == abi.LIGHT["directional"] and fs.lights[0]["strength"] == [5.0, 5.0, 5.0]
    assert fs.aggregate == abi.AGG["group"] and fs.light_sampler == abi.LS["uniform"]
    assert abs(fs.cameras[0]["fov"] - math.radians(45.0)) < 1e-6
    assert fs.textures[0]["value"] == pytest.approx([0.5, 0.7, 1.0])


def test_q23_integer_literals_are_not_floats():
    d = scenes.c1b()
    d["surfaces"][0]["emissive"] = [45, 45, 45]       # ints: get_float3_or silently falls back to black
    fs = loader.load_scene(d, scenes.BASE_PATH)
    assert fs.surfaces[fs.names["surface"]["light_white"]]["emissive"] == [0.0, 0.0, 0.0]
    assert not any(l["tag"] == abi.LIGHT["shape"] for l in fs.lights)
    d = scenes.c1()
    d["primitives"][0]["radius"] = 1                   # required float given as int: hard error
    with pytest.raises(loader.SceneError, match="should be float"):
        loader.load_scene(d, scenes.BASE_PATH)


def test_q13_medium_sigma_s_read_from_sigma_a():
    fs = scenes.load("c4")
    m = fs.mediums[0]
    assert m["sigma_s"] == m["sigma_a"] == pytest.approx([0.9, 1.1, 1.6])
    assert any("sigma_s" in w for w in fs.warnings)    # the real key is reported as unused, like the reference


def test_q8_env_light_index_and_light_order():
    d = scenes.c3()
    d["lights"] = [{"type": "point", "name": "zz", "position": [0.0, 3.0, 0.0], "strength": [1.0, 1.0, 1.0]},
                   {"type": "point", "name": "aa", "position": [1.0, 3.0, 0.0], "strength": [1.0, 1.0, 1.0]}]
    fs = loader.load_scene(d, scenes.BASE_PATH)
    assert fs.env_light_index == 2                      # number of lights registered before the environment
    assert [l["tag"] for l in fs.lights] == [abi.LIGHT["env"], abi.LIGHT["point"], abi.LIGHT["point"]]  # "$env" < "aa" < "zz"
    assert fs.lights[1]["position"][0] == 1.0


def test_shape_lights_follow_instance_order_and_comments_are_skipped():
    d = scenes.c2()
    d["instances"][0]["#note"] = "keys starting with # are comments (core/loader.rs:396)"
    fs = loader.load_scene(d, scenes.BASE_PATH)
    shape = [l for l in fs.lights if l["tag"] == abi.LIGHT["shape"]]
    assert len(shape) == 1 and fs.instances[shape[0]["instance"]] is fs.instances[fs.names["instance"]["light"]]
    assert not any("#note" in w for w in fs.warnings)
    names = sorted(fs.names["instance"], key=fs.names["instance"].get)
    assert names == sorted(names)                       # Q9: HashMap order pinned to name order


def test_texture_wrappers_and_unknown_types():
    d = scenes.c1()
    d["textures"] = [{"type": "scalar", "name": "a", "value": [0.5, 0.5, 0.5], "is_srgb": True, "tiling": [2.0, 2.0]},
                     {"type": "scalar", "name": "b", "value": [0.1, 0.2, 0.3]},
                     {"type": "mul", "name": "tex_blue", "t1": "a", "t2": "b"}]
    fs = loader.load_scene(d, scenes.BASE_PATH)
    tags = [t["tag"] for t in fs.textures]
    assert tags == [abi.TEX["scalar"], abi.TEX["srgb"], abi.TEX["input_modifier"], abi.TEX["scalar"], abi.TEX["mul"]]
    assert fs.names["texture"]["a"] == 2 and fs.textures[4]["child"] == [2, 3]
    assert fs.textures[2]["tiling"] == [2.0, 2.0, 1.0]
    d["textures"].append({"type": "voronoi", "name": "x"})
    with pytest.raises(loader.SceneError, match="unknown type"):
        loader.load_scene(d, scenes.BASE_PATH)
    d = scenes.c1(); d["aggregate"] = "kdtree"
    with pytest.raises(loader.SceneError, match="Unknown aggregate"):
        loader.load_scene(d, scenes.BASE_PATH)
    d = scenes.c1(); del d["mediums"]
    with pytest.raises(loader.SceneError, match="no 'mediums'"):
        loader.load_scene(d, scenes.BASE_PATH)


def test_external_json_sections(tmp_path):
    d = scenes.c1()
    (tmp_path / "tex.json").write_text(json.dumps(d["textures"]))
    d["textures"] = "tex.json"
    (tmp_path / "scene.json").write_text(json.dumps(d))
    fs = loader.load_scene(str(tmp_path / "scene.json"))
    assert fs.textures[0]["value"] == pytest.approx([0.5, 0.7, 1.0])


def test_obj_single_index_and_tangents(tmp_path):
    obj = "v -1 0 -1\nv -1 0 1\nv 1 0 1\nv 1 0 -1\nvn 0 1 0\nvt 0 0\nvt 0 1\nvt 1 1\nvt 1 0\nf 1/1/1 2/2/1 3/3/1 4/4/1\n"
    p = tmp_path / "q.obj"; p.write_text(obj)
    pos, nrm, uv, idx = meshio.load_obj(str(p))
    assert pos.shape == (4, 3) and idx.tolist() == [[0, 1, 2], [0, 2, 3]]        # triangulated fan
    tan, bit = meshio.calc_tangents(pos, uv, idx)
    assert np.allclose(tan, [[1, 0, 0]] * 4) and np.allclose(bit, [[0, 0, 1]] * 4)  # d/du = +x, d/dv = +z
    # a vertex used with two different normals is duplicated (tobj single_index)
    obj2 = "v 0 0 0\nv 1 0 0\nv 0 1 0\nv 0 0 1\nvn 0 0 1\nvn 1 0 0\nf 1//1 2//1 3//1\nf 1//2 3//2 4//2\n"
    p.write_text(obj2)
    pos, nrm, uv, idx = meshio.load_obj(str(p))
    assert len(pos) == 6 and np.allclose(uv, 0)
    tan, bit = meshio.calc_tangents(pos, uv, idx)
    assert np.allclose(tan, [[1, 0, 0]] * 6) and np.allclose(bit, [[0, 1, 0]] * 6)  # no UVs: MeshVertex defaults


def test_bunny_asset_matches_reference_counts():
    pos, nrm, uv, idx = meshio.load_mesh(os.path.join(scenes.ASSETS, "bunny_5k.npz"))
    assert pos.shape == (2502, 3) and idx.shape == (5000, 3) and int(idx.max()) == 2501
    assert np.allclose(np.linalg.norm(nrm, axis=1), 1.0, atol=1e-3)


def test_renderer_json_and_interface():
    r = loader.load_renderer(scenes.PT_JSON)
    assert r == dict(type="pt", max_depth=8, sampler=dict(kind=abi.SAMPLER["recurrence"], spp=256, div_x=1, div_y=1), filter_radius=0.5)
    with pytest.raises(loader.SceneError, match="shoule be integer"):
        loader.load_renderer(dict(scenes.PT_JSON, max_depth=8.0))
    rr = renderer.load_renderer(scenes.PT_JSON)
    assert isinstance(rr, renderer.WavefrontPathTracer) and rr.pixel_sampler.total_spp() == 256
    with pytest.raises(loader.SceneError, match="unknown type"):
        renderer.create_renderer("bdpt", 8, renderer.PixelSampler(), renderer.BoxFilter())
    with pytest.raises(abi.SptError):
        renderer.create_renderer("pt", 8, renderer.PixelSampler(), renderer.BoxFilter())
    j = renderer.PixelSampler("jittered", division_x=4, division_y=3)
    assert j.total_spp() == 12


def test_all_configs_flatten():
    for name in ("c1", "c1b", "c2", "c3", "c4"):
        fs = scenes.load(name)
        d = fs.to_desc()
        assert d.n_instances == len(fs.instances) and d.n_lights == len(fs.lights) >= 1
    fs = scenes.load("c5", nx=5, nz=4)
    assert len(fs.instances) == 22 and len(fs.meshes) == 3
