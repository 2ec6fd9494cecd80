"""The library's C++ loaders (Scene::from_file scene.rs:74-76, Mesh::from_file mesh.rs:63-76) against
the independent pure-Python readers in _pyload.py; no GPU needed (loaders are host code)."""
import ctypes as C
import glob
import json
import os

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
from _pyload import GOLDEN, mesh_fixture, parse_obj, scene_from_json

REF_ASSETS = "/root/reference/assets"


def desc_arrays(lowered):
    d = lowered.desc
    sph = [(d.spheres[i].center.tuple(), d.spheres[i].radius, d.spheres[i].material) for i in range(d.n_spheres)]
    obj = [(d.objects[i].kind, d.objects[i].index) for i in range(d.n_objects)]
    mat = [(d.materials[i].kind, d.materials[i].color.tuple(), d.materials[i].param, d.materials[i].first,
            d.materials[i].second) for i in range(d.n_materials)]
    bg = (d.background.kind, d.background.color.tuple(), d.background.bottom.tuple(), d.background.width, d.background.height)
    vol = [(d.volumes[i].boundary.kind, d.volumes[i].boundary.index, d.volumes[i].neg_inv_density) for i in range(d.n_volumes)]
    return sph, obj, mat, bg, vol


@pytest.mark.parametrize("name", ["simple.json", "benchmark.json"])
def test_json_loader_matches_python_reader(name):
    path = os.path.join(GOLDEN, "scenes", name)
    a, b = rt.Scene.from_file(path), scene_from_json(path)
    assert desc_arrays(a.lower()) == desc_arrays(b.lower())
    assert a.settings == b.settings and a.settings.clamp_indirect == 10.0  # defaulted (scene.rs:55)
    assert a.camera.lookfrom.tuple() == b.camera.lookfrom.tuple() and a.camera.aspect_ratio == b.camera.aspect_ratio
    p = rt.make_params(a, 24, 16, seed=4)
    assert np.array_equal(orc.OracleScene(a).render(p)[1], orc.OracleScene(b).render(p)[1])


def test_json_all_variants_roundtrip(tmp_path):
    doc = {
        "camera": {"lookfrom": {"x": 1, "y": 2, "z": 3}, "lookat": {"x": 0, "y": 0, "z": -1},
                   "vertical": {"x": 0, "y": 1, "z": 0}, "vertical_fov": 40.5, "aspect_ratio": 1.5},
        "settings": {"samples_per_pixel": 7, "max_ray_depth": 9, "clamp_indirect": 3.5, "enable_multithreading": False,
                     "enable_bvh_tree": True, "mode": "Clay"},
        "background": {"type": "SkyMap", "image": [{"r": i, "g": 2 * i, "b": 0.5} for i in range(6)], "width": 3, "height": 2},
        "objects": [{"type": "Sphere", "center": {"x": 0, "y": 0, "z": -1}, "radius": 0.5, "material": 1},
                    {"type": "Volume", "boundary": {"type": "Sphere", "center": {"x": 0, "y": 0, "z": 0}, "radius": 9.0, "material": 2},
                     "neg_inv_density": -10.0}],
        "materials": [{"type": "Metal", "albedo": {"r": 1, "g": 1, "b": 1}, "fuzz": 0.1},
                      {"type": "MixMaterial", "factor": 0.3,
                       "first": {"type": "Dielectric", "ir": 1.5},
                       "second": {"type": "MixMaterial", "factor": 0.6, "first": {"type": "Emission", "color": {"r": 2, "g": 2, "b": 2}},
                                  "second": {"type": "Lambertian", "albedo": {"r": 0.1, "g": 0.2, "b": 0.3}}}},
                      {"type": "Isotropic", "color": {"r": 0.9, "g": 0.9, "b": 0.9}}],
    }
    path = tmp_path / "all.json"
    path.write_text(json.dumps(doc))
    a, b = rt.Scene.from_file(str(path)), scene_from_json(str(path))
    assert desc_arrays(a.lower()) == desc_arrays(b.lower())
    assert a.settings == b.settings and a.settings.mode == rt.RenderMode.Clay and a.settings.clamp_indirect == 3.5
    d = a.lower().desc
    assert d.n_materials == 7 and d.materials[1].kind == A.MAT_MIX  # handles 0..2 keep their indices
    assert np.allclose(np.ctypeslib.as_array(d.background.texels, (18,))[:6], [0, 0, 0.5, 1, 2, 0.5])


@pytest.mark.parametrize("doc,status", [
    ("{ not json", A.RTR_INVALID_ARG),
    ('{"camera": {}}', A.RTR_INVALID_ARG),                         # serde: missing field
])
def test_json_errors(tmp_path, doc, status):
    p = tmp_path / "bad.json"
    p.write_text(doc)
    with pytest.raises(rt.RtrError) as e:
        rt.Scene.from_file(str(p))
    assert e.value.status == status


def test_json_mesh_and_missing_file(tmp_path):
    with pytest.raises(rt.RtrError) as e:
        rt.Scene.from_file(str(tmp_path / "nope.json"))  # .unwrap() on read_to_string (scene.rs:75)
    assert e.value.status == A.RTR_IO_ERROR
    doc = json.load(open(os.path.join(GOLDEN, "scenes", "benchmark.json")))
    doc["objects"].append({"type": "Mesh"})
    p = tmp_path / "mesh.json"
    p.write_text(json.dumps(doc))
    with pytest.raises(rt.RtrError) as e:
        rt.Scene.from_file(str(p))  # Mesh::deserialize panics (mesh.rs:51-60)
    assert "Mesh" in str(e.value)


CUBE_OBJ = """# unit cube, two styles of face corner
o cube
v -1 -1 -1
v 1 -1 -1
v 1 1 -1
v -1 1 -1
v -1 -1 1
v 1 -1 1
v 1 1 1
v -1 1 1
vt 0 0
vn 0 0 -1
vn 0 0 1
vn 0 -1 0
s off
f 1//1 3//1 2//1
f 1/1/1 4/1/1 3/1/1
f 5//2 6//2 7//2
f -4//2 -2//2 -1//2
f 1//3 2//3 6//3
"""


def test_obj_loader_matches_python_reader(tmp_path):
    p = tmp_path / "cube.obj"
    p.write_text(CUBE_OBJ)
    pos, nrm, idx = rt.load_obj(str(p))
    ppos, pnrm, pidx = parse_obj(str(p))
    assert np.array_equal(pos, ppos) and np.array_equal(nrm, pnrm) and np.array_equal(idx, pidx)
    assert len(idx) == 15 and len(pos) == 11  # (position, normal) pairs are de-duplicated in first-use order
    assert idx[0] == idx[3] and idx[1] == idx[5]


def test_obj_loader_errors(tmp_path):
    quad = tmp_path / "quad.obj"
    quad.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nv 0 1 0\nvn 0 0 1\nf 1//1 2//1 3//1 4//1\n")
    with pytest.raises(rt.RtrError) as e:
        rt.load_obj(str(quad))
    assert "triangulated" in str(e.value)
    non = tmp_path / "nonormal.obj"
    non.write_text("v 0 0 0\nv 1 0 0\nv 1 1 0\nf 1 2 3\n")
    with pytest.raises(rt.RtrError) as e:
        rt.load_obj(str(non))
    assert "normal" in str(e.value)
    with pytest.raises(rt.RtrError) as e:
        rt.load_obj(str(tmp_path / "missing.obj"))
    assert e.value.status == A.RTR_IO_ERROR


@pytest.mark.skipif(not os.path.isdir(REF_ASSETS), reason="reference checkout not present (GPU box)")
def test_obj_loader_on_reference_assets_matches_fixtures():
    for path in sorted(glob.glob(os.path.join(REF_ASSETS, "*.obj"))):
        name = os.path.splitext(os.path.basename(path))[0]
        if not os.path.exists(os.path.join(GOLDEN, "meshes", name + ".npz")):
            continue
        pos, nrm, idx = rt.load_obj(path)
        fp, fn, fi = mesh_fixture(name)
        assert np.array_equal(pos, fp) and np.array_equal(nrm, fn) and np.array_equal(idx, fi), name


def test_write_png_round_trip(tmp_path):
    """Headless output (replaces the window blit of RenderTarget.data): PNG written by the host library decodes back."""
    import cv2
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (37, 53, 4), dtype=np.uint8)
    path = str(tmp_path / "out.png")
    rt.write_png(path, img)
    back = cv2.imread(path, cv2.IMREAD_UNCHANGED)  # BGRA
    assert back is not None and back.shape == img.shape
    assert np.array_equal(back[..., [2, 1, 0, 3]], img)
    target = rt.RenderTarget(8, 4)
    target.data[:] = 200
    target.save_png(path)
    assert np.all(cv2.imread(path, cv2.IMREAD_UNCHANGED) == 200)
    with pytest.raises(rt.RtrError):
        rt.write_png(str(tmp_path / "no_such_dir" / "x.png"), img)
