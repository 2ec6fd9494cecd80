"""The C++ host side (include/rtr.hpp + examples/rtr_cli.cpp): the crate's scene API and its callers in a compiled
language over the C ABI.  CPU tests: it builds, lowers scenes to the same rtr_scene_desc bytes as the Python mirror,
and "panics" (exit 101) where the crate would.  GPU tests: its renders are byte-identical to the Python path's."""
import ctypes as C
import json
import os
import shutil
import subprocess

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
import _scenes
from _pyload import GOLDEN, mesh_fixture

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "raytracingrust_b200", "rtr_cli")


@pytest.fixture(scope="module")
def cli():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "raytracingrust_b200", "csrc")])
    assert os.path.exists(CLI)
    return CLI


def run(cli, *args, ok=True):
    p = subprocess.run([cli, *map(str, args)], capture_output=True, text=True, timeout=600)
    if ok:
        assert p.returncode == 0, p.stderr
        return json.loads(p.stdout.strip().splitlines()[-1])
    return p


def fnv1a(data: bytes) -> str:
    h = 1469598103934665603
    for chunk in np.frombuffer(data, np.uint8).tolist():
        h = ((h ^ chunk) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"


def lowered_fnv(scene):
    low = scene.lower() if hasattr(scene, "lower") else scene
    d = low.desc

    def raw(ptr, n, ctype):
        return C.string_at(ptr, n * C.sizeof(ctype)) if n else b""
    return {"objects": d.n_objects, "spheres": d.n_spheres, "meshes": d.n_meshes, "volumes": d.n_volumes,
            "materials": d.n_materials, "fnv_objects": fnv1a(raw(d.objects, d.n_objects, A.ObjectRef)),
            "fnv_spheres": fnv1a(raw(d.spheres, d.n_spheres, A.SphereS)),
            "fnv_materials": fnv1a(raw(d.materials, d.n_materials, A.MaterialS))}


@pytest.fixture(scope="module")
def assets(tmp_path_factory):
    """scenes/ and assets/ laid out as the crate's benches expect them (paths relative to the crate root)."""
    root = tmp_path_factory.mktemp("crate")
    shutil.copytree(os.path.join(GOLDEN, "scenes"), root / "scenes")
    (root / "assets").mkdir()
    pos, nrm, idx = mesh_fixture("monkey")
    with open(root / "assets" / "monkey.obj", "w") as f:
        for p in pos:
            f.write("v %r %r %r\n" % tuple(float(x) for x in p))
        for n in nrm:
            f.write("vn %r %r %r\n" % tuple(float(x) for x in n))
        for t in idx.reshape(-1, 3) + 1:
            f.write("f %d//%d %d//%d %d//%d\n" % (t[0], t[0], t[1], t[1], t[2], t[2]))
    return str(root)


@pytest.mark.parametrize("name", ["simple", "benchmark"])
def test_from_file_lowers_like_the_python_mirror(cli, name):
    path = os.path.join(GOLDEN, "scenes", name + ".json")
    got = run(cli, "lower", path)
    want = lowered_fnv(getattr(_scenes, name)())
    for k, v in want.items():
        assert got[k] == v, k
    s = getattr(_scenes, name)()
    assert got["spp"] == s.settings.samples_per_pixel and got["depth"] == s.settings.max_ray_depth


def test_bench_scenes_lower_like_the_python_recipes(cli, assets):
    got = run(cli, "lower", "bvh_stress")  # benches/bvh_tree_stress.rs:11-26
    want = lowered_fnv(_scenes.grid_spheres(10))
    for k, v in want.items():
        assert got[k] == v, k
    assert got["depth"] == 50
    got = run(cli, "lower", "mesh", "--assets", assets)  # benches/mesh.rs:10-19
    assert got["meshes"] == 1 and got["triangles"] == len(mesh_fixture("monkey")[2]) // 3
    assert got["spp"] == 10 and got["depth"] == 50
    assert got["fnv_materials"] == lowered_fnv(_scenes.mesh_bench())["fnv_materials"]


def test_missing_files_panic_like_the_crate(cli, tmp_path):
    p = run(cli, "lower", str(tmp_path / "nope.json"), ok=False)  # Scene::from_file unwrap (scene.rs:75)
    assert p.returncode == 101 and "panic" in p.stderr
    p = run(cli, "lower", "mesh", "--assets", str(tmp_path), ok=False)  # Mesh::from_file expect (mesh.rs:65)
    assert p.returncode == 101 and "panic" in p.stderr
    p = run(cli, "frobnicate", ok=False)
    assert p.returncode == 2


# ------------------------------------------------------------------------------------------------ GPU
def _python_render(scene, W, H, seed):
    ctx = rt.Context(0)
    rgba, _, st = ctx.upload(scene).render(rt.make_params(scene, W, H, seed=seed))
    return rgba, st


@pytest.mark.gpu
def test_cli_render_is_byte_identical_to_the_python_path(cli, tmp_path):
    png, exr = tmp_path / "simple.png", tmp_path / "simple.exr"
    got = run(cli, "render", os.path.join(GOLDEN, "scenes", "simple.json"), 100, 100, png, "--seed", 3, "--exr", exr)
    rgba, st = _python_render(_scenes.simple(), 100, 100, 3)
    assert got["fnv1a"] == fnv1a(rgba.tobytes())
    assert got["rays"] == st["rays"] and got["samples"] == 100 * 100 * 5
    cv2 = pytest.importorskip("cv2")
    back = cv2.imread(str(png), cv2.IMREAD_UNCHANGED)
    assert np.array_equal(back[:, :, [2, 1, 0, 3]], rgba)
    lin = rt.load_image(str(exr))  # linear radiance; the RGBA8 image is sqrt(lin) * 255 (render.rs:103-110)
    assert np.abs(np.minimum(np.sqrt(lin) * 255.0, 255.0).astype(np.uint8).astype(int) - rgba[:, :, :3].astype(int)).max() <= 1


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["simple", "mesh", "bvh_stress"])
def test_cli_benches_match_the_python_recipes(cli, assets, which):
    got = run(cli, "bench", which, "--assets", assets, "--iters", 2)
    scene, W, H = {"simple": (_scenes.simple(), 100, 100), "mesh": (_scenes.mesh_bench(), 40, 40),
                   "bvh_stress": (_scenes.grid_spheres(10), 100, 100)}[which]
    rgba, st = _python_render(scene, W, H, 0)
    assert got["fnv1a"] == fnv1a(rgba.tobytes())
    assert got["rays"] == st["rays"]
    assert got["render_call_ms"] > 0 and got["rerender_ms"] > 0
