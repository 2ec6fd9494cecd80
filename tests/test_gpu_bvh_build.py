"""Device BVH build (csrc/bvh_build.cu) against the host rebuild of BvhNode::from_list (bvh.rs:59-144)
and against the oracle: same closest hits, same tie winners, same images, same work counters on the reference walk."""
import os
from contextlib import contextmanager

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _parity import assert_hits_identical

pytestmark = pytest.mark.gpu


@contextmanager
def build_mode(mode):
    old = os.environ.get("RTR_BVH_BUILD")
    os.environ["RTR_BVH_BUILD"] = mode  # read by rtr_scene_upload
    try:
        yield
    finally:
        os.environ.pop("RTR_BVH_BUILD") if old is None else os.environ.__setitem__("RTR_BVH_BUILD", old)


def upload(ctx, scene, mode):
    with build_mode(mode):
        return ctx.upload(scene)


def two_spheres():
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    s.add_object(rt.Sphere(rt.Vec3(-0.6, 0, -2), 0.5, m))
    s.add_object(rt.Sphere(rt.Vec3(0.6, 0, -2), 0.5, m))
    return s


def duplicates():
    """Identical centroids everywhere: the stable sort must keep insertion order, and equal-t ties go to the last leaf."""
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    for k in range(37):
        s.add_object(rt.Sphere(rt.Vec3(0.0 if k % 3 else -0.0, 0, -2), 0.5, m))
    return s


SCENES = {"simple": _scenes.simple, "benchmark": _scenes.benchmark, "grid8000": lambda: _scenes.grid_spheres(10),
          "random20k": lambda: _scenes.random_spheres(20000), "two": two_spheres, "three": lambda: _scenes.grid_spheres(1),
          "duplicates": duplicates, "odd_1001": lambda: _scenes.random_spheres(1001, seed=5)}


@pytest.mark.parametrize("name", sorted(SCENES))
def test_device_build_equals_host_build_and_oracle(ctx, name):
    scene = SCENES[name]()
    rays = np.concatenate([_scenes.camera_rays_np(scene, 96, 64), _scenes.random_rays(6000, 3, 60.0 if "random" in name or "odd" in name else 3.0)])
    dev, host = upload(ctx, scene, "device"), upload(ctx, scene, "host")
    ora = orc.OracleScene(scene).trace_batch(rays, seed=4)
    for flags in (0, A.FLAG_EXHAUSTIVE, A.FLAG_REFERENCE_TREE, A.FLAG_REFERENCE_TREE | A.FLAG_EXHAUSTIVE):
        hd = dev.trace_batch(rays, seed=4, flags=flags)
        assert_hits_identical(hd, ora, f"{name} device build flags={flags}")
        assert np.array_equal(hd, host.trace_batch(rays, seed=4, flags=flags))
    assert np.array_equal(dev.trace_batch(rays, enable_bvh_tree=False), host.trace_batch(rays, enable_bvh_tree=False))
    # the walk of the reference topology does the same work on both (and what the oracle's counting walk does)
    _, sd = dev.trace_batch(rays, want_stats=True, flags=A.FLAG_REFERENCE_TREE)
    _, sh = host.trace_batch(rays, want_stats=True, flags=A.FLAG_REFERENCE_TREE)
    counts, _ = orc.OracleScene(scene).count_batch(rays)
    for k_dev, k_ora in (("nodes_visited", "nodes"), ("sphere_tests", "spheres")):
        assert abs(sd[k_dev] - sh[k_dev]) <= 0.002 * sh[k_dev] + 4, (k_dev, sd[k_dev], sh[k_dev])
        assert abs(sd[k_dev] - counts[k_ora]) <= 0.02 * counts[k_ora] + 16, (k_dev, sd[k_dev], counts[k_ora])
    # images: identical accumulation buffers
    p = rt.make_params(scene, 96, 64, seed=2)
    assert np.array_equal(dev.render(p, want_accum=True)[1], host.render(p, want_accum=True)[1])


def test_device_build_is_the_default_for_large_sphere_scenes_and_much_faster(ctx):
    scene = _scenes.random_spheres(200_000)
    auto, host = ctx.upload(scene), upload(ctx, scene, "host")
    assert auto.info()["n_nodes"] == 200_000          # n-1 Splits + the root's pseudo parent, one shared tree
    assert host.info()["n_nodes"] > 200_000           # reference tree + traversal tree
    assert auto.info()["build_ms"] < 0.5 * host.info()["build_ms"]
    rays = _scenes.random_rays(1 << 16, 9, 60.0)
    assert np.array_equal(auto.trace_batch(rays), host.trace_batch(rays))
    small = _scenes.simple()  # 102 spheres: stays on the host in auto mode
    assert ctx.upload(small).info()["n_nodes"] == upload(ctx, small, "host").info()["n_nodes"]


def test_device_build_rejects_nan_like_the_reference_panics(ctx):
    s = two_spheres()
    s.add_object(rt.Sphere(rt.Vec3(float("nan"), 0, -2), 0.5, rt.MaterialHandle(0)))
    with pytest.raises(rt.RtrError):
        upload(ctx, s, "device")


# ---- triangle meshes (build_mesh_on_device): mesh.build_bvh() of every mesh on the device, scene level on the host ----
def random_triangle_mesh(n_tris, seed, material, flat_axis=None):
    """Small triangles scattered in a box; `flat_axis` makes every triangle axis-aligned in one plane family, which produces
    zero-extent boxes (epsilon_expand, aabb.rs:57-78) and many equal centroid coordinates (stable-sort ties)."""
    rng = np.random.default_rng(seed)
    centers = rng.uniform(-2.0, 2.0, (n_tris, 1, 3)).astype(np.float32)
    if flat_axis is not None:
        centers[:, :, flat_axis] = np.round(centers[:, :, flat_axis] * 2.0) / 2.0
    pos = centers + rng.uniform(-0.15, 0.15, (n_tris, 3, 3)).astype(np.float32)
    if flat_axis is not None:
        pos[:, :, flat_axis] = centers[:, :, flat_axis]
    pos = pos.reshape(-1, 3)
    nrm = rng.normal(size=pos.shape).astype(np.float32)
    mesh = rt.Mesh.from_buffers(pos, nrm, np.arange(3 * n_tris, dtype=np.uint32), material)
    mesh.build_bvh()
    return mesh


def meshes_spheres_and_media():
    s = rt.Scene.default()
    s.settings.max_ray_depth = 6
    lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4)))
    glass = s.add_material(rt.Dielectric(1.5))
    iso = s.add_material(rt.Isotropic(rt.Color(0.9, 0.9, 0.9)))
    s.add_object(rt.Sphere(rt.Vec3(0, -102.5, -1), 100.0, lam))
    s.add_object(random_triangle_mesh(3001, 1, lam))
    s.add_object(rt.Sphere(rt.Vec3(0.3, 0.2, 1.0), 0.4, glass))
    s.add_object(random_triangle_mesh(2, 2, glass))                      # the smallest tree: one Split, two leaves
    s.add_object(rt.Volume(random_triangle_mesh(500, 3, iso), 1.5))       # a mesh-bounded medium: nested boundary walks
    s.add_object(random_triangle_mesh(777, 4, lam, flat_axis=1))          # flat boxes, tied centroids
    s.add_object(random_triangle_mesh(1, 5, lam))                         # a single triangle stays a host leaf
    return s


MESH_SCENES = {"mesh_hdri": lambda: _scenes.mesh_hdri(64, 32), "volumes": lambda: _scenes.volumes(64, 32), "monkey": _scenes.mesh_bench,
               "mixed": meshes_spheres_and_media}


@pytest.mark.parametrize("mode", ["device", "hybrid"])
@pytest.mark.parametrize("name", sorted(MESH_SCENES))
def test_device_mesh_build_equals_host_build_and_oracle(ctx, name, mode):
    """device: reference tree, records and traversal all from the device; hybrid: the host adds its SAH traversal tree over
    the gates the device built (RTR_BVH_BUILD, capi.cpp prepare_scene)."""
    scene = MESH_SCENES[name]()
    rays = np.concatenate([_scenes.camera_rays_np(scene, 128, 96), _scenes.random_rays(20000, 3, 3.0)])
    dev, host = upload(ctx, scene, mode), upload(ctx, scene, "host")
    if mode == "device":
        assert dev.info()["n_nodes"] < host.info()["n_nodes"]  # one shared tree per mesh instead of reference + traversal tree
    ora = orc.OracleScene(scene).trace_batch(rays, seed=4)
    assert (ora["object"] != A.RTR_NO_HIT).mean() > 0.05
    for flags in (0, A.FLAG_EXHAUSTIVE, A.FLAG_REFERENCE_TREE, A.FLAG_REFERENCE_TREE | A.FLAG_EXHAUSTIVE, A.FLAG_COUNT_WORK):
        hd = dev.trace_batch(rays, seed=4, flags=flags)
        assert_hits_identical(hd, ora, f"{name} device mesh build flags={flags}", 4 if name in ("volumes", "mixed") else 0)
        assert hd.tobytes() == host.trace_batch(rays, seed=4, flags=flags).tobytes(), flags
    assert dev.trace_batch(rays, enable_bvh_tree=False, seed=4).tobytes() == host.trace_batch(rays, enable_bvh_tree=False, seed=4).tobytes()
    # the reference topology is the same tree: the same work on its walk
    _, sd = dev.trace_batch(rays, want_stats=True, flags=A.FLAG_REFERENCE_TREE, seed=4)
    _, sh = host.trace_batch(rays, want_stats=True, flags=A.FLAG_REFERENCE_TREE, seed=4)
    for k in ("nodes_visited", "tri_tests", "sphere_tests"):
        assert abs(sd[k] - sh[k]) <= 0.002 * sh[k] + 4, (k, sd[k], sh[k])
    for kernel in (A.KERNEL_MEGAKERNEL, A.KERNEL_WAVEFRONT):
        p = rt.make_params(scene, 160, 120, seed=2, kernel=kernel)
        assert np.array_equal(dev.render(p, want_accum=True)[1], host.render(p, want_accum=True)[1]), kernel


def test_device_mesh_build_is_faster_to_upload_on_a_large_mesh(ctx):
    s = rt.Scene.default()
    lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4)))
    s.add_object(random_triangle_mesh(200_000, 7, lam))
    dev, hyb, host = upload(ctx, s, "device"), upload(ctx, s, "hybrid"), upload(ctx, s, "host")
    assert dev.info()["build_ms"] < 0.5 * host.info()["build_ms"], (dev.info()["build_ms"], host.info()["build_ms"])
    assert hyb.info()["build_ms"] < 0.7 * host.info()["build_ms"], (hyb.info()["build_ms"], host.info()["build_ms"])
    rays = _scenes.random_rays(1 << 16, 9, 3.0)
    assert dev.trace_batch(rays).tobytes() == host.trace_batch(rays).tobytes()
    assert hyb.trace_batch(rays).tobytes() == host.trace_batch(rays).tobytes()
    # auto: a mesh this large takes the hybrid way (capi.cpp prepare_scene), small ones stay on the host
    auto = ctx.upload(s)
    assert auto.info()["n_nodes"] == hyb.info()["n_nodes"] != host.info()["n_nodes"]
    assert auto.trace_batch(rays).tobytes() == host.trace_batch(rays).tobytes()
    small = _scenes.mesh_bench()
    assert ctx.upload(small).info()["n_nodes"] == upload(ctx, small, "host").info()["n_nodes"]


def test_device_mesh_build_rejects_nan_like_the_reference_panics(ctx):
    s = rt.Scene.default()
    lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4)))
    pos = np.zeros((9, 3), np.float32); pos[1, 0] = pos[2, 1] = pos[7, 0] = pos[8, 1] = 1.0
    pos[6:9, 0] += 5.0    # x is the widest centroid axis, so the root sorts on it
    pos[3:6, 0] = np.nan  # f32::min / max skip a NaN operand (aabb.rs:41-54): only an all-NaN coordinate reaches the sort key
    mesh = rt.Mesh.from_buffers(pos, np.ones_like(pos), np.arange(9, dtype=np.uint32), lam)
    mesh.build_bvh()
    s.add_object(mesh)
    with pytest.raises(rt.RtrError):
        upload(ctx, s, "device")
    with pytest.raises(rt.RtrError):
        upload(ctx, s, "host")
