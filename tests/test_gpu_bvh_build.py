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
