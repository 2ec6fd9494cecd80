"""Parity level 1 — closest hit for fixed ray batches: rtr_trace_batch vs the oracle's reference-shaped
(unculled, both-children) walk.  IDs exact, t/point/normal bit-identical."""
import os

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _parity import assert_hits_identical
from _pyload import mesh_fixture

pytestmark = pytest.mark.gpu

SCENES = {
    "simple": _scenes.simple, "benchmark": _scenes.benchmark, "monkey": _scenes.mesh_bench,
    "grid8000": lambda: _scenes.grid_spheres(10), "mesh_hdri": lambda: _scenes.mesh_hdri(64, 32),
    "volumes": lambda: _scenes.volumes(64, 32), "random20k": lambda: _scenes.random_spheres(20000),
}


def ray_set(scene, n_cam=96, seed=3, scale=3.0):
    cam = _scenes.camera_rays_np(scene, n_cam, (n_cam * 2) // 3)
    return np.concatenate([cam, _scenes.random_rays(6000, seed, scale)])


@pytest.mark.parametrize("name", sorted(SCENES))
def test_bvh_closest_hit(ctx, name):
    scene = SCENES[name]()
    rays = ray_set(scene, scale=60.0 if name == "random20k" else 3.0)
    ds, os_ = ctx.upload(scene), orc.OracleScene(scene)
    ora = os_.trace_batch(rays, enable_bvh_tree=True, seed=11)
    dev, st = ds.trace_batch(rays, enable_bvh_tree=True, seed=11, want_stats=True)
    # volumes draw ln(xi) (volume.rs:57-58): CUDA logf vs glibc logf may differ by an ulp in t
    ulp = 4 if name == "volumes" else 0
    n_hit = assert_hits_identical(dev, ora, f"{name} culled", ulp)
    assert_hits_identical(ds.trace_batch(rays, enable_bvh_tree=True, seed=11, flags=A.FLAG_EXHAUSTIVE), ora, f"{name} exhaustive", ulp)
    # the same two walks over the reference's own topology (the default walks the gate-box traversal tree)
    for flags in (A.FLAG_REFERENCE_TREE, A.FLAG_REFERENCE_TREE | A.FLAG_EXHAUSTIVE):
        assert_hits_identical(ds.trace_batch(rays, enable_bvh_tree=True, seed=11, flags=flags), ora, f"{name} reference tree {flags}", ulp)
    assert n_hit > 100 and st["rays"] == len(rays)


@pytest.mark.parametrize("name", ["simple", "benchmark", "monkey", "mesh_hdri", "volumes"])
def test_linear_scan_closest_hit(ctx, name):
    """enable_bvh_tree = false: Hittable for Scene (scene.rs:141-156), including its mesh quirk (Q9)."""
    scene = SCENES[name]()
    rays = ray_set(scene)
    ora = orc.OracleScene(scene).trace_batch(rays, enable_bvh_tree=False, seed=2)
    ds = ctx.upload(scene)
    for flags in (0, A.FLAG_REFERENCE_TREE, A.FLAG_EXHAUSTIVE):
        dev = ds.trace_batch(rays, enable_bvh_tree=False, seed=2, flags=flags)
        assert_hits_identical(dev, ora, f"{name} linear flags={flags}", 4 if name == "volumes" else 0)


def test_work_counters_match_counting_walk(ctx):
    """On the reference topology (RTR_FLAG_REFERENCE_TREE) the device's culled walk does the algorithmic work
    SURVEY 8(d) defines (same ordered traversal, up to near/far ties and the culling slack): counters within 2 %
    of the oracle's counting walk.  The default traversal tree tests the same primitives' gates with fewer boxes."""
    for name in ("simple", "monkey", "grid8000"):
        scene = SCENES[name]()
        rays = ray_set(scene)
        counts, _ = orc.OracleScene(scene).count_batch(rays)
        _, st = ctx.upload(scene).trace_batch(rays, want_stats=True, flags=A.FLAG_REFERENCE_TREE)
        for k_dev, k_ora in (("nodes_visited", "nodes"), ("tri_tests", "tris"), ("sphere_tests", "spheres")):
            if counts[k_ora]:
                assert abs(st[k_dev] - counts[k_ora]) <= 0.02 * counts[k_ora] + 16, (name, k_dev, st[k_dev], counts[k_ora])


def test_ties_touching_spheres_and_duplicates(ctx):
    """Q2: exact-t ties resolve to the LAST leaf in the reference's DFS order, whatever order the GPU visits."""
    s = rt.Scene.default()
    m = [s.add_material(rt.Lambertian(rt.Color(0.1 * i, 0, 0))) for i in range(1, 7)]
    for i in range(6):  # six coincident spheres among others
        s.add_object(rt.Sphere(rt.Vec3(0, 0, -3), 1.0, m[i]))
        s.add_object(rt.Sphere(rt.Vec3(2.5 * (i - 2.5), 1.0, -4), 0.7, m[i]))
    rays = ray_set(s)
    ora = orc.OracleScene(s).trace_batch(rays)
    ds = ctx.upload(s)
    assert_hits_identical(ds.trace_batch(rays), ora, "duplicates culled")
    assert_hits_identical(ds.trace_batch(rays, flags=A.FLAG_EXHAUSTIVE), ora, "duplicates exhaustive")
    on_dupes = ora["object"][(ora["object"] != A.RTR_NO_HIT) & (ora["object"] % 2 == 0)]
    assert len(on_dupes) > 500 and set(on_dupes.tolist()) == {10}  # only the last coincident sphere is ever returned


def test_duplicate_triangles_tie(ctx):
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    p, n, i = mesh_fixture("cube")
    mesh = rt.Mesh.from_buffers(p + np.array((0, 0, -4), np.float32), n, np.concatenate([i, i, i[::-1].copy()]), m)
    mesh.build_bvh()
    s.add_object(mesh)
    rays = ray_set(s)
    ora = orc.OracleScene(s).trace_batch(rays)
    assert_hits_identical(ctx.upload(s).trace_batch(rays), ora, "duplicate triangles")
    assert (ora["primitive"][ora["object"] == 0] >= 12).all()  # a later copy always wins the tie


def test_tmin_variants(ctx):
    """tmin is a run-time parameter (Volume::hit passes -inf and hit1.t + 1e-5, volume.rs:36-37)."""
    scene = SCENES["monkey"]()
    rays = ray_set(scene)
    ds, os_ = ctx.upload(scene), orc.OracleScene(scene)
    for tmin in (1e-5, 0.5, -1.0, float("-inf")):
        assert_hits_identical(ds.trace_batch(rays, tmin=tmin), os_.trace_batch(rays, tmin=tmin), f"tmin={tmin}")


def test_degenerate_rays(ctx):
    """Axis-parallel directions (1/0 = inf in the slab test, aabb.rs:86), zero components, huge/tiny |d|."""
    scene = SCENES["simple"]()
    r = np.zeros(12, A.RAY_DTYPE)
    r["origin"] = [(0, 0, 0)] * 6 + [(0, 5, -1)] * 3 + [(0, 0, 0)] * 3
    r["direction"] = [(0, 0, -1), (0, -1, 0), (1, 0, 0), (0, -0.0, -1), (-0.0, 0, -1), (0, 1, 0),
                      (0, -1, 0), (0, -1e-20, 0), (0, -1e20, 0), (1e-30, -1e-30, -1e-30), (0, 0, 0), (3, -1, -7)]
    assert_hits_identical(ctx.upload(scene).trace_batch(r), orc.OracleScene(scene).trace_batch(r), "degenerate")


@pytest.mark.timeout(900)
def test_full_size_culled_equals_exhaustive_1m_spheres(ctx):
    """BASELINE config C3 at full size (1 M spheres, 2^22 primary + 2^22 secondary rays): too big for the CPU
    oracle, so the size-independent property is checked on the device itself — the culled, ordered walk returns
    exactly what the reference-shaped exhaustive walk returns — plus an oracle spot check on 2^15 rays."""
    scene = _scenes.random_spheres(1_000_000)
    ds = ctx.upload(scene)
    info = ds.info()
    assert info["n_nodes"] >= 999_999
    p = rt.make_params(scene, 2048, 2048, seed=0)
    pix = np.arange(2048 * 2048, dtype=np.uint32)
    primary = ctx.camera_rays(p, pix, np.zeros_like(pix))
    h = ds.trace_batch(primary)
    hx = ds.trace_batch(primary, flags=A.FLAG_EXHAUSTIVE)
    os.environ["RTR_BVH_BUILD"] = "host"  # the default above was the device build: the host rebuild must agree
    try:
        assert np.array_equal(h, ctx.upload(scene).trace_batch(primary))
    finally:
        os.environ.pop("RTR_BVH_BUILD")
    assert np.array_equal(h, hx)
    # ... and the gate-box traversal tree returns exactly what a walk of the reference's own topology returns
    assert np.array_equal(h, ds.trace_batch(primary, flags=A.FLAG_REFERENCE_TREE))
    hit = h["object"] != A.RTR_NO_HIT
    assert hit.mean() > 0.3
    rng = np.random.default_rng(1)
    u = rng.normal(size=(len(primary), 3)).astype(np.float32)
    u /= np.linalg.norm(u, axis=1, keepdims=True)
    secondary = np.zeros(len(primary), A.RAY_DTYPE)
    secondary["origin"] = np.where(hit[:, None], h["point"], rng.uniform(-50, 50, (len(primary), 3)).astype(np.float32))
    secondary["direction"] = np.where(hit[:, None], h["normal"] + u, u)
    h2 = ds.trace_batch(secondary)
    assert np.array_equal(h2, ds.trace_batch(secondary, flags=A.FLAG_EXHAUSTIVE))
    assert np.array_equal(h2, ds.trace_batch(secondary, flags=A.FLAG_REFERENCE_TREE | A.FLAG_EXHAUSTIVE))
    sel = rng.choice(len(primary), 1 << 15, replace=False)
    os_ = orc.OracleScene(scene)
    assert_hits_identical(h[sel], os_.trace_batch(primary[sel]), "1M primary sample")
    assert_hits_identical(h2[sel], os_.trace_batch(secondary[sel]), "1M secondary sample")
