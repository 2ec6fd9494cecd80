"""The all-sphere BVH-off kernel (csrc/scan.cu: merged Philox site + warp-cooperative rejection loop) against the
generic megakernel (RTR_SCAN_KERNEL=0) — bit-identical — and against the oracle with the same Philox stream.
`Hittable for Scene` lib/core/scene.rs:141-156; materials lib/materials.rs; sample.rs:44-52."""
import os

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _parity import assert_images_close

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def generic_ctx():
    """A context whose all-sphere BVH-off renders go through the generic megakernel."""
    old = os.environ.get("RTR_SCAN_KERNEL")
    os.environ["RTR_SCAN_KERNEL"] = "0"
    try:
        c = rt.Context(0)
    finally:
        if old is None:
            del os.environ["RTR_SCAN_KERNEL"]
        else:
            os.environ["RTR_SCAN_KERNEL"] = old
    yield c
    c.close()


def every_material(background="uniform", depth=6):
    """Spheres carrying each of the six material kinds (incl. a nested Mix and an Isotropic surface), BVH off."""
    s = rt.Scene.default()
    s.settings = rt.RenderSettings(samples_per_pixel=8, max_ray_depth=depth, enable_bvh_tree=False)
    s.camera = rt.Camera(rt.Vec3(0, 1.0, 4.0), rt.Vec3(0, 0.5, 0), rt.Vec3(0, 1, 0), 60.0, 1.5)
    if background == "gradient":
        s.background = rt.GradientBackground(rt.Color(0.3, 0.5, 1.0), rt.Color(1.0, 0.9, 0.8))
    elif background == "skymap":
        s.background = rt.SkyMap(_scenes.synthetic_hdri(128, 64, 3))
    mats = [
        rt.Lambertian(rt.Color(0.6, 0.6, 0.5)),
        rt.Metal(rt.Color(0.9, 0.8, 0.7), 0.4),
        rt.Dielectric(1.5),
        rt.Emission(rt.Color(1.0, 0.7, 0.4), 6.0),
        rt.MixMaterial(rt.Metal(rt.Color(0.9, 0.9, 0.9), 0.0),
                       rt.MixMaterial(rt.Lambertian(rt.Color(0.2, 0.7, 0.3)), rt.Dielectric(1.3), 0.3), 0.5),
        rt.Isotropic(rt.Color(0.8, 0.3, 0.8)),
        rt.Metal(rt.Color(0.7, 0.7, 0.9), 1.0),  # fuzz 1: many absorbed rays (materials.rs:64-69)
    ]
    handles = [s.add_material(m) for m in mats]
    s.add_object(rt.Sphere(rt.Vec3(0, -100.5, 0), 100.0, handles[0]))
    for i, h in enumerate(handles):
        s.add_object(rt.Sphere(rt.Vec3(-3.0 + i, 0.0 + 0.15 * (i % 3), -0.3 * (i % 2)), 0.5, h))
    s.add_object(rt.Sphere(rt.Vec3(0.0, 0.0, 1.0), -0.3, handles[2]))  # negative radius: inward normals (objects.rs:47)
    return s


def both(ctx, generic_ctx, scene, W, H, seed, **kw):
    p = rt.make_params(scene, W, H, seed=seed, kernel=A.KERNEL_MEGAKERNEL, **kw)
    return ctx.upload(scene).render(p, want_accum=True), generic_ctx.upload(scene).render(p, want_accum=True), p


@pytest.mark.parametrize("background", ["uniform", "gradient", "skymap"])
@pytest.mark.parametrize("depth", [1, 2, 6, 50])
def test_scan_kernel_bit_identical_to_generic_and_matches_oracle(ctx, generic_ctx, background, depth):
    scene = every_material(background, depth)
    (rgba_s, acc_s, st_s), (rgba_g, acc_g, st_g), p = both(ctx, generic_ctx, scene, 96, 64, seed=11)
    assert np.array_equal(acc_s.view(np.uint32), acc_g.view(np.uint32)) and np.array_equal(rgba_s, rgba_g)
    assert st_s["rays"] == st_g["rays"] and st_s["samples"] == st_g["samples"]
    rgba_o, acc_o, st_o = orc.OracleScene(scene).render(p)
    off = 5e-3 if background == "skymap" else 0.0  # acosf/atan2f texel flips (CUDA libm vs glibc)
    assert_images_close(rgba_s, acc_s, rgba_o, acc_o, f"{background} depth {depth}", frac_pixels_off=off)
    if off == 0.0:
        assert st_s["rays"] == st_o["rays"]


@pytest.mark.parametrize("mode", [rt.RenderMode.Clay, rt.RenderMode.Normal, rt.RenderMode.Random])
def test_scan_kernel_render_modes(ctx, generic_ctx, mode):
    scene = every_material()
    scene.settings.mode = mode
    (rgba_s, acc_s, st_s), (rgba_g, acc_g, st_g), p = both(ctx, generic_ctx, scene, 80, 48, seed=5)
    assert np.array_equal(acc_s.view(np.uint32), acc_g.view(np.uint32)) and st_s["rays"] == st_g["rays"]
    rgba_o, acc_o, _ = orc.OracleScene(scene).render(p)
    assert_images_close(rgba_s, acc_s, rgba_o, acc_o, f"mode {mode}")


def test_scan_kernel_benchmark_scene_chunks_slices_and_counters(ctx, generic_ctx):
    """scenes/benchmark.json as shipped: ragged image (not a multiple of the 8x4 tile), a small image whose samples
    are split into chunks, a sample slice with progressive accumulation, and the work counters."""
    scene = _scenes.benchmark()
    assert not scene.settings.enable_bvh_tree
    scene.settings.samples_per_pixel = 24
    for W, H in ((101, 37), (13, 7)):
        (rgba_s, acc_s, st_s), (rgba_g, acc_g, st_g), p = both(ctx, generic_ctx, scene, W, H, seed=9, flags=A.FLAG_COUNT_WORK)
        assert np.array_equal(acc_s.view(np.uint32), acc_g.view(np.uint32)) and np.array_equal(rgba_s, rgba_g)
        for k in ("rays", "sphere_tests", "env_lookups", "nodes_visited", "tri_tests"):
            assert st_s[k] == st_g[k], k
        assert st_s["sphere_tests"] == st_s["rays"] * len(scene.objects)
    scene.settings.samples_per_pixel = 2000  # 13x7 pixels: chunked samples, summed in order
    (_, acc_s, _), (_, acc_g, _), _ = both(ctx, generic_ctx, scene, 13, 7, seed=2)
    assert np.array_equal(acc_s.view(np.uint32), acc_g.view(np.uint32))
    scene.settings.samples_per_pixel = 12
    ds = ctx.upload(scene)
    full = ds.render(rt.make_params(scene, 64, 40, seed=8), want_accum=True)[1]
    parts = [ds.render(rt.make_params(scene, 64, 40, seed=8, sample_begin=b, sample_count=c), want_accum=True)[1]
             for b, c in ((0, 7), (7, 5))]
    np.testing.assert_allclose(sum(parts), full, rtol=2e-6, atol=1e-6)


def test_scan_kernel_is_the_default_for_benchmark_json(ctx):
    scene = _scenes.benchmark()
    scene.settings.samples_per_pixel = 4
    p = rt.make_params(scene, 64, 36, seed=1)
    rgba_o, acc_o, st_o = orc.OracleScene(scene).render(p)
    rgba_d, acc_d, st_d = ctx.upload(scene).render(p, want_accum=True)
    assert st_d["kernel_used"] == A.KERNEL_MEGAKERNEL and st_d["rays"] == st_o["rays"]
    assert_images_close(rgba_d, acc_d, rgba_o, acc_o, "benchmark.json default kernel")
