"""Edge cases through the C ABI: empty / degenerate scenes and targets, limits, error statuses."""
import ctypes as C

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _parity import assert_hits_identical, assert_images_close
from test_oracle_unit import quad_scene

pytestmark = pytest.mark.gpu


def both(ctx, scene, W, H, seed=0, **kw):
    p = rt.make_params(scene, W, H, seed=seed, **kw)
    return ctx.upload(scene).render(p, want_accum=True), orc.OracleScene(scene).render(p)


def test_empty_scene_is_all_background(ctx):
    s = rt.Scene.default()  # BvhNode::None (bvh.rs:61)
    for bvh in (True, False):
        s.settings.enable_bvh_tree = bvh
        (rgba_d, acc_d, st), (rgba_o, acc_o, _) = both(ctx, s, 33, 17)
        assert_images_close(rgba_d, acc_d, rgba_o, acc_o, "empty")
        assert st["rays"] == 33 * 17 * 5 and np.all(rgba_d[..., :3] == int(np.sqrt(np.float32(0.8)) * 255))
    h = ctx.upload(s).trace_batch(_scenes.random_rays(100, 1))
    assert np.all(h["object"] == A.RTR_NO_HIT)


def test_single_object_scenes(ctx):
    """n == 1: BvhNode::Object with no box at all (bvh.rs:62)."""
    for scene in (_scenes.mesh_bench("cube"), quad_scene()):
        rays = np.concatenate([_scenes.camera_rays_np(scene, 32, 32), _scenes.random_rays(2000, 6)])
        assert_hits_identical(ctx.upload(scene).trace_batch(rays), orc.OracleScene(scene).trace_batch(rays), "single")
    s = rt.Scene.default()
    m = s.add_material(rt.Metal(rt.Color(0.9, 0.9, 0.9), 0.2))
    s.add_object(rt.Sphere(rt.Vec3(0, 0, -2), 1.0, m))
    (rgba_d, acc_d, _), (rgba_o, acc_o, _) = both(ctx, s, 40, 40)
    assert_images_close(rgba_d, acc_d, rgba_o, acc_o, "one sphere")


def test_mesh_without_build_bvh_is_invisible_but_bounds_count(ctx):
    s = quad_scene(built=False)
    m = s.add_material(rt.Lambertian(rt.Color(0.3, 0.3, 0.3)))
    s.add_object(rt.Sphere(rt.Vec3(0, -1, -2), 0.5, m))
    rays = np.concatenate([_scenes.camera_rays_np(s, 32, 32), _scenes.random_rays(2000, 7)])
    dev, ora = ctx.upload(s).trace_batch(rays), orc.OracleScene(s).trace_batch(rays)
    assert_hits_identical(dev, ora, "unbuilt mesh")
    assert not np.any(dev["object"] == 0)


def test_zero_sized_targets_and_zero_samples(ctx):
    s = _scenes.simple()
    ds = ctx.upload(s)
    for W, H in ((0, 10), (10, 0), (0, 0)):
        rgba, acc, st = ds.render(rt.make_params(s, W, H))
        assert rgba.size == 0 and st["rays"] == 0
    s.settings.samples_per_pixel = 0  # sqrt(0/0) = NaN -> `as u8` -> 0 (color.rs:34-41)
    (rgba_d, acc_d, st), (rgba_o, acc_o, _) = both(ctx, s, 8, 8)
    assert st["rays"] == 0 and np.array_equal(rgba_d, rgba_o) and np.all(rgba_d[..., :3] == 0) and np.all(rgba_d[..., 3] == 255)


def test_depth_limits(ctx):
    s = _scenes.benchmark()
    for depth in (0, 1, 255):
        s.settings.max_ray_depth = depth
        (rgba_d, acc_d, st), (rgba_o, acc_o, st_o) = both(ctx, s, 48, 27, seed=depth)
        assert_images_close(rgba_d, acc_d, rgba_o, acc_o, f"depth {depth}")
        assert st["rays"] == st_o["rays"]
    assert st["rays"] > 48 * 27 * 5


def test_one_pixel_wide_target_divides_by_zero_like_the_reference(ctx):
    """W = 1: u = (x + xi) / (W - 1) = inf (render.rs:88) -> NaN rays -> NaN colour -> 0."""
    s = _scenes.simple()
    (rgba_d, acc_d, _), (rgba_o, acc_o, _) = both(ctx, s, 1, 4)
    assert np.array_equal(rgba_d, rgba_o)
    assert np.array_equal(np.isnan(acc_d), np.isnan(acc_o))


def test_clamp_and_saturation(ctx):
    s = rt.Scene.default()
    m = s.add_material(rt.Emission(rt.Color(1, 0.5, 0.01), 50.0))
    s.add_object(rt.Sphere(rt.Vec3(0, 0, -2), 1.2, m))
    s.settings.clamp_indirect = 2.5
    (rgba_d, acc_d, _), (rgba_o, acc_o, _) = both(ctx, s, 9, 9)
    assert np.array_equal(acc_d, acc_o) and np.array_equal(rgba_d, rgba_o)
    assert tuple(acc_d[4, 4, :3]) == (12.5, 12.5, 2.5) and tuple(rgba_d[4, 4]) == (255, 255, int(np.sqrt(np.float32(0.5)) * 255), 255)


def test_error_statuses(ctx):
    s = _scenes.simple()
    ds = ctx.upload(s)
    p = rt.make_params(s, 8, 8)
    p.settings.mode = 9
    with pytest.raises(rt.RtrError) as e:
        ds.render(p)
    assert e.value.status == A.RTR_INVALID_ARG
    p = rt.make_params(s, 8, 8)
    p.settings.max_ray_depth = 256
    with pytest.raises(rt.RtrError):
        ds.render(p)
    with pytest.raises(rt.RtrError) as e:
        rt.Context(99)
    assert e.value.status == A.RTR_INVALID_ARG
    # descriptor validation happens in the library too (not only in the Python mirror)
    low = s.lower()
    low.desc.spheres[3].material = 1000
    with pytest.raises(rt.RtrError) as e:
        ctx.upload(low)
    assert e.value.status == A.RTR_INVALID_ARG and "material" in str(e.value)
    low = s.lower()
    vol = (A.VolumeS * 1)(A.VolumeS(A.ObjectRef(A.OBJ_VOLUME, 0), -1.0))
    low.desc.volumes, low.desc.n_volumes = vol, 1
    with pytest.raises(rt.RtrError) as e:
        ctx.upload(low)  # a volume bounded by itself: Rust's ownership tree cannot express it, the index-based ABI can
    assert e.value.status == A.RTR_INVALID_ARG and "cyclic" in str(e.value)
    other = rt.Context(0)
    with pytest.raises(rt.RtrError):
        rt.lib().rtr_render  # noqa: B018
        other_scene = other.upload(s)
        st = A.Stats()
        rc = rt.lib().rtr_render(ctx._h, other_scene._h, C.byref(rt.make_params(s, 8, 8)), None, None, C.byref(st))
        assert rc == A.RTR_INVALID_ARG
        raise rt.RtrError(rc, "scene belongs to another context")


def test_scene_info(ctx):
    info = ctx.upload(_scenes.simple()).info()
    assert 102 < info["n_nodes"] <= 2 * 102 and 7 <= info["max_depth"] <= 16 and info["h2d_bytes"] > 102 * 64


def nested_volume_scene():
    """Volume whose boundary is a Volume (lib/volume.rs:20-23 allows any Box<dyn Object>), among ordinary objects."""
    s = rt.Scene.default()
    s.settings.max_ray_depth = 8
    lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4)))
    iso = s.add_material(rt.Isotropic(rt.Color(0.9, 0.9, 0.9)))
    s.add_object(rt.Sphere(rt.Vec3(0, -100.5, -1), 100.0, lam))
    s.add_object(rt.Sphere(rt.Vec3(-0.6, 0, -1.2), 0.5, lam))
    s.add_object(rt.Volume(rt.Sphere(rt.Vec3(0.7, 0.1, -1.0), 0.45, iso), 3.0))
    s.add_object(rt.Volume(rt.Volume(rt.Sphere(rt.Vec3(0, 0.3, -1.0), 0.8, iso), 5.0), 2.0))           # two levels
    s.add_object(rt.Volume(rt.Volume(rt.Volume(rt.Sphere(rt.Vec3(0, 0, -2), 3.0, iso), 1.0), 1.0), 1.0))  # three
    return s


def test_volume_bounded_by_volume_matches_oracle(ctx):
    """Such a volume never hits in the reference (inner query with tmin = -inf: t1 >= t2, volume.rs:36-49) but it is a
    leaf of the scene BVH through its bounds: hits and images must match the oracle, which recurses as the Rust does."""
    scene = nested_volume_scene()
    ds, os_ = ctx.upload(scene), orc.OracleScene(scene)
    rays = _scenes.random_rays(40000, 3, scale=2.5)
    for bvh in (True, False):
        ora = os_.trace_batch(rays, enable_bvh_tree=bvh, seed=9)
        assert not np.isin(ora["object"], [3, 4]).any() and np.isin(2, ora["object"])  # the plain volume hits, the nested never
        for flags in (0, A.FLAG_EXHAUSTIVE, A.FLAG_REFERENCE_TREE):
            assert_hits_identical(ds.trace_batch(rays, enable_bvh_tree=bvh, seed=9, flags=flags), ora, f"nested volume bvh={bvh} flags={flags}", 4)
    for kernel in (A.KERNEL_MEGAKERNEL, A.KERNEL_WAVEFRONT):
        p = rt.make_params(scene, 96, 64, seed=2, kernel=kernel)
        rgba_o, acc_o, _ = os_.render(p)
        rgba_d, acc_d, _ = ds.render(p, want_accum=True)
        assert_images_close(rgba_d, acc_d, rgba_o, acc_o, f"nested volume kernel {kernel}", frac_pixels_off=3e-2)
