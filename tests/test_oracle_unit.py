"""Pins the CPU oracle.  The reference ships no golden vectors (parity unpinned), so
these are the analytic self-checks SURVEY §8(c) lists, each tied to the reference line it probes."""
import math

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _pyload import mesh_fixture

f32 = np.float32


def rays_of(origins, dirs):
    r = np.zeros(len(origins), A.RAY_DTYPE)
    r["origin"], r["direction"] = origins, dirs
    return r


def one_sphere(center=(0, 0, -2), radius=0.5, material=None):
    s = rt.Scene.default()
    m = s.add_material(material or rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    s.add_object(rt.Sphere(rt.Vec3(*center), radius, m))
    return s


# ---------------------------------------------------------------- RNG substitute
def test_philox_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    assert [hex(x) for x in orc.philox([0] * 4, [0] * 2)] == ["0x6627e8d5", "0xe169c58d", "0xbc57ac4c", "0x9b00dbd8"]
    assert [hex(x) for x in orc.philox([0xFFFFFFFF] * 4, [0xFFFFFFFF] * 2)] == ["0x408f276d", "0x41c83b0e", "0xa20bc7c6", "0x6d5451fd"]
    assert [hex(x) for x in orc.philox([0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344], [0xA4093822, 0x299F31D0])] == \
        ["0xd16cfe09", "0x94fdcceb", "0x5001e420", "0x24126ea1"]


def test_jitter_is_inclusive_unit_interval():
    """render.rs:81,88-89: xi in [0,1]; u = (x+xi)/(W-1) may exceed 1 by < 1 px (Q10)."""
    s = one_sphere()
    p = rt.make_params(s, 16, 16, seed=5)
    pix = np.repeat(np.arange(256, dtype=np.uint32), 64)
    smp = np.tile(np.arange(64, dtype=np.uint32), 256)
    r = orc.camera_rays(p, pix, smp)
    # default camera: origin 0, looking -z, fov 90, aspect 1: dir = (-1 + 2u, 1 - 2v, -1)
    u = (r["direction"][:, 0] + 1) / 2 * 15
    v = (1 - r["direction"][:, 1]) / 2 * 15
    jx, jy = u - (pix % 16), v - (pix // 16)
    assert jx.min() >= -1e-5 and jx.max() <= 1 + 1e-5 and jy.min() >= -1e-5 and jy.max() <= 1 + 1e-5
    assert abs(jx.mean() - 0.5) < 0.01 and abs(jy.mean() - 0.5) < 0.01
    assert np.all(r["direction"][:, 2] == -1.0) and np.all(r["origin"] == 0)


# ---------------------------------------------------------------- camera (camera.rs:19-56)
def test_camera_basis_upper_left_and_aspect():
    s = one_sphere()
    s.camera.aspect_ratio = 1.2  # independent of W/H (benchmark.json)
    p = rt.make_params(s, 3, 3, seed=0)
    r = orc.camera_rays(p, [0, 8], [0, 0])
    # pixel 0 is the TOP-left: x<0, y>0 ; pixel 8 bottom-right
    assert r["direction"][0, 0] < 0 and r["direction"][0, 1] > 0
    assert r["direction"][1, 0] > 0 and r["direction"][1, 1] < 0
    assert abs(r["direction"][:, 0]).max() <= 1.2 * 1.5 and not np.allclose(np.linalg.norm(r["direction"], axis=1), 1)


# ---------------------------------------------------------------- Sphere::hit (objects.rs:23-50)
def test_sphere_closed_form_and_unnormalised_direction():
    o = orc.OracleScene(one_sphere())
    rays = rays_of([(0, 0, 0), (0, 0, 0), (0, 0, -2), (0, 0, 0), (0, 0.6, 0)],
                   [(0, 0, -1), (0, 0, -4), (0, 0, -1), (0, 0, 1), (0, 0, -1)])
    h = o.trace_batch(rays)
    assert h["t"][0] == f32(1.5) and tuple(h["normal"][0]) == (0, 0, 1) and h["front_face"][0] == 1
    assert h["t"][1] == f32(1.5 / 4)                      # t scales with 1/|d| (Q6)
    assert h["t"][2] == f32(0.5) and h["front_face"][2] == 0 and tuple(h["normal"][2]) == (0, 0, 1)  # inside: flipped
    assert h["object"][3] == A.RTR_NO_HIT and h["object"][4] == A.RTR_NO_HIT


def test_sphere_tmin_excludes_self_hit():
    o = orc.OracleScene(one_sphere())
    h = o.trace_batch(rays_of([(0, 0, -1.5)], [(0, 0, 1)]), tmin=1e-5)  # starts on the surface, leaving
    assert h["object"][0] == A.RTR_NO_HIT


# ---------------------------------------------------------------- BVH semantics (bvh.rs:146-199)
def test_equal_t_last_dfs_leaf_wins():
    """merge_optionals keeps left only if strictly smaller (Q2): coincident spheres -> later object."""
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    for _ in range(3):
        s.add_object(rt.Sphere(rt.Vec3(0, 0, -2), 0.5, m))
    o = orc.OracleScene(s)
    r = rays_of([(0, 0, 0)], [(0, 0, -1)])
    assert o.trace_batch(r, enable_bvh_tree=True)["object"][0] == 2
    assert o.trace_batch(r, enable_bvh_tree=False)["object"][0] == 2  # linear scan: tmax < root is strict


def test_bvh_and_linear_scan_agree_on_spheres():
    s = _scenes.simple()
    o = orc.OracleScene(s)
    rays = np.concatenate([_scenes.camera_rays_np(s, 64, 64), _scenes.random_rays(2000, 1)])
    a, b = o.trace_batch(rays, enable_bvh_tree=True), o.trace_batch(rays, enable_bvh_tree=False)
    # a ray can numerically hit a sphere yet miss one of its ancestors' boxes; those are the only differences
    diff = np.flatnonzero(a["object"] != b["object"])
    assert len(diff) <= 2e-3 * len(rays)
    assert np.all(b["object"][diff] != A.RTR_NO_HIT)


def test_counting_walk_finds_same_t():
    for s in (_scenes.simple(), _scenes.mesh_bench(), _scenes.grid_spheres(4)):
        o = orc.OracleScene(s)
        rays = np.concatenate([_scenes.camera_rays_np(s, 48, 48), _scenes.random_rays(2000, 2)])
        h = o.trace_batch(rays)
        counts, t = o.count_batch(rays)
        hit = h["object"] != A.RTR_NO_HIT
        assert np.array_equal(np.isfinite(t), hit)
        assert np.array_equal(t[hit], h["t"][hit])
        assert counts["rays"] == len(rays) and counts["nodes"] > 0


# ---------------------------------------------------------------- Triangle::hit (mesh.rs:147-196)
def quad_scene(normals=((0, 1, 0),) * 4, built=True):
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    pos = [(-1, 0, -1), (1, 0, -1), (1, 0, 1), (-1, 0, 1)]
    mesh = rt.Mesh.from_buffers(pos, normals, [0, 2, 1, 0, 3, 2], m)  # wound so the face normal is +y
    if built:
        mesh.build_bvh()
    s.add_object(mesh)
    return s


def test_triangle_hit_tmax_ignored_and_flat_normal():
    o = orc.OracleScene(quad_scene())
    h = o.trace_batch(rays_of([(0.3, 2, 0.2), (0.3, -2, 0.2), (3, 2, 0)], [(0, -1, 0), (0, 1, 0), (0, -1, 0)]))
    assert h["t"][0] == 2 and h["front_face"][0] == 1 and tuple(h["normal"][0]) == (0, 1, 0)
    assert h["t"][1] == 2 and h["front_face"][1] == 0 and tuple(h["normal"][1]) == (0, -1, 0)  # flipped to face the ray
    assert h["object"][2] == A.RTR_NO_HIT


def test_smooth_normals_are_dead_code_unless_exactly_parallel():
    """mesh.rs:185-189 (Q5): the interpolated normal survives only when |n x fn| == 0 and n.fn > 0."""
    tilted = [(0.1, 1, 0)] * 4
    h = orc.OracleScene(quad_scene(tilted)).trace_batch(rays_of([(0.3, 2, 0.2)], [(0, -1, 0)]))
    assert tuple(h["normal"][0]) == (0, 1, 0)          # replaced by the face normal
    scaled = [(0, 2, 0)] * 4
    h = orc.OracleScene(quad_scene(scaled)).trace_batch(rays_of([(0.3, 2, 0.2)], [(0, -1, 0)]))
    assert h["normal"][0][1] == pytest.approx(2.0, rel=1e-6) and h["normal"][0][1] != 1.0  # kept, unnormalised
    flipped = [(0, -1, 0)] * 4
    h = orc.OracleScene(quad_scene(flipped)).trace_batch(rays_of([(0.3, 2, 0.2)], [(0, -1, 0)]))
    assert tuple(h["normal"][0]) == (0, 1, 0)          # angle = pi > 0 -> face normal


def test_mesh_without_build_bvh_is_invisible():
    h = orc.OracleScene(quad_scene(built=False)).trace_batch(rays_of([(0.3, 2, 0.2)], [(0, -1, 0)]))
    assert h["object"][0] == A.RTR_NO_HIT


def test_determinant_epsilon_scales_with_direction_length():
    """mesh.rs:166: |a| < tmin rejects; a scales with |d| (Q4)."""
    o = orc.OracleScene(quad_scene())
    ray = rays_of([(0.3, 2, 0.2)], [(0, -1e-6, 0)])
    assert o.trace_batch(ray, tmin=1e-5)["object"][0] == A.RTR_NO_HIT   # a = 4 * 1e-6 < 1e-5
    assert o.trace_batch(ray, tmin=1e-7)["object"][0] == 0


# ---------------------------------------------------------------- materials (materials.rs)
def scatter_in(n, material, normal=(0, 0, 1), direction=(0, 0, -1), front=1):
    q = np.zeros(n, A.SCATTER_IN_DTYPE)
    q["origin"], q["direction"] = (0, 0, 1), direction
    q["point"], q["normal"], q["t"], q["front_face"], q["material"] = (0, 0, 0), normal, 1.0, front, material
    q["pixel"], q["sample"], q["bounce"] = np.arange(n), 7, 1
    return q


def materials_scene():
    s = rt.Scene.default()
    s.add_material(rt.Lambertian(rt.Color(0.1, 0.2, 0.3)))            # 0
    s.add_material(rt.Metal(rt.Color(0.9, 0.8, 0.7), 0.0))            # 1
    s.add_material(rt.Metal(rt.Color(0.9, 0.8, 0.7), 0.5))            # 2
    s.add_material(rt.Emission(rt.Color(1.0, 0.5, 0.25), 4.0))        # 3
    s.add_material(rt.Dielectric(1.5))                                # 4
    s.add_material(rt.Isotropic(rt.Color(0.3, 0.3, 0.9)))             # 5
    s.add_material(rt.MixMaterial(rt.Emission(rt.Color(1, 0, 0), 1.0), rt.Emission(rt.Color(0, 1, 0), 1.0), 0.25))  # 6
    s.add_material(rt.MixMaterial(rt.MixMaterial(rt.Lambertian(rt.Color(1, 1, 1)), rt.Dielectric(1.3), 0.5),
                                  rt.Metal(rt.Color(0.5, 0.5, 0.5), 0.1), 0.3))  # 7 (nested)
    s.add_object(rt.Sphere(rt.Vec3(0, 0, -2), 0.5, rt.MaterialHandle(0)))
    return s


def test_lambertian_scatter_is_normal_plus_unit_vector():
    o = orc.OracleScene(materials_scene())
    out = o.scatter_batch(scatter_in(4000, 0))
    assert np.all(out["scattered"] == 1) and np.allclose(out["attenuation"], (0.1, 0.2, 0.3))
    d = out["direction"] - np.array((0, 0, 1), f32)
    assert np.allclose(np.linalg.norm(d, axis=1), 1, atol=1e-6)
    assert np.abs(d.mean(axis=0)).max() < 0.05  # uniform on the sphere


def test_metal_mirror_and_absorption():
    o = orc.OracleScene(materials_scene())
    q = scatter_in(8, 1, direction=(1, 0, -1))
    out = o.scatter_batch(q)
    inv = 1 / math.sqrt(2)
    assert np.allclose(out["direction"], (inv, 0, inv), atol=1e-6) and np.all(out["scattered"] == 1)
    # fuzz pushing the direction under the surface -> (black, None) (materials.rs:65-69)
    out = o.scatter_batch(scatter_in(4000, 2, direction=(1, 0, -0.05)))
    absorbed = out["scattered"] == 0
    assert 0.05 < absorbed.mean() < 0.6
    assert np.all(out["attenuation"][absorbed] == 0) and np.allclose(out["attenuation"][~absorbed], (0.9, 0.8, 0.7))


def test_emission_premultiplied_and_terminal():
    out = orc.OracleScene(materials_scene()).scatter_batch(scatter_in(3, 3))
    assert np.all(out["scattered"] == 0) and np.allclose(out["attenuation"], (4.0, 2.0, 1.0))


def test_dielectric_schlick_endpoints_and_tir():
    o = orc.OracleScene(materials_scene())
    # normal incidence from outside: reflect with probability r0 = ((1-1/1.5)/(1+1/1.5))^2 = 0.04
    out = o.scatter_batch(scatter_in(20000, 4))
    refl = out["direction"][:, 2] > 0
    assert abs(refl.mean() - 0.04) < 0.006
    assert np.allclose(out["direction"][~refl], (0, 0, -1), atol=1e-6) and np.all(out["attenuation"] == 1)
    # grazing (cos -> 0): reflectance -> 1
    out = o.scatter_batch(scatter_in(2000, 4, direction=(1, 0, -1e-4)))
    assert (out["direction"][:, 2] > 0).mean() > 0.99
    # inside (ratio 1.5) beyond the critical angle: always reflect, no draw (Q11)
    out = o.scatter_batch(scatter_in(2000, 4, direction=(1, 0, -0.5), front=0))
    assert np.all(out["direction"][:, 2] > 0)


def test_mix_material_factor_picks_second_below_factor():
    """materials.rs:151: gen_range(0..1) >= factor -> first, else second."""
    out = orc.OracleScene(materials_scene()).scatter_batch(scatter_in(20000, 6))
    second = out["attenuation"][:, 1] == 1
    assert abs(second.mean() - 0.25) < 0.01
    assert np.all(out["attenuation"][~second] == (1, 0, 0))


def test_isotropic_direction_inside_unit_ball_unnormalised():
    out = orc.OracleScene(materials_scene()).scatter_batch(scatter_in(4000, 5))
    n = np.linalg.norm(out["direction"], axis=1)
    assert n.max() <= 1 + 1e-6 and abs((n ** 3).mean() - 0.5) < 0.03 and np.all(out["scattered"] == 1)


# ---------------------------------------------------------------- backgrounds (backgrounds.rs)
def test_gradient_background():
    s = one_sphere()
    s.background = rt.GradientBackground(rt.Color(1, 0, 0), rt.Color(0, 0, 1))
    c = orc.OracleScene(s).background_batch(rays_of([(0, 0, 0)] * 3, [(0, 5, 0), (0, -2, 0), (3, 0, 0)]))
    # ray.rs:24-25: t = 0.5 (d^.y + 1); top*(1-t) + bottom*t  -> straight up gives BOTTOM colour
    assert np.allclose(c, [(0, 0, 1), (1, 0, 0), (0.5, 0, 0.5)])


def test_skymap_texel_indexing_at_axis_directions():
    """backgrounds.rs:85-92 with to_spherical_coords (math.rs:24-29); nearest texel, no filtering."""
    W, H = 8, 4
    img = np.arange(W * H * 3, dtype=f32).reshape(H, W, 3)
    s = one_sphere()
    s.background = rt.SkyMap(img)
    dirs = [(1, 0, 0), (-1, 0, 0), (0, 0, 1), (0, 0, -1), (0, 1, 0), (0, -1, 0)]
    c = orc.OracleScene(s).background_batch(rays_of([(0, 0, 0)] * 6, dirs))

    def expect(d):
        x, y, z = (f32(v) for v in d)
        sx = f32(math.acos(-y)); sy = f32(f32(math.atan2(-z, x)) + f32(math.pi))
        u = f32(sx / f32(math.pi)); v = f32(sy / f32(2 * f32(math.pi)))
        xi = int(f32(v * f32(W))) % W
        yi = H - 1 - int(f32(u * f32(H))) % H
        return img[yi, xi]
    for d, got in zip(dirs, c):
        assert np.array_equal(got, expect(d)), d
    assert np.array_equal(c[4], img[H - 1, 4])  # straight up (u=1) wraps to the bottom row


# ---------------------------------------------------------------- integrator (render.rs)
def furnace(albedo, bg, depth):
    s = one_sphere(center=(0, 0, 0), radius=1.0, material=rt.Lambertian(rt.Color(albedo, albedo, albedo)))
    s.background = rt.UniformBackground(rt.Color(bg, bg, bg))
    s.settings = rt.RenderSettings(samples_per_pixel=4, max_ray_depth=depth, clamp_indirect=100.0)
    return s


def test_closed_furnace_hits_depth_limit_black():
    """Camera inside a closed Lambertian sphere: every path reaches max depth -> black (render.rs:34-36)."""
    s = furnace(0.8, 1.0, 5)
    rgba, acc, st = orc.OracleScene(s).render(rt.make_params(s, 32, 32, seed=1))
    # a path can only end early through a numerical self-hit from outside (t > 1e-5, "shadow acne",
    # same arithmetic as objects.rs:39-44) that scatters outward; rare, but not impossible
    assert (acc[..., :3] / 4).mean() < 0.01 and np.all(rgba[..., 3] == 255)  # sky is 1.0
    assert 0.995 * 32 * 32 * 4 * 5 <= st["rays"] <= 32 * 32 * 4 * 5


def test_convex_sphere_single_bounce_radiance():
    """Outside a Lambertian sphere under a uniform sky: a hit returns albedo*sky (convex: second ray escapes)."""
    s = one_sphere(center=(0, 0, -2), radius=1.2, material=rt.Lambertian(rt.Color(0.5, 0.25, 1.0)))
    s.background = rt.UniformBackground(rt.Color(0.8, 0.8, 0.8))
    s.settings = rt.RenderSettings(samples_per_pixel=16, max_ray_depth=6)
    rgba, acc, st = orc.OracleScene(s).render(rt.make_params(s, 9, 9, seed=2))
    centre = acc[4, 4, :3] / 16
    assert np.allclose(centre, (0.4, 0.2, 0.8), atol=1e-6)
    assert np.allclose(acc[0, 0, :3] / 16, 0.8)
    assert tuple(rgba[4, 4]) == (int(math.sqrt(f32(0.4)) * 255), int(math.sqrt(f32(0.2)) * 255), int(math.sqrt(f32(0.8)) * 255), 255)


def test_clamp_and_u8_saturation():
    """render.rs:95-97 per-sample clamp; color.rs:34-41 saturating `as u8` (Q12)."""
    s = one_sphere(material=rt.Emission(rt.Color(1, 1, 1), 50.0))
    s.settings = rt.RenderSettings(samples_per_pixel=2, clamp_indirect=10.0)
    rgba, acc, _ = orc.OracleScene(s).render(rt.make_params(s, 9, 9, seed=0))
    assert np.all(acc[4, 4, :3] == 20.0) and tuple(rgba[4, 4]) == (255, 255, 255, 255)


def test_render_modes():
    s = one_sphere(radius=1.2)
    for mode, check in ((rt.RenderMode.Normal, lambda a: abs(np.linalg.norm(2 * a[4, 4, :3] - 1) - 1) < 1e-5 and a[4, 4, 2] > 0.9),
                        (rt.RenderMode.Random, lambda a: np.all(a[4, 4, :3] == 0)),
                        (rt.RenderMode.Clay, lambda a: np.allclose(a[4, 4, :3], 0.8 * 0.8, atol=1e-6))):
        s.settings = rt.RenderSettings(samples_per_pixel=1, mode=mode)
        _, acc, _ = orc.OracleScene(s).render(rt.make_params(s, 9, 9, seed=0))
        assert check(acc), mode


def test_render_is_deterministic_and_seed_dependent():
    s = _scenes.simple()
    o = orc.OracleScene(s)
    a = o.render(rt.make_params(s, 40, 30, seed=9))[1]
    b = o.render(rt.make_params(s, 40, 30, seed=9), threads=1)[1]
    c = o.render(rt.make_params(s, 40, 30, seed=10))[1]
    assert np.array_equal(a, b) and not np.array_equal(a, c)


def test_sample_slices_add_up():
    """Philox is keyed by the global sample index: slices are a partition of the full render."""
    s = _scenes.benchmark()
    s.settings.samples_per_pixel = 8
    o = orc.OracleScene(s)
    full = o.render(rt.make_params(s, 32, 24, seed=3))[1]
    a = o.render(rt.make_params(s, 32, 24, seed=3, sample_begin=0, sample_count=3))[1]
    b = o.render(rt.make_params(s, 32, 24, seed=3, sample_begin=3, sample_count=5))[1]
    np.testing.assert_allclose(a + b, full, rtol=1e-6)
    assert np.all(full[..., 3] == 8)


# ---------------------------------------------------------------- Volume::hit (volume.rs:35-72)
def test_sphere_bounded_volume_density_law():
    """P(hit) = 1 - exp(-density * chord) for a chord through the medium."""
    s = rt.Scene.default()
    m = s.add_material(rt.Isotropic(rt.Color(1, 1, 1)))
    s.add_object(rt.Volume(rt.Sphere(rt.Vec3(0, 0, -3), 1.0, m), 0.7))
    o = orc.OracleScene(s)
    n = 40000
    h = o.trace_batch(rays_of([(0, 0, 0)] * n, [(0, 0, -2)] * n))  # |d| = 2: distances are in world units
    hit = h["object"] != A.RTR_NO_HIT
    assert abs(hit.mean() - (1 - math.exp(-0.7 * 2.0))) < 0.01
    assert np.all(h["t"][hit] >= 1.0) and np.all(h["t"][hit] <= 2.0)
    # constant outward normal (1,0,0) (volume.rs:68); d.x = 0 -> d.n = 0 is not < 0 -> back face -> flipped
    assert np.all(h["normal"][hit] == (-1, 0, 0)) and np.all(h["front_face"][hit] == 0)


def test_mesh_bounded_volume_vanishes_from_outside():
    """Q7b: hit2's tmin' = hit1.t + 1e-5 doubles as the determinant epsilon and rejects every triangle."""
    s = rt.Scene.default()
    m = s.add_material(rt.Isotropic(rt.Color(1, 1, 1)))
    p, n, i = mesh_fixture("cube")
    mesh = rt.Mesh.from_buffers(p + np.array((0, 0, -5), f32), n, i, m)
    mesh.build_bvh()
    s.add_object(rt.Volume(mesh, 50.0))
    o = orc.OracleScene(s)
    k = 2000
    outside = o.trace_batch(rays_of([(0.1, 0.2, 0)] * k, [(0, 0, -1)] * k))
    assert np.all(outside["object"] == A.RTR_NO_HIT)
    inside = o.trace_batch(rays_of([(0.1, 0.2, -5)] * k, [(0, 0, -1)] * k))
    assert (inside["object"] != A.RTR_NO_HIT).mean() > 0.95


def test_volume_bounded_by_volume_never_hits_but_is_a_bvh_leaf():
    """lib/volume.rs:20-23 allows any Box<dyn Object> as a boundary.  With a Volume there, the outer query
    boundary.hit(ray, -inf, +inf) runs the inner Volume::hit with tmin = -inf: its second boundary query starts at
    t1 + (-inf) and repeats the first, so t1 >= t2 -> None (volume.rs:36-49).  Adding such an object must not change any hit."""
    def scene(nested):
        s = rt.Scene.default()
        lam = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
        iso = s.add_material(rt.Isotropic(rt.Color(1, 1, 1)))
        s.add_object(rt.Sphere(rt.Vec3(0, -100.5, -1), 100.0, lam))
        s.add_object(rt.Sphere(rt.Vec3(-0.6, 0, -1.2), 0.5, lam))
        s.add_object(rt.Volume(rt.Sphere(rt.Vec3(0.7, 0.1, -1.0), 0.45, iso), 3.0))
        if nested:
            s.add_object(rt.Volume(rt.Volume(rt.Sphere(rt.Vec3(0, 0.3, -1.0), 0.8, iso), 5.0), 2.0))
        return s
    rng = np.random.default_rng(5)
    rays = np.zeros(20000, A.RAY_DTYPE)
    rays["origin"] = rng.uniform(-2, 2, (len(rays), 3))
    rays["direction"] = rng.normal(size=(len(rays), 3))
    for bvh in (True, False):
        a = orc.OracleScene(scene(False)).trace_batch(rays, enable_bvh_tree=bvh, seed=3)
        b = orc.OracleScene(scene(True)).trace_batch(rays, enable_bvh_tree=bvh, seed=3)
        assert np.array_equal(a, b) and (a["object"] == 2).sum() > 100
    # a boundary chain that loops cannot be built in Rust; through the C description it is refused
    low = scene(True).lower()
    low.desc.volumes[low.desc.volumes[low.desc.objects[3].index].boundary.index].boundary = A.ObjectRef(A.OBJ_VOLUME, low.desc.objects[3].index)
    with pytest.raises(ValueError):
        orc.OracleScene(low)
