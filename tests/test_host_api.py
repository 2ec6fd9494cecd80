"""Host-side mirror of the reference's scene API (raytracingrust_b200/scene.py): names, defaults,
handle semantics and error behaviour (scene.rs:42-114, materials.rs, volume.rs:26-31)."""
import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from raytracingrust_b200.distributed import sample_slice
from oracle import orc


def test_scene_default_matches_reference():
    s = rt.Scene.default()  # scene.rs:42-65
    assert (s.settings.samples_per_pixel, s.settings.max_ray_depth, s.settings.clamp_indirect) == (5, 6, 10.0)
    assert s.settings.enable_multithreading and s.settings.enable_bvh_tree and s.settings.mode == rt.RenderMode.Full
    assert s.camera.lookat.tuple() == (0, 0, -1) and s.camera.vertical_fov == 90.0 and s.camera.aspect_ratio == 1.0
    assert isinstance(s.background, rt.UniformBackground) and s.background.color.r == pytest.approx(0.8)
    assert rt.Color().g == 1.0  # Color::default is green (color.rs:43-47)


def test_handles_are_indices():
    s = rt.Scene.default()
    a = s.add_material(rt.Dielectric(1.5))
    b = s.add_material(rt.MixMaterial(rt.Metal(rt.Color(0.1, 0.1, 0.1), 0.0), rt.Lambertian(rt.Color(0.4, 0.4, 0.4)), 0.5))
    c = s.add_material(rt.Lambertian(rt.Color(1, 1, 1)))
    assert (a.index, b.index, c.index) == (0, 1, 2) and s.material(b).factor == 0.5
    h = s.add_object(rt.Sphere(rt.Vec3(0, 0, -1), 0.5, c))
    assert h.index == 0 and s.object(h).radius == 0.5
    d = s.lower().desc
    # MixMaterial children are appended after the handles, so MaterialHandle(i) == materials[i]
    assert d.n_materials == 5 and d.materials[1].kind == A.MAT_MIX and d.materials[2].kind == A.MAT_LAMBERTIAN
    assert {d.materials[1].first, d.materials[1].second} == {3, 4}


def test_emission_and_volume_constructors():
    e = rt.Emission(rt.Color(1.0, 0.5, 0.25), 50.0)  # pre-multiplied (materials.rs:79-83, Q13)
    assert (e.color.r, e.color.g, e.color.b) == (50.0, 25.0, 12.5)
    v = rt.Volume(rt.Sphere(rt.Vec3(0, 0, 0), 1.0, rt.MaterialHandle(0)), 0.3)
    assert v.neg_inv_density == float(np.float32(-1.0) / np.float32(0.3))  # volume.rs:26-31


def test_bad_handles_and_nesting_raise():
    s = rt.Scene.default()
    s.add_object(rt.Sphere(rt.Vec3(0, 0, -1), 0.5, rt.MaterialHandle(3)))
    with pytest.raises(IndexError):
        s.lower()  # the reference would panic indexing materials
    s = rt.Scene.default()
    m = s.add_material(rt.Isotropic(rt.Color(1, 1, 1)))
    s.add_object(rt.Volume(rt.Volume(rt.Sphere(rt.Vec3(0, 0, 0), 1.0, m), 1.0), 1.0))
    low = s.lower()  # Box<dyn Object>: a Volume may bound a Volume (volume.rs:20-23)
    assert low.desc.n_volumes == 2 and low.desc.n_objects == 1
    outer = low.desc.volumes[low.desc.objects[0].index]
    assert outer.boundary.kind == A.OBJ_VOLUME and low.desc.volumes[outer.boundary.index].boundary.kind == A.OBJ_SPHERE
    with pytest.raises(IndexError):
        rt.Mesh.from_buffers(np.zeros((3, 3)), np.zeros((3, 3)), [0, 1, 3], m)
    with pytest.raises(ValueError):
        rt.RenderSettings(max_ray_depth=256).to_c()  # u8 in the reference


def test_oracle_rejects_what_the_reference_panics_on():
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(1, 1, 1)))
    s.add_object(rt.Sphere(rt.Vec3(0, 0, float("nan")), 1.0, m))  # z is the default sort axis
    s.add_object(rt.Sphere(rt.Vec3(0, 0, 0), 1.0, m))
    with pytest.raises(ValueError):
        orc.OracleScene(s)  # partial_cmp().unwrap() on a NaN centroid (bvh.rs:99)


def test_render_target_zero_initialised():
    t = rt.RenderTarget(7, 5)  # render.rs:21-30
    assert t.data.shape == (7 * 5 * 4,) and not t.data.any() and t.request_redraw is False


def test_bulk_spheres_lowering_order():
    s = rt.Scene.default()
    m = s.add_material(rt.Lambertian(rt.Color(1, 1, 1)))
    s.add_object(rt.Sphere(rt.Vec3(9, 9, 9), 1.0, m))
    s.add_spheres_bulk(np.arange(12).reshape(4, 3), np.arange(4) + 1, m)
    d = s.lower().desc
    assert d.n_objects == 5 and d.n_spheres == 5
    assert [d.objects[i].index for i in range(5)] == [0, 1, 2, 3, 4]
    assert d.spheres[3].center.tuple() == (6, 7, 8) and d.spheres[3].radius == 3


@pytest.mark.parametrize("spp,world", [(1024, 8), (10, 4), (3, 8), (0, 2), (7, 1)])
def test_sample_slices_partition(spp, world):
    got = [sample_slice(spp, r, world) for r in range(world)]
    assert sum(c for _, c in got) == spp
    pos = 0
    for b, c in got:
        assert b == pos
        pos += c
    assert max(c for _, c in got) - min(c for _, c in got) <= 1
