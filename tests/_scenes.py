"""Scene recipes shared by tests and bench.py."""
import os

import numpy as np

import raytracingrust_b200 as rt
from _pyload import GOLDEN, mesh_fixture, scene_from_json


def simple():
    return scene_from_json(os.path.join(GOLDEN, "scenes", "simple.json"))


def benchmark():
    return scene_from_json(os.path.join(GOLDEN, "scenes", "benchmark.json"))


def mesh_bench(name="monkey"):
    """benches/mesh.rs:10-21: Dielectric(1.5) mesh, camera at the origin, depth 50, spp 10."""
    s = rt.Scene.default()
    s.settings.max_ray_depth = 50
    s.settings.samples_per_pixel = 10
    m = s.add_material(rt.Dielectric(1.5))
    mesh = rt.Mesh.from_buffers(*mesh_fixture(name), m)
    mesh.build_bvh()
    s.add_object(mesh)
    return s


def grid_spheres(n=10):
    """benches/bvh_tree_stress.rs:11-28 (n=10 -> 8000 touching spheres)."""
    s = rt.Scene.default()
    s.camera.lookfrom.y = -20.0
    s.settings.max_ray_depth = 50
    m = s.add_material(rt.Dielectric(1.5))
    r = np.arange(-n, n, dtype=np.float32)
    c = np.stack(np.meshgrid(r, r, r, indexing="ij"), -1).reshape(-1, 3)
    s.add_spheres_bulk(c, np.full(len(c), 0.5, np.float32), m)
    return s


def random_spheres(n, seed=42):
    """C3(ii): n spheres, centres U([-50,50]^3), r U(0.05,0.5), one Lambertian."""
    rng = np.random.default_rng(seed)
    s = rt.Scene.default()
    s.camera = rt.Camera(rt.Vec3(0, 0, 120), rt.Vec3(0, 0, 0), rt.Vec3(0, 1, 0), 60.0, 1.0)
    m = s.add_material(rt.Lambertian(rt.Color(0.7, 0.7, 0.7)))
    c = rng.uniform(-50, 50, (n, 3)).astype(np.float32)
    r = rng.uniform(0.05, 0.5, n).astype(np.float32)
    s.add_spheres_bulk(c, r, m)
    return s


def synthetic_hdri(w, h, seed):
    """Seeded stand-in for the missing assets/*.exr: sky gradient + sun + softboxes."""
    rng = np.random.default_rng(seed)
    v = np.linspace(0, 1, h, dtype=np.float32)[:, None]
    u = np.linspace(0, 1, w, dtype=np.float32)[None, :]
    img = np.zeros((h, w, 3), np.float32)
    img[..., 0] = 0.25 + 0.5 * v
    img[..., 1] = 0.35 + 0.45 * v
    img[..., 2] = 0.55 + 0.4 * v
    for _ in range(6):
        cu, cv, rad, pw = rng.uniform(0, 1), rng.uniform(0.1, 0.9), rng.uniform(0.01, 0.06), rng.uniform(2, 30)
        d2 = (u - cu) ** 2 + (v - cv) ** 2
        img += (pw * np.exp(-d2 / (rad * rad)))[..., None].astype(np.float32) * rng.uniform(0.6, 1.0, 3).astype(np.float32)
    img += rng.uniform(0, 0.02, img.shape).astype(np.float32)
    return img


def mesh_hdri(env_w=1024, env_h=512):
    """C4: house + ground plane + glass monkey under a synthetic studio HDRI (src/gui.rs:56-90 recipe)."""
    s = rt.Scene.default()
    s.camera = rt.Camera(rt.Vec3(-3, 2, 4), rt.Vec3(0, 2, -1), rt.Vec3(0, 1, 0), 90.0, 1.0)
    s.settings.max_ray_depth = 6
    s.background = rt.SkyMap(synthetic_hdri(env_w, env_h, 7))
    lam = s.add_material(rt.Lambertian(rt.Color(0.9, 0.2, 0.2)))
    glass = s.add_material(rt.Dielectric(1.5))
    ground = s.add_material(rt.MixMaterial(rt.Metal(rt.Color(0.1, 0.1, 0.1), 0.0), rt.Lambertian(rt.Color(0.4, 0.4, 0.4)), 0.5))
    for name, mat, off in (("house", lam, (0, 0, 0)), ("plane", ground, (0, 0, 0)), ("fancy_monkey", glass, (-2.5, 1.0, 1.5))):
        p, n, i = mesh_fixture(name)
        mesh = rt.Mesh.from_buffers(p + np.array(off, np.float32), n, i, mat)
        mesh.build_bvh()
        s.add_object(mesh)
    return s


def volumes(env_w=512, env_h=256):
    """C5: mesh-bounded media (camera inside the cylinder, Q7b) + sphere fog over a plane."""
    s = rt.Scene.default()
    s.camera = rt.Camera(rt.Vec3(0, 1, 0.5), rt.Vec3(0, 1, -3), rt.Vec3(0, 1, 0), 90.0, 1.0)
    s.settings.max_ray_depth = 16
    s.background = rt.SkyMap(synthetic_hdri(env_w, env_h, 11))
    ground = s.add_material(rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    iso1 = s.add_material(rt.Isotropic(rt.Color(0.9, 0.9, 0.9)))
    iso2 = s.add_material(rt.Isotropic(rt.Color(0.2, 0.4, 0.9)))
    fog = s.add_material(rt.Isotropic(rt.Color(1.0, 1.0, 1.0)))
    p, n, i = mesh_fixture("plane")
    m = rt.Mesh.from_buffers(p, n, i, ground); m.build_bvh(); s.add_object(m)
    p, n, i = mesh_fixture("donut")
    m = rt.Mesh.from_buffers(p + np.array((0, 1, -3), np.float32), n, i, iso1); m.build_bvh()
    s.add_object(rt.Volume(m, 2.0))
    p, n, i = mesh_fixture("cylinder")
    m = rt.Mesh.from_buffers(p, n, i, iso2); m.build_bvh()
    s.add_object(rt.Volume(m, 0.5))
    s.add_object(rt.Volume(rt.Sphere(rt.Vec3(0, 0, 0), 100.0, fog), 0.01))
    s.add_object(rt.Sphere(rt.Vec3(1.5, 0.5, -2.0), 0.5, ground))
    return s


def camera_rays_np(scene, w, h, seed=0):
    """One jittered camera ray per pixel via the oracle (level-1 parity inputs)."""
    from oracle import orc
    p = rt.make_params(scene, w, h, seed=seed)
    pix = np.arange(w * h, dtype=np.uint32)
    return orc.camera_rays(p, pix, np.zeros_like(pix))


def random_rays(n, seed, scale=3.0):
    rng = np.random.default_rng(seed)
    r = np.zeros(n, rt._abi.RAY_DTYPE)
    r["origin"] = rng.uniform(-scale, scale, (n, 3)).astype(np.float32)
    r["direction"] = rng.normal(size=(n, 3)).astype(np.float32) * rng.uniform(0.1, 3.0, (n, 1)).astype(np.float32)
    return r
