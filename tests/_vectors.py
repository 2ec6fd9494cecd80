"""Runs an implementation (the oracle, or the CUDA library) against tests/golden/vectors/*.npz — the outputs of the
independent numpy restatement tests/golden/make_vectors.py — and returns per-check mismatch reports.

`impl` is anything with:  trace(scene, rays, tmin, seed) -> HIT_DTYPE array;  scatter(scene, inp, seed) -> SCATTER_OUT;
background(scene, rays) -> (n, 3);  camera(params, pixels, samples) -> RAY_DTYPE;  optionally aabb(boxes, rays, tmin, tmax)."""
import os

import numpy as np

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A

VECTORS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "vectors")


def load(name):
    return np.load(os.path.join(VECTORS, name + ".npz"))


def rays_of(o, d):
    r = np.zeros(len(o), A.RAY_DTYPE)
    r["origin"], r["direction"] = o, d
    return r


def same_bits(a, b):
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def ulp_diff(a, b):
    """distance in float32 steps (same-sign finite values)"""
    ia = np.ascontiguousarray(a, np.float32).view(np.int32).astype(np.int64)
    ib = np.ascontiguousarray(b, np.float32).view(np.int32).astype(np.int64)
    return np.abs(ia - ib)


def one_object_scene(obj_factory, material=None):
    s = rt.Scene.default()
    m = s.add_material(material or rt.Lambertian(rt.Color(0.5, 0.5, 0.5)))
    s.add_object(obj_factory(m))
    return s


def check_hits(got, z, sel, what, ulp=0, with_normal=True):
    """got: HIT_DTYPE for the rays `sel` of vector file z.  Hit/miss exact; t / point / normal bit-identical (ulp=0)."""
    hit = z["hit"][sel]
    ghit = got["object"] != A.RTR_NO_HIT
    assert np.array_equal(ghit, hit), f"{what}: hit/miss differs on {np.count_nonzero(ghit != hit)} of {len(hit)} rays"
    fields = [("t", z["t"][sel][hit]), ("point", z["point"][sel][hit])]
    if with_normal:
        fields.append(("normal", z["normal"][sel][hit]))
    for f, want in fields:
        have = got[f][hit]
        if ulp == 0:
            ok = same_bits(have, want)
            assert ok.all(), f"{what}: {f} differs in {np.count_nonzero(~ok)} of {ok.size} values (max ulp {ulp_diff(have, want).max()})"
        else:
            tol = ulp * 1.2e-7 * (np.abs(np.asarray(want, np.float64)).reshape(len(want), -1).max(1) + 1.0)
            err = np.abs(np.asarray(have, np.float64) - np.asarray(want, np.float64)).reshape(len(want), -1).max(1)
            assert (err <= tol).all(), f"{what}: {f} off by up to {(err / tol).max():.2f} x tolerance"
    if with_normal and "front_face" in z.files:
        assert np.array_equal(got["front_face"][hit].astype(bool), z["front_face"][sel][hit]), f"{what}: front_face"
    return int(hit.sum())


def run_sphere(impl):
    z = load("sphere")
    per, n_hit = int(z["per"]), 0
    block = 0
    for s in z["spheres"]:
        scene = one_object_scene(lambda m: rt.Sphere(rt.Vec3(*map(float, s[:3])), float(s[3]), m))
        for _ in range(2):  # two tmin values per sphere
            sel = slice(block * per, (block + 1) * per)
            tmin = float(z["tmin"][sel][0])
            got = impl.trace(scene, rays_of(z["origin"][sel], z["direction"][sel]), tmin, 0)
            n_hit += check_hits(got, z, sel, f"sphere {s} tmin {tmin}")
            block += 1
    return n_hit


def run_triangle(impl):
    z = load("triangle")
    per, n_hit = int(z["per"]), 0
    for i in range(len(z["p0"])):
        pos = np.stack([z["p0"][i], z["p1"][i], z["p2"][i]])
        scene = one_object_scene(lambda m: _mesh(pos, z["vn"][i], m))
        sel = slice(i * per, (i + 1) * per)
        got = impl.trace(scene, rays_of(z["origin"][sel], z["direction"][sel]), float(z["tmin"]), 0)
        n_hit += check_hits(got, z, sel, f"triangle {i}")
    return n_hit


def _mesh(pos, nrm, m):
    mesh = rt.Mesh.from_buffers(pos, nrm, np.arange(3, dtype=np.uint32), m)
    mesh.build_bvh()
    return mesh


def run_volume(impl):
    z = load("volume")
    per, n_hit = int(z["per"]), 0
    for i, (s, rho) in enumerate(zip(z["spheres"], z["density"])):
        scene = one_object_scene(lambda m: rt.Volume(rt.Sphere(rt.Vec3(*map(float, s[:3])), float(s[3]), m), float(rho)),
                                 rt.Isotropic(rt.Color(0.9, 0.9, 0.9)))
        sel = slice(i * per, (i + 1) * per)
        got = impl.trace(scene, rays_of(z["origin"][sel], z["direction"][sel]), float(z["tmin"]), int(z["seed"]))
        # ln(): libm-dependent (numpy / glibc / CUDA), a few ulp of t; hit/miss can flip when hit_distance grazes the exit
        hit = z["hit"][sel]
        ghit = got["object"] != A.RTR_NO_HIT
        flips = np.count_nonzero(ghit != hit)
        assert flips <= 2, f"volume {i}: {flips} hit/miss flips"
        both = hit & ghit
        want, have = z["t"][sel][both].astype(np.float64), got["t"][both].astype(np.float64)
        assert (np.abs(want - have) <= 4 * 1.2e-7 * (np.abs(want) + 1.0)).all(), f"volume {i}: t beyond 4 ulp"
        assert (got["normal"][both] == np.where(got["front_face"][both][:, None] != 0, 1.0, -1.0) * np.array([1.0, 0, 0], np.float32)).all()
        n_hit += int(both.sum())
    return n_hit


def materials_scene(z):
    s = rt.Scene.default()
    kinds, colors, params = z["mat_kind"], z["mat_color"], z["mat_param"]
    first, second = z["mat_first"], z["mat_second"]

    def build(i):
        c = rt.Color(*map(float, colors[i]))
        k, p = int(kinds[i]), float(params[i])
        if k == A.MAT_LAMBERTIAN: return rt.Lambertian(c)
        if k == A.MAT_METAL: return rt.Metal(c, p)
        if k == A.MAT_EMISSION: return rt.Emission(c, 1.0)  # the vectors hold color * strength
        if k == A.MAT_DIELECTRIC: return rt.Dielectric(p)
        if k == A.MAT_ISOTROPIC: return rt.Isotropic(c)
        return rt.MixMaterial(build(int(first[i])), build(int(second[i])), p)

    handles = [s.add_material(build(i)) for i in range(len(kinds))]
    assert [h.index for h in handles] == list(range(len(kinds)))
    s.add_object(rt.Sphere(rt.Vec3(0, 0, -1), 0.5, handles[0]))
    return s


def run_scatter(impl):
    z = load("scatter")
    scene = materials_scene(z)
    n = len(z["material"])
    q = np.zeros(n, A.SCATTER_IN_DTYPE)
    q["origin"] = z["point"] - z["direction"]
    q["direction"], q["point"], q["normal"] = z["direction"], z["point"], z["normal"]
    q["t"], q["front_face"], q["material"] = 1.0, z["front_face"], z["material"]
    q["pixel"], q["sample"], q["bounce"] = z["pixel"], z["sample"], z["bounce"]
    got = impl.scatter(scene, q, int(z["seed"]))
    some = z["scattered"]
    assert np.array_equal(got["scattered"].astype(bool), some), "scatter: Some/None differs"
    ok = same_bits(got["attenuation"], z["attenuation"])
    assert ok.all(), f"scatter: attenuation differs in {np.count_nonzero(~ok)} values"
    ok = same_bits(got["direction"][some], z["out_direction"][some])
    bad = np.flatnonzero(~ok.all(1))
    assert ok.all(), (f"scatter: scattered direction differs for {len(bad)} events, materials "
                      f"{np.unique(z['material'][some][bad])}, max ulp {ulp_diff(got['direction'][some], z['out_direction'][some]).max()}")
    assert same_bits(got["origin"][some], z["point"][some]).all(), "scatter: scattered origin is not the hit point"
    return int(some.sum())


def run_background(impl):
    z = load("background")
    rays = rays_of(np.zeros_like(z["direction"]), z["direction"])
    s = rt.Scene.default()
    s.background = rt.SkyMap(z["texels"])
    got = impl.background(s, rays)
    want = z["sky_rgb"]
    off = ~(got == want).all(1)
    # acos / atan2 are libm-dependent: a direction within an ulp of a texel border may land in the neighbouring texel
    assert off.mean() < 2e-3, f"SkyMap: {off.sum()} of {len(off)} lookups differ"
    H, W = z["texels"].shape[:2]
    flat = z["texels"].reshape(-1, 3)
    for i in np.flatnonzero(off):  # ... and only there
        idx = int(z["texel_index"][i]); x, y = idx % W, idx // W
        near = [flat[((x + dx) % W) + min(max(y + dy, 0), H - 1) * W] for dx in (-1, 0, 1) for dy in (-1, 0, 1)]
        assert any((got[i] == t).all() for t in near), f"SkyMap: lookup {i} is not an adjacent texel"
    s.background = rt.GradientBackground(rt.Color(*map(float, z["top"])), rt.Color(*map(float, z["bottom"])))
    ok = same_bits(impl.background(s, rays), z["gradient_rgb"])
    assert ok.all(), f"GradientBackground differs in {np.count_nonzero(~ok)} values"
    return len(rays)


def run_camera(impl):
    z = load("camera")
    n = 0
    for i in range(int(z["n_cameras"])):
        c = z[f"cam{i}"]
        W, H = map(int, z[f"size{i}"])
        s = rt.Scene.default()
        s.camera = rt.Camera(rt.Vec3(*map(float, c[0:3])), rt.Vec3(*map(float, c[3:6])), rt.Vec3(*map(float, c[6:9])), float(c[9]), float(c[10]))
        p = rt.make_params(s, W, H, seed=int(z["seed"]))
        got = impl.camera(p, z[f"pixel{i}"], z[f"sample{i}"])
        assert same_bits(got["origin"], z[f"origin{i}"]).all(), f"camera {i}: origin"
        # tan(): one libm call per camera; allow it one ulp, which moves every direction component by <= 2 ulp of the basis
        d_ulp = ulp_diff(got["direction"], z[f"direction{i}"])
        exact = same_bits(got["direction"], z[f"direction{i}"]).all()
        assert exact or d_ulp.max() <= 64, f"camera {i}: direction off by {d_ulp.max()} ulp"
        n += len(got)
    return n


def run_aabb(impl):
    z = load("aabb")
    rays = rays_of(z["origin"], z["direction"])
    boxes = np.concatenate([z["bmin"], z["bmax"]], 1)
    n = 0
    for k in "abc":
        got = impl.aabb(boxes, rays, float(z[f"tmin_{k}"]), float(z[f"tmax_{k}"]))
        assert np.array_equal(got, z[f"hit_{k}"]), f"AABB::hit set {k}: {np.count_nonzero(got != z[f'hit_{k}'])} differ"
        n += int(got.sum())
    return n
