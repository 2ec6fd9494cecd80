#!/usr/bin/env python
"""Second opinion for the oracle: an INDEPENDENT float32 numpy restatement of the reference's level-1 / level-2
routines, written from the .rs files alone (not from oracle/oracle.cpp), plus the generator of tests/golden/vectors/*.npz.

    python tests/golden/make_vectors.py          # rewrites tests/golden/vectors/*.npz (inputs, Philox draws, outputs)

tests/test_second_opinion.py checks oracle/ (CPU) against these files, tests/test_gpu_second_opinion.py the CUDA path.
Each function cites the reference lines it restates.  numpy float32 arithmetic is IEEE single without contraction, i.e.
what rustc emits for these expressions.

Assumptions about third-party crates that are NOT under /root/reference (no Cargo.lock; ranges from Cargo.toml:42-61).
A maintainer with the crate sources can check each in minutes:
  cgmath 0.18  src/vector.rs   Vector3::dot      = (x*ox + y*oy) + z*oz          (mul_element_wise().sum(), sum = x + y + z)
               src/structure.rs InnerSpace::magnitude2 = dot(self, self); magnitude = sqrt(magnitude2)
               src/structure.rs InnerSpace::normalize  = self * (1 / magnitude)   (normalize_to(one) = self * (one / magnitude))
               src/vector.rs   Vector3::cross    = (y*oz - z*oy, z*ox - x*oz, x*oy - y*ox)
               src/vector.rs   Vector3::angle    = Rad::atan2(cross(self, other).magnitude(), dot(self, other))
  rand 0.8     src/distributions/uniform.rs  UniformFloat<f32>: value1_2 = f32 with exponent 0 and mantissa (u32 >> 9),
               value0_1 = value1_2 - 1;  gen_range(lo..hi) = value1_2 * (hi - lo) + (lo - (hi - lo));
               Uniform::new_inclusive(lo, hi) / gen_range(lo..=hi) = value0_1 * ((hi - lo) / (1 - 2^-23)) + lo
  compiler-rt  __powisf2(a, 5) = a * (a*a)*(a*a)                                  (f32::powi)
The reference's ChaCha12 stream (thread_rng, OS-seeded) cannot be reproduced; oracle and kernels draw Philox4x32-10 words
named by (pixel, sample, bounce | purpose << 8, iteration).  That naming is this repo's convention; the words are inputs
of the vectors below (stored in the .npz files), the float mapping on top of them is rand's.
"""
import os

import numpy as np

F = np.float32
INF = F(np.inf)
PI = F(3.14159265358979323846)  # std::f32::consts::PI (types.rs)
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "vectors")

# ---------------------------------------------------------------------------------------------- cgmath (see header)


def f32(a):
    return np.asarray(a, dtype=F)


def dot(a, b):
    return (a[..., 0] * b[..., 0] + a[..., 1] * b[..., 1]) + a[..., 2] * b[..., 2]


def cross(a, b):
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1],
                     a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], -1)


def magnitude(a):
    return np.sqrt(dot(a, a))


def normalize(a):
    return a * (F(1.0) / magnitude(a))[..., None]


def scale(s, v):
    """scalar (array) times vector"""
    return np.asarray(s, F)[..., None] * v


# ---------------------------------------------------------------------------------------------- rand 0.8 float mapping
def value0_1(bits):
    return (np.asarray(bits, np.uint32) >> np.uint32(9)).astype(F) * F(1.0 / 8388608.0)  # == value1_2 - 1, exactly


def gen_range_01(bits):        # gen_range(0.0..1.0): value1_2 * 1 + (0 - 1)
    return (value0_1(bits) + F(1.0)) * F(1.0) + F(-1.0)


def gen_range_m11(bits):       # gen_range(-1.0..1.0): value1_2 * 2 + (-1 - 2)
    return (value0_1(bits) + F(1.0)) * F(2.0) + F(-3.0)


def gen_range_01_inclusive(bits):  # Uniform::from(0.0..=1.0) / gen_range(0.0..=1.0)
    s = F(1.0) / (F(1.0) - F(1.0 / 8388608.0))
    return value0_1(bits) * s + F(0.0)


# ---------------------------------------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011)
P_JITTER, P_SPHERE, P_MIX, P_DIELECTRIC, P_VOLUME = range(5)


def philox4x32_10(ctr, key):
    """ctr (n, 4) uint32, key (2,) uint32 -> (n, 4) uint32.  Known answers: tests/test_second_opinion.py."""
    c = [np.asarray(ctr[..., i], np.uint64) for i in range(4)]
    k0, k1 = np.uint64(key[0]), np.uint64(key[1])
    M0, M1, W0, W1, MASK = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57), np.uint64(0x9E3779B9), np.uint64(0xBB67AE85), np.uint64(0xFFFFFFFF)
    for _ in range(10):
        p0, p1 = M0 * c[0], M1 * c[2]
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c = [hi1 ^ c[1] ^ k0, lo1, hi0 ^ c[3] ^ k1, lo0]
        k0, k1 = (k0 + W0) & MASK, (k1 + W1) & MASK
    return np.stack(c, -1).astype(np.uint32)


def draw(seed, pixel, sample, bounce, purpose, iteration):
    """This repo's naming of the random words (see header); all arguments broadcast to (n,)."""
    pixel, sample, bounce, iteration = np.broadcast_arrays(*(np.asarray(v, np.uint32) for v in (pixel, sample, bounce, iteration)))
    ctr = np.stack([pixel, sample, (bounce & np.uint32(0xFF)) | np.uint32(purpose << 8), iteration], -1)
    return philox4x32_10(ctr, (seed & 0xFFFFFFFF, seed >> 32))


# ---------------------------------------------------------------------------------------------- ray.rs
def ray_at(o, d, t):  # ray.rs:15-17
    return o + d * np.asarray(t, F)[..., None]


def hit_record(d, outward):  # HitRecord::new, ray.rs:39-54 -> (front_face, normal)
    front = dot(d, outward) < F(0.0)
    return front, np.where(front[..., None], outward, -outward)


# ---------------------------------------------------------------------------------------------- aabb.rs:81-137
def aabb_hit(bmin, bmax, o, d, tmin, tmax):
    n = len(o)
    tmin, tmax = np.full(n, tmin, F), np.full(n, tmax, F)
    alive = np.ones(n, bool)
    with np.errstate(all="ignore"):
        for ax in range(3):
            inv_d = F(1.0) / d[:, ax]
            t0 = (bmin[:, ax] - o[:, ax]) * inv_d
            t1 = (bmax[:, ax] - o[:, ax]) * inv_d
            swap = inv_d < F(0.0)
            t0, t1 = np.where(swap, t1, t0), np.where(swap, t0, t1)
            tmin = np.where(t0 > tmin, t0, tmin)
            tmax = np.where(t1 < tmax, t1, tmax)
            alive &= ~(tmax <= tmin)  # `if tmax <= tmin { return false }` after each axis
    return alive


# ---------------------------------------------------------------------------------------------- objects.rs:23-50
def sphere_hit(center, radius, o, d, tmin, tmax):
    """-> hit, t, point, outward normal"""
    with np.errstate(all="ignore"):
        oc = o - center
        a = dot(d, d)
        half_b = dot(oc, d)
        c = dot(oc, oc) - radius * radius
        disc = half_b * half_b - a * c
        ok = ~(disc < F(0.0))
        sqrtd = np.sqrt(disc)
        root = (-half_b - sqrtd) / a
        bad1 = (root < tmin) | (tmax < root)
        root2 = (-half_b + sqrtd) / a
        bad2 = (root2 < tmin) | (tmax < root2)
        root = np.where(bad1, root2, root)
        ok &= ~(bad1 & bad2)
        point = ray_at(o, d, root)
        normal = (point - center) / np.asarray(radius, F)[..., None]
    return ok, root, point, normal


# ---------------------------------------------------------------------------------------------- mesh.rs:78-96, 147-196
def face_normal(p0, p1, p2):  # Mesh::from_buffers, mesh.rs:85-90
    return normalize(cross(p2 - p1, p0 - p1))


def triangle_hit(p0, p1, p2, n0, n1, n2, fn, o, d, tmin):
    """Triangle::hit (tmax is ignored, mesh.rs:151) -> hit, t, point, outward normal"""
    with np.errstate(all="ignore"):
        e1, e2 = p1 - p0, p2 - p0
        h = cross(d, e2)
        a = dot(e1, h)
        ok = ~((a > -tmin) & (a < tmin))
        f = F(1.0) / a
        s = o - p0
        u = f * dot(s, h)
        ok &= ~((u < F(0.0)) | (u > F(1.0)))
        q = cross(s, e1)
        v = f * dot(d, q)
        ok &= ~((v < F(0.0)) | (u + v > F(1.0)))
        t = f * dot(e2, q)
        normal = scale(u, n1) + scale(v, n2) + scale(F(1.0) - u - v, n0)
        angle = np.arctan2(magnitude(cross(normal, fn)), dot(normal, fn))  # normal.angle(self.normal).0
        normal = np.where((angle > F(0.0))[..., None], fn, normal)         # degrees_to_radians(0.0) == 0
        ok &= t > tmin
    return ok, t, ray_at(o, d, t), normal


# ---------------------------------------------------------------------------------------------- volume.rs:35-72 (sphere boundary)
def volume_hit_sphere(center, radius, neg_inv_density, o, d, tmin, tmax, word):
    """Volume::hit with a Sphere boundary; `word` = the u32 behind thread_rng().gen_range(0.0..=1.0) (volume.rs:57).
    -> hit, t, point (the normal is the constant (1, 0, 0), volume.rs:68)"""
    with np.errstate(all="ignore"):
        ok1, t1, _, _ = sphere_hit(center, radius, o, d, -INF, INF)              # :36
        ok2, t2, _, _ = sphere_hit(center, radius, o, d, t1 + F(tmin), INF)      # :37
        ok = ok1 & ok2
        t1 = np.where(t1 < F(tmin), F(tmin), t1)                                  # :39-41
        t2 = np.where(t2 > F(tmax), F(tmax), t2)                                  # :43-45
        ok &= ~(t1 >= t2)                                                         # :47-49
        t1 = np.where(t1 < F(0.0), F(0.0), t1)                                    # :51-53
        ray_length = magnitude(d)
        inside = (t2 - t1) * ray_length
        random_float = gen_range_01_inclusive(word)
        hit_distance = F(neg_inv_density) * np.log(random_float)                 # libm-dependent: compared within 4 ulp
        ok &= ~(hit_distance > inside)
        t = t1 + hit_distance / ray_length
    return ok, t, ray_at(o, d, t)


# ---------------------------------------------------------------------------------------------- math.rs
def near_zero(v):  # math.rs:6-9
    e = F(1.0e-8)
    return (np.abs(v[..., 0]) < e) & (np.abs(v[..., 1]) < e) & (np.abs(v[..., 2]) < e)


def reflect(v, n):  # math.rs:12-14: vector - 2.0 * vector.dot(normal) * normal
    return v - scale(F(2.0) * dot(v, n), n)


def refract(uv, n, eta):  # math.rs:16-21
    cos_theta = np.minimum(dot(n, -uv), F(1.0))
    out_perp = scale(eta, uv + scale(cos_theta, n))
    out_parallel = scale(-np.sqrt(np.abs(F(1.0) - dot(out_perp, out_perp))), n)
    return out_perp + out_parallel


def reflectance(cosine, idx):  # materials.rs:103-108
    r0 = (F(1.0) - idx) / (F(1.0) + idx)
    r0 = r0 * r0
    a = F(1.0) - cosine
    a2 = a * a
    return r0 + (F(1.0) - r0) * (a * (a2 * a2))  # powi(a, 5), see header


# ---------------------------------------------------------------------------------------------- sample.rs:44-70
def unit_sphere_volume(seed, pixel, sample, bounce, max_iter=64):
    """The rejection loop of UnitSphereVolumeSampler: candidate k = (gen_range(-1..1) x3) of word k.  -> vec, words used"""
    n = len(pixel)
    out = np.zeros((n, 3), F)
    done = np.zeros(n, bool)
    used = np.zeros(n, np.uint32)
    for k in range(max_iter):
        w = draw(seed, pixel, sample, bounce, P_SPHERE, k)
        v = np.stack([gen_range_m11(w[:, 0]), gen_range_m11(w[:, 1]), gen_range_m11(w[:, 2])], -1)
        take = ~done & (dot(v, v) <= F(1.0))
        out[take] = v[take]
        used[take] = k + 1
        done |= take
        if done.all():
            break
    assert done.all()
    return out, used


# ---------------------------------------------------------------------------------------------- materials.rs:31-156, volume.rs:82-87
LAMBERTIAN, METAL, EMISSION, DIELECTRIC, MIX, ISOTROPIC = range(6)


def scatter(materials, mat_index, ray_d, point, normal, front_face, seed, pixel, sample, bounce):
    """Material::scatter for a batch.  materials: list of dicts(kind, color, param, first, second) as rtr_material holds
    them (Emission's color is already color * strength, materials.rs:79-83).  -> attenuation, some, dir"""
    n = len(mat_index)
    idx = np.asarray(mat_index).copy()
    kinds = np.array([m["kind"] for m in materials])
    # MixMaterial (materials.rs:150-156): gen_range(0.0..1.0) >= factor -> first, else second; one word per nesting level
    for level in range(8):
        is_mix = kinds[idx] == MIX
        if not is_mix.any():
            break
        w = gen_range_01(draw(seed, pixel, sample, bounce, P_MIX, level)[:, 0])
        factor = np.array([m["param"] for m in materials], F)[idx]
        first = np.array([m["first"] for m in materials])[idx]
        second = np.array([m["second"] for m in materials])[idx]
        idx = np.where(is_mix, np.where(w >= factor, first, second), idx)
    kind = kinds[idx]
    color = np.array([m["color"] for m in materials], F)[idx]
    param = np.array([m["param"] for m in materials], F)[idx]
    att = color.copy()
    some = np.ones(n, bool)
    direction = np.zeros((n, 3), F)
    ball, _ = unit_sphere_volume(seed, pixel, sample, bounce)
    with np.errstate(all="ignore"):
        surface = normalize(ball)                                           # sample_unit_sphere_surface, sample.rs:66-70
        # Lambertian (materials.rs:31-41)
        m = kind == LAMBERTIAN
        dl = normal + surface
        dl = np.where(near_zero(dl)[:, None], normal, dl)
        direction[m] = dl[m]
        # Metal (:58-70)
        m = kind == METAL
        reflected = normalize(reflect(ray_d, normal))
        dm = reflected + scale(param, surface)
        okm = dot(dm, normal) > F(0.0)
        direction[m] = dm[m]
        some[m] = okm[m]
        att[m & ~okm] = F(0.0)
        # Emission (:88-90)
        m = kind == EMISSION
        some[m] = False
        # Dielectric (:112-138)
        m = kind == DIELECTRIC
        ratio = np.where(front_face, F(1.0) / param, param)
        unit = normalize(ray_d)
        cos_theta = np.minimum(dot(normal, -unit), F(1.0))
        sin_theta = np.sqrt(F(1.0) - cos_theta * cos_theta)
        cannot = (ratio * sin_theta) > F(1.0)
        xi = gen_range_01(draw(seed, pixel, sample, bounce, P_DIELECTRIC, 0)[:, 0])
        refl = cannot | (reflectance(cos_theta, ratio) > xi)
        dd = np.where(refl[:, None], reflect(unit, normal), refract(unit, normal, ratio))
        direction[m] = dd[m]
        att[m] = F(1.0)
        # Isotropic (volume.rs:82-87)
        m = kind == ISOTROPIC
        direction[m] = ball[m]
    return att, some, direction


# ---------------------------------------------------------------------------------------------- backgrounds.rs
def gradient_sample(top, bottom, d):  # :46-48 + ray.rs:19-26
    t = F(0.5) * (normalize(d)[:, 1] + F(1.0))
    return f32(top)[None, :] * (F(1.0) - t)[:, None] + f32(bottom)[None, :] * t[:, None]


def skymap_texel(width, height, d):
    """SkyMap::sample (:85-92) + to_spherical_coords (math.rs:24-29) -> texel index x + y * width (acos / atan2: libm)"""
    with np.errstate(all="ignore"):
        v = normalize(d)
        sx = np.arccos(-v[:, 1])
        sy = np.arctan2(-v[:, 2], v[:, 0]) + PI
        u = sx / PI
        vv = sy / (F(2.0) * PI)
        as_usize = lambda a: np.where(np.isnan(a) | (a < 0), 0, np.minimum(a, F(1.8e19))).astype(np.uint64)  # `as usize` saturates
        x = as_usize(vv * F(width)) % np.uint64(width)
        y = np.uint64(height - 1) - as_usize(u * F(height)) % np.uint64(height)
    return (x + y * np.uint64(width)).astype(np.int64)


# ---------------------------------------------------------------------------------------------- camera.rs:19-56, render.rs:88-92
def camera_basis(lookfrom, lookat, vup, vfov, aspect):
    lookfrom, lookat, vup = f32(lookfrom), f32(lookat), f32(vup)
    theta = F(vfov) * PI / F(180.0)                 # degrees_to_radians, math.rs:32-34
    h = np.tan(theta / F(2.0), dtype=F)
    viewport_height = F(2.0) * h
    viewport_width = F(aspect) * viewport_height
    w = normalize(lookfrom - lookat)
    u = normalize(cross(vup, w))
    v = cross(w, u)
    horizontal = viewport_width * u
    vertical = viewport_height * v
    llc = lookfrom - horizontal / F(2.0) + vertical / F(2.0) - w
    return lookfrom, horizontal, vertical, llc


def camera_rays(basis, W, H, pixel, sample, seed):
    origin, horizontal, vertical, llc = basis
    x, y = (pixel % W).astype(F), (pixel // W).astype(F)
    w = draw(seed, pixel, sample, 0, P_JITTER, 0)
    s = (x + gen_range_01_inclusive(w[:, 0])) / F(W - 1)
    t = (y + gen_range_01_inclusive(w[:, 1])) / F(H - 1)
    d = llc[None, :] + horizontal[None, :] * s[:, None] - vertical[None, :] * t[:, None] - origin[None, :]
    return np.broadcast_to(origin, d.shape).copy(), d


# ============================================================================================== generator
def unit_vectors(rng, n):
    v = rng.normal(size=(n, 3)).astype(F)
    return (v / np.linalg.norm(v, axis=1, keepdims=True)).astype(F)


def rays_towards(rng, n, target, spread, far=6.0):
    """origins around `target` at distance ~far, directions towards points within `spread` of it (unnormalised)"""
    o = (f32(target) + unit_vectors(rng, n) * rng.uniform(0.2, far, (n, 1))).astype(F)
    aim = (f32(target) + rng.normal(size=(n, 3)) * spread).astype(F)
    d = ((aim - o) * rng.uniform(0.2, 3.0, (n, 1))).astype(F)
    return o, d


def gen_aabb(rng):
    n = 16000
    c = rng.uniform(-3, 3, (n, 3)).astype(F)
    half = rng.uniform(0.0, 2.0, (n, 3)).astype(F)
    half[rng.random((n, 3)) < 0.05] = 0.0          # flat boxes
    bmin, bmax = (c - half).astype(F), (c + half).astype(F)
    o, d = rays_towards(rng, n, (0, 0, 0), 3.0)
    o[: n // 8] = c[: n // 8]                        # origin inside the box
    d[rng.random((n, 3)) < 0.04] = 0.0              # axis-parallel rays: 1/0 = inf, 0 * inf = NaN
    d[rng.random((n, 3)) < 0.02] = -0.0
    out = {}
    for name, tmin, tmax in (("a", F(1e-5), INF), ("b", -INF, INF), ("c", F(0.5), F(4.0))):
        out[f"hit_{name}"] = aabb_hit(bmin, bmax, o, d, tmin, tmax)
        out[f"tmin_{name}"], out[f"tmax_{name}"] = tmin, tmax
    return dict(bmin=bmin, bmax=bmax, origin=o, direction=d, **out)


def gen_sphere(rng):
    spheres = f32([[0, 0, -1, 0.5], [0, -100.5, -1, 100.0], [1.5, 0.25, -2.0, 0.011], [-1, 0.5, -2.5, 1.0],
                   [0.3, 0.2, -1.2, -0.45], [10, 15.8, -1, 10.0]])  # negative radius: the hollow-sphere trick
    per = 3000
    O, D, T, HIT, TT, P, N, FF = [], [], [], [], [], [], [], []
    for s in spheres:
        o, d = rays_towards(rng, per, s[:3], abs(s[3]) * 1.2, far=abs(s[3]) * 3 + 3)
        k = per // 6  # origins exactly on the surface (a scattered ray leaving it: c == 0 up to rounding)
        on = (s[:3] + unit_vectors(rng, k) * abs(s[3])).astype(F)
        o[:k] = on
        o[k: 2 * k] = (s[:3] + unit_vectors(rng, k) * abs(s[3]) * rng.uniform(0, 0.9, (k, 1))).astype(F)  # inside
        for tmin in (F(1e-5), F(1e-3)):
            ok, t, p, nrm = sphere_hit(s[:3][None, :], s[3], o, d, tmin, INF)
            ff, nn = hit_record(d, nrm)
            O.append(o); D.append(d); T.append(np.full(per, tmin, F)); HIT.append(ok); TT.append(t); P.append(p); N.append(nn); FF.append(ff)
    return dict(spheres=spheres, per=per, origin=np.concatenate(O), direction=np.concatenate(D), tmin=np.concatenate(T),
                hit=np.concatenate(HIT), t=np.concatenate(TT), point=np.concatenate(P), normal=np.concatenate(N),
                front_face=np.concatenate(FF))


def gen_triangle(rng):
    tris, per = 5, 4000
    P0 = rng.uniform(-1, 1, (tris, 3)).astype(F)
    P1 = (P0 + rng.uniform(-1.5, 1.5, (tris, 3))).astype(F)
    P2 = (P0 + rng.uniform(-1.5, 1.5, (tris, 3))).astype(F)
    FN = face_normal(P0, P1, P2)
    VN = np.stack([unit_vectors(rng, tris) for _ in range(3)], 1)   # (tris, 3 vertices, 3)
    VN[0] = FN[0]                                                    # all three vertex normals == face normal
    VN[1] = -FN[1]                                                   # opposite: angle = pi
    VN[2] = (FN[2][None, :] + 0.2 * VN[2]).astype(F)                # smooth-shading-like
    O, D, HIT, TT, P, N, FF = [], [], [], [], [], [], []
    for i in range(tris):
        cen = ((P0[i] + P1[i] + P2[i]) / F(3.0)).astype(F)
        o, d = rays_towards(rng, per, cen, 0.9)
        ok, t, p, nrm = triangle_hit(P0[i][None], P1[i][None], P2[i][None], VN[i, 0][None], VN[i, 1][None], VN[i, 2][None],
                                     FN[i][None], o, d, F(1e-5))
        ff, nn = hit_record(d, nrm)
        O.append(o); D.append(d); HIT.append(ok); TT.append(t); P.append(p); N.append(nn); FF.append(ff)
    return dict(p0=P0, p1=P1, p2=P2, vn=VN.astype(F), face_normal=FN, per=per, tmin=F(1e-5), origin=np.concatenate(O),
                direction=np.concatenate(D), hit=np.concatenate(HIT), t=np.concatenate(TT), point=np.concatenate(P),
                normal=np.concatenate(N), front_face=np.concatenate(FF))


def gen_volume(rng, seed=20241105):
    spheres = f32([[0, 1, -3, 1.0], [0, 0, 0, 100.0], [0.5, 0.2, -1.0, 0.3]])
    density = f32([2.0, 0.01, 8.0])
    per = 5000
    O, D, HIT, TT, P, W = [], [], [], [], [], []
    for s, rho in zip(spheres, density):
        o, d = rays_towards(rng, per, s[:3], min(abs(s[3]), 3.0), far=min(abs(s[3]) * 2, 8.0))
        o[: per // 4] = (s[:3] + unit_vectors(rng, per // 4) * s[3] * rng.uniform(0, 0.95, (per // 4, 1))).astype(F)  # camera inside (Q7b)
        word = draw(seed, np.arange(per), 0, 0, P_VOLUME, 0)[:, 0]   # ray i of a batch: pixel = i, volume index 0
        ok, t, p = volume_hit_sphere(s[:3][None, :], s[3], F(-1.0) / rho, o, d, F(1e-5), INF, word)
        O.append(o); D.append(d); HIT.append(ok); TT.append(t); P.append(p); W.append(word)
    return dict(spheres=spheres, density=density, per=per, seed=np.uint64(seed), tmin=F(1e-5), origin=np.concatenate(O),
                direction=np.concatenate(D), word=np.concatenate(W), hit=np.concatenate(HIT), t=np.concatenate(TT),
                point=np.concatenate(P))


MATERIALS = [
    dict(kind=LAMBERTIAN, color=(0.9, 0.2, 0.2), param=0.0, first=0, second=0),
    dict(kind=METAL, color=(0.8, 0.8, 0.9), param=0.3, first=0, second=0),
    dict(kind=METAL, color=(0.1, 0.1, 0.1), param=0.0, first=0, second=0),
    dict(kind=EMISSION, color=(4.0, 3.0, 2.0), param=0.0, first=0, second=0),
    dict(kind=DIELECTRIC, color=(0.0, 0.0, 0.0), param=1.5, first=0, second=0),
    dict(kind=ISOTROPIC, color=(0.2, 0.4, 0.9), param=0.0, first=0, second=0),
    dict(kind=MIX, color=(0.0, 0.0, 0.0), param=0.5, first=2, second=0),
    dict(kind=MIX, color=(0.0, 0.0, 0.0), param=0.3, first=6, second=4),   # nested Mix
    dict(kind=DIELECTRIC, color=(0.0, 0.0, 0.0), param=2.4, first=0, second=0),
]


def gen_scatter(rng, seed=77):
    n = 30000
    mat = rng.integers(0, len(MATERIALS), n)
    normal = unit_vectors(rng, n)
    d = (unit_vectors(rng, n) * rng.uniform(0.2, 3.0, (n, 1))).astype(F)
    flip = dot(d, normal) > 0                      # HitRecord::new makes the normal face the ray
    normal = np.where(flip[:, None], -normal, normal).astype(F)
    graze = rng.random(n) < 0.1                    # grazing incidence: total internal reflection for the glass
    d[graze] = (d[graze] - scale(dot(d[graze], normal[graze]) * F(0.97), normal[graze])).astype(F)
    point = rng.uniform(-2, 2, (n, 3)).astype(F)
    front = rng.random(n) < 0.5
    pixel = rng.integers(0, 1 << 21, n).astype(np.uint32)
    sample = rng.integers(0, 4096, n).astype(np.uint32)
    bounce = rng.integers(0, 50, n).astype(np.uint32)
    att, some, out_d = scatter(MATERIALS, mat, d, point, normal, front, seed, pixel, sample, bounce)
    _, used = unit_sphere_volume(seed, pixel, sample, bounce)
    return dict(material=mat.astype(np.uint32), direction=d, point=point, normal=normal, front_face=front, pixel=pixel,
                sample=sample, bounce=bounce, seed=np.uint64(seed), attenuation=att, scattered=some, out_direction=out_d,
                rejection_words_used=used,
                mat_kind=np.array([m["kind"] for m in MATERIALS], np.uint32), mat_color=f32([m["color"] for m in MATERIALS]),
                mat_param=f32([m["param"] for m in MATERIALS]), mat_first=np.array([m["first"] for m in MATERIALS], np.uint32),
                mat_second=np.array([m["second"] for m in MATERIALS], np.uint32))


def gen_background(rng):
    W, H, n = 64, 32, 20000
    texels = rng.uniform(0, 4, (H, W, 3)).astype(F)
    d = (unit_vectors(rng, n) * rng.uniform(0.1, 5.0, (n, 1))).astype(F)
    d[:6] = f32([[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]])
    idx = skymap_texel(W, H, d)
    top, bottom = f32([0.5, 0.7, 1.0]), f32([1.0, 1.0, 1.0])
    return dict(texels=texels, direction=d, texel_index=idx, sky_rgb=texels.reshape(-1, 3)[idx], top=top, bottom=bottom,
                gradient_rgb=gradient_sample(top, bottom, d))


def gen_camera(rng, seed=5):
    cams = [dict(lookfrom=(0, 0, 0), lookat=(0, 0, -1), vup=(0, 1, 0), vfov=90.0, aspect=1.0, W=100, H=100),
            dict(lookfrom=(-3, 2, 4), lookat=(0, 2, -1), vup=(0, 1, 0), vfov=90.0, aspect=1.0, W=1920, H=1080),
            dict(lookfrom=(13, 2, 3), lookat=(0, 0.5, 0), vup=(0.1, 1, 0), vfov=20.0, aspect=1.2, W=7, H=5)]
    out = dict(seed=np.uint64(seed), n_cameras=len(cams))
    for i, c in enumerate(cams):
        n = 3000
        pixel = rng.integers(0, c["W"] * c["H"], n).astype(np.uint32)
        sample = rng.integers(0, 5000, n).astype(np.uint32)
        o, d = camera_rays(camera_basis(c["lookfrom"], c["lookat"], c["vup"], c["vfov"], c["aspect"]), c["W"], c["H"], pixel, sample, seed)
        out.update({f"cam{i}": f32(list(c["lookfrom"]) + list(c["lookat"]) + list(c["vup"]) + [c["vfov"], c["aspect"]]),
                    f"size{i}": np.array([c["W"], c["H"]], np.uint32), f"pixel{i}": pixel, f"sample{i}": sample,
                    f"origin{i}": o, f"direction{i}": d})
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    for name, fn in (("aabb", gen_aabb), ("sphere", gen_sphere), ("triangle", gen_triangle), ("volume", gen_volume),
                     ("scatter", gen_scatter), ("background", gen_background), ("camera", gen_camera)):
        data = fn(np.random.default_rng(sum(map(ord, name))))
        np.savez_compressed(os.path.join(OUT, name + ".npz"), **data)
        print(name, {k: (v.shape if hasattr(v, "shape") else v) for k, v in list(data.items())[:4]}, "...")


if __name__ == "__main__":
    main()
