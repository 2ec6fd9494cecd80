/*
 * oracle.cpp — CPU restatement of the RayTracingRust render hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under raytracingrust_b200/ links, imports or executes
 * this file; it is called from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * `--impl reference` leg as the checker and the timed CPU baseline.
 *
 * PARITY UNPINNED: the reference (a Rust crate) ships no tests, golden vectors or expected
 * images, and cannot be built here (no cargo/rustc, crates un-vendored).  Fidelity rests on
 * line-by-line correspondence with the cited reference code plus analytic self-checks in
 * tests/test_oracle_*.py.  Citations are file:line relative to the reference checkout.
 *
 * Arithmetic: IEEE f32 throughout, compiled with -ffp-contract=off (Rust never contracts to
 * FMA).  cgmath 0.18 semantics restated from its published source: dot = (x*x'+y*y')+z*z',
 * normalize = v * (1/|v|), angle = atan2(|a x b|, a.b).
 * Randomness: the reference draws from rand 0.8 thread_rng (OS-seeded ChaCha12), which is not
 * reproducible by design.  Substitute: Philox4x32-10 keyed by the render seed, counter =
 * (pixel, sample, bounce | purpose<<8, iteration), so every draw is identified by what it is
 * for, not by call order.  Float mapping follows rand 0.8 UniformFloat: 23 mantissa bits
 * (u32 >> 9) -> [0,1), then *scale + low.
 */
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstring>
#include <limits>
#include <memory>
#include <string>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

#include "../include/rtr.h"

namespace orc {

typedef float Float;
static const Float PI = 3.14159265358979323846f;  // std::f32::consts::PI (types.rs:5)
static const Float INF = std::numeric_limits<float>::infinity();

// ---------------------------------------------------------------- Vec3 (cgmath 0.18)
struct Vec3 {
    Float x, y, z;
};
static inline Vec3 V(Float x, Float y, Float z) { return Vec3{x, y, z}; }
static inline Vec3 V(const rtr_vec3& v) { return Vec3{v.x, v.y, v.z}; }
static inline rtr_vec3 R(const Vec3& v) { return rtr_vec3{v.x, v.y, v.z}; }
static inline Vec3 operator+(Vec3 a, Vec3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline Vec3 operator-(Vec3 a, Vec3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline Vec3 operator-(Vec3 a) { return V(-a.x, -a.y, -a.z); }
static inline Vec3 operator*(Vec3 a, Float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline Vec3 operator*(Float s, Vec3 a) { return V(s * a.x, s * a.y, s * a.z); }
static inline Vec3 operator/(Vec3 a, Float s) { return V(a.x / s, a.y / s, a.z / s); }
static inline Float dot(Vec3 a, Vec3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
static inline Vec3 cross(Vec3 a, Vec3 b) {
    return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline Float magnitude2(Vec3 a) { return dot(a, a); }
static inline Float magnitude(Vec3 a) { return std::sqrt(dot(a, a)); }
static inline Vec3 normalize(Vec3 a) { return a * (1.0f / magnitude(a)); }
static inline Float angle(Vec3 a, Vec3 b) { return std::atan2(magnitude(cross(a, b)), dot(a, b)); }

// ---------------------------------------------------------------- Color (color.rs)
struct Color {
    Float r, g, b;
};
static inline Color C(Float r, Float g, Float b) { return Color{r, g, b}; }
static inline Color C(const rtr_vec3& v) { return Color{v.x, v.y, v.z}; }
static inline Color operator+(Color a, Color b) { return C(a.r + b.r, a.g + b.g, a.b + b.b); }  // :60-69
static inline Color operator*(Color a, Color b) { return C(a.r * b.r, a.g * b.g, a.b * b.b); }  // :71-80
static inline Color operator*(Color a, Float s) { return C(a.r * s, a.g * s, a.b * s); }        // :49-58

// Rust `f32 as u8`: saturating, NaN -> 0 (color.rs:34-41)
static inline uint8_t f32_as_u8(Float v) {
    if (!(v == v)) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)v;  // truncation toward zero
}
// Rust `f32 as usize`: saturating, NaN -> 0 (backgrounds.rs:89-90)
static inline uint64_t f32_as_usize(Float v) {
    if (!(v == v)) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 18446744073709551616.0f) return UINT64_MAX;
    return (uint64_t)v;
}
// f32::clamp(min,max): NaN stays NaN (render.rs:95-97)
static inline Float f32_clamp(Float v, Float lo, Float hi) {
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;
}

// ---------------------------------------------------------------- Philox4x32-10
struct U4 {
    uint32_t x, y, z, w;
};
static inline U4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0,
                               uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
    for (int i = 0; i < 10; ++i) {
        uint64_t p0 = (uint64_t)M0 * c0, p1 = (uint64_t)M1 * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
        uint32_t n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
        uint32_t n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += W0; k1 += W1;
    }
    return U4{c0, c1, c2, c3};
}

enum Purpose { P_JITTER = 0, P_SPHERE = 1, P_MIX = 2, P_DIELECTRIC = 3, P_VOLUME = 4 };

struct RngKey {
    uint64_t seed;
    uint32_t pixel, sample, bounce;
};
static inline U4 draw(const RngKey& k, uint32_t purpose, uint32_t iter) {
    return philox4x32_10(k.pixel, k.sample, (k.bounce & 0xFFu) | (purpose << 8), iter,
                         (uint32_t)k.seed, (uint32_t)(k.seed >> 32));
}
// rand 0.8 UniformFloat<f32>: value0_1 = (bits>>9) * 2^-23
static inline Float u01(uint32_t bits) { return (Float)(bits >> 9) * (1.0f / 8388608.0f); }
// gen_range(0.0..1.0): value0_1 * 1.0 + 0.0
static inline Float gen_range_01(uint32_t bits) { return u01(bits); }
// gen_range(-1.0..1.0): value0_1 * 2.0 + -1.0 (sample.rs:27-33)
static inline Float gen_range_m11(uint32_t bits) { return u01(bits) * 2.0f + -1.0f; }
// Uniform::from(0.0..=1.0) / gen_range(0.0..=1.0): scale = 1/(1-2^-23) = 1+2^-23 in f32
static inline Float gen_range_01_inclusive(uint32_t bits) {
    const Float scale = 1.0f / (1.0f - 1.0f / 8388608.0f);
    return u01(bits) * scale + 0.0f;
}

// sample.rs:44-52 (ball, rejection) — iteration k uses Philox(purpose SPHERE, iter k).xyz
static inline Vec3 sample_unit_sphere_volume(const RngKey& key) {
    for (uint32_t k = 0;; ++k) {
        U4 r = draw(key, P_SPHERE, k);
        Vec3 v = V(gen_range_m11(r.x), gen_range_m11(r.y), gen_range_m11(r.z));
        if (magnitude2(v) <= 1.0f) return v;
    }
}
// sample.rs:66-70
static inline Vec3 sample_unit_sphere_surface(const RngKey& key) {
    return normalize(sample_unit_sphere_volume(key));
}

// ---------------------------------------------------------------- Ray / HitRecord (ray.rs)
struct Ray {
    Vec3 origin, direction;
    Vec3 at(Float t) const { return origin + direction * t; }  // :15-17
};

struct HitRecord {
    Vec3 point, normal;
    Float t;
    bool front_face;
    uint32_t material;
    // which primitive produced it (not in the reference; needed for ID parity)
    uint32_t object, primitive;
    bool some;
};
static inline HitRecord none() {
    HitRecord h;
    std::memset(&h, 0, sizeof(h));
    h.some = false;
    h.object = RTR_NO_HIT;
    return h;
}
// HitRecord::new (ray.rs:39-54)
static inline HitRecord hit_new(Vec3 point, Vec3 outward_normal, Float t, const Ray& ray,
                                uint32_t material) {
    HitRecord h;
    h.front_face = dot(ray.direction, outward_normal) < 0.0f;
    h.normal = h.front_face ? outward_normal : -outward_normal;
    h.point = point;
    h.t = t;
    h.material = material;
    h.object = 0;
    h.primitive = 0;
    h.some = true;
    return h;
}

// ---------------------------------------------------------------- AABB (aabb.rs)
struct AABB {
    Vec3 min, max;
};
static inline AABB aabb_default() { return AABB{V(0, 0, 0), V(0, 0, 0)}; }  // :140-147
static inline AABB from_point(Vec3 p) { return AABB{p, p}; }                // :21-26
static inline Vec3 centroid(const AABB& b) { return (b.min + b.max) / 2.0f; }  // :28-30
static inline AABB surround(const AABB& a, const AABB& b) {                 // :41-54
    // f32::min/max: if one is NaN return the other
    return AABB{V(std::fmin(a.min.x, b.min.x), std::fmin(a.min.y, b.min.y), std::fmin(a.min.z, b.min.z)),
                V(std::fmax(a.max.x, b.max.x), std::fmax(a.max.y, b.max.y), std::fmax(a.max.z, b.max.z))};
}
static inline AABB epsilon_expand(const AABB& bounds, Float eps) {  // :57-78
    Vec3 dim = V(bounds.max.x - bounds.min.x, bounds.max.y - bounds.min.y, bounds.max.z - bounds.min.z);
    Vec3 c = centroid(bounds);
    AABB r = bounds;
    if (dim.x < eps) { r.min.x = c.x - eps; r.max.x = c.x + eps; }
    if (dim.y < eps) { r.min.y = c.y - eps; r.max.y = c.y + eps; }
    if (dim.z < eps) { r.min.z = c.z - eps; r.max.z = c.z + eps; }
    return r;
}
// AABB::hit (aabb.rs:81-137) — literal: 3 divides, per-axis early out
static inline bool aabb_hit(const AABB& b, const Ray& ray, Float tmin, Float tmax) {
    {
        Float inv_d = 1.0f / ray.direction.x;
        Float t0 = (b.min.x - ray.origin.x) * inv_d;
        Float t1 = (b.max.x - ray.origin.x) * inv_d;
        if (inv_d < 0.0f) std::swap(t0, t1);
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
        if (tmax <= tmin) return false;
    }
    {
        Float inv_d = 1.0f / ray.direction.y;
        Float t0 = (b.min.y - ray.origin.y) * inv_d;
        Float t1 = (b.max.y - ray.origin.y) * inv_d;
        if (inv_d < 0.0f) std::swap(t0, t1);
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
        if (tmax <= tmin) return false;
    }
    {
        Float inv_d = 1.0f / ray.direction.z;
        Float t0 = (b.min.z - ray.origin.z) * inv_d;
        Float t1 = (b.max.z - ray.origin.z) * inv_d;
        if (inv_d < 0.0f) std::swap(t0, t1);
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
        if (tmax <= tmin) return false;
    }
    return true;
}
// same test, also returning the entry distance (final tmin) — used only by the counting walk
static inline bool aabb_hit_enter(const AABB& b, const Ray& ray, Float tmin, Float tmax, Float* enter) {
    const Float o[3] = {ray.origin.x, ray.origin.y, ray.origin.z};
    const Float d[3] = {ray.direction.x, ray.direction.y, ray.direction.z};
    const Float mn[3] = {b.min.x, b.min.y, b.min.z}, mx[3] = {b.max.x, b.max.y, b.max.z};
    for (int a = 0; a < 3; ++a) {
        Float inv_d = 1.0f / d[a];
        Float t0 = (mn[a] - o[a]) * inv_d, t1 = (mx[a] - o[a]) * inv_d;
        if (inv_d < 0.0f) std::swap(t0, t1);
        if (t0 > tmin) tmin = t0;
        if (t1 < tmax) tmax = t1;
        if (tmax <= tmin) return false;
    }
    *enter = tmin;
    return true;
}

// ---------------------------------------------------------------- math.rs
static inline bool near_zero(Vec3 v) {  // :6-9
    const Float EPS = 1.0e-8f;
    return std::fabs(v.x) < EPS && std::fabs(v.y) < EPS && std::fabs(v.z) < EPS;
}
static inline Vec3 reflect(Vec3 v, Vec3 n) { return v - 2.0f * dot(v, n) * n; }  // :12-14
static inline Vec3 refract(Vec3 uv, Vec3 n, Float etai_over_etat) {              // :16-21
    Float cos_theta = std::fmin(dot(n, -uv), 1.0f);
    Vec3 out_perp = etai_over_etat * (uv + cos_theta * n);
    Vec3 out_parallel = -std::sqrt(std::fabs(1.0f - magnitude2(out_perp))) * n;
    return out_perp + out_parallel;
}
static inline Float degrees_to_radians(Float d) { return d * PI / 180.0f; }  // :32-34
// f32::powi(x, 5) lowers to compiler-rt __powisf2: square-and-multiply, r = x; r *= (x^2)^2
static inline Float powi5(Float a) {
    Float r = a;            // n=5, odd: r = 1*a
    Float a2 = a * a;       // n=2
    Float a4 = a2 * a2;     // n=1
    return r * a4;
}

// ---------------------------------------------------------------- scene model
struct Scene;

// bvh.rs:18-28 — BoundsCollection
struct BoundsCollection {
    virtual ~BoundsCollection() {}
    virtual HitRecord hit(uint32_t handle, const Ray& ray, Float tmin, Float tmax, const RngKey& key) const = 0;
    virtual AABB bounds(uint32_t handle) const = 0;
    virtual std::vector<uint32_t> objects() const = 0;
};

struct Counters {
    uint64_t rays = 0, nodes = 0, tris = 0, spheres = 0, env = 0;
    bool work = false;  // also run the counting walk for every ray
};

// bvh.rs:52-56 — enum BvhNode { None, Object(u32), Split(AABB, Box, Box) }
struct BvhNode {
    enum Kind { None, Object, Split } kind = None;
    uint32_t handle = 0;
    AABB bounds;
    std::unique_ptr<BvhNode> left, right;
};

// merge_optionals (bvh.rs:177-199): left only if strictly smaller; ties / NaN -> right
static inline HitRecord merge_optionals(const HitRecord& l, const HitRecord& r) {
    if (l.some && r.some) return (l.t < r.t) ? l : r;
    if (l.some) return l;
    if (r.some) return r;
    return none();
}

// BvhNode::from_list (bvh.rs:59-144)
static std::unique_ptr<BvhNode> from_list(std::vector<uint32_t>& objects, const BoundsCollection& scene,
                                          bool* nan_centroid) {
    std::unique_ptr<BvhNode> node(new BvhNode());
    if (objects.size() == 0) {
        node->kind = BvhNode::None;
        return node;
    }
    if (objects.size() == 1) {
        node->kind = BvhNode::Object;
        node->handle = objects[0];
        return node;
    }
    // :68-72
    AABB bounds = scene.bounds(objects[0]);
    for (size_t i = 1; i < objects.size(); ++i) bounds = surround(bounds, scene.bounds(objects[i]));
    // :75-79
    AABB cents = from_point(centroid(scene.bounds(objects[0])));
    for (size_t i = 1; i < objects.size(); ++i)
        cents = surround(cents, from_point(centroid(scene.bounds(objects[i]))));
    Vec3 spread = cents.max - cents.min;
    int axis = (spread.x > spread.y && spread.x > spread.z) ? 0
             : (spread.y > spread.x && spread.y > spread.z) ? 1 : 2;  // :81-88
    // :91-122 — Rust sort_by is stable; keys cached here (same values the comparator recomputes)
    std::vector<Float> key(objects.size());
    for (size_t i = 0; i < objects.size(); ++i) {
        Vec3 c = centroid(scene.bounds(objects[i]));
        key[i] = axis == 0 ? c.x : axis == 1 ? c.y : c.z;
        if (!(key[i] == key[i])) *nan_centroid = true;  // partial_cmp().unwrap() would panic
    }
    std::vector<uint32_t> order(objects.size());
    for (size_t i = 0; i < order.size(); ++i) order[i] = (uint32_t)i;
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return key[a] < key[b]; });
    // :125-134
    std::vector<uint32_t> left_list, right_list;
    for (size_t i = 0; i < order.size(); ++i) {
        if (i < order.size() / 2) left_list.push_back(objects[order[i]]);
        else right_list.push_back(objects[order[i]]);
    }
    node->kind = BvhNode::Split;
    node->bounds = bounds;
    node->left = from_list(left_list, scene, nan_centroid);
    node->right = from_list(right_list, scene, nan_centroid);
    return node;
}

// BvhNode::hit (bvh.rs:146-173): both children, same tmin/tmax, no ordering
static HitRecord bvh_hit(const BvhNode* n, const Ray& ray, Float tmin, Float tmax,
                         const BoundsCollection& scene, const RngKey& key) {
    switch (n->kind) {
        case BvhNode::Object: return scene.hit(n->handle, ray, tmin, tmax, key);
        case BvhNode::Split:
            if (aabb_hit(n->bounds, ray, tmin, tmax)) {
                HitRecord l = bvh_hit(n->left.get(), ray, tmin, tmax, scene, key);
                HitRecord r = bvh_hit(n->right.get(), ray, tmin, tmax, scene, key);
                return merge_optionals(l, r);
            }
            return none();
        default: return none();
    }
}

// Object trait (traits.rs:17-20) + Hittable + Bounded
struct Object {
    virtual ~Object() {}
    virtual HitRecord hit(const Ray& ray, Float tmin, Float tmax, const RngKey& key) const = 0;
    virtual AABB bounds() const = 0;
    virtual uint32_t material() const = 0;
    virtual int kind() const = 0;
};

// objects.rs:15-60
struct Sphere : Object {
    Vec3 center;
    Float radius;
    uint32_t mat;
    HitRecord hit(const Ray& ray, Float tmin, Float tmax, const RngKey&) const override {
        Vec3 oc = ray.origin - center;
        Float a = magnitude2(ray.direction);
        Float half_b = dot(oc, ray.direction);
        Float c = magnitude2(oc) - radius * radius;
        Float discriminant = half_b * half_b - a * c;
        if (discriminant < 0.0f) return none();
        Float sqrtd = std::sqrt(discriminant);
        Float root = (-half_b - sqrtd) / a;
        if (root < tmin || tmax < root) {
            root = (-half_b + sqrtd) / a;
            if (root < tmin || tmax < root) return none();
        }
        Vec3 point = ray.at(root);
        Vec3 normal = (point - center) / radius;
        return hit_new(point, normal, root, ray, mat);
    }
    AABB bounds() const override {
        return AABB{center - V(radius, radius, radius), center + V(radius, radius, radius)};
    }
    uint32_t material() const override { return mat; }
    int kind() const override { return RTR_OBJ_SPHERE; }
};

// mesh.rs:19-38
struct Vertex {
    Vec3 position, normal;
};
struct Triangle {
    uint32_t vertices[3];
    Vec3 normal;
};

struct Mesh : Object, BoundsCollection {
    std::vector<Vertex> vertices;
    std::vector<Triangle> triangles;
    AABB mesh_bounds;
    std::unique_ptr<BvhNode> bvh_root;
    uint32_t mat;

    // Mesh::from_buffers (mesh.rs:78-117)
    void from_buffers(const float* pos, const float* nrm, uint32_t nv, const uint32_t* idx, uint32_t ni) {
        vertices.resize(nv);
        for (uint32_t i = 0; i < nv; ++i) {
            vertices[i].position = V(pos[3 * i], pos[3 * i + 1], pos[3 * i + 2]);
            vertices[i].normal = V(nrm[3 * i], nrm[3 * i + 1], nrm[3 * i + 2]);
        }
        for (uint32_t i = 0; i + 2 < ni; i += 3) {  // chunks_exact(3)
            Vec3 e1 = vertices[idx[i]].position - vertices[idx[i + 1]].position;
            Vec3 e2 = vertices[idx[i + 2]].position - vertices[idx[i + 1]].position;
            Triangle t;
            t.vertices[0] = idx[i]; t.vertices[1] = idx[i + 1]; t.vertices[2] = idx[i + 2];
            t.normal = normalize(cross(e2, e1));
            triangles.push_back(t);
        }
        AABB b = aabb_default();  // :99 — seeded with the origin (Q8)
        for (const Vertex& v : vertices) b = surround(b, AABB{v.position, v.position});
        mesh_bounds = b;
    }
    // Triangle::hit (mesh.rs:147-196)
    HitRecord tri_hit(const Triangle& tr, const Ray& ray, Float tmin) const {
        const Vertex& v0 = vertices[tr.vertices[0]];
        const Vertex& v1 = vertices[tr.vertices[1]];
        const Vertex& v2 = vertices[tr.vertices[2]];
        Vec3 e1 = v1.position - v0.position;
        Vec3 e2 = v2.position - v0.position;
        Vec3 h = cross(ray.direction, e2);
        Float a = dot(e1, h);
        if (a > -tmin && a < tmin) return none();
        Float f = 1.0f / a;
        Vec3 s = ray.origin - v0.position;
        Float u = f * dot(s, h);
        if (u < 0.0f || u > 1.0f) return none();
        Vec3 q = cross(s, e1);
        Float v = f * dot(ray.direction, q);
        if (v < 0.0f || u + v > 1.0f) return none();
        Float t = f * dot(e2, q);
        Vec3 normal = u * v1.normal + v * v2.normal + (1.0f - u - v) * v0.normal;
        if (angle(normal, tr.normal) > degrees_to_radians(0.0f)) normal = tr.normal;
        if (t > tmin) return hit_new(ray.at(t), normal, t, ray, mat);
        return none();
    }
    // BoundsCollection for Mesh (mesh.rs:200-232)
    HitRecord hit(uint32_t handle, const Ray& ray, Float tmin, Float, const RngKey&) const override {
        HitRecord h = tri_hit(triangles[handle], ray, tmin);
        if (h.some) h.primitive = handle;
        return h;
    }
    AABB bounds(uint32_t handle) const override {
        const Triangle& t = triangles[handle];
        return epsilon_expand(
            surround(from_point(vertices[t.vertices[0]].position),
                     surround(from_point(vertices[t.vertices[1]].position),
                              from_point(vertices[t.vertices[2]].position))),
            0.01f);
    }
    std::vector<uint32_t> objects() const override {
        std::vector<uint32_t> v(triangles.size());
        for (size_t i = 0; i < v.size(); ++i) v[i] = (uint32_t)i;
        return v;
    }
    // Hittable for Mesh (mesh.rs:124-127)
    HitRecord hit(const Ray& ray, Float tmin, Float tmax, const RngKey& key) const override {
        return bvh_hit(bvh_root.get(), ray, tmin, tmax, *this, key);
    }
    AABB bounds() const override { return mesh_bounds; }
    uint32_t material() const override { return mat; }
    int kind() const override { return RTR_OBJ_MESH; }
};

// volume.rs:19-72
struct Volume : Object {
    std::unique_ptr<Object> boundary;
    Float neg_inv_density;
    uint32_t volume_index;  // RNG iteration slot for this volume's draw
    HitRecord hit(const Ray& ray, Float tmin, Float tmax, const RngKey& key) const override {
        HitRecord hit1 = boundary->hit(ray, -INF, INF, key);
        if (!hit1.some) return none();
        HitRecord hit2 = boundary->hit(ray, hit1.t + tmin, INF, key);
        if (!hit2.some) return none();
        if (hit1.t < tmin) hit1.t = tmin;
        if (hit2.t > tmax) hit2.t = tmax;
        if (hit1.t >= hit2.t) return none();
        if (hit1.t < 0.0f) hit1.t = 0.0f;
        Float ray_length = magnitude(ray.direction);
        Float distance_inside_boundary = (hit2.t - hit1.t) * ray_length;
        Float random_float = gen_range_01_inclusive(draw(key, P_VOLUME, volume_index).x);  // :57
        Float hit_distance = neg_inv_density * std::log(random_float);
        if (hit_distance > distance_inside_boundary) return none();
        Float t = hit1.t + hit_distance / ray_length;
        return hit_new(ray.at(t), V(1.0f, 0.0f, 0.0f), t, ray, boundary->material());
    }
    AABB bounds() const override { return boundary->bounds(); }
    uint32_t material() const override { return boundary->material(); }
    int kind() const override { return RTR_OBJ_VOLUME; }
};

// ---------------------------------------------------------------- materials (materials.rs, volume.rs:75-87)
struct Scattered {
    Color attenuation;
    bool some;
    Ray ray;
};

struct Scene : BoundsCollection {
    std::vector<std::unique_ptr<Object>> objs;
    std::vector<rtr_material> materials;
    rtr_background background;
    std::vector<float> texels;
    std::unique_ptr<BvhNode> root;  // scene.build_bvh() (scene.rs:112-114)
    bool nan_centroid = false;

    // BoundsCollection for Scene (scene.rs:117-139)
    HitRecord hit(uint32_t handle, const Ray& ray, Float tmin, Float tmax, const RngKey& key) const override {
        HitRecord h = objs[handle]->hit(ray, tmin, tmax, key);
        if (h.some) h.object = handle;
        return h;
    }
    AABB bounds(uint32_t handle) const override { return objs[handle]->bounds(); }
    std::vector<uint32_t> objects() const override {
        std::vector<uint32_t> v(objs.size());
        for (size_t i = 0; i < v.size(); ++i) v[i] = (uint32_t)i;
        return v;
    }
    // Hittable for BvhTree (bvh.rs:42-50)
    HitRecord hit_bvh(const Ray& ray, Float tmin, Float tmax, const RngKey& key) const {
        return bvh_hit(root.get(), ray, tmin, tmax, *this, key);
    }
    // Hittable for Scene — linear scan (scene.rs:141-156)
    HitRecord hit_linear(const Ray& ray, Float tmin, Float tmax, const RngKey& key) const {
        HitRecord result = none();
        Float closest_so_far = tmax;
        for (size_t i = 0; i < objs.size(); ++i) {
            HitRecord h = objs[i]->hit(ray, tmin, closest_so_far, key);
            if (h.some) {
                h.object = (uint32_t)i;
                closest_so_far = h.t;
                result = h;
            }
        }
        return result;
    }
    HitRecord world_hit(bool bvh, const Ray& ray, Float tmin, Float tmax, const RngKey& key) const {
        return bvh ? hit_bvh(ray, tmin, tmax, key) : hit_linear(ray, tmin, tmax, key);
    }
};

// Dielectric::reflectance (materials.rs:103-108)
static inline Float reflectance(Float cosine, Float idx) {
    Float r0 = (1.0f - idx) / (1.0f + idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * powi5(1.0f - cosine);
}

static inline Scattered lambertian_scatter(Color albedo, const HitRecord& hit, const RngKey& key) {  // :31-41
    Vec3 dir = hit.normal + sample_unit_sphere_surface(key);
    if (near_zero(dir)) dir = hit.normal;
    return Scattered{albedo, true, Ray{hit.point, dir}};
}

static Scattered scatter(const Scene& sc, uint32_t mat_index, const Ray& ray, const HitRecord& hit,
                         const RngKey& key, uint32_t mix_level) {
    const rtr_material& m = sc.materials[mat_index];
    switch (m.kind) {
        case RTR_MAT_LAMBERTIAN: return lambertian_scatter(C(m.color), hit, key);
        case RTR_MAT_METAL: {  // :58-70
            Vec3 reflected = normalize(reflect(ray.direction, hit.normal));
            Ray scattered{hit.point, reflected + m.param * sample_unit_sphere_surface(key)};
            if (dot(scattered.direction, hit.normal) > 0.0f) return Scattered{C(m.color), true, scattered};
            return Scattered{C(0, 0, 0), false, scattered};
        }
        case RTR_MAT_EMISSION: return Scattered{C(m.color), false, Ray{hit.point, V(0, 0, 0)}};  // :88-90
        case RTR_MAT_DIELECTRIC: {  // :113-138
            Float ir = m.param;
            Float refraction_ratio = hit.front_face ? 1.0f / ir : ir;
            Vec3 unit_direction = normalize(ray.direction);
            Float cos_theta = std::fmin(dot(hit.normal, -unit_direction), 1.0f);
            Float sin_theta = std::sqrt(1.0f - cos_theta * cos_theta);
            bool cannot_refract = (refraction_ratio * sin_theta) > 1.0f;
            Vec3 direction;
            // short-circuit: the draw only happens when refraction is possible (Q11)
            if (cannot_refract ||
                reflectance(cos_theta, refraction_ratio) > gen_range_01(draw(key, P_DIELECTRIC, 0).x))
                direction = reflect(unit_direction, hit.normal);
            else
                direction = refract(unit_direction, hit.normal, refraction_ratio);
            return Scattered{C(1, 1, 1), true, Ray{hit.point, direction}};
        }
        case RTR_MAT_MIX: {  // :150-156
            if (gen_range_01(draw(key, P_MIX, mix_level).x) >= m.param)
                return scatter(sc, m.first, ray, hit, key, mix_level + 1);
            return scatter(sc, m.second, ray, hit, key, mix_level + 1);
        }
        case RTR_MAT_ISOTROPIC:  // volume.rs:82-87
            return Scattered{C(m.color), true, Ray{hit.point, sample_unit_sphere_volume(key)}};
    }
    return Scattered{C(0, 0, 0), false, Ray{hit.point, V(0, 0, 0)}};
}

// ---------------------------------------------------------------- backgrounds (backgrounds.rs)
static inline Vec3 to_spherical_coords_xy(Vec3 v, Float* sx, Float* sy) {  // math.rs:24-29
    *sx = std::acos(-v.y);
    *sy = std::atan2(-v.z, v.x) + PI;
    return v;
}
static inline Color background_sample(const Scene& sc, const Ray& ray, Counters* cnt) {
    const rtr_background& bg = sc.background;
    switch (bg.kind) {
        case RTR_BG_UNIFORM: return C(bg.color);  // :27-29
        case RTR_BG_GRADIENT: {                   // :46-48 -> ray.rs:19-26
            Float t = 0.5f * (normalize(ray.direction).y + 1.0f);
            return C(bg.color) * (1.0f - t) + C(bg.bottom) * t;
        }
        default: {  // SkyMap::sample :85-92
            Float sx, sy;
            to_spherical_coords_xy(normalize(ray.direction), &sx, &sy);
            Float u = sx / PI;
            Float v = sy / (2.0f * PI);
            uint64_t W = bg.width, H = bg.height;
            uint64_t x = f32_as_usize(v * (Float)W) % W;
            uint64_t y = H - 1 - f32_as_usize(u * (Float)H) % H;
            if (cnt) cnt->env++;
            const float* p = &sc.texels[3 * (x + y * W)];
            return C(p[0], p[1], p[2]);
        }
    }
}

// ---------------------------------------------------------------- camera (camera.rs)
struct RayOrigin {
    Vec3 origin, horizontal, vertical, lower_left_corner;
    Ray get_ray(Float s, Float t) const {  // :19-24
        return Ray{origin, lower_left_corner + horizontal * s - vertical * t - origin};
    }
};
static RayOrigin ray_origin(const rtr_camera& cam) {  // :37-56
    Float theta = degrees_to_radians(cam.vertical_fov);
    Float h = std::tan(theta / 2.0f);
    Float viewport_height = 2.0f * h;
    Float viewport_width = cam.aspect_ratio * viewport_height;
    Vec3 w = normalize(V(cam.lookfrom) - V(cam.lookat));
    Vec3 u = normalize(cross(V(cam.vertical), w));
    Vec3 v = cross(w, u);
    Vec3 horizontal = viewport_width * u;
    Vec3 vertical = viewport_height * v;
    RayOrigin ro;
    ro.origin = V(cam.lookfrom);
    ro.horizontal = horizontal;
    ro.vertical = vertical;
    ro.lower_left_corner = V(cam.lookfrom) - horizontal / 2.0f + vertical / 2.0f - w;
    return ro;
}

// ---------------------------------------------------------------- integrator (render.rs)
static void count_ray_work(const Scene& sc, bool bvh, const Ray& ray, Float tmin, const RngKey& key, Counters& cnt);
// trace_ray (render.rs:32-59), literal recursion
static Color trace_ray(const Scene& sc, const rtr_settings& st, const Ray& ray, uint32_t depth,
                       RngKey key, Counters* cnt) {
    if (depth >= st.max_ray_depth) return C(0, 0, 0);
    key.bounce = depth;
    if (cnt) cnt->rays++;
    if (cnt && cnt->work) count_ray_work(sc, st.enable_bvh_tree != 0, ray, 0.00001f, key, *cnt);
    HitRecord hit = sc.world_hit(st.enable_bvh_tree != 0, ray, 0.00001f, INF, key);
    if (hit.some) {
        Scattered s;
        switch (st.mode) {
            case RTR_MODE_FULL: s = scatter(sc, hit.material, ray, hit, key, 0); break;
            case RTR_MODE_CLAY: s = lambertian_scatter(C(0.8f, 0.8f, 0.8f), hit, key); break;
            case RTR_MODE_NORMAL: {
                Vec3 n = 0.5f * (normalize(hit.normal) + V(1.0f, 1.0f, 1.0f));
                return C(n.x, n.y, n.z);
            }
            default: return C(0, 0, 0);  // Random
        }
        if (s.some) return s.attenuation * trace_ray(sc, st, s.ray, depth + 1, key, cnt);
        return s.attenuation;
    }
    return background_sample(sc, ray, cnt);
}

// one sample of the loop body render.rs:86-100 (returns the clamped colour)
static inline Ray camera_sample_ray(const RayOrigin& ro, uint32_t W, uint32_t H, uint32_t x, uint32_t y,
                                    const RngKey& key) {
    U4 j = draw(key, P_JITTER, 0);
    Float u = ((Float)x + gen_range_01_inclusive(j.x)) / (Float)(W - 1);
    Float v = ((Float)y + gen_range_01_inclusive(j.y)) / (Float)(H - 1);
    return ro.get_ray(u, v);
}
static inline Color sample_pixel(const Scene& sc, const rtr_settings& st, const RayOrigin& ro, uint32_t W,
                                 uint32_t H, uint32_t index, uint32_t s, uint64_t seed, Counters* cnt) {
    uint32_t x = index % W, y = index / W;  // :76-77
    RngKey key{seed, index, s, 0};
    Ray ray = camera_sample_ray(ro, W, H, x, y, key);
    Color c = trace_ray(sc, st, ray, 0, key, cnt);
    c.r = f32_clamp(c.r, 0.0f, st.clamp_indirect);  // :95-97
    c.g = f32_clamp(c.g, 0.0f, st.clamp_indirect);
    c.b = f32_clamp(c.b, 0.0f, st.clamp_indirect);
    return c;
}

// ---------------------------------------------------------------- counting walk
// Ordered, t-culled traversal of the SAME topology, used only to count the algorithmic work a
// culling tracer needs for a ray set: interior boxes tested, triangle tests, sphere tests.
// It returns the same closest t as the reference walk (asserted in tests) but is not the oracle.
struct CountedHit {
    bool some;
    Float t;
};
static CountedHit counted_object_hit(const Object* o, const Ray& ray, Float tmin, Float tmax,
                                     const RngKey& key, Counters& cnt);

static void counted_walk(const BvhNode* n, const BoundsCollection& coll, const Mesh* mesh,
                         const Scene* scene, const Ray& ray, Float tmin, const RngKey& key,
                         Counters& cnt, CountedHit& best) {
    if (n->kind == BvhNode::None) return;
    if (n->kind == BvhNode::Object) {
        CountedHit h;
        if (mesh) {
            cnt.tris++;
            HitRecord r = mesh->tri_hit(mesh->triangles[n->handle], ray, tmin);
            h = CountedHit{r.some, r.t};
        } else {
            h = counted_object_hit(scene->objs[n->handle].get(), ray, tmin, INF, key, cnt);
        }
        if (h.some && (!best.some || h.t <= best.t)) best = h;
        return;
    }
    // Split: own box was tested by the caller; test the children's boxes, near one first
    const BvhNode* ch[2] = {n->left.get(), n->right.get()};
    bool pierced[2];
    Float enter[2];
    for (int i = 0; i < 2; ++i) {
        if (ch[i]->kind == BvhNode::Split) {
            cnt.nodes++;
            pierced[i] = aabb_hit_enter(ch[i]->bounds, ray, tmin, INF, &enter[i]);
        } else {
            pierced[i] = true;
            enter[i] = -INF;
        }
    }
    int first = (enter[1] < enter[0]) ? 1 : 0;
    for (int k = 0; k < 2; ++k) {
        int i = k == 0 ? first : 1 - first;
        if (!pierced[i]) continue;
        if (best.some && enter[i] > best.t) continue;  // culled
        counted_walk(ch[i], coll, mesh, scene, ray, tmin, key, cnt, best);
    }
}
static CountedHit counted_root(const BvhNode* root, const BoundsCollection& coll, const Mesh* mesh,
                               const Scene* scene, const Ray& ray, Float tmin, const RngKey& key,
                               Counters& cnt) {
    CountedHit best{false, 0.0f};
    if (root->kind == BvhNode::Split) {
        cnt.nodes++;
        Float e;
        if (!aabb_hit_enter(root->bounds, ray, tmin, INF, &e)) return best;
    }
    counted_walk(root, coll, mesh, scene, ray, tmin, key, cnt, best);
    return best;
}
static CountedHit counted_object_hit(const Object* o, const Ray& ray, Float tmin, Float tmax,
                                     const RngKey& key, Counters& cnt) {
    switch (o->kind()) {
        case RTR_OBJ_SPHERE: {
            cnt.spheres++;
            HitRecord r = o->hit(ray, tmin, tmax, key);
            return CountedHit{r.some, r.t};
        }
        case RTR_OBJ_MESH: {
            const Mesh* m = static_cast<const Mesh*>(o);
            return counted_root(m->bvh_root.get(), *m, m, nullptr, ray, tmin, key, cnt);
        }
        default: {
            const Volume* v = static_cast<const Volume*>(o);
            CountedHit h1 = counted_object_hit(v->boundary.get(), ray, -INF, INF, key, cnt);
            if (!h1.some) return h1;
            counted_object_hit(v->boundary.get(), ray, h1.t + tmin, INF, key, cnt);
            HitRecord r = o->hit(ray, tmin, tmax, key);
            return CountedHit{r.some, r.t};
        }
    }
}

static void count_ray_work(const Scene& sc, bool bvh, const Ray& ray, Float tmin, const RngKey& key, Counters& cnt) {
    if (bvh) {
        counted_root(sc.root.get(), sc, nullptr, &sc, ray, tmin, key, cnt);
    } else {
        Float closest = INF;
        for (const auto& o : sc.objs) {
            CountedHit h = counted_object_hit(o.get(), ray, tmin, closest, key, cnt);
            if (h.some) closest = h.t;
        }
    }
}

}  // namespace orc

// ================================================================ C API (ctypes)
using namespace orc;

struct orc_scene {
    Scene scene;
    std::string error;
};

static std::unique_ptr<Object> make_object(const rtr_scene_desc* d, rtr_object_ref ref, uint32_t* volume_counter,
                                           bool allow_volume, std::string* err, bool* nan_centroid) {
    if (ref.kind == RTR_OBJ_SPHERE) {
        if (ref.index >= d->n_spheres) { *err = "sphere index out of range"; return nullptr; }
        const rtr_sphere& s = d->spheres[ref.index];
        std::unique_ptr<Sphere> sp(new Sphere());
        sp->center = V(s.center); sp->radius = s.radius; sp->mat = s.material;
        return sp;
    }
    if (ref.kind == RTR_OBJ_MESH) {
        if (ref.index >= d->n_meshes) { *err = "mesh index out of range"; return nullptr; }
        const rtr_mesh& m = d->meshes[ref.index];
        for (uint32_t i = 0; i < m.n_indices; ++i)
            if (m.indices[i] >= m.n_vertices) { *err = "mesh vertex index out of range"; return nullptr; }
        std::unique_ptr<Mesh> me(new Mesh());
        me->mat = m.material;
        me->from_buffers(m.positions, m.normals, m.n_vertices, m.indices, m.n_indices);
        std::vector<uint32_t> objs = me->objects();
        me->bvh_root = from_list(objs, *me, nan_centroid);  // mesh.build_bvh() (mesh.rs:119-121)
        return me;
    }
    if (ref.kind == RTR_OBJ_VOLUME && allow_volume) {
        if (ref.index >= d->n_volumes) { *err = "volume index out of range"; return nullptr; }
        const rtr_volume& v = d->volumes[ref.index];
        std::unique_ptr<Volume> vo(new Volume());
        // boundary: Box<dyn Object> (volume.rs:20-23) — a Sphere, a Mesh or again a Volume.  The plain-C description is
        // index based, so unlike Rust's ownership tree it could loop: follow the chain first.
        {
            rtr_object_ref b = v.boundary;
            for (uint32_t steps = 0; b.kind == RTR_OBJ_VOLUME; ++steps) {
                if (b.index >= d->n_volumes || steps > d->n_volumes) { *err = "volume boundary chain is cyclic or out of range"; return nullptr; }
                b = d->volumes[b.index].boundary;
            }
        }
        vo->boundary = make_object(d, v.boundary, volume_counter, true, err, nan_centroid);
        if (!vo->boundary) { if (err->empty()) *err = "bad volume boundary"; return nullptr; }
        vo->neg_inv_density = v.neg_inv_density;  // as stored by Volume::new (volume.rs:26-31)
        vo->volume_index = ref.index;
        return vo;
    }
    *err = "unsupported object kind";
    return nullptr;
}

extern "C" {

const char* orc_build_info(void) {
#ifdef _OPENMP
    return "oracle: C++ restatement, -ffp-contract=off, OpenMP";
#else
    return "oracle: C++ restatement, -ffp-contract=off, serial";
#endif
}

void orc_philox(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
    U4 r = philox4x32_10(ctr[0], ctr[1], ctr[2], ctr[3], key[0], key[1]);
    out[0] = r.x; out[1] = r.y; out[2] = r.z; out[3] = r.w;
}

int orc_scene_create(const rtr_scene_desc* d, orc_scene** out, char* errbuf, size_t errlen) {
    std::unique_ptr<orc_scene> os(new orc_scene());
    Scene& sc = os->scene;
    std::string err;
    uint32_t vc = 0;
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        const rtr_material& m = d->materials[i];
        if (m.kind > RTR_MAT_ISOTROPIC) err = "bad material kind";
        if (m.kind == RTR_MAT_MIX && (m.first >= d->n_materials || m.second >= d->n_materials)) err = "mix child out of range";
        sc.materials.push_back(m);
    }
    for (uint32_t i = 0; i < d->n_objects && err.empty(); ++i) {
        std::unique_ptr<Object> o = make_object(d, d->objects[i], &vc, true, &err, &sc.nan_centroid);
        if (!o) break;
        if (o->material() >= d->n_materials) { err = "material handle out of range"; break; }
        sc.objs.push_back(std::move(o));
    }
    sc.background = d->background;
    if (d->background.kind == RTR_BG_SKYMAP) {
        if (!d->background.texels || d->background.width == 0 || d->background.height == 0) err = "skymap without texels";
        else sc.texels.assign(d->background.texels, d->background.texels + (size_t)3 * d->background.width * d->background.height);
        sc.background.texels = nullptr;
    }
    if (err.empty()) {
        std::vector<uint32_t> objs = sc.objects();
        sc.root = from_list(objs, sc, &sc.nan_centroid);
        if (sc.nan_centroid) err = "NaN centroid (the reference panics in partial_cmp().unwrap())";
    }
    if (!err.empty()) {
        if (errbuf && errlen) std::snprintf(errbuf, errlen, "%s", err.c_str());
        return RTR_INVALID_ARG;
    }
    *out = os.release();
    return RTR_OK;
}
void orc_scene_destroy(orc_scene* s) { delete s; }

static void fill_hit(const HitRecord& h, rtr_hit* o) {
    if (!h.some) {
        std::memset(o, 0, sizeof(*o));
        o->object = RTR_NO_HIT;
        return;
    }
    o->object = h.object; o->primitive = h.primitive; o->t = h.t;
    o->point = R(h.point); o->normal = R(h.normal);
    o->front_face = h.front_face ? 1u : 0u; o->material = h.material;
}

// AABB::hit (aabb.rs:81-137) on its own, for the second-opinion vectors (tests/golden/vectors): boxes = n x (min.xyz, max.xyz)
int orc_aabb_hit_batch(const float* boxes, const rtr_ray* rays, size_t n, float tmin, float tmax, uint8_t* out) {
    for (size_t i = 0; i < n; ++i) {
        const float* b = boxes + 6 * i;
        AABB box{V(b[0], b[1], b[2]), V(b[3], b[4], b[5])};
        Ray ray{V(rays[i].origin.x, rays[i].origin.y, rays[i].origin.z), V(rays[i].direction.x, rays[i].direction.y, rays[i].direction.z)};
        out[i] = aabb_hit(box, ray, tmin, tmax) ? 1 : 0;
    }
    return RTR_OK;
}

// world.hit(ray, tmin, +INF) for a ray batch (render.rs:38)
int orc_trace_batch(const orc_scene* s, const rtr_ray* rays, size_t n, float tmin, uint32_t enable_bvh_tree,
                    uint64_t seed, rtr_hit* hits_out) {
#pragma omp parallel for schedule(dynamic, 256)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r{V(rays[i].origin), V(rays[i].direction)};
        RngKey key{seed, (uint32_t)i, 0, 0};
        HitRecord h = s->scene.world_hit(enable_bvh_tree != 0, r, tmin, INF, key);
        fill_hit(h, &hits_out[i]);
    }
    return RTR_OK;
}

// counts for the same batch with an ordered, culled walk; closest_t_out optional
int orc_count_batch(const orc_scene* s, const rtr_ray* rays, size_t n, float tmin, uint64_t seed,
                    uint64_t counts_out[4], float* closest_t_out) {
    uint64_t nodes = 0, tris = 0, sph = 0;
#pragma omp parallel for schedule(dynamic, 256) reduction(+ : nodes, tris, sph)
    for (long long i = 0; i < (long long)n; ++i) {
        Ray r{V(rays[i].origin), V(rays[i].direction)};
        RngKey key{seed, (uint32_t)i, 0, 0};
        Counters c;
        CountedHit h = counted_root(s->scene.root.get(), s->scene, nullptr, &s->scene, r, tmin, key, c);
        if (closest_t_out) closest_t_out[i] = h.some ? h.t : INF;
        nodes += c.nodes; tris += c.tris; sph += c.spheres;
    }
    counts_out[0] = n; counts_out[1] = nodes; counts_out[2] = tris; counts_out[3] = sph;
    return RTR_OK;
}

int orc_scatter_batch(const orc_scene* s, const rtr_scatter_in* in, size_t n, uint32_t mode, uint64_t seed,
                      rtr_scatter_out* out) {
    for (size_t i = 0; i < n; ++i) {
        Ray r{V(in[i].ray.origin), V(in[i].ray.direction)};
        HitRecord h = none();
        h.some = true; h.point = V(in[i].point); h.normal = V(in[i].normal); h.t = in[i].t;
        h.front_face = in[i].front_face != 0; h.material = in[i].material;
        if (mode == RTR_MODE_FULL && h.material >= s->scene.materials.size()) return RTR_INVALID_ARG;
        RngKey key{seed, in[i].pixel, in[i].sample, in[i].bounce};
        Scattered sc = mode == RTR_MODE_CLAY ? lambertian_scatter(C(0.8f, 0.8f, 0.8f), h, key)
                                             : scatter(s->scene, h.material, r, h, key, 0);
        out[i].attenuation = rtr_vec3{sc.attenuation.r, sc.attenuation.g, sc.attenuation.b};
        out[i].scattered = sc.some ? 1u : 0u;
        out[i].ray.origin = R(sc.ray.origin);
        out[i].ray.direction = sc.some ? R(sc.ray.direction) : rtr_vec3{0, 0, 0};
    }
    return RTR_OK;
}

int orc_background_batch(const orc_scene* s, const rtr_ray* rays, size_t n, float* rgb_out) {
    for (size_t i = 0; i < n; ++i) {
        Ray r{V(rays[i].origin), V(rays[i].direction)};
        Color c = background_sample(s->scene, r, nullptr);
        rgb_out[3 * i] = c.r; rgb_out[3 * i + 1] = c.g; rgb_out[3 * i + 2] = c.b;
    }
    return RTR_OK;
}

int orc_camera_rays(const rtr_render_params* p, const uint32_t* pixels, const uint32_t* samples, size_t n,
                    rtr_ray* rays_out) {
    if (p->width < 1 || p->height < 1) return RTR_INVALID_ARG;
    RayOrigin ro = ray_origin(p->camera);
    for (size_t i = 0; i < n; ++i) {
        uint32_t idx = pixels[i];
        RngKey key{p->seed, idx, samples[i], 0};
        Ray r = camera_sample_ray(ro, p->width, p->height, idx % p->width, idx / p->width, key);
        rays_out[i].origin = R(r.origin); rays_out[i].direction = R(r.direction);
    }
    return RTR_OK;
}

// render() (render.rs:61-127).  threads: 0 = all (enable_multithreading), 1 = serial.
// accum_out (optional): W*H*4 floats = sequential per-pixel sum (rgb) and sample count (a).
// count_work != 0 additionally runs the counting walk for every ray (slower; fills node/tri/sphere).
int orc_render(const orc_scene* s, const rtr_render_params* p, int threads, int count_work,
               uint8_t* rgba8_out, float* accum_out, rtr_stats* stats) {
    const Scene& sc = s->scene;
    const rtr_settings st = p->settings;
    const uint32_t W = p->width, H = p->height;
    if (W == 0 || H == 0) return RTR_INVALID_ARG;
    const uint32_t s0 = p->sample_begin;
    const uint32_t sn = p->sample_count ? p->sample_count : st.samples_per_pixel;
    RayOrigin ro = ray_origin(p->camera);
    uint64_t rays = 0, env = 0, nodes = 0, tris = 0, sph = 0;
    int nthreads = threads > 0 ? threads : 1;
#ifdef _OPENMP
    if (threads <= 0) nthreads = omp_get_max_threads();
#endif
    auto t0 = std::chrono::steady_clock::now();
#pragma omp parallel for schedule(dynamic, 16) reduction(+ : rays, env, nodes, tris, sph) num_threads(nthreads)
    for (long long index = 0; index < (long long)W * H; ++index) {
        Color color = C(0, 0, 0);
        Counters cnt;
        cnt.work = count_work != 0;
        for (uint32_t k = 0; k < sn; ++k) {
            Color c = sample_pixel(sc, st, ro, W, H, (uint32_t)index, s0 + k, p->seed, &cnt);
            color = color + c;  // :99
        }
        rays += cnt.rays; env += cnt.env; nodes += cnt.nodes; tris += cnt.tris; sph += cnt.spheres;
        if (accum_out) {
            accum_out[4 * index] = color.r; accum_out[4 * index + 1] = color.g;
            accum_out[4 * index + 2] = color.b; accum_out[4 * index + 3] = (float)sn;
        }
        if (rgba8_out) {
            Float spp = (Float)sn;  // :103-107 (== samples_per_pixel for a full render)
            Color g = C(std::sqrt(color.r / spp), std::sqrt(color.g / spp), std::sqrt(color.b / spp));
            rgba8_out[4 * index] = f32_as_u8(g.r * 255.0f);  // color.rs:34-41
            rgba8_out[4 * index + 1] = f32_as_u8(g.g * 255.0f);
            rgba8_out[4 * index + 2] = f32_as_u8(g.b * 255.0f);
            rgba8_out[4 * index + 3] = 255;
        }
    }
    auto t1 = std::chrono::steady_clock::now();
    if (stats) {
        std::memset(stats, 0, sizeof(*stats));
        stats->device_ms = std::chrono::duration<double, std::milli>(t1 - t0).count();
        stats->rays = rays; stats->samples = (uint64_t)W * H * sn; stats->env_lookups = env;
        stats->kernel_launches = (uint32_t)nthreads;  // host threads used
        stats->nodes_visited = nodes; stats->tri_tests = tris; stats->sphere_tests = sph;
    }
    return RTR_OK;
}

}  // extern "C"
