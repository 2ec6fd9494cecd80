// device_core.cuh — device routines of the render hot path (sm_100a).
//
// Every arithmetic expression below reproduces the reference's f32 evaluation order so that,
// compiled with -fmad=false (no FMA contraction; Rust never contracts), -prec-div=true and
// -prec-sqrt=true (nvcc defaults), results are bit-identical to the CPU oracle except for
// logf/acosf/atan2f (CUDA libm vs glibc, <= 2 ulp).  Reference citations are file:line.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/rtr.h"
#include "flat.h"
#include "kernels.h"

namespace rtr {

#define RTR_INF __int_as_float(0x7f800000)
#define RTR_PI 3.14159265358979323846f  // std::f32::consts::PI (types.rs:5)
#define RTR_STACK_CAP 96                // (ref, enter) pairs per thread; the host flattens every scene to fit (capi.cpp stack_limit)

// ------------------------------------------------------------------ vectors (cgmath 0.18 order)  [bucket:vecmath]
// struct v3 { float x, y, z; } comes from kernels.h
__device__ __forceinline__ v3 mk(float x, float y, float z) { v3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ v3 operator+(v3 a, v3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ v3 operator-(v3 a, v3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ v3 operator-(v3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ v3 operator*(v3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ v3 operator*(float s, v3 a) { return mk(s * a.x, s * a.y, s * a.z); }
__device__ __forceinline__ v3 operator*(v3 a, v3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ v3 operator/(v3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
__device__ __forceinline__ float dot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ v3 cross(v3 a, v3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float magnitude(v3 a) { return sqrtf(dot(a, a)); }
__device__ __forceinline__ v3 normalize(v3 a) { return a * (1.0f / magnitude(a)); }
__device__ __forceinline__ v3 xyz(const float4& q) { return mk(q.x, q.y, q.z); }

struct Ray { v3 o, d; };
__device__ __forceinline__ v3 ray_at(const Ray& r, float t) { return r.o + r.d * t; }  // ray.rs:15-17

// 128-bit global loads, one LDG.E.128 each.  Written as volatile inline PTX for two reasons seen in the SASS:
//  * through a by-value struct the compiler treats scene pointers as GENERIC and, worse, splits a float4
//    load whose components are used under different conditions into up to four narrower LD.E (10 load
//    instructions per 64-B node instead of 4) — and the kernels are bound by L1 wavefronts (ncu:
//    l1tex__data_pipe_lsu_wavefronts 73 % of peak), where every extra divergent load instruction costs
//    one wavefront per distinct cache line;
//  * with ld.global.nc (__ldg) nvcc 12.9 hoisted a leaf load above the ref-kind test that guards it
//    (reading spheres[node index], out of bounds); a volatile asm is never speculated.
__device__ __forceinline__ float4 ldg4(const F4* p) {
    float4 v;
    asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
// 256-bit global load (LDG.E.ENL2.256, new with sm_100): one instruction fetches half a 64-B node / triangle
// record, halving the L1 wavefronts of a divergent fetch again.  p must be 32-byte aligned.
__device__ __forceinline__ void ldg8(const F4* p, float4& a, float4& b) {
    asm volatile("ld.global.v8.f32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(p));
}
__device__ __forceinline__ uint4 ldgu4(const U4* p) {
    uint4 v;
    asm volatile("ld.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}

// ------------------------------------------------------------------ Philox4x32-10, keyed draws  [bucket:rng]
__device__ __forceinline__ uint4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                               uint32_t k0, uint32_t k1) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        uint32_t h0 = __umulhi(M0, c0), l0 = M0 * c0;
        uint32_t h1 = __umulhi(M1, c2), l1 = M1 * c2;
        uint32_t n0 = h1 ^ c1 ^ k0, n2 = h0 ^ c3 ^ k1;
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
        k0 += W0; k1 += W1;
    }
    return make_uint4(c0, c1, c2, c3);
}

enum { P_JITTER = 0, P_SPHERE = 1, P_MIX = 2, P_DIELECTRIC = 3, P_VOLUME = 4 };

struct RngKey { uint32_t k0, k1, pixel, sample, bounce; };
__device__ __forceinline__ uint4 draw(const RngKey& k, uint32_t purpose, uint32_t iter) {
    return philox4x32_10(k.pixel, k.sample, (k.bounce & 0xFFu) | (purpose << 8), iter, k.k0, k.k1);
}
// rand 0.8 UniformFloat<f32>: 23 mantissa bits -> [0,1)
__device__ __forceinline__ float u01(uint32_t b) { return (float)(b >> 9) * (1.0f / 8388608.0f); }
__device__ __forceinline__ float gen_range_01(uint32_t b) { return u01(b); }
__device__ __forceinline__ float gen_range_m11(uint32_t b) { return u01(b) * 2.0f + -1.0f; }
__device__ __forceinline__ float gen_range_01_inclusive(uint32_t b) {
    const float scale = 1.0f / (1.0f - 1.0f / 8388608.0f);
    return u01(b) * scale + 0.0f;
}
// sample.rs:44-52
__device__ __forceinline__ v3 sample_unit_sphere_volume(const RngKey& key) {
    for (uint32_t k = 0;; ++k) {
        uint4 r = draw(key, P_SPHERE, k);
        v3 v = mk(gen_range_m11(r.x), gen_range_m11(r.y), gen_range_m11(r.z));
        if (dot(v, v) <= 1.0f) return v;
    }
}
__device__ __forceinline__ v3 sample_unit_sphere_surface(const RngKey& key) {  // sample.rs:66-70
    return normalize(sample_unit_sphere_volume(key));
}

// ------------------------------------------------------------------ work counters (optional)  [bucket:other]
struct Counters { uint32_t nodes, tris, spheres, env; };

// ------------------------------------------------------------------ AABB::hit (aabb.rs:81-137)  [bucket:slab]
// Same t0/t1/swap/update sequence; the per-axis early-outs are folded into the final test, which
// is equivalent because tmin only grows and tmax only shrinks.  `if (t0 > tmin) tmin = t0` is written
// fmaxf(tmin, t0): identical for every input (a NaN t0 — 0 * inf — leaves tmin unchanged in both forms;
// tmin/tmax themselves are never NaN) and one FMNMX instead of FSETP + FSEL.  inv = 1/d is hoisted per ray
// (same IEEE quotient the reference recomputes at every node).  Returns pierced; enter = final tmin.
struct RayInv { float ix, iy, iz; };
__device__ __forceinline__ bool slab(const float4& mn, const float4& mx, const Ray& r, const RayInv& inv,
                                     float tmin, float tmax, float& enter) {
    float t0 = (mn.x - r.o.x) * inv.ix, t1 = (mx.x - r.o.x) * inv.ix;
    if (inv.ix < 0.0f) { float s = t0; t0 = t1; t1 = s; }
    tmin = fmaxf(tmin, t0);
    tmax = fminf(tmax, t1);
    t0 = (mn.y - r.o.y) * inv.iy; t1 = (mx.y - r.o.y) * inv.iy;
    if (inv.iy < 0.0f) { float s = t0; t0 = t1; t1 = s; }
    tmin = fmaxf(tmin, t0);
    tmax = fminf(tmax, t1);
    t0 = (mn.z - r.o.z) * inv.iz; t1 = (mx.z - r.o.z) * inv.iz;
    if (inv.iz < 0.0f) { float s = t0; t0 = t1; t1 = s; }
    tmin = fmaxf(tmin, t0);
    tmax = fminf(tmax, t1);
    enter = tmin;
    return !(tmax <= tmin);
}

// ------------------------------------------------------------------ primitives  [bucket:primitive_tests]
// Sphere::hit (objects.rs:23-50), t only; point/normal are rebuilt for the winner in finish_hit.
// Three shortcuts keep the reference's result bit for bit while skipping its IEEE sqrt / divisions (each ~10 SASS
// instructions, and a ~45-instruction slow path whenever the numerator is exactly 0 — ncu on benchmark.json: 7 % of all
// warp-instructions, because a ray leaving the r=100 ground sphere has c == 0 exactly, so -half_b -/+ sqrtd == 0):
//  (1) origin outside and moving away (c > 0, half_b > 0): fl(a*c) >= 0 so disc <= fl(half_b^2) and, sqrt being
//      monotonic with sqrt(fl(x^2)) == |x|, sqrtd <= half_b: both numerators are <= 0, both roots <= 0 < tmin -> miss;
//  (2) origin inside (c < 0): disc >= fl(half_b^2), sqrtd >= |half_b|, the first numerator is <= 0 -> the first
//      root fails `root < tmin` without being computed;
//  (3) a zero numerator over a > 0 is that same signed zero.
// All three need tmin > 0, a > 0, a non-NaN discriminant and half_b^2 neither overflowed nor denormal (the
// sqrt(fl(x^2)) == |x| identity); otherwise the plain sequence runs (tests/test_sphere_shortcuts.py sweeps both forms).
// SHORTCUTS is on in the linear scan (scan.cu: +3.2 % on benchmark.json, where rays leaving the ground sphere meet
// cases 1-3 at every bounce) and off in the BVH walks, whose candidates have already passed a box test: there the
// extra compares and branches cost 4 % (simple.json, 1 M spheres; same-box A/B).
template <bool SHORTCUTS>
__device__ __forceinline__ float div_or_zero(float num, float a) {
    if (!SHORTCUTS) return num / a;
    float q = num;
    if (!(num == 0.0f && a > 0.0f)) {
        // the volatile move pins the division inside this branch: written as a select, nvcc divides
        // unconditionally (FCHK + slow-path call included) and picks the result afterwards
        float n;
        asm volatile("mov.f32 %0, %1;" : "=f"(n) : "f"(num));
        q = n / a;
    }
    return q;
}
template <bool SHORTCUTS = false>
__device__ __forceinline__ bool sphere_t(const float4& s, const Ray& r, float tmin, float tmax, float& t) {
    v3 oc = r.o - xyz(s);
    float a = dot(r.d, r.d);
    float half_b = dot(oc, r.d);
    float c = dot(oc, oc) - s.w * s.w;
    float hb2 = half_b * half_b;
    float disc = hb2 - a * c;
    if (disc < 0.0f) return false;
    const bool tame = SHORTCUTS && disc >= 0.0f && a > 0.0f && tmin > 0.0f && hb2 >= 1.17549435e-38f && hb2 < RTR_INF;
    if (tame && c > 0.0f && half_b > 0.0f) return false;
    float sqrtd = sqrtf(disc);
    bool second = tame && c < 0.0f;
    float root = 0.0f;
    if (!second) {
        root = div_or_zero<SHORTCUTS>(-half_b - sqrtd, a);
        second = root < tmin || tmax < root;
    }
    if (second) {
        root = div_or_zero<SHORTCUTS>(-half_b + sqrtd, a);
        if (root < tmin || tmax < root) return false;
    }
    t = root;
    return true;
}

// Triangle::hit (mesh.rs:147-196): tmin doubles as the determinant epsilon (Q4); tmax ignored
__device__ __forceinline__ bool triangle_t(const float4& q0, const float4& q1, const float4& q2, const Ray& r,
                                           float tmin, float& t, float& u, float& v) {
    v3 e1 = xyz(q1), e2 = xyz(q2);
    v3 h = cross(r.d, e2);
    float a = dot(e1, h);
    if (a > -tmin && a < tmin) return false;
    float f = 1.0f / a;
    v3 s = r.o - xyz(q0);
    u = f * dot(s, h);
    if (u < 0.0f || u > 1.0f) return false;
    v3 q = cross(s, e1);
    v = f * dot(r.d, q);
    if (v < 0.0f || u + v > 1.0f) return false;
    t = f * dot(e2, q);
    return t > tmin;
}

struct LiteHit {
    float t;
    uint32_t ref;  // REF_EMPTY_VALUE = miss
    float u, v;    // barycentrics of a triangle winner
};

// candidate update: min t; among equal t the larger DFS leaf rank wins (bvh.rs:183-187, Q2)
__device__ __forceinline__ uint32_t leaf_rank(const SceneView& sc, uint32_t ref) {
    uint32_t k = RTR_REF_KIND(ref), i = RTR_REF_INDEX(ref);
    if (k == REF_SPHERE) return sc.sphere_info[i].w;
    if (k == REF_TRI) return __float_as_uint(sc.tris[(size_t)i * 4 + 3].w);
    return sc.volumes[i].rank;
}
__device__ __forceinline__ void offer(const SceneView& sc, LiteHit& best, float t, uint32_t ref, float u, float v) {
    bool take = t < best.t;
    if (!take && t == best.t) take = best.ref == REF_EMPTY_VALUE || leaf_rank(sc, ref) > leaf_rank(sc, best.ref);
    if (take) { best.t = t; best.ref = ref; best.u = u; best.v = v; }
}

// t-culling threshold.  [bucket:stack_pop]  A subtree is skipped when its box entry lies beyond the current best by
// more than this slack; the slack absorbs the few-ulp (sphere quadratic: worse) disagreement
// between a primitive's computed t and its box's computed entry.  RTR_FLAG_EXHAUSTIVE disables
// culling altogether (reference-shaped walk) and is used to validate the culled walk at scale.
__device__ __forceinline__ float cull_limit(float best_t) {
    return best_t + fmaxf(fabsf(best_t) * (1.0f / 256.0f), 1.0e-6f);
}

struct StackEntry { uint32_t ref; float enter; };

// Traversal stacks.  LocalStack: per-thread array in local memory.  ShortStack: the first SHORT entries of every
// thread live in shared memory, laid out [entry][thread] so a warp's access is conflict-free and costs two
// wavefronts whatever the lanes' depths are (in local memory lanes at different depths hit different lines:
// ncu showed local traffic at 40 % of the L1 wavefronts of wf_extend); deeper entries spill to local memory.
struct LocalStack {
    StackEntry* p;
    __device__ __forceinline__ void put(int sp, uint32_t ref, float enter) const { p[sp].ref = ref; p[sp].enter = enter; }
    __device__ __forceinline__ StackEntry get(int sp) const { return p[sp]; }
    __device__ __forceinline__ StackEntry* nested(int sp) const { return p + sp; }  // scratch for a nested query
};
template <int SHORT, int STRIDE>
struct ShortStack {
    uint2* sh;        // &smem[0][threadIdx.x]
    StackEntry* loc;  // spill area
    __device__ __forceinline__ void put(int sp, uint32_t ref, float enter) const {
        if (sp < SHORT) sh[sp * STRIDE] = make_uint2(ref, __float_as_uint(enter));
        else { loc[sp - SHORT].ref = ref; loc[sp - SHORT].enter = enter; }
    }
    __device__ __forceinline__ StackEntry get(int sp) const {
        StackEntry e;
        if (sp < SHORT) { const uint2 v = sh[sp * STRIDE]; e.ref = v.x; e.enter = __uint_as_float(v.y); }
        else e = loc[sp - SHORT];
        return e;
    }
    __device__ __forceinline__ StackEntry* nested(int sp) const { return loc + (sp > SHORT ? sp - SHORT : 0); }
};

// Closest hit below `entry`; boxes are tested against (tmin, tmax), primitives as the reference does.  NESTED = a volume boundary / mesh sub-query:
// volume leaves cannot occur there.  Volume leaves of the outer walk are handled by the caller
// through `on_volume`.
// Volume::hit is force-inlined into its callers: as a real call it cost the walk kernels ~400 bytes of spills around
// the call site (C5: wavefront 1031 -> 1615, megakernel 697 -> 844 Mrays/s, same-box A/B), and inlined its nested walks
// can use the 256-bit loads too (ptxas 12.9 segfaults on a NON-inlined function holding ld.global.v8 that two kernels call).
#ifndef RTR_VOLUME_INLINE
#define RTR_VOLUME_INLINE __forceinline__
#endif
#ifndef RTR_NESTED_WIDE
#define RTR_NESTED_WIDE 1  // 256-bit loads in the nested (volume boundary) walks; set 0 together with RTR_VOLUME_INLINE=__noinline__
#endif
static __device__ RTR_VOLUME_INLINE float volume_hit(const SceneView& sc, const VolumeRec& vol, const Ray& r, const RayInv& inv,
                            float tmin, float tmax, const RngKey& key, StackEntry* stack, bool exhaustive,
                            Counters* cnt);

#define RTR_REF_DONE 0xDFFFFFFFu  // kind 6: traversal finished (never stored in a node)

// pop the next stack entry that survives t-culling
template <bool EXHAUSTIVE, typename STACK>
__device__ __forceinline__ uint32_t pop_next(const STACK& stack, int& sp, float best_t, int floor = 0) {
    const float lim = cull_limit(best_t);
    while (sp > floor) {  // floor > 0: a nested walk stacked on top of a suspended outer walk (wf_extend, volumes)
        --sp;
        const StackEntry e = stack.get(sp);
        if (EXHAUSTIVE || !(e.enter > lim)) return e.ref;
    }
    return RTR_REF_DONE;
}

// Traversal is split into two resumable steps over the state (cur, sp, best):  [bucket:node_step]
//   node_step  cur is an interior node: slab-test both children (one 64-B fetch), descend into the
//              nearer pierced child, push the farther, or pop
//   leaf_step  cur is a primitive: test it as the reference does, offer the hit, pop
// cur == RTR_REF_DONE: finished.  Callers schedule the steps: walk() (nested queries, batch kernel)
// runs the classic "while-while" order; the render kernels pick, per warp, whichever step has more
// lanes waiting and refill finished lanes in between (kernels.cu, wavefront.cu).
// WIDE: fetch with 256-bit loads (everywhere by default; RTR_NESTED_WIDE=0 narrows the nested walks again, which a
// non-inlined volume_hit needs: see RTR_VOLUME_INLINE above).
template <bool EXHAUSTIVE, bool COUNT, bool WIDE = true, typename STACK>
__device__ __forceinline__ void node_step(const SceneView& sc, uint32_t& cur, int& sp, const LiteHit& best, const Ray& r,
                                          const RayInv& inv, float tmin, float tmax, const STACK& stack, Counters* cnt,
                                          int floor = 0) {
    const F4* q = sc.nodes + (size_t)RTR_REF_INDEX(cur) * 4;
    float4 q0, q1, q2, q3;
    if (WIDE) { ldg8(q, q0, q1); ldg8(q + 2, q2, q3); }
    else { q0 = ldg4(q); q1 = ldg4(q + 1); q2 = ldg4(q + 2); q3 = ldg4(q + 3); }
    uint32_t lref = __float_as_uint(q0.w), rref = __float_as_uint(q1.w);
    float le = -RTR_INF, re = -RTR_INF;
    bool lh = lref != REF_EMPTY_VALUE, rh = rref != REF_EMPTY_VALUE;
    // traversal tree: every child slot carries a box (a leaf's slot holds its gate, flat.h); reference
    // topology: leaves carry none (bvh.rs:157-159) and are tested as soon as the parent's box was pierced
    // (a paired leaf stands for the 2-leaf Split it folds, so its box is that Split's and is tested in both modes)
    // Only the COUNT instantiations can walk the reference topology (capi.cpp selects them for RTR_FLAG_REFERENCE_TREE):
    // the others drop the per-child uniform branch.
    const bool boxed_leaves = COUNT ? sc.reference_tree == 0u : true;
    const uint32_t lk = RTR_REF_KIND(lref), rk = RTR_REF_KIND(rref);
    if (lh && (boxed_leaves || lk == REF_NODE || lk == REF_TRI2 || lk == REF_SPHERE2)) { lh = slab(q0, q1, r, inv, tmin, tmax, le); if (COUNT) cnt->nodes++; }
    if (rh && (boxed_leaves || rk == REF_NODE || rk == REF_TRI2 || rk == REF_SPHERE2)) { rh = slab(q2, q3, r, inv, tmin, tmax, re); if (COUNT) cnt->nodes++; }
    if (!EXHAUSTIVE) {
        const float lim = cull_limit(best.t);
        lh = lh && !(le > lim);
        rh = rh && !(re > lim);
    }
    if (lh && rh) {
        if (re < le) { uint32_t tr = lref; lref = rref; rref = tr; float te = le; le = re; re = te; }
        stack.put(sp, rref, re); ++sp;
        cur = lref;
    } else if (lh) {
        cur = lref;
    } else if (rh) {
        cur = rref;
    } else {
        cur = pop_next<EXHAUSTIVE>(stack, sp, best.t, floor);
    }
}

// 4-wide node step (flat.h: nodes4).  [bucket:node_step]  One 128-B fetch, four slab tests — each against the very box the
// 2-wide traversal tree tests for that slot, so the set of leaves reached is the same — then the pierced children are
// ordered by entry distance with a 5-exchange network: the nearest becomes `cur`, the others go onto the stack far to
// near.  Visiting order is a speed matter only (offer() breaks ties by DFS rank, culling keeps its slack).
__device__ __forceinline__ void order2(float& ea, uint32_t& ra, float& eb, uint32_t& rb) {
    const bool sw = eb < ea;
    const float e0 = sw ? eb : ea, e1 = sw ? ea : eb;
    const uint32_t r0 = sw ? rb : ra, r1 = sw ? ra : rb;
    ea = e0; eb = e1; ra = r0; rb = r1;
}
template <bool COUNT, typename STACK>
__device__ __forceinline__ void node4_step(const SceneView& sc, uint32_t& cur, int& sp, const LiteHit& best, const Ray& r,
                                           const RayInv& inv, float tmin, float tmax, const STACK& stack, Counters* cnt, int floor = 0) {
    const F4* q = sc.nodes4 + (size_t)RTR_REF_INDEX(cur) * 8;
    float4 q0, q1, q2, q3, q4, q5, q6, q7;
    ldg8(q, q0, q1); ldg8(q + 2, q2, q3); ldg8(q + 4, q4, q5); ldg8(q + 6, q6, q7);
    uint32_t r0 = __float_as_uint(q0.w), r1 = __float_as_uint(q1.w), r2 = __float_as_uint(q2.w), r3 = __float_as_uint(q3.w);
    if (COUNT) cnt->nodes += (r0 != REF_EMPTY_VALUE) + (r1 != REF_EMPTY_VALUE) + (r2 != REF_EMPTY_VALUE) + (r3 != REF_EMPTY_VALUE);  // boxes tested
    const float lim = cull_limit(best.t);
    float e0, e1, e2, e3;
    const bool h0 = slab(q0, q1, r, inv, tmin, tmax, e0) && r0 != REF_EMPTY_VALUE && !(e0 > lim);
    const bool h1 = slab(q2, q3, r, inv, tmin, tmax, e1) && r1 != REF_EMPTY_VALUE && !(e1 > lim);
    const bool h2 = slab(q4, q5, r, inv, tmin, tmax, e2) && r2 != REF_EMPTY_VALUE && !(e2 > lim);
    const bool h3 = slab(q6, q7, r, inv, tmin, tmax, e3) && r3 != REF_EMPTY_VALUE && !(e3 > lim);
    const int n = (int)h0 + (int)h1 + (int)h2 + (int)h3;
    if (n == 0) { cur = pop_next<false>(stack, sp, best.t, floor); return; }
    // a pierced box has a finite entry (or -inf under a -inf tmin): +inf sorts the others behind every pierced one
    e0 = h0 ? e0 : RTR_INF; e1 = h1 ? e1 : RTR_INF; e2 = h2 ? e2 : RTR_INF; e3 = h3 ? e3 : RTR_INF;
    order2(e0, r0, e1, r1); order2(e2, r2, e3, r3); order2(e0, r0, e2, r2); order2(e1, r1, e3, r3); order2(e1, r1, e2, r2);
    if (n > 3) { stack.put(sp, r3, e3); ++sp; }
    if (n > 2) { stack.put(sp, r2, e2); ++sp; }
    if (n > 1) { stack.put(sp, r1, e1); ++sp; }
    cur = r0;
}

// [bucket:leaf_step]
// the primitive test of a leaf step, without the pop: `cur` is read only
template <bool EXHAUSTIVE, bool COUNT, bool NESTED, bool VOLUMES = true, typename STACK>
__device__ __forceinline__ void leaf_test(const SceneView& sc, const uint32_t cur, const int sp, LiteHit& best, const Ray& r,
                                          const RayInv& inv, float tmin, const RngKey& key, const STACK& stack,
                                          Counters* cnt) {
    const uint32_t kind = RTR_REF_KIND(cur);
    // A single leaf and a paired leaf share their first test: lanes holding either kind run it together, the second
    // primitive of a pair follows (handled as separate branches the two kinds serialised: three tests' worth of
    // instructions per leaf step instead of two).
    if (kind == REF_TRI || kind == REF_TRI2) {
        const uint32_t i = RTR_REF_INDEX(cur);
        const F4* q = sc.tris + (size_t)i * 4;
        float4 a0, a1;
        if (!NESTED || RTR_NESTED_WIDE) ldg8(q, a0, a1);
        else { a0 = ldg4(q); a1 = ldg4(q + 1); }
        const float4 a2 = ldg4(q + 2);
        const bool pair = kind == REF_TRI2;  // the sibling behind the same gate (traversal tree)
        float4 b0 = a0, b1 = a1, b2 = a2;
        if (pair) {  // fetched before the first test runs, so the two records' latencies overlap
            if (!NESTED || RTR_NESTED_WIDE) ldg8(q + 4, b0, b1);
            else { b0 = ldg4(q + 4); b1 = ldg4(q + 5); }
            b2 = ldg4(q + 6);
        }
        float t, u, v;
        if (COUNT) cnt->tris++;
        if (triangle_t(a0, a1, a2, r, tmin, t, u, v)) offer(sc, best, t, RTR_MAKE_REF(REF_TRI, i), u, v);
        if (pair) {
            if (COUNT) cnt->tris++;
            if (triangle_t(b0, b1, b2, r, tmin, t, u, v)) offer(sc, best, t, RTR_MAKE_REF(REF_TRI, i + 1), u, v);
        }
    } else if (kind == REF_SPHERE || kind == REF_SPHERE2) {
        const uint32_t i = RTR_REF_INDEX(cur);
        const bool pair = kind == REF_SPHERE2;
        const float4 s0 = ldg4(sc.spheres + i);
        float4 s1 = s0;
        if (pair) s1 = ldg4(sc.spheres + i + 1);
        float t;
        if (COUNT) cnt->spheres++;
        if (sphere_t(s0, r, tmin, RTR_INF, t)) offer(sc, best, t, RTR_MAKE_REF(REF_SPHERE, i), 0.0f, 0.0f);
        if (pair) {
            if (COUNT) cnt->spheres++;
            if (sphere_t(s1, r, tmin, RTR_INF, t)) offer(sc, best, t, RTR_MAKE_REF(REF_SPHERE, i + 1), 0.0f, 0.0f);
        }
    } else if (VOLUMES && !NESTED && kind == REF_VOLUME) {
        const VolumeRec& vol = sc.volumes[RTR_REF_INDEX(cur)];
        float t = volume_hit(sc, vol, r, inv, tmin, RTR_INF, key, stack.nested(sp), EXHAUSTIVE, COUNT ? cnt : nullptr);
        if (t == t) offer(sc, best, t, cur, 0.0f, 0.0f);
    }
}
template <bool EXHAUSTIVE, bool COUNT, bool NESTED, bool VOLUMES = true, typename STACK>
__device__ __forceinline__ void leaf_step(const SceneView& sc, uint32_t& cur, int& sp, LiteHit& best, const Ray& r,
                                          const RayInv& inv, float tmin, const RngKey& key, const STACK& stack,
                                          Counters* cnt, int floor = 0) {
    leaf_test<EXHAUSTIVE, COUNT, NESTED, VOLUMES>(sc, cur, sp, best, r, inv, tmin, key, stack, cnt);
    cur = pop_next<EXHAUSTIVE>(stack, sp, best.t, floor);
}

// one "while-while" round: all lanes descend interior nodes together, then test their pending leaf together  [bucket:nested_walk]
template <bool EXHAUSTIVE, bool COUNT, bool NESTED>
__device__ __forceinline__ void walk_round(const SceneView& sc, uint32_t& cur, int& sp, LiteHit& best, const Ray& r,
                                           const RayInv& inv, float tmin, float tmax, const RngKey& key,
                                           const LocalStack& stack, Counters* cnt) {
    while (RTR_REF_KIND(cur) == REF_NODE) node_step<EXHAUSTIVE, COUNT, !NESTED || RTR_NESTED_WIDE>(sc, cur, sp, best, r, inv, tmin, tmax, stack, cnt);
    if (cur == RTR_REF_DONE) return;
    leaf_step<EXHAUSTIVE, COUNT, NESTED>(sc, cur, sp, best, r, inv, tmin, key, stack, cnt);
}

__device__ __forceinline__ uint32_t walk_entry(uint32_t entry) { return entry == REF_EMPTY_VALUE ? RTR_REF_DONE : entry; }

template <bool EXHAUSTIVE, bool COUNT, bool NESTED>
__device__ __forceinline__ void walk(const SceneView& sc, uint32_t entry, const Ray& r, const RayInv& inv,
                                     float tmin, float tmax, const RngKey& key, StackEntry* stack, LiteHit& best,
                                     Counters* cnt) {
    uint32_t cur = walk_entry(entry);
    int sp = 0;
    LocalStack ls; ls.p = stack;
    while (cur != RTR_REF_DONE) walk_round<EXHAUSTIVE, COUNT, NESTED>(sc, cur, sp, best, r, inv, tmin, tmax, key, ls, cnt);
}

// boundary.hit(ray, tmin, +INF) of a volume boundary: only t is consumed (volume.rs:36-37)
__device__ __forceinline__ bool boundary_t(const SceneView& sc, uint32_t entry, const Ray& r, const RayInv& inv,
                                           float tmin, StackEntry* stack, bool exhaustive, Counters* cnt, float& t) {
    LiteHit best; best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
    RngKey nokey; nokey.k0 = nokey.k1 = nokey.pixel = nokey.sample = nokey.bounce = 0;
    Counters dummy; dummy.nodes = dummy.tris = dummy.spheres = dummy.env = 0;
    Counters* c = cnt ? cnt : &dummy;
    if (exhaustive) walk<true, true, true>(sc, entry, r, inv, tmin, RTR_INF, nokey, stack, best, c);
    else walk<false, true, true>(sc, entry, r, inv, tmin, RTR_INF, nokey, stack, best, c);
    t = best.t;
    return best.ref != REF_EMPTY_VALUE;
}

// The rest of Volume::hit once both boundary crossings t1, t2 are known (volume.rs:39-72).  Returns t, or NaN for None.
__device__ __forceinline__ float volume_sample(const VolumeRec& vol, const Ray& r, float t1, float t2, float tmin, float tmax,
                                               const RngKey& key) {
    const float NONE = __int_as_float(0x7fc00000);
    if (t1 < tmin) t1 = tmin;
    if (t2 > tmax) t2 = tmax;
    if (t1 >= t2) return NONE;
    if (t1 < 0.0f) t1 = 0.0f;
    float ray_length = magnitude(r.d);
    float distance_inside_boundary = (t2 - t1) * ray_length;
    float random_float = gen_range_01_inclusive(draw(key, P_VOLUME, vol.index).x);  // volume.rs:57
    float hit_distance = vol.neg_inv_density * logf(random_float);
    if (hit_distance > distance_inside_boundary) return NONE;
    return t1 + hit_distance / ray_length;
}

// Volume::hit (volume.rs:35-72).  Returns t, or NaN for None.
static __device__ RTR_VOLUME_INLINE float volume_hit(const SceneView& sc, const VolumeRec& vol, const Ray& r, const RayInv& inv,
                                         float tmin, float tmax, const RngKey& key, StackEntry* stack,
                                         bool exhaustive, Counters* cnt) {
    const float NONE = __int_as_float(0x7fc00000);
    float t1, t2;
    const uint32_t entry = sc.reference_tree ? vol.entry_reference : vol.entry;
    if (!boundary_t(sc, entry, r, inv, -RTR_INF, stack, exhaustive, cnt, t1)) return NONE;
    if (!boundary_t(sc, entry, r, inv, t1 + tmin, stack, exhaustive, cnt, t2)) return NONE;
    return volume_sample(vol, r, t1, t2, tmin, tmax, key);
}

// world.hit(ray, tmin, +INF) with the BVH (bvh.rs:47-49) or the linear scan (scene.rs:141-156)
template <bool EXHAUSTIVE, bool COUNT>
__device__ __forceinline__ LiteHit world_hit(const SceneView& sc, bool bvh, const Ray& r, float tmin,
                                             const RngKey& key, StackEntry* stack, Counters* cnt) {
    LiteHit best; best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
    RayInv inv; inv.ix = inv.iy = inv.iz = 0.0f;
    if (bvh || sc.linear_needs_inv) { inv.ix = 1.0f / r.d.x; inv.iy = 1.0f / r.d.y; inv.iz = 1.0f / r.d.z; }
    if (bvh) {
        walk<EXHAUSTIVE, COUNT, false>(sc, sc.root_entry, r, inv, tmin, RTR_INF, key, stack, best, cnt);
        return best;
    }
    float closest = RTR_INF;  // closest_so_far (scene.rs:144)
    if (sc.linear_spheres != nullptr) {  // all spheres: same scan, same order, without the per-object dispatch
        uint32_t winner = 0xFFFFFFFFu;
        for (uint32_t i = 0; i < sc.n_objects; ++i) {
            float t;
            if (COUNT) cnt->spheres++;
            if (sphere_t(ldg4(sc.linear_spheres + i), r, tmin, closest, t)) { closest = t; winner = i; }
        }
        if (winner != 0xFFFFFFFFu) { best.t = closest; best.ref = sc.objects[winner].entry; }
        return best;
    }
    for (uint32_t i = 0; i < sc.n_objects; ++i) {
        const uint32_t entry = sc.reference_tree ? sc.objects[i].entry_reference : sc.objects[i].entry;
        const uint32_t kind = RTR_REF_KIND(entry);
        if (kind == REF_SPHERE) {
            float t;
            if (COUNT) cnt->spheres++;
            if (sphere_t(ldg4(sc.spheres + RTR_REF_INDEX(entry)), r, tmin, closest, t)) {
                closest = t; best.t = t; best.ref = entry;
            }
        } else if (kind == REF_VOLUME) {
            float t = volume_hit(sc, sc.volumes[RTR_REF_INDEX(entry)], r, inv, tmin, closest, key, stack, EXHAUSTIVE,
                                 COUNT ? cnt : nullptr);
            if (t == t) { closest = t; best.t = t; best.ref = entry; }
        } else if (kind != REF_EMPTY) {
            // mesh: the boxes honour tmax = closest_so_far (bvh.rs:161) but Triangle::hit ignores it
            // (mesh.rs:151), and any hit returned unconditionally replaces the result (Q9)
            LiteHit m; m.t = RTR_INF; m.ref = REF_EMPTY_VALUE; m.u = m.v = 0.0f;
            walk<EXHAUSTIVE, COUNT, true>(sc, entry, r, inv, tmin, closest, key, stack, m, cnt);
            if (m.ref != REF_EMPTY_VALUE) { closest = m.t; best = m; }
        }
    }
    return best;
}

// ------------------------------------------------------------------ HitRecord for the winner  [bucket:finish_hit]
struct Hit {
    v3 point, normal;
    float t;
    bool front_face;
    uint32_t material, object, primitive;
};

// Rebuilds what the reference's hit() returned for the winning primitive:
// objects.rs:45-50, mesh.rs:184-192, volume.rs:66-71, then HitRecord::new (ray.rs:39-54)
__device__ __forceinline__ Hit finish_hit(const SceneView& sc, const Ray& r, const LiteHit& lh) {
    Hit h;
    const uint32_t kind = RTR_REF_KIND(lh.ref), i = RTR_REF_INDEX(lh.ref);
    v3 n;
    h.t = lh.t;
    h.point = ray_at(r, lh.t);
    if (kind == REF_SPHERE) {
        const float4 s = ldg4(sc.spheres + i);
        const uint4 info = ldgu4(sc.sphere_info + i);
        n = (h.point - xyz(s)) / s.w;
        h.object = info.x; h.primitive = 0; h.material = info.z;
    } else if (kind == REF_TRI) {
        const F4* q = sc.tris + (size_t)i * 4;
        const F4* vn = sc.tri_vn + (size_t)i * 3;
        const v3 fn = mk(q[0].w, q[1].w, q[2].w);
        const v3 n0 = xyz(ldg4(vn)), n1 = xyz(ldg4(vn + 1)), n2 = xyz(ldg4(vn + 2));
        const uint4 info = ldgu4(reinterpret_cast<const U4*>(q + 3));
        n = lh.u * n1 + lh.v * n2 + (1.0f - lh.u - lh.v) * n0;  // mesh.rs:185
        // `normal.angle(face).0 > 0` (mesh.rs:187) with angle = atan2(|n x f|, n.f):
        // atan2(y>=0, x) > 0  <=>  y > 0 || (y == 0 && x negative (incl. -0)); NaN -> false
        const float y = magnitude(cross(n, fn)), x = dot(n, fn);
        if (y > 0.0f || (y == 0.0f && (__float_as_uint(x) >> 31) && x == x)) n = fn;
        h.object = info.x; h.primitive = info.y; h.material = info.z;
    } else {
        const VolumeRec& vol = sc.volumes[i];
        n = mk(1.0f, 0.0f, 0.0f);
        h.object = vol.object; h.primitive = 0; h.material = vol.material;
    }
    h.front_face = dot(r.d, n) < 0.0f;
    h.normal = h.front_face ? n : -n;
    return h;
}

// ------------------------------------------------------------------ materials  [bucket:scatter]
__device__ __forceinline__ bool near_zero(v3 v) {  // math.rs:6-9
    const float EPS = 1.0e-8f;
    return fabsf(v.x) < EPS && fabsf(v.y) < EPS && fabsf(v.z) < EPS;
}
__device__ __forceinline__ v3 reflect(v3 v, v3 n) { return v - 2.0f * dot(v, n) * n; }  // math.rs:12-14
__device__ __forceinline__ v3 refract(v3 uv, v3 n, float eta) {                         // math.rs:16-21
    float cos_theta = fminf(dot(n, -uv), 1.0f);
    v3 out_perp = eta * (uv + cos_theta * n);
    v3 out_parallel = -sqrtf(fabsf(1.0f - dot(out_perp, out_perp))) * n;
    return out_perp + out_parallel;
}
__device__ __forceinline__ float powi5(float a) { float a2 = a * a; float a4 = a2 * a2; return a * a4; }
__device__ __forceinline__ float reflectance(float cosine, float idx) {  // materials.rs:103-108
    float r0 = (1.0f - idx) / (1.0f + idx);
    r0 = r0 * r0;
    return r0 + (1.0f - r0) * powi5(1.0f - cosine);
}

struct Scatter { v3 attenuation; bool some; Ray ray; };

// Material::scatter for Scene.material(handle) (mode Full, render.rs:40) or the fresh Lambertian(0.8) of mode Clay
// (render.rs:41), in three parts so that callers can put the random draws where the warp is converged:
//   scatter_prepare   resolve MixMaterial (materials.rs:150-156), Dielectric prologue (:114-124), say which draw is needed
//   (draw)            unit-ball rejection loop for Lambertian / Metal / Isotropic (sample.rs:44-52), one uniform for a
//                     Dielectric that can refract (no draw under total internal reflection, Q11)
//   scatter_finish    the per-material arithmetic, a few instructions each
// scatter_event() strings them together with the plain sequential loop (megakernel refill, wf_shade, scatter_batch);
// scan.cu runs the draw as ONE site for the whole warp followed by the warp-cooperative tail (coop_ball_tail).  The
// same arrangement in wf_shade measured neutral (+1.3 % C4, -1.3 % C1: that kernel waits on memory, not on issue slots).
struct ScatterPrep {
    uint32_t kind;
    v3 color;
    float param, ratio, cos_theta;
    v3 unit;
    bool cannot_refract, want_ball, want_uniform;
};
__device__ __forceinline__ ScatterPrep scatter_prepare(const SceneView& sc, bool clay, uint32_t material, const Ray& ray,
                                                       const Hit& hit, const RngKey& key) {
    ScatterPrep p;
    p.kind = RTR_MAT_LAMBERTIAN;
    p.color = mk(0.8f, 0.8f, 0.8f);
    p.param = 0.0f;
    if (!clay) {
        MaterialRec m = sc.materials[material];
        for (uint32_t level = 0; m.kind == RTR_MAT_MIX; ++level) {  // MixMaterial (materials.rs:150-156)
            const bool first = gen_range_01(draw(key, P_MIX, level).x) >= m.param;
            m = sc.materials[first ? m.first : m.second];
        }
        p.kind = m.kind; p.color = mk(m.r, m.g, m.b); p.param = m.param;
    }
    // Dielectric prologue (materials.rs:114-124)
    p.ratio = 1.0f; p.cos_theta = 0.0f;
    p.unit = mk(0.0f, 0.0f, 0.0f);
    p.cannot_refract = false;
    // one normalize site for the Dielectric's unit direction (:120) and the Metal's unit reflection (:59)
    if (p.kind == RTR_MAT_DIELECTRIC || p.kind == RTR_MAT_METAL)
        p.unit = normalize(p.kind == RTR_MAT_METAL ? reflect(ray.d, hit.normal) : ray.d);
    if (p.kind == RTR_MAT_DIELECTRIC) {
        p.ratio = hit.front_face ? 1.0f / p.param : p.param;
        p.cos_theta = fminf(dot(hit.normal, -p.unit), 1.0f);
        const float sin_theta = sqrtf(1.0f - p.cos_theta * p.cos_theta);
        p.cannot_refract = (p.ratio * sin_theta) > 1.0f;
    }
    p.want_ball = p.kind == RTR_MAT_LAMBERTIAN || p.kind == RTR_MAT_METAL || p.kind == RTR_MAT_ISOTROPIC;
    p.want_uniform = p.kind == RTR_MAT_DIELECTRIC && !p.cannot_refract;
    return p;
}
__device__ __forceinline__ Scatter scatter_finish(const ScatterPrep& p, v3 ball, float xi, const Ray& ray, const Hit& hit) {
    Scatter s;
    s.ray.o = hit.point;
    s.ray.d = mk(0.0f, 0.0f, 0.0f);
    s.attenuation = p.color;
    s.some = true;
    v3 nball = ball;  // unit_sphere_surface() of Lambertian and Metal (sample.rs:66-70): one normalize site for both
    if (p.kind == RTR_MAT_LAMBERTIAN || p.kind == RTR_MAT_METAL) nball = normalize(ball);
    if (p.kind == RTR_MAT_LAMBERTIAN) {  // materials.rs:31-41
        v3 dir = hit.normal + nball;
        if (near_zero(dir)) dir = hit.normal;
        s.ray.d = dir;
    } else if (p.kind == RTR_MAT_METAL) {  // :58-70 (p.unit = the unit reflection, scatter_prepare)
        s.ray.d = p.unit + p.param * nball;
        s.some = dot(s.ray.d, hit.normal) > 0.0f;
        if (!s.some) s.attenuation = mk(0.0f, 0.0f, 0.0f);
    } else if (p.kind == RTR_MAT_DIELECTRIC) {  // :125-138
        const bool refl = p.cannot_refract || reflectance(p.cos_theta, p.ratio) > xi;
        s.ray.d = refl ? reflect(p.unit, hit.normal) : refract(p.unit, hit.normal, p.ratio);
        s.attenuation = mk(1.0f, 1.0f, 1.0f);
    } else if (p.kind == RTR_MAT_ISOTROPIC) {  // volume.rs:82-87
        s.ray.d = ball;
    } else {  // Emission (:88-90)
        s.some = false;
    }
    return s;
}
__device__ __forceinline__ Scatter scatter_event(const SceneView& sc, bool clay, uint32_t material, const Ray& ray,
                                                 const Hit& hit, const RngKey& key) {
    const ScatterPrep p = scatter_prepare(sc, clay, material, ray, hit, key);
    // the one draw site: lanes holding different materials still converge on the expensive part
    v3 ball = mk(0.0f, 0.0f, 0.0f);
    float xi = 0.0f;
    if (p.want_ball || p.want_uniform) {
        const uint32_t purpose = p.want_ball ? P_SPHERE : P_DIELECTRIC;
        for (uint32_t k = 0;; ++k) {
            const uint4 r = draw(key, purpose, k);
            if (!p.want_ball) { xi = gen_range_01(r.x); break; }
            ball = mk(gen_range_m11(r.x), gen_range_m11(r.y), gen_range_m11(r.z));
            if (dot(ball, ball) <= 1.0f) break;
        }
    }
    return scatter_finish(p, ball, xi, ray, hit);
}

// ------------------------------------------------------------------ warp-cooperative rejection tail  [bucket:rng]
// Philox4x32-10 with the key schedule taken from the kernel parameters: the ten round keys (k + i*W) are launch
// constants, so each round is 2 IMAD.WIDE + 2 LOP3 with a constant-bank operand (the in-kernel schedule costs two
// extra uniform adds per round, a quarter of the generator's instructions).
__device__ __forceinline__ uint4 philox_rk(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, const PhiloxKeys& pk) {
    const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u;
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const uint32_t h0 = __umulhi(M0, c0), l0 = M0 * c0;
        const uint32_t h1 = __umulhi(M1, c2), l1 = M1 * c2;
        const uint32_t n0 = h1 ^ c1 ^ pk.rk0[i], n2 = h0 ^ c3 ^ pk.rk1[i];
        c0 = n0; c1 = l1; c2 = n2; c3 = l0;
    }
    return make_uint4(c0, c1, c2, c3);
}
__device__ __forceinline__ uint32_t warp_lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}
// Continues the unit-ball rejection loop (sample.rs:44-52) of the lanes whose candidate 0 was rejected (`failed`;
// 47.6 % of them).  Counter-based RNG: any lane can compute candidate k of any lane's stream, so all 32 lanes serve
// the failing ones, `per` candidates each (the largest power of two with failing * per <= 32), and a failing lane
// takes the FIRST accepted k of its segment — exactly the sequential loop's result, in two or three full-width
// rounds instead of five or six thin ones.  Must be called by the whole warp; `tab` = 32 uint4 of shared memory per
// warp; (pixel, sample, depth) name the lane's stream (draw(key, P_SPHERE, k)).
__device__ __forceinline__ void coop_ball_tail(bool failed, uint32_t pixel, uint32_t sample, uint32_t depth, uint4* tab,
                                               const PhiloxKeys& pk, v3& ball) {
    const uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt = warp_lanemask_lt();
    uint32_t fail = __ballot_sync(FULL, failed);
    uint32_t kbase = 1u;
    while (fail != 0u) {
        const uint32_t nf = (uint32_t)__popc(fail);
        const uint32_t shift = 5u - (nf <= 1u ? 0u : 32u - (uint32_t)__clz((int)(nf - 1u)));
        const uint32_t per = 1u << shift;
        const uint32_t my_j = (uint32_t)__popc(fail & lt);
        if (failed) tab[my_j] = make_uint4(pixel, sample, depth, 0u);
        __syncwarp();
        const uint32_t j = lane >> shift;
        const bool serving = j < nf;
        const uint4 served = tab[serving ? j : 0u];
        const uint4 r = philox_rk(served.x, served.y, (served.z & 0xFFu) | ((uint32_t)P_SPHERE << 8),
                                  kbase + (lane & (per - 1u)), pk);
        const v3 cand = mk(gen_range_m11(r.x), gen_range_m11(r.y), gen_range_m11(r.z));
        const uint32_t okm = __ballot_sync(FULL, serving && dot(cand, cand) <= 1.0f);
        const uint32_t seg_mask = per == 32u ? FULL : ((1u << per) - 1u);
        const uint32_t seg = failed ? (okm >> (my_j << shift)) & seg_mask : 0u;
        const uint32_t from = seg != 0u ? (my_j << shift) + (uint32_t)(__ffs((int)seg) - 1) : lane;
        const float bx = __shfl_sync(FULL, cand.x, from);
        const float by = __shfl_sync(FULL, cand.y, from);
        const float bz = __shfl_sync(FULL, cand.z, from);
        if (seg != 0u) { ball = mk(bx, by, bz); failed = false; }
        fail = __ballot_sync(FULL, failed);
        kbase += per;
        __syncwarp();
    }
}

// ------------------------------------------------------------------ backgrounds (backgrounds.rs)  [bucket:background]
// Rust `f32 as usize` saturates (NaN -> 0); __float2ull_rz saturates the same way, NaN -> 0x8000...
__device__ __forceinline__ unsigned long long f32_as_usize(float v) {
    return (v == v) ? __float2ull_rz(v) : 0ull;
}
__device__ __forceinline__ v3 background_sample(const SceneView& sc, const Ray& ray, Counters* cnt) {
    const BackgroundRec& bg = sc.background;
    if (bg.kind == RTR_BG_UNIFORM) return mk(bg.r, bg.g, bg.b);  // :27-29
    if (bg.kind == RTR_BG_GRADIENT) {                            // :46-48, ray.rs:19-26
        float t = 0.5f * (normalize(ray.d).y + 1.0f);
        return mk(bg.r, bg.g, bg.b) * (1.0f - t) + mk(bg.br, bg.bg, bg.bb) * t;
    }
    // SkyMap::sample :85-92, to_spherical_coords math.rs:24-29
    const v3 d = normalize(ray.d);
    const float sx = acosf(-d.y);
    const float sy = atan2f(-d.z, d.x) + RTR_PI;
    const float u = sx / RTR_PI;
    const float v = sy / (2.0f * RTR_PI);
    const unsigned long long W = bg.width, H = bg.height;
    const unsigned long long x = f32_as_usize(v * (float)W) % W;
    const unsigned long long y = H - 1ull - f32_as_usize(u * (float)H) % H;
    if (cnt) cnt->env++;
    const float4 texel = ldg4(sc.env + (x + y * W));
    return mk(texel.x, texel.y, texel.z);
}

// ------------------------------------------------------------------ camera (camera.rs:19-24)  [bucket:camera]
__device__ __forceinline__ Ray camera_sample_ray(const CameraBasis& cam, uint32_t W, uint32_t H, uint32_t x,
                                                 uint32_t y, const RngKey& key) {
    const uint4 j = draw(key, P_JITTER, 0);
    const float u = ((float)x + gen_range_01_inclusive(j.x)) / (float)(W - 1);  // render.rs:88-89
    const float v = ((float)y + gen_range_01_inclusive(j.y)) / (float)(H - 1);
    Ray r;
    r.o = cam.origin;
    r.d = cam.lower_left_corner + cam.horizontal * u - cam.vertical * v - cam.origin;
    return r;
}

// f32::clamp(0, hi): NaN stays NaN (render.rs:95-97)
__device__ __forceinline__ float f32_clamp(float v, float lo, float hi) {
    if (v < lo) return lo;
    if (v > hi) return hi;
    return v;
}
// Rust `f32 as u8`: saturating, NaN -> 0 (color.rs:34-41)
__device__ __forceinline__ uint8_t f32_as_u8(float v) {
    if (!(v == v) || v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)__float2uint_rz(v);
}

}  // namespace rtr
