// bvh_build.cu — device build of the reference BVH for a scene made of spheres.
//
// render() rebuilds the BVH on every call (lib/core/render.rs:67) and BvhNode::from_list (lib/core/bvh.rs:59-144)
// is O(N log^2 N) comparator calls on the host; at 1 M primitives the host rebuild dominates a render by two
// orders of magnitude.  This file reproduces the SAME tree on the GPU:
//
//  * from_list splits n handles into the first n/2 and the rest after a STABLE sort on the centroid of the axis
//    picked from the centroid box (bvh.rs:75-134).  The SHAPE of the tree therefore depends on N alone: the node
//    for positions [lo, hi) has children [lo, mid) and [mid, hi), mid = lo + (hi - lo)/2.  Only the permutation
//    of the primitives is data dependent.
//  * Level by level, every segment of the current level picks its axis (segmented min/max of centroids, strict
//    '>' rule), and ONE stable LSD radix sort of all positions on the 64-bit key (segment start << 32 | ordered
//    float) sorts every segment of the level at once (cub::DeviceRadixSort — a library primitive; everything
//    else here is hand written).  ceil(log2 N) levels.
//  * Boxes are unions (min/max are exact and associative, so any reduction order gives the reference's bits).
//
// The same builder serves a triangle mesh (mesh.build_bvh(), mesh.rs:119-121): its leaves are the triangles' boxes
// (mesh.rs:200-213) and its records land in index ranges the host flattener reserved (host_flatten.h DeviceMesh).
//
// Output = the flat.h layout, with every leaf slot carrying its gate box and 2-leaf Splits folded into paired
// leaves, so ONE node array serves both the default traversal and RTR_FLAG_REFERENCE_TREE (a paired leaf is just
// a compact encoding of the Split it replaces: one box test, then both primitives).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <cstring>
#include <vector>

#include "bvh_build.h"

namespace rtr {

namespace {

__device__ __forceinline__ uint32_t ordered(float f) {  // monotone float -> uint; -0 and +0 compare equal (partial_cmp)
    if (f == 0.0f) f = 0.0f;
    const uint32_t u = __float_as_uint(f);
    return u ^ ((u >> 31) ? 0xFFFFFFFFu : 0x80000000u);
}
__device__ __forceinline__ float unordered(uint32_t o) {
    const uint32_t u = (o >> 31) ? (o ^ 0x80000000u) : ~o;
    return __uint_as_float(u);
}

struct BuildArrays {
    // inputs, by object handle (Scene.objects order)
    const float4* sph;       // (center, radius)
    const uint32_t* mat;
    uint32_t n;
    // per handle
    float* cx; float* cy; float* cz;  // centroid of Sphere::bounds (objects.rs:53-60, aabb.rs:28-30)
    // per position
    uint32_t* perm; uint32_t* perm_alt;
    unsigned long long* key; unsigned long long* key_alt;
    uint32_t* lo; uint32_t* hi; uint32_t* idx;   // current segment of the position and that segment's node index
    // per node (pre-order index; a subtree with n leaves owns n-1 consecutive indices)
    uint32_t* cmin; uint32_t* cmax;              // 3 x n ordered-float centroid bounds
    uint32_t* node_lo; uint32_t* node_hi; uint32_t* node_depth;
    float4* bmin; float4* bmax;                  // node boxes
    float4* pmin; float4* pmax;                  // leaf boxes by position
    uint32_t* flags;                             // [0]: NaN centroid seen
    // what the leaves are and where this collection's records go (all zero for a sphere scene)
    uint32_t leaf_kind, pair_kind;               // REF_SPHERE / REF_SPHERE2 or REF_TRI / REF_TRI2
    uint32_t slot_base, node_base, node4_base;
    float4* hmin; float4* hmax;                  // triangle meshes: leaf boxes by handle
};

__global__ void k_init(BuildArrays a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float4 s = a.sph[i];
    // Sphere::bounds = center -/+ radius; AABB::centroid = (min + max) / 2.0
    const float x0 = s.x - s.w, x1 = s.x + s.w, y0 = s.y - s.w, y1 = s.y + s.w, z0 = s.z - s.w, z1 = s.z + s.w;
    const float cx = (x0 + x1) / 2.0f, cy = (y0 + y1) / 2.0f, cz = (z0 + z1) / 2.0f;
    a.cx[i] = cx; a.cy[i] = cy; a.cz[i] = cz;
    a.perm[i] = i; a.lo[i] = 0u; a.hi[i] = a.n; a.idx[i] = 0u;
    for (int k = 0; k < 3; ++k) { a.cmin[k * a.n + i] = 0xFFFFFFFFu; a.cmax[k * a.n + i] = 0u; }
}

// centroid box of every segment of the level (bvh.rs:75-79): warp-aggregated when a warp sits in one segment
__global__ void k_centroid_bounds(BuildArrays a) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = p < a.n;
    const uint32_t lo = in ? a.lo[p] : 0u, hi = in ? a.hi[p] : 0u;
    const bool active = in && hi - lo >= 2u;
    uint32_t node = active ? a.idx[p] : 0xFFFFFFFFu;
    uint32_t mn[3] = {0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu}, mx[3] = {0u, 0u, 0u};
    if (active) {  // f32::min / max skip a NaN operand (aabb.rs:41-54): a NaN coordinate does not move the centroid box
        const uint32_t h = a.perm[p];
        const float c[3] = {a.cx[h], a.cy[h], a.cz[h]};
        for (int k = 0; k < 3; ++k) if (c[k] == c[k]) mn[k] = mx[k] = ordered(c[k]);
    }
    const uint32_t first = __shfl_sync(0xffffffffu, node, 0);
    const bool uniform = __all_sync(0xffffffffu, node == first);
    if (uniform) {
        if (first == 0xFFFFFFFFu) return;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mn[k] = __reduce_min_sync(0xffffffffu, mn[k]);
            mx[k] = __reduce_max_sync(0xffffffffu, mx[k]);
        }
        if ((threadIdx.x & 31u) == 0u)
            for (int k = 0; k < 3; ++k) { atomicMin(&a.cmin[k * a.n + node], mn[k]); atomicMax(&a.cmax[k * a.n + node], mx[k]); }
    } else if (active) {
        for (int k = 0; k < 3; ++k) { atomicMin(&a.cmin[k * a.n + node], mn[k]); atomicMax(&a.cmax[k * a.n + node], mx[k]); }
    }
}

// sort key of the level: (segment start, centroid on the segment's axis) — axis rule of bvh.rs:81-88
__global__ void k_keys(BuildArrays a) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n) return;
    const uint32_t lo = a.lo[p], hi = a.hi[p];
    uint32_t low = 0u;
    if (hi - lo >= 2u) {
        const uint32_t node = a.idx[p], n = a.n;
        const float sx = unordered(a.cmax[node]) - unordered(a.cmin[node]);
        const float sy = unordered(a.cmax[n + node]) - unordered(a.cmin[n + node]);
        const float sz = unordered(a.cmax[2 * n + node]) - unordered(a.cmin[2 * n + node]);
        const int axis = (sx > sy && sx > sz) ? 0 : (sy > sx && sy > sz) ? 1 : 2;
        const uint32_t h = a.perm[p];
        const float key = axis == 0 ? a.cx[h] : axis == 1 ? a.cy[h] : a.cz[h];
        if (!(key == key)) a.flags[0] = 1u;  // partial_cmp(..).unwrap() on a NaN sort key: the reference panics (bvh.rs:99)
        low = ordered(key);
    }
    a.key[p] = ((unsigned long long)lo << 32) | low;
}

// after the sort: the segment's first n/2 go left, the rest right (bvh.rs:125-134); record the node
__global__ void k_advance(BuildArrays a, uint32_t depth) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n) return;
    const uint32_t lo = a.lo[p], hi = a.hi[p];
    if (hi - lo < 2u) return;
    const uint32_t node = a.idx[p], mid = lo + (hi - lo) / 2u;
    if (p == lo) { a.node_lo[node] = lo; a.node_hi[node] = hi; a.node_depth[node] = depth; }
    if (p < mid) { a.hi[p] = mid; a.idx[p] = node + 1u; }
    else { a.lo[p] = mid; a.idx[p] = node + (mid - lo); }
}

// leaves in DFS order: slot = position, rank = position (bvh.rs:183-187 tie rule)
__global__ void k_emit_leaves(BuildArrays a, F4* spheres, PrimInfo* info, ObjectRec* objects, F4* linear) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n) return;
    const uint32_t h = a.perm[p];
    const float4 s = a.sph[h];
    spheres[p] = F4{s.x, s.y, s.z, s.w};
    info[p] = PrimInfo{h, 0u, a.mat[h], p};
    a.pmin[p] = make_float4(s.x - s.w, s.y - s.w, s.z - s.w, 0.0f);
    a.pmax[p] = make_float4(s.x + s.w, s.y + s.w, s.z + s.w, 0.0f);
    ObjectRec o;
    o.entry = o.entry_reference = RTR_MAKE_REF(REF_SPHERE, p);
    o.entry4 = o.entry; o._pad = 0u;
    objects[h] = o;
    linear[h] = F4{s.x, s.y, s.z, s.w};
}

// ---- triangle meshes ------------------------------------------------------------------------------------------
struct MeshInput {
    const float* pos; const float* nrm; const uint32_t* idx;  // the crate's vertex / index buffers, on the device
    uint32_t object, material, rank_base;
};
// f32::min / f32::max as host_flatten.cpp writes them (aabb.rs:41-54): the same expression, so the same bits
__device__ __forceinline__ float fmin_(float x, float y) { return x <= y ? x : x > y ? y : (y != y ? x : y); }
__device__ __forceinline__ float fmax_(float x, float y) { return x >= y ? x : x < y ? y : (y != y ? x : y); }
__device__ __forceinline__ float3 vtx(const float* p, uint32_t i) { return make_float3(p[3 * (size_t)i], p[3 * (size_t)i + 1], p[3 * (size_t)i + 2]); }

// Triangle::bounds (mesh.rs:200-213): box of the three vertices, epsilon_expand(0.01) (aabb.rs:57-78); its centroid
__global__ void k_init_tri(BuildArrays a, MeshInput m) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float3 p0 = vtx(m.pos, m.idx[3 * (size_t)i]), p1 = vtx(m.pos, m.idx[3 * (size_t)i + 1]), p2 = vtx(m.pos, m.idx[3 * (size_t)i + 2]);
    float mnx = fmin_(p0.x, fmin_(p1.x, p2.x)), mny = fmin_(p0.y, fmin_(p1.y, p2.y)), mnz = fmin_(p0.z, fmin_(p1.z, p2.z));
    float mxx = fmax_(p0.x, fmax_(p1.x, p2.x)), mxy = fmax_(p0.y, fmax_(p1.y, p2.y)), mxz = fmax_(p0.z, fmax_(p1.z, p2.z));
    const float eps = 0.01f;
    const float ex = (mnx + mxx) / 2.0f, ey = (mny + mxy) / 2.0f, ez = (mnz + mxz) / 2.0f;
    if (mxx - mnx < eps) { mnx = ex - eps; mxx = ex + eps; }
    if (mxy - mny < eps) { mny = ey - eps; mxy = ey + eps; }
    if (mxz - mnz < eps) { mnz = ez - eps; mxz = ez + eps; }
    a.hmin[i] = make_float4(mnx, mny, mnz, 0.0f);
    a.hmax[i] = make_float4(mxx, mxy, mxz, 0.0f);
    const float cx = (mnx + mxx) / 2.0f, cy = (mny + mxy) / 2.0f, cz = (mnz + mxz) / 2.0f;
    a.cx[i] = cx; a.cy[i] = cy; a.cz[i] = cz;
    a.perm[i] = i; a.lo[i] = 0u; a.hi[i] = a.n; a.idx[i] = 0u;
    for (int k = 0; k < 3; ++k) { a.cmin[k * a.n + i] = 0xFFFFFFFFu; a.cmax[k * a.n + i] = 0u; }
}

// triangle records in DFS leaf order (flat.h `tris` / `tri_vn`), as host_flatten.cpp MeshColl::leaf writes them
__global__ void k_emit_leaves_tri(BuildArrays a, MeshInput m, F4* tris, F4* tri_vn) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= a.n) return;
    const uint32_t h = a.perm[p];
    const uint32_t i0 = m.idx[3 * (size_t)h], i1 = m.idx[3 * (size_t)h + 1], i2 = m.idx[3 * (size_t)h + 2];
    const float3 p0 = vtx(m.pos, i0), p1 = vtx(m.pos, i1), p2 = vtx(m.pos, i2);
    // face normal as Mesh::from_buffers computes it (mesh.rs:85-90): normalize(cross(p2 - p1, p0 - p1))
    const float ax = p2.x - p1.x, ay = p2.y - p1.y, az = p2.z - p1.z, bx = p0.x - p1.x, by = p0.y - p1.y, bz = p0.z - p1.z;
    const float nx = ay * bz - az * by, ny = az * bx - ax * bz, nz = ax * by - ay * bx;
    const float s = 1.0f / sqrtf((nx * nx + ny * ny) + nz * nz);
    const size_t slot = (size_t)a.slot_base + p;
    F4* q = tris + slot * 4;
    q[0] = F4{p0.x, p0.y, p0.z, nx * s};
    q[1] = F4{p1.x - p0.x, p1.y - p0.y, p1.z - p0.z, ny * s};  // edges as Triangle::hit computes them (mesh.rs:160-161)
    q[2] = F4{p2.x - p0.x, p2.y - p0.y, p2.z - p0.z, nz * s};
    q[3] = F4{__uint_as_float(m.object), __uint_as_float(h), __uint_as_float(m.material), __uint_as_float(m.rank_base + p)};
    const uint32_t iv[3] = {i0, i1, i2};
    for (int k = 0; k < 3; ++k) {
        const float3 nn = vtx(m.nrm, iv[k]);
        tri_vn[slot * 3 + k] = F4{nn.x, nn.y, nn.z, 0.0f};
    }
    a.pmin[p] = a.hmin[h]; a.pmax[p] = a.hmax[h];
}

__device__ __forceinline__ void child_box(const BuildArrays& a, uint32_t lo, uint32_t hi, uint32_t node, float4& mn, float4& mx) {
    if (hi - lo == 1u) { mn = a.pmin[lo]; mx = a.pmax[lo]; }
    else { mn = a.bmin[node]; mx = a.bmax[node]; }
}

// bounds of a Split = union of its children's (bvh.rs:68-72), one tree level per launch, deepest first
__global__ void k_node_boxes(BuildArrays a, uint32_t depth) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node + 1u >= a.n || a.node_depth[node] != depth) return;
    const uint32_t lo = a.node_lo[node], hi = a.node_hi[node], mid = lo + (hi - lo) / 2u;
    float4 lmn, lmx, rmn, rmx;
    child_box(a, lo, mid, node + 1u, lmn, lmx);
    child_box(a, mid, hi, node + (mid - lo), rmn, rmx);
    a.bmin[node] = make_float4(fminf(lmn.x, rmn.x), fminf(lmn.y, rmn.y), fminf(lmn.z, rmn.z), 0.0f);
    a.bmax[node] = make_float4(fmaxf(lmx.x, rmx.x), fmaxf(lmx.y, rmx.y), fmaxf(lmx.z, rmx.z), 0.0f);
}

// child reference + the box its slot carries: a Split's own box, or (leaf) the gate = the parent's box
__device__ __forceinline__ uint32_t child_ref(const BuildArrays& a, uint32_t lo, uint32_t hi, uint32_t node, uint32_t parent,
                                              float4& mn, float4& mx) {
    const uint32_t n = hi - lo;
    if (n == 1u) { mn = a.bmin[parent]; mx = a.bmax[parent]; return RTR_MAKE_REF(a.leaf_kind, a.slot_base + lo); }
    mn = a.bmin[node]; mx = a.bmax[node];
    return n == 2u ? RTR_MAKE_REF(a.pair_kind, a.slot_base + lo) : RTR_MAKE_REF(REF_NODE, a.node_base + node);
}

__global__ void k_emit_nodes(BuildArrays a, F4* nodes) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node + 1u >= a.n) return;
    const uint32_t lo = a.node_lo[node], hi = a.node_hi[node], mid = lo + (hi - lo) / 2u;
    float4 lmn, lmx, rmn, rmx;
    const uint32_t l = child_ref(a, lo, mid, node + 1u, node, lmn, lmx);
    const uint32_t r = child_ref(a, mid, hi, node + (mid - lo), node, rmn, rmx);
    F4* q = nodes + (size_t)(a.node_base + node) * 4;
    q[0] = F4{lmn.x, lmn.y, lmn.z, __uint_as_float(l)};
    q[1] = F4{lmx.x, lmx.y, lmx.z, __uint_as_float(r)};
    q[2] = F4{rmn.x, rmn.y, rmn.z, 0.0f};
    q[3] = F4{rmx.x, rmx.y, rmx.z, 0.0f};
    if (node == 0u) {  // pseudo parent of the root: the root's own box is tested on entry (flat.h); index n-1
        F4* e = nodes + (size_t)(a.node_base + a.n - 1u) * 4;
        const uint32_t root = a.n == 2u ? RTR_MAKE_REF(a.pair_kind, a.slot_base) : RTR_MAKE_REF(REF_NODE, a.node_base);
        const float4 mn = a.bmin[0], mx = a.bmax[0];
        e[0] = F4{mn.x, mn.y, mn.z, __uint_as_float(root)};
        e[1] = F4{mx.x, mx.y, mx.z, __uint_as_float(REF_EMPTY_VALUE)};
        e[2] = F4{0.0f, 0.0f, 0.0f, 0.0f};
        e[3] = F4{0.0f, 0.0f, 0.0f, 0.0f};
    }
}

// Gated entries for a host-built traversal tree (hybrid build, host_flatten.h build_traversal_over_gates): per leaf
// position {gate.min, bits(ref)} {gate.max, 0} — a single leaf is gated by its parent's box, the two leaves of a 2-leaf
// Split are ONE paired entry gated by that Split's own box (the second position holds EMPTY).
__global__ void k_emit_gated(BuildArrays a, F4* gated) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node + 1u >= a.n) return;
    const uint32_t lo = a.node_lo[node], hi = a.node_hi[node], mid = lo + (hi - lo) / 2u;
    if (hi - lo < 3u) return;  // a 2-leaf Split is written by its parent
    const uint32_t clo[2] = {lo, mid}, chi[2] = {mid, hi}, cnode[2] = {node + 1u, node + (mid - lo)};
    for (int k = 0; k < 2; ++k) {
        const uint32_t n = chi[k] - clo[k];
        if (n > 2u) continue;
        float4 mn, mx;
        const uint32_t ref = child_ref(a, clo[k], chi[k], cnode[k], node, mn, mx);
        gated[2 * (size_t)clo[k]] = F4{mn.x, mn.y, mn.z, __uint_as_float(ref)};
        gated[2 * (size_t)clo[k] + 1] = F4{mx.x, mx.y, mx.z, 0.0f};
        if (n == 2u) {
            gated[2 * (size_t)clo[k] + 2] = F4{0.0f, 0.0f, 0.0f, __uint_as_float(REF_EMPTY_VALUE)};
            gated[2 * (size_t)clo[k] + 3] = F4{0.0f, 0.0f, 0.0f, 0.0f};
        }
    }
}

// ---- 4-wide traversal tree (flat.h nodes4) ----------------------------------------------------------------------
// The tree's shape depends on n alone and every interior child of a Split can be replaced by its own two children
// without overflowing four slots, so the 4-wide nodes are exactly the Splits at even depth that hold three or more
// leaves (plus the root); each slot keeps the box its 2-wide slot carries (k_emit_nodes / child_ref).
__global__ void k_flag4(BuildArrays a, uint32_t* flag) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= a.n) return;
    flag[node] = node + 1u < a.n && (a.node_depth[node] & 1u) == 0u && (a.node_hi[node] - a.node_lo[node] >= 3u || node == 0u) ? 1u : 0u;
}

struct Slots4 { float4 mn[4], mx[4]; uint32_t ref[4]; int n; };
__device__ __forceinline__ void slot4(const BuildArrays& a, const uint32_t* idx4, Slots4& s, uint32_t lo, uint32_t hi, uint32_t node, uint32_t parent) {
    float4 mn, mx;
    uint32_t ref = child_ref(a, lo, hi, node, parent, mn, mx);
    if (RTR_REF_KIND(ref) == REF_NODE) ref = RTR_MAKE_REF(REF_NODE, a.node4_base + idx4[node]);
    s.mn[s.n] = mn; s.mx[s.n] = mx; s.ref[s.n] = ref; ++s.n;
}
__device__ __forceinline__ void expand4(const BuildArrays& a, const uint32_t* idx4, Slots4& s, uint32_t lo, uint32_t hi, uint32_t node, uint32_t parent) {
    if (hi - lo < 3u) { slot4(a, idx4, s, lo, hi, node, parent); return; }  // a leaf or a paired leaf: one slot
    const uint32_t mid = lo + (hi - lo) / 2u;
    slot4(a, idx4, s, lo, mid, node + 1u, node);
    slot4(a, idx4, s, mid, hi, node + (mid - lo), node);
}
__global__ void k_emit_nodes4(BuildArrays a, const uint32_t* flag, const uint32_t* idx4, F4* nodes4) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node >= a.n || flag[node] == 0u) return;
    const uint32_t lo = a.node_lo[node], hi = a.node_hi[node], mid = lo + (hi - lo) / 2u;
    Slots4 s;
    s.n = 0;
    expand4(a, idx4, s, lo, mid, node + 1u, node);
    expand4(a, idx4, s, mid, hi, node + (mid - lo), node);
    for (int i = s.n; i < 4; ++i) { s.mn[i] = s.mx[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); s.ref[i] = REF_EMPTY_VALUE; }
    F4* q = nodes4 + (size_t)(a.node4_base + idx4[node]) * 8;
    for (int i = 0; i < 4; ++i) {
        q[2 * i] = F4{s.mn[i].x, s.mn[i].y, s.mn[i].z, 0.0f};
        q[2 * i + 1] = F4{s.mx[i].x, s.mx[i].y, s.mx[i].z, 0.0f};
    }
    for (int i = 0; i < 4; ++i) q[i].w = __uint_as_float(s.ref[i]);  // flat.h: the refs live in q0.w .. q3.w
}

}  // namespace

bool device_build_applies(const rtr_scene_desc& d) {
    if (d.n_objects < 2) return false;
    if (!d.objects || !d.spheres) return false;
    for (uint32_t i = 0; i < d.n_objects; ++i)
        if (d.objects[i].kind != RTR_OBJ_SPHERE || d.objects[i].index >= d.n_spheres) return false;
    return true;
}

namespace {

// Scratch + launches shared by the two front ends.  Everything is stream-ordered on `st`; scratch comes from `pool`.
struct Builder {
    cudaStream_t st;
    cudaMemPool_t pool;
    cudaError_t e = cudaSuccess;
    std::vector<void*> scratch;
    BuildArrays a;
    uint32_t n = 0, levels = 0;
    unsigned grid = 1;
    void* d_temp = nullptr;
    size_t temp_bytes = 0;
    int end_bit = 32;
    uint32_t* flag4 = nullptr;
    uint32_t* idx4 = nullptr;

    void* alloc(size_t bytes) {
        void* p = nullptr;
        if (e != cudaSuccess) return nullptr;
        e = cudaMallocFromPoolAsync(&p, bytes ? bytes : 16, pool, st);  // stream-ordered, reused across uploads
        if (e != cudaSuccess) return nullptr;
        scratch.push_back(p);
        return p;
    }
    void step(cudaError_t r) { if (e == cudaSuccess) e = r; }
    void cleanup() { for (void* p : scratch) cudaFreeAsync(p, st); scratch.clear(); }

    void setup(uint32_t count, bool mesh) {
        n = count;
        std::memset(&a, 0, sizeof(a));
        a.n = n;
        a.leaf_kind = mesh ? REF_TRI : REF_SPHERE; a.pair_kind = mesh ? REF_TRI2 : REF_SPHERE2;
        grid = (n + 255u) / 256u;
        a.cx = (float*)alloc(n * 4); a.cy = (float*)alloc(n * 4); a.cz = (float*)alloc(n * 4);
        a.perm = (uint32_t*)alloc(n * 4); a.perm_alt = (uint32_t*)alloc(n * 4);
        a.key = (unsigned long long*)alloc((size_t)n * 8); a.key_alt = (unsigned long long*)alloc((size_t)n * 8);
        a.lo = (uint32_t*)alloc(n * 4); a.hi = (uint32_t*)alloc(n * 4); a.idx = (uint32_t*)alloc(n * 4);
        a.cmin = (uint32_t*)alloc((size_t)n * 12); a.cmax = (uint32_t*)alloc((size_t)n * 12);
        a.node_lo = (uint32_t*)alloc(n * 4); a.node_hi = (uint32_t*)alloc(n * 4); a.node_depth = (uint32_t*)alloc(n * 4);
        a.bmin = (float4*)alloc(n * sizeof(float4)); a.bmax = (float4*)alloc(n * sizeof(float4));
        a.pmin = (float4*)alloc(n * sizeof(float4)); a.pmax = (float4*)alloc(n * sizeof(float4));
        if (mesh) { a.hmin = (float4*)alloc(n * sizeof(float4)); a.hmax = (float4*)alloc(n * sizeof(float4)); }
        a.flags = (uint32_t*)alloc(16);
        // one sort plan for every level: 64-bit keys, the bits above (32 + bits(n)) are zero
        end_bit = 32;
        while (end_bit < 64 && (1ull << (end_bit - 32)) < n) ++end_bit;
        if (e == cudaSuccess) {
            cub::DoubleBuffer<unsigned long long> kb(a.key, a.key_alt);
            cub::DoubleBuffer<uint32_t> vb(a.perm, a.perm_alt);
            e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, kb, vb, (int)n, 0, end_bit, st);
            d_temp = alloc(temp_bytes);
        }
        levels = 0;
        while ((1ull << levels) < n) ++levels;  // ceil(log2 n): after this many splits every segment is one leaf
        if (e == cudaSuccess) step(cudaMemsetAsync(a.flags, 0, 16, st));
    }
    // after the front end's init kernel: the level-by-level sorts (bvh.rs:75-134)
    void sort_levels() {
        for (uint32_t L = 0; L < levels && e == cudaSuccess; ++L) {
            k_centroid_bounds<<<grid, 256, 0, st>>>(a);
            k_keys<<<grid, 256, 0, st>>>(a);
            cub::DoubleBuffer<unsigned long long> kb(a.key, a.key_alt);
            cub::DoubleBuffer<uint32_t> vb(a.perm, a.perm_alt);
            step(cub::DeviceRadixSort::SortPairs(d_temp, temp_bytes, kb, vb, (int)n, 0, end_bit, st));
            if (kb.Current() != a.key) { unsigned long long* t = a.key; a.key = a.key_alt; a.key_alt = t; }
            if (vb.Current() != a.perm) { uint32_t* t = a.perm; a.perm = a.perm_alt; a.perm_alt = t; }
            k_advance<<<grid, 256, 0, st>>>(a, L);
            step(cudaGetLastError());
        }
    }
    // after the front end's leaf kernel (which fills pmin / pmax): Split boxes bottom-up, then the node records
    void emit_nodes(F4* nodes) {
        for (uint32_t L = levels; L-- > 0 && e == cudaSuccess;) k_node_boxes<<<grid, 256, 0, st>>>(a, L);
        if (e == cudaSuccess) k_emit_nodes<<<grid, 256, 0, st>>>(a, nodes);
        step(cudaGetLastError());
    }
    // 4-wide tree, part 1: flag its nodes and number them in pre-order
    void number_nodes4() {
        flag4 = (uint32_t*)alloc((size_t)n * 4);
        idx4 = (uint32_t*)alloc((size_t)n * 4 + 4);
        size_t scan_bytes = 0;
        step(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, flag4, idx4, (int)n, st));
        void* scan_temp = alloc(scan_bytes);
        if (e == cudaSuccess) {
            k_flag4<<<grid, 256, 0, st>>>(a, flag4);
            step(cub::DeviceScan::ExclusiveSum(scan_temp, scan_bytes, flag4, idx4, (int)n, st));
        }
    }
    // part 2: the node records (the caller sized nodes4 from the count of part 1, or knows it in advance)
    void emit_nodes4(F4* nodes4) {
        if (e == cudaSuccess) k_emit_nodes4<<<grid, 256, 0, st>>>(a, flag4, idx4, nodes4);
        step(cudaGetLastError());
    }
};

}  // namespace

cudaError_t build_sphere_scene_on_device(const rtr_scene_desc& d, cudaStream_t st, cudaMemPool_t pool, DeviceBvh* out, std::vector<void*>& allocs,
                                         bool want_nodes4) {
    const uint32_t n = d.n_objects;
    Builder b;
    b.st = st; b.pool = pool;
    auto keep = [&](size_t bytes) -> void* {
        void* p = nullptr;
        if (b.e != cudaSuccess) return nullptr;
        b.e = cudaMallocFromPoolAsync(&p, bytes ? bytes : 16, pool, st);
        if (b.e != cudaSuccess) return nullptr;
        allocs.push_back(p);
        out->device_bytes += bytes;
        return p;
    };
    // inputs by object handle (Scene.objects order; an object is a reference into spheres[])
    std::vector<float4> h_sph(n);
    std::vector<uint32_t> h_mat(n);
    for (uint32_t i = 0; i < n; ++i) {
        const rtr_sphere& s = d.spheres[d.objects[i].index];
        h_sph[i] = make_float4(s.center.x, s.center.y, s.center.z, s.radius);
        h_mat[i] = s.material;
    }
    b.setup(n, false);
    BuildArrays& a = b.a;
    float4* d_sph = (float4*)b.alloc(n * sizeof(float4));
    uint32_t* d_mat = (uint32_t*)b.alloc(n * sizeof(uint32_t));
    a.sph = d_sph; a.mat = d_mat;
    out->nodes = (F4*)keep((size_t)n * 4 * sizeof(F4));
    out->spheres = (F4*)keep(n * sizeof(F4));
    out->sphere_info = (PrimInfo*)keep(n * sizeof(PrimInfo));
    out->objects = (ObjectRec*)keep(n * sizeof(ObjectRec));
    out->linear_spheres = (F4*)keep(n * sizeof(F4));
    if (b.e != cudaSuccess) { b.cleanup(); return b.e; }
    b.step(cudaMemcpyAsync(d_sph, h_sph.data(), n * sizeof(float4), cudaMemcpyHostToDevice, st));
    b.step(cudaMemcpyAsync(d_mat, h_mat.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    out->h2d_bytes += n * (sizeof(float4) + sizeof(uint32_t));
    k_init<<<b.grid, 256, 0, st>>>(a);
    b.sort_levels();
    k_emit_leaves<<<b.grid, 256, 0, st>>>(a, out->spheres, out->sphere_info, out->objects, out->linear_spheres);
    b.emit_nodes(out->nodes);
    // 4-wide tree: the node count comes back with the NaN flag (one synchronisation)
    if (want_nodes4) b.number_nodes4();
    uint32_t h_flags[4] = {0, 0, 0, 0};
    uint32_t h_last[2] = {0, 0};  // flag and exclusive index of the last Split (node n-2): their sum is the node count
    b.step(cudaMemcpyAsync(h_flags, a.flags, 16, cudaMemcpyDeviceToHost, st));
    if (b.idx4 && b.e == cudaSuccess) {
        b.step(cudaMemcpyAsync(&h_last[0], b.flag4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
        b.step(cudaMemcpyAsync(&h_last[1], b.idx4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
    }
    b.step(cudaStreamSynchronize(st));
    if (b.idx4 && b.e == cudaSuccess) {
        out->n_nodes4 = h_last[0] + h_last[1];
        out->nodes4 = (F4*)keep((size_t)out->n_nodes4 * 8 * sizeof(F4));
        b.emit_nodes4(out->nodes4);
        if (b.e == cudaSuccess) {
            out->root_entry4 = RTR_MAKE_REF(REF_NODE, 0u);
            out->max_depth4 = b.levels / 2u + 2u;
        }
    }
    b.cleanup();  // stream-ordered frees: after the kernel above
    out->nan_centroid = h_flags[0] != 0;
    out->n_nodes = n;
    out->root_entry = RTR_MAKE_REF(REF_NODE, n - 1u);
    out->max_depth = b.levels + 2;
    return b.e;
}

cudaError_t build_mesh_on_device(const rtr_mesh& mesh, const DeviceMesh& dm, cudaStream_t st, cudaMemPool_t pool, F4* nodes, F4* nodes4,
                                 F4* tris, F4* tri_vn, uint64_t* h2d_bytes, bool* nan_centroid, bool* count_mismatch, std::vector<F4>* gated_out) {
    const uint32_t n = dm.n_tris;
    Builder b;
    b.st = st; b.pool = pool;
    b.setup(n, true);
    BuildArrays& a = b.a;
    a.slot_base = dm.slot_base; a.node_base = dm.node_base; a.node4_base = dm.node4_base;
    const size_t vbytes = (size_t)mesh.n_vertices * 3 * sizeof(float), ibytes = (size_t)n * 3 * sizeof(uint32_t);
    float* d_pos = (float*)b.alloc(vbytes);
    float* d_nrm = (float*)b.alloc(vbytes);
    uint32_t* d_idx = (uint32_t*)b.alloc(ibytes);
    if (b.e != cudaSuccess) { b.cleanup(); return b.e; }
    b.step(cudaMemcpyAsync(d_pos, mesh.positions, vbytes, cudaMemcpyHostToDevice, st));
    b.step(cudaMemcpyAsync(d_nrm, mesh.normals, vbytes, cudaMemcpyHostToDevice, st));
    b.step(cudaMemcpyAsync(d_idx, mesh.indices, ibytes, cudaMemcpyHostToDevice, st));
    *h2d_bytes += 2 * vbytes + ibytes;
    MeshInput in;
    in.pos = d_pos; in.nrm = d_nrm; in.idx = d_idx;
    in.object = dm.object; in.material = dm.material; in.rank_base = dm.rank_base;
    k_init_tri<<<b.grid, 256, 0, st>>>(a, in);
    b.sort_levels();
    k_emit_leaves_tri<<<b.grid, 256, 0, st>>>(a, in, tris, tri_vn);
    b.emit_nodes(nodes);
    if (nodes4) { b.number_nodes4(); b.emit_nodes4(nodes4); }  // the host reserved count_nodes4(n) records (host_flatten.h)
    if (gated_out) {  // hybrid: the gates go back to the host, which builds the SAH traversal tree over them
        F4* d_gated = (F4*)b.alloc((size_t)n * 2 * sizeof(F4));
        gated_out->resize((size_t)n * 2);
        if (b.e == cudaSuccess) k_emit_gated<<<b.grid, 256, 0, st>>>(a, d_gated);
        b.step(cudaGetLastError());
        b.step(cudaMemcpyAsync(gated_out->data(), d_gated, (size_t)n * 2 * sizeof(F4), cudaMemcpyDeviceToHost, st));
    }
    uint32_t h_flags[4] = {0, 0, 0, 0};
    uint32_t h_last[2] = {0, 0};
    b.step(cudaMemcpyAsync(h_flags, a.flags, 16, cudaMemcpyDeviceToHost, st));
    if (b.idx4 && b.e == cudaSuccess) {
        b.step(cudaMemcpyAsync(&h_last[0], b.flag4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
        b.step(cudaMemcpyAsync(&h_last[1], b.idx4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
    }
    b.step(cudaStreamSynchronize(st));  // the caller's mesh buffers may die at return
    b.cleanup();
    *nan_centroid = h_flags[0] != 0;
    *count_mismatch = b.idx4 != nullptr && b.e == cudaSuccess && h_last[0] + h_last[1] != count_nodes4(n);
    return b.e;
}

}  // namespace rtr
