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
// Output = the flat.h layout, with every leaf slot carrying its gate box and 2-leaf Splits folded into paired
// leaves, so ONE node array serves both the default traversal and RTR_FLAG_REFERENCE_TREE (a paired leaf is just
// a compact encoding of the Split it replaces: one box test, then both primitives).
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

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
};

__global__ void k_init(BuildArrays a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= a.n) return;
    const float4 s = a.sph[i];
    // Sphere::bounds = center -/+ radius; AABB::centroid = (min + max) / 2.0
    const float x0 = s.x - s.w, x1 = s.x + s.w, y0 = s.y - s.w, y1 = s.y + s.w, z0 = s.z - s.w, z1 = s.z + s.w;
    const float cx = (x0 + x1) / 2.0f, cy = (y0 + y1) / 2.0f, cz = (z0 + z1) / 2.0f;
    a.cx[i] = cx; a.cy[i] = cy; a.cz[i] = cz;
    if (!(cx == cx) || !(cy == cy) || !(cz == cz)) a.flags[0] = 1u;  // the reference panics (bvh.rs:99)
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
    if (active) {
        const uint32_t h = a.perm[p];
        mn[0] = mx[0] = ordered(a.cx[h]); mn[1] = mx[1] = ordered(a.cy[h]); mn[2] = mx[2] = ordered(a.cz[h]);
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
        low = ordered(axis == 0 ? a.cx[h] : axis == 1 ? a.cy[h] : a.cz[h]);
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
    if (n == 1u) { mn = a.bmin[parent]; mx = a.bmax[parent]; return RTR_MAKE_REF(REF_SPHERE, lo); }
    mn = a.bmin[node]; mx = a.bmax[node];
    return n == 2u ? RTR_MAKE_REF(REF_SPHERE2, lo) : RTR_MAKE_REF(REF_NODE, node);
}

__global__ void k_emit_nodes(BuildArrays a, F4* nodes) {
    const uint32_t node = blockIdx.x * blockDim.x + threadIdx.x;
    if (node + 1u >= a.n) return;
    const uint32_t lo = a.node_lo[node], hi = a.node_hi[node], mid = lo + (hi - lo) / 2u;
    float4 lmn, lmx, rmn, rmx;
    const uint32_t l = child_ref(a, lo, mid, node + 1u, node, lmn, lmx);
    const uint32_t r = child_ref(a, mid, hi, node + (mid - lo), node, rmn, rmx);
    F4* q = nodes + (size_t)node * 4;
    q[0] = F4{lmn.x, lmn.y, lmn.z, __uint_as_float(l)};
    q[1] = F4{lmx.x, lmx.y, lmx.z, __uint_as_float(r)};
    q[2] = F4{rmn.x, rmn.y, rmn.z, 0.0f};
    q[3] = F4{rmx.x, rmx.y, rmx.z, 0.0f};
    if (node == 0u) {  // pseudo parent of the root: the root's own box is tested on entry (flat.h); index n-1
        F4* e = nodes + (size_t)(a.n - 1u) * 4;
        const uint32_t root = a.n == 2u ? RTR_MAKE_REF(REF_SPHERE2, 0u) : RTR_MAKE_REF(REF_NODE, 0u);
        const float4 mn = a.bmin[0], mx = a.bmax[0];
        e[0] = F4{mn.x, mn.y, mn.z, __uint_as_float(root)};
        e[1] = F4{mx.x, mx.y, mx.z, __uint_as_float(REF_EMPTY_VALUE)};
        e[2] = F4{0.0f, 0.0f, 0.0f, 0.0f};
        e[3] = F4{0.0f, 0.0f, 0.0f, 0.0f};
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
    if (RTR_REF_KIND(ref) == REF_NODE) ref = RTR_MAKE_REF(REF_NODE, idx4[node]);
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
    F4* q = nodes4 + (size_t)idx4[node] * 8;
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

cudaError_t build_sphere_scene_on_device(const rtr_scene_desc& d, cudaStream_t st, cudaMemPool_t pool, DeviceBvh* out, std::vector<void*>& allocs,
                                         bool want_nodes4) {
    const uint32_t n = d.n_objects;
    cudaError_t e = cudaSuccess;
    std::vector<void*> scratch;
    auto alloc = [&](size_t bytes, bool keep) -> void* {
        void* p = nullptr;
        if (e != cudaSuccess) return nullptr;
        e = cudaMallocFromPoolAsync(&p, bytes ? bytes : 16, pool, st);  // stream-ordered, reused across uploads
        if (e != cudaSuccess) return nullptr;
        (keep ? allocs : scratch).push_back(p);
        if (keep) out->device_bytes += bytes;
        return p;
    };
    auto cleanup = [&]() { for (void* p : scratch) cudaFreeAsync(p, st); };

    // inputs by object handle (Scene.objects order; an object is a reference into spheres[])
    std::vector<float4> h_sph(n);
    std::vector<uint32_t> h_mat(n);
    for (uint32_t i = 0; i < n; ++i) {
        const rtr_sphere& s = d.spheres[d.objects[i].index];
        h_sph[i] = make_float4(s.center.x, s.center.y, s.center.z, s.radius);
        h_mat[i] = s.material;
    }
    BuildArrays a;
    a.n = n;
    float4* d_sph = (float4*)alloc(n * sizeof(float4), false);
    uint32_t* d_mat = (uint32_t*)alloc(n * sizeof(uint32_t), false);
    a.sph = d_sph; a.mat = d_mat;
    a.cx = (float*)alloc(n * 4, false); a.cy = (float*)alloc(n * 4, false); a.cz = (float*)alloc(n * 4, false);
    a.perm = (uint32_t*)alloc(n * 4, false); a.perm_alt = (uint32_t*)alloc(n * 4, false);
    a.key = (unsigned long long*)alloc(n * 8, false); a.key_alt = (unsigned long long*)alloc(n * 8, false);
    a.lo = (uint32_t*)alloc(n * 4, false); a.hi = (uint32_t*)alloc(n * 4, false); a.idx = (uint32_t*)alloc(n * 4, false);
    a.cmin = (uint32_t*)alloc((size_t)n * 12, false); a.cmax = (uint32_t*)alloc((size_t)n * 12, false);
    a.node_lo = (uint32_t*)alloc(n * 4, false); a.node_hi = (uint32_t*)alloc(n * 4, false); a.node_depth = (uint32_t*)alloc(n * 4, false);
    a.bmin = (float4*)alloc(n * sizeof(float4), false); a.bmax = (float4*)alloc(n * sizeof(float4), false);
    a.pmin = (float4*)alloc(n * sizeof(float4), false); a.pmax = (float4*)alloc(n * sizeof(float4), false);
    a.flags = (uint32_t*)alloc(16, false);
    out->nodes = (F4*)alloc((size_t)n * 4 * sizeof(F4), true);
    out->spheres = (F4*)alloc(n * sizeof(F4), true);
    out->sphere_info = (PrimInfo*)alloc(n * sizeof(PrimInfo), true);
    out->objects = (ObjectRec*)alloc(n * sizeof(ObjectRec), true);
    out->linear_spheres = (F4*)alloc(n * sizeof(F4), true);
    // one sort plan for every level: 64-bit keys, the bits above (32 + bits(n)) are zero
    int end_bit = 32;
    while (end_bit < 64 && (1ull << (end_bit - 32)) < n) ++end_bit;
    size_t temp_bytes = 0;
    void* d_temp = nullptr;
    if (e == cudaSuccess) {
        cub::DoubleBuffer<unsigned long long> kb(a.key, a.key_alt);
        cub::DoubleBuffer<uint32_t> vb(a.perm, a.perm_alt);
        e = cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, kb, vb, (int)n, 0, end_bit, st);
        d_temp = alloc(temp_bytes, false);
    }
    if (e != cudaSuccess) { cleanup(); return e; }
    auto step = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    step(cudaMemcpyAsync(d_sph, h_sph.data(), n * sizeof(float4), cudaMemcpyHostToDevice, st));
    step(cudaMemcpyAsync(d_mat, h_mat.data(), n * sizeof(uint32_t), cudaMemcpyHostToDevice, st));
    step(cudaMemsetAsync(a.flags, 0, 16, st));
    out->h2d_bytes += n * (sizeof(float4) + sizeof(uint32_t));
    const unsigned grid = (n + 255u) / 256u;
    k_init<<<grid, 256, 0, st>>>(a);
    uint32_t levels = 0;
    while ((1ull << levels) < n) ++levels;  // ceil(log2 n): after this many splits every segment is one leaf
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
    k_emit_leaves<<<grid, 256, 0, st>>>(a, out->spheres, out->sphere_info, out->objects, out->linear_spheres);
    for (uint32_t L = levels; L-- > 0 && e == cudaSuccess;) k_node_boxes<<<grid, 256, 0, st>>>(a, L);
    k_emit_nodes<<<grid, 256, 0, st>>>(a, out->nodes);
    step(cudaGetLastError());
    // 4-wide tree: flag its nodes, number them in pre-order; the count comes back with the NaN flag (one synchronisation)
    uint32_t* flag4 = nullptr;
    uint32_t* idx4 = nullptr;
    if (want_nodes4 && e == cudaSuccess) {
        flag4 = (uint32_t*)alloc((size_t)n * 4, false);
        idx4 = (uint32_t*)alloc((size_t)n * 4 + 4, false);
        size_t scan_bytes = 0;
        step(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, flag4, idx4, (int)n, st));
        void* scan_temp = alloc(scan_bytes, false);
        if (e == cudaSuccess) {
            k_flag4<<<grid, 256, 0, st>>>(a, flag4);
            step(cub::DeviceScan::ExclusiveSum(scan_temp, scan_bytes, flag4, idx4, (int)n, st));
        }
    }
    uint32_t h_flags[4] = {0, 0, 0, 0};
    uint32_t h_last[2] = {0, 0};  // flag and exclusive index of the last Split (node n-2): their sum is the node count
    step(cudaMemcpyAsync(h_flags, a.flags, 16, cudaMemcpyDeviceToHost, st));
    if (idx4 && e == cudaSuccess) {
        step(cudaMemcpyAsync(&h_last[0], flag4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
        step(cudaMemcpyAsync(&h_last[1], idx4 + (n - 2u), 4, cudaMemcpyDeviceToHost, st));
    }
    step(cudaStreamSynchronize(st));
    if (idx4 && e == cudaSuccess) {
        out->n_nodes4 = h_last[0] + h_last[1];
        out->nodes4 = (F4*)alloc((size_t)out->n_nodes4 * 8 * sizeof(F4), true);
        if (e == cudaSuccess) {
            k_emit_nodes4<<<grid, 256, 0, st>>>(a, flag4, idx4, out->nodes4);
            step(cudaGetLastError());
            out->root_entry4 = RTR_MAKE_REF(REF_NODE, 0u);
            out->max_depth4 = levels / 2u + 2u;
        }
    }
    cleanup();  // stream-ordered frees: after the kernel above
    out->nan_centroid = h_flags[0] != 0;
    out->n_nodes = n;
    out->root_entry = RTR_MAKE_REF(REF_NODE, n - 1u);
    out->max_depth = levels + 2;
    return e;
}

}  // namespace rtr
