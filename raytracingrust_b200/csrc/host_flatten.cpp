// host_flatten.cpp — rebuild the reference BVH topology and flatten it (flat.h).
//
// Host-side mirror of:
//   BvhNode::from_list           lib/core/bvh.rs:59-144   (median split, stable sort, n/2)
//   AABB surround/centroid/...   lib/utils/aabb.rs:21-78
//   Sphere::bounds               lib/objects.rs:53-60
//   Mesh::from_buffers / bounds  lib/core/mesh.rs:78-117,200-213
//   Volume::bounds / ::new       lib/volume.rs:26-31,89-93
// The topology decides exact-t ties (leaf DFS rank), so it is reproduced, not improved.
// Compiled with -ffp-contract=off: every box coordinate must equal the reference's f32 value.
#include "host_flatten.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <system_error>
#include <thread>

namespace rtr {

namespace {

struct V3 { float x, y, z; };
struct Box { V3 min, max; };

inline V3 v3(const float* p) { return V3{p[0], p[1], p[2]}; }
inline V3 sub(V3 a, V3 b) { return V3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 cross(V3 a, V3 b) { return V3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline V3 normalize(V3 a) { float s = 1.0f / std::sqrt(dot(a, a)); return V3{a.x * s, a.y * s, a.z * s}; }

// f32::min / f32::max (aabb.rs:41-54) ignore a NaN operand.  Written inline (as libm calls they were a quarter of the flatten
// time); +0 against -0 returns the first argument, as glibc does — which zero a union carries is unspecified in the
// reference too (LLVM minnum) and unobservable: the slab test only compares.
inline float fmin_(float x, float y) { return x <= y ? x : x > y ? y : (y != y ? x : y); }
inline float fmax_(float x, float y) { return x >= y ? x : x < y ? y : (y != y ? x : y); }
inline Box surround(const Box& a, const Box& b) {  // aabb.rs:41-54
    return Box{V3{fmin_(a.min.x, b.min.x), fmin_(a.min.y, b.min.y), fmin_(a.min.z, b.min.z)},
               V3{fmax_(a.max.x, b.max.x), fmax_(a.max.y, b.max.y), fmax_(a.max.z, b.max.z)}};
}
inline V3 centroid(const Box& b) {  // aabb.rs:28-30
    return V3{(b.min.x + b.max.x) / 2.0f, (b.min.y + b.max.y) / 2.0f, (b.min.z + b.max.z) / 2.0f};
}
inline Box epsilon_expand(Box b, float eps) {  // aabb.rs:57-78
    V3 c = centroid(b);
    Box r = b;
    if (b.max.x - b.min.x < eps) { r.min.x = c.x - eps; r.max.x = c.x + eps; }
    if (b.max.y - b.min.y < eps) { r.min.y = c.y - eps; r.max.y = c.y + eps; }
    if (b.max.z - b.min.z < eps) { r.min.z = c.z - eps; r.max.z = c.z + eps; }
    return r;
}

struct Built {
    uint32_t ref;   // child reference as stored in the parent (REFERENCE topology)
    Box box;        // own box (valid iff split)
    bool split;
    uint32_t depth; // levels below and including this node
    // the same subtree in the TRAVERSAL tree (see build_traversal): equal to `ref` for a primitive
    uint32_t fast_ref = 0xFFFFFFFFu;
    uint32_t fast_depth = 0;
    bool leaf = false;  // came from Collection::leaf (BvhNode::Object), not from a split of this collection
    Built() : ref(0xFFFFFFFFu), box{}, split(false), depth(0) {}
    Built(uint32_t r, Box b, bool s, uint32_t d) : ref(r), box(b), split(s), depth(d), fast_ref(r), fast_depth(d) {}
};

// one leaf of a collection as the traversal tree sees it: what to visit, and the box that GATES the visit
struct GatedLeaf {
    uint32_t ref; Box gate; uint32_t depth;
    V3 c;  // centroid of the gate (cached: the SAH build reads it once per level and axis)
    GatedLeaf() : ref(0xFFFFFFFFu), gate{}, depth(0), c{0, 0, 0} {}
    GatedLeaf(uint32_t r, const Box& g, uint32_t d) : ref(r), gate(g), depth(d), c(centroid(g)) {}
};

// Anything BvhNode::from_list can be built over (bvh.rs:18-28)
struct Collection {
    virtual ~Collection() {}
    virtual Box bounds(uint32_t handle) const = 0;
    virtual Built leaf(uint32_t handle) = 0;  // emit the leaf, return its child reference
};

struct Flattener {
    HostFlat& out;
    const rtr_scene_desc& d;
    std::string& err;
    uint32_t rank = 0;
    uint32_t slack = 8;  // levels the SAH traversal tree may exceed the balanced depth by (0 = always balanced)

    Flattener(HostFlat& o, const rtr_scene_desc& desc, std::string& e) : out(o), d(desc), err(e) {}

    // host-built nodes are numbered after the device-built ones (HostFlat::device_nodes); out.nodes holds only the former
    uint32_t alloc_node() {
        uint32_t idx = (uint32_t)(out.nodes.size() / 4);
        out.nodes.resize(out.nodes.size() + 4, F4{0, 0, 0, 0});
        return out.device_nodes + idx;
    }
    static float bits(uint32_t u) { float f; std::memcpy(&f, &u, 4); return f; }
    void write_node_raw(uint32_t idx, uint32_t lref, const Box& lb, uint32_t rref, const Box& rb) {
        F4* q = &out.nodes[(size_t)(idx - out.device_nodes) * 4];
        q[0] = F4{lb.min.x, lb.min.y, lb.min.z, bits(lref)};
        q[1] = F4{lb.max.x, lb.max.y, lb.max.z, bits(rref)};
        q[2] = F4{rb.min.x, rb.min.y, rb.min.z, 0.0f};
        q[3] = F4{rb.max.x, rb.max.y, rb.max.z, 0.0f};
    }
    void write_node(uint32_t idx, const Built& l, const Built& r) {
        const Box zero{{0, 0, 0}, {0, 0, 0}};
        write_node_raw(idx, l.ref, l.split ? l.box : zero, r.ref, r.split ? r.box : zero);
    }

    // ---- traversal tree ------------------------------------------------------------------------------
    // The reference tests a leaf of a collection exactly when the box of the Split that holds it is pierced
    // (bvh.rs:157-163; leaves carry no box, Q3) — ancestors above that Split contain it, and the slab test is
    // monotone under containment for a fixed ray, so "parent box pierced" already implies every ancestor's.
    // That parent box is the leaf's GATE.  Any tree whose interior boxes are unions of the gates below them
    // and whose leaf slots carry the gate therefore visits the same candidate set as the reference topology,
    // and ties are still decided by the reference's DFS leaf rank stored with each primitive.  This frees the
    // topology: the traversal tree is a binned-SAH tree over the gates (the reference's object-median split on
    // centroids costs ~2x the node visits on the mesh scenes), depth-capped so the stack bound holds.
    static float half_area(const Box& b) {
        const float dx = b.max.x - b.min.x, dy = b.max.y - b.min.y, dz = b.max.z - b.min.z;
        return dx * dy + dy * dz + dz * dx;
    }
    struct FastOut { uint32_t ref; Box box; uint32_t depth; };
    // helper threads a build may use besides its caller's: RTR_HOST_THREADS - 1, by default the host's cores (at most 15)
    static int spare_threads() {
        if (const char* e = std::getenv("RTR_HOST_THREADS")) return std::max(0, std::min(63, std::atoi(e) - 1));
        const unsigned hw = std::thread::hardware_concurrency();
        return (int)std::min(15u, hw > 1 ? hw - 1 : 0u);
    }
    // min/max for the SAH BINS only: plain compare-select (minss / maxss).  The bins feed the cost heuristic; the boxes
    // written into the nodes are exact unions of the gates (surround), so NaN / signed-zero subtleties do not matter here.
    static Box surround_fast(const Box& a, const Box& b) {
        return Box{V3{a.min.x < b.min.x ? a.min.x : b.min.x, a.min.y < b.min.y ? a.min.y : b.min.y, a.min.z < b.min.z ? a.min.z : b.min.z},
                   V3{a.max.x > b.max.x ? a.max.x : b.max.x, a.max.y > b.max.y ? a.max.y : b.max.y, a.max.z > b.max.z ? a.max.z : b.max.z}};
    }
    std::vector<GatedLeaf> part_scratch;  // stable partition buffer, as long as the collection: a range partitions through its own part
    // A binary tree over n entries has n - 1 nodes, and the serial recursion numbers them in pre-order: node `idx` of the
    // range [lo, hi) split at mid has its left subtree from idx + 1 and its right subtree from idx + (mid - lo).  With the
    // indices known in advance the subtrees are independent: the top of a large tree is built by several threads, and
    // the arrays come out byte for byte as the serial build writes them.
    FastOut build_traversal(std::vector<GatedLeaf>& L, size_t lo, size_t hi, uint32_t depth_left) {
        const size_t n = hi - lo;
        if (n == 1) return FastOut{L[lo].ref, L[lo].gate, L[lo].depth};
        const uint32_t first = out.device_nodes + (uint32_t)(out.nodes.size() / 4);
        out.nodes.resize(out.nodes.size() + 4 * (n - 1), F4{0, 0, 0, 0});
        if (part_scratch.size() < L.size()) part_scratch.resize(L.size());
        std::atomic<int> spare(spare_threads());
        return build_traversal_at(L, lo, hi, depth_left, first, spare);
    }
    FastOut build_traversal_at(std::vector<GatedLeaf>& L, size_t lo, size_t hi, uint32_t depth_left, uint32_t idx, std::atomic<int>& spare) {
        const size_t n = hi - lo;
        if (n == 1) return FastOut{L[lo].ref, L[lo].gate, L[lo].depth};
        Box cb{L[lo].c, L[lo].c};  // centroid bounds
        for (size_t i = lo + 1; i < hi; ++i) cb = surround_fast(cb, Box{L[i].c, L[i].c});
        size_t mid = lo + n / 2;
        uint32_t levels_needed = 0;
        while ((size_t(1) << levels_needed) < n) ++levels_needed;
        const bool must_balance = levels_needed + 1 >= depth_left;
        const V3 ext = sub(cb.max, cb.min);
        const float emax = std::fmax(ext.x, std::fmax(ext.y, ext.z));
        bool split_done = false;
        constexpr int NB = 16;
        auto bin_of = [](float c, float mn, float scale) {
            int k = (int)((c - mn) * scale);
            return k < 0 ? 0 : k >= NB ? NB - 1 : k;
        };
        if (!must_balance && emax > 0.0f && n > 2) {
            const float e3[3] = {ext.x, ext.y, ext.z}, mn3[3] = {cb.min.x, cb.min.y, cb.min.z};
            float scale3[3];
            for (int axis = 0; axis < 3; ++axis) scale3[axis] = e3[axis] > 0.0f ? (float)NB / e3[axis] : 0.0f;
            Box bb[3][NB] = {}; size_t cnt[3][NB] = {};
            for (size_t i = lo; i < hi; ++i) {  // one pass bins the element on every axis
                const float c3[3] = {L[i].c.x, L[i].c.y, L[i].c.z};
                for (int axis = 0; axis < 3; ++axis) {
                    if (!(e3[axis] > 0.0f)) continue;
                    const int k = bin_of(c3[axis], mn3[axis], scale3[axis]);
                    bb[axis][k] = cnt[axis][k] ? surround_fast(bb[axis][k], L[i].gate) : L[i].gate;
                    ++cnt[axis][k];
                }
            }
            float best_cost = INFINITY;
            int best_axis = -1, best_plane = -1;
            for (int axis = 0; axis < 3; ++axis) {
                if (!(e3[axis] > 0.0f)) continue;
                float right_area[NB]; size_t right_cnt[NB];
                Box acc{}; size_t c = 0;
                for (int k = NB - 1; k > 0; --k) {
                    if (cnt[axis][k]) { acc = c ? surround_fast(acc, bb[axis][k]) : bb[axis][k]; c += cnt[axis][k]; }
                    right_area[k] = c ? half_area(acc) : 0.0f; right_cnt[k] = c;
                }
                c = 0;
                for (int k = 0; k < NB - 1; ++k) {
                    if (cnt[axis][k]) { acc = c ? surround_fast(acc, bb[axis][k]) : bb[axis][k]; c += cnt[axis][k]; }
                    if (c == 0 || right_cnt[k + 1] == 0) continue;
                    const float cost = half_area(acc) * (float)c + right_area[k + 1] * (float)right_cnt[k + 1];
                    if (cost < best_cost) { best_cost = cost; best_axis = axis; best_plane = k; }
                }
            }
            if (best_axis >= 0) {
                // stable partition into [bins <= plane][the rest] through the shared scratch buffer
                const float mn = mn3[best_axis], scale = scale3[best_axis];
                GatedLeaf* scratch = part_scratch.data() + lo;
                size_t nl = 0, nr = 0;
                for (size_t i = lo; i < hi; ++i) {
                    const float c = best_axis == 0 ? L[i].c.x : best_axis == 1 ? L[i].c.y : L[i].c.z;
                    if (bin_of(c, mn, scale) <= best_plane) L[lo + nl++] = L[i];
                    else scratch[nr++] = L[i];
                }
                for (size_t i = 0; i < nr; ++i) L[lo + nl + i] = scratch[i];
                const size_t m = lo + nl;
                if (m > lo && m < hi) { mid = m; split_done = true; }
            }
        }
        if (!split_done && emax > 0.0f) {  // balanced fallback: median of the widest centroid axis
            const int axis = (ext.x >= ext.y && ext.x >= ext.z) ? 0 : (ext.y >= ext.z ? 1 : 2);
            std::nth_element(L.begin() + lo, L.begin() + mid, L.begin() + hi, [&](const GatedLeaf& a, const GatedLeaf& b) {
                return (axis == 0 ? a.c.x : axis == 1 ? a.c.y : a.c.z) < (axis == 0 ? b.c.x : axis == 1 ? b.c.y : b.c.z);
            });
        }
        const uint32_t lidx = idx + 1u, ridx = idx + (uint32_t)(mid - lo);
        FastOut l, r;
        bool forked = false;
        if (n >= 16384 && spare.fetch_sub(1) > 0) {
            try {
                std::thread t([&] { l = build_traversal_at(L, lo, mid, depth_left - 1, lidx, spare); });
                forked = true;
                r = build_traversal_at(L, mid, hi, depth_left - 1, ridx, spare);
                t.join();
            } catch (const std::system_error&) {  // no thread to be had: this range stays on the caller's
                if (forked) throw;
            }
            spare.fetch_add(1);
        } else if (n >= 16384) {
            spare.fetch_add(1);
        }
        if (!forked) {
            l = build_traversal_at(L, lo, mid, depth_left - 1, lidx, spare);
            r = build_traversal_at(L, mid, hi, depth_left - 1, ridx, spare);
        }
        write_node_raw(idx, l.ref, l.box, r.ref, r.box);
        return FastOut{RTR_MAKE_REF(REF_NODE, idx), surround(l.box, r.box), 1 + std::max(l.depth, r.depth)};
    }

    // reference tree + traversal tree of one collection (scene.build_bvh / mesh.build_bvh)
    Built build_collection(std::vector<uint32_t>& handles, Collection& coll) {
        const size_t n = handles.size();
        std::vector<GatedLeaf> gated;
        gated.reserve(n);
        Built root = build(handles, coll, &gated);
        if (!err.empty() || n < 2) return root;  // None / a single Object: both trees are that leaf
        if (gated.size() == 1) {  // two sibling primitives: one paired leaf gated by the root's own box
            root.fast_ref = gated[0].ref;
            root.fast_depth = 1;
            return root;
        }
        uint32_t levels = 0;
        while ((size_t(1) << levels) < n) ++levels;
        const FastOut f = build_traversal(gated, 0, gated.size(), levels + slack);
        root.fast_ref = f.ref;
        root.fast_depth = f.depth;
        return root;
    }

    // BvhNode::from_list (bvh.rs:59-144) over the handles list[lo, hi), in place.  The reference asks the collection for
    // every handle's bounds at every level (bvh.rs:68-79, 91-122) and sorts a fresh Vec per node; here bounds and centroids
    // are computed once per collection, the list is one array whose ranges are the nodes, and the stable sort (Rust's
    // slice::sort_by is a stable merge sort; with a strict weak order the result is unique) runs on (key, handle) pairs
    // through one preallocated buffer.  Same topology, leaves, ranks and gate boxes as the per-node version it replaces (checked by
    // hashing the flattened arrays of mesh, sphere and volume scenes); with build_traversal's one-pass binning a 28 K-triangle
    // mesh flattens in 60 ms instead of 150 ms in this container.
    struct KV { float k; uint32_t h; };
    struct BuildCtx {
        Collection* coll = nullptr;
        std::vector<GatedLeaf>* gated = nullptr;
        std::vector<Box> hb;        // bounds(handle)
        std::vector<V3> hc;         // bounds(handle).centroid()
        std::vector<uint32_t> list; // the handles; a node is a range of it
        std::vector<KV> kv, tmp;    // sort scratch
    };
    static void stable_sort_kv(KV* a, KV* tmp, size_t n) {
        if (n <= 24) {  // insertion sort: stable, strict '<'
            for (size_t i = 1; i < n; ++i) {
                const KV v = a[i];
                size_t j = i;
                while (j > 0 && v.k < a[j - 1].k) { a[j] = a[j - 1]; --j; }
                a[j] = v;
            }
            return;
        }
        const size_t h = n / 2;
        stable_sort_kv(a, tmp, h);
        stable_sort_kv(a + h, tmp, n - h);
        size_t i = 0, j = h, o = 0;
        while (i < h && j < n) tmp[o++] = (a[j].k < a[i].k) ? a[j++] : a[i++];  // ties keep the left (earlier) element first
        while (i < h) tmp[o++] = a[i++];
        while (j < n) tmp[o++] = a[j++];
        std::memcpy(a, tmp, n * sizeof(KV));
    }
    Built build(std::vector<uint32_t>& handles, Collection& coll, std::vector<GatedLeaf>* gated) {
        BuildCtx c;
        c.coll = &coll; c.gated = gated;
        const size_t n = handles.size();
        uint32_t max_h = 0;
        for (uint32_t h : handles) max_h = std::max(max_h, h);
        c.hb.resize(n ? (size_t)max_h + 1 : 0); c.hc.resize(c.hb.size());
        for (uint32_t h : handles) { c.hb[h] = coll.bounds(h); c.hc[h] = centroid(c.hb[h]); }
        c.list.swap(handles);
        c.kv.resize(n); c.tmp.resize(n);
        // Phase 1, the O(n log^2 n) part: every node's stable sort.  The tree's shape is known beforehand (a node is a
        // range, its children the two halves), a node's sort touches its own range only, so the subtrees of the upper
        // levels are sorted by several threads.  Phase 2 (build_range) then walks the ranges in order, serially, and
        // emits nodes, leaves, ranks and gates exactly as before.
        std::atomic<int> spare(spare_threads());
        std::atomic<bool> nan_key(false);
        sort_range(c, 0, n, spare, nan_key);
        if (nan_key.load()) { err = "NaN centroid (the reference panics at bvh.rs:99)"; return Built{REF_EMPTY_VALUE, Box{}, false, 0}; }
        return build_range(c, 0, n);
    }
    template <typename F, typename G>
    static void fork_join(size_t n, std::atomic<int>& spare, F&& left, G&& right, size_t min_n = 4096) {  // left on another thread when one is spare
        bool forked = false;
        if (n >= min_n) {
            if (spare.fetch_sub(1) > 0) {
                try {
                    std::thread t(left);
                    forked = true;
                    right();
                    t.join();
                } catch (const std::system_error&) {
                    if (forked) throw;
                }
            }
            spare.fetch_add(1);
        }
        if (!forked) { left(); right(); }
    }
    static void stable_sort_kv_mt(KV* a, KV* tmp, size_t n, std::atomic<int>& spare) {  // the same merge sort, halves in parallel
        if (n < 65536) { stable_sort_kv(a, tmp, n); return; }
        const size_t h = n / 2;
        fork_join(n, spare, [&] { stable_sort_kv_mt(a, tmp, h, spare); }, [&] { stable_sort_kv_mt(a + h, tmp + h, n - h, spare); }, 65536);
        size_t i = 0, j = h, o = 0;
        while (i < h && j < n) tmp[o++] = (a[j].k < a[i].k) ? a[j++] : a[i++];
        while (i < h) tmp[o++] = a[i++];
        while (j < n) tmp[o++] = a[j++];
        std::memcpy(a, tmp, n * sizeof(KV));
    }
    void sort_range(BuildCtx& c, size_t lo, size_t hi, std::atomic<int>& spare, std::atomic<bool>& nan_key) {
        const size_t n = hi - lo;
        if (n < 2) return;
        const uint32_t* L = c.list.data() + lo;
        Box cents{c.hc[L[0]], c.hc[L[0]]};                           // :75-79
        for (size_t i = 1; i < n; ++i) cents = surround(cents, Box{c.hc[L[i]], c.hc[L[i]]});
        const V3 spread = sub(cents.max, cents.min);
        const int axis = (spread.x > spread.y && spread.x > spread.z) ? 0
                       : (spread.y > spread.x && spread.y > spread.z) ? 1 : 2;  // :81-88
        KV* kv = c.kv.data() + lo;
        for (size_t i = 0; i < n; ++i) {
            const V3& ce = c.hc[L[i]];
            kv[i].k = axis == 0 ? ce.x : axis == 1 ? ce.y : ce.z;
            kv[i].h = L[i];
            if (!(kv[i].k == kv[i].k)) { nan_key.store(true); return; }  // partial_cmp(..).unwrap() panics (bvh.rs:99)
        }
        stable_sort_kv_mt(kv, c.tmp.data() + lo, n, spare);
        for (size_t i = 0; i < n; ++i) c.list[lo + i] = kv[i].h;    // first n/2 left, the rest right (:125-134)
        const size_t mid = lo + n / 2;
        fork_join(n, spare, [&] { sort_range(c, lo, mid, spare, nan_key); }, [&] { sort_range(c, mid, hi, spare, nan_key); });
    }
    Built build_range(BuildCtx& c, size_t lo, size_t hi) {
        const size_t n = hi - lo;
        if (n == 0) return Built{REF_EMPTY_VALUE, Box{}, false, 0};  // BvhNode::None
        if (n == 1) { Built b = c.coll->leaf(c.list[lo]); b.leaf = true; return b; }  // BvhNode::Object
        const size_t mid = lo + n / 2;
        uint32_t idx = alloc_node();  // pre-order index: a subtree is contiguous in memory
        Built l = build_range(c, lo, mid);
        if (!err.empty()) return l;
        Built r = build_range(c, mid, hi);
        if (!err.empty()) return r;
        // bounds of the Split (:68-72): the reference folds surround() over the whole list at every node; min / max are
        // associative, so the union of the two halves' bounds is the same box (up to the sign of a zero, which nothing
        // observes) at O(n) instead of O(n log n) for the tree
        const Box bounds = surround(mid - lo == 1 ? c.hb[c.list[lo]] : l.box, hi - mid == 1 ? c.hb[c.list[mid]] : r.box);
        write_node(idx, l, r);
        std::vector<GatedLeaf>* gated = c.gated;
        // a leaf child of this Split is gated by this Split's box
        // (two sibling primitives share the gate and sit in consecutive slots: one paired leaf, one box test)
        const uint32_t lk = RTR_REF_KIND(l.ref), rk = RTR_REF_KIND(r.ref);
        if (l.leaf && r.leaf && !l.split && !r.split && lk == rk && (lk == REF_TRI || lk == REF_SPHERE) &&
            RTR_REF_INDEX(r.ref) == RTR_REF_INDEX(l.ref) + 1) {
            gated->push_back(GatedLeaf{RTR_MAKE_REF(lk == REF_TRI ? REF_TRI2 : REF_SPHERE2, RTR_REF_INDEX(l.ref)), bounds, 1u});
        } else {
            if (l.leaf) gated->push_back(GatedLeaf{traversal_entry(l), bounds, l.fast_depth + (l.split ? 1u : 0u)});
            if (r.leaf) gated->push_back(GatedLeaf{traversal_entry(r), bounds, r.fast_depth + (r.split ? 1u : 0u)});
        }
        return Built{RTR_MAKE_REF(REF_NODE, idx), bounds, true, 1 + std::max(l.depth, r.depth)};
    }

    // wrap a root so its own box is tested on entry (see flat.h)
    uint32_t make_entry(const Built& root, uint32_t* depth) {
        *depth = root.depth;
        if (!root.split) return root.ref;
        uint32_t idx = alloc_node();
        write_node(idx, root, Built{REF_EMPTY_VALUE, Box{}, false, 0});
        *depth = root.depth + 1;
        return RTR_MAKE_REF(REF_NODE, idx);
    }

    // the same for the traversal tree: a nested collection (mesh) still tests its own root box on entry
    uint32_t traversal_entry(const Built& b) {
        if (!b.split) return b.fast_ref;
        uint32_t idx = alloc_node();
        write_node_raw(idx, b.fast_ref, b.box, REF_EMPTY_VALUE, Box{{0, 0, 0}, {0, 0, 0}});
        return RTR_MAKE_REF(REF_NODE, idx);
    }

    // ---- 4-wide traversal tree (flat.h) ------------------------------------------------------------------
    // Folds the 2-wide traversal tree below `entry`: a node takes its two child slots and, while it has a free slot,
    // replaces the interior child with the largest box by that child's own two slots (single-child pseudo parents are
    // looked through).  Each slot keeps the box the 2-wide tree tests for it.  Returns the 4-wide ref (a leaf ref as is).
    struct Slot { uint32_t ref; Box box; };
    void slots_of(uint32_t node, std::vector<Slot>& out) const {  // the (at most two) child slots of a 2-wide node
        const F4* q = &this->out.nodes[(size_t)(node - this->out.device_nodes) * 4];
        uint32_t l, r;
        std::memcpy(&l, &q[0].w, 4); std::memcpy(&r, &q[1].w, 4);
        if (l != REF_EMPTY_VALUE) out.push_back(Slot{l, Box{V3{q[0].x, q[0].y, q[0].z}, V3{q[1].x, q[1].y, q[1].z}}});
        if (r != REF_EMPTY_VALUE) out.push_back(Slot{r, Box{V3{q[2].x, q[2].y, q[2].z}, V3{q[3].x, q[3].y, q[3].z}}});
    }
    static uint32_t levels_of(uint32_t n) { uint32_t l = 0; while ((1ull << l) < n) ++l; return l; }
    const DeviceMesh* device_mesh_at(uint32_t node) const {  // the device-built mesh whose root Split this is
        for (const DeviceMesh& m : out.device_meshes) if (m.node_base == node || (m.hybrid && m.sah_base == node)) return &m;
        return nullptr;
    }
    uint32_t collapse4(uint32_t entry, uint32_t* depth) {
        if (entry == REF_EMPTY_VALUE || RTR_REF_KIND(entry) != REF_NODE) { *depth = entry == REF_EMPTY_VALUE ? 0u : 1u; return entry; }
        if (RTR_REF_INDEX(entry) < out.device_nodes) {  // the root Split of a device-built mesh: its 4-wide root
            const DeviceMesh* m = device_mesh_at(RTR_REF_INDEX(entry));
            *depth = m->hybrid ? m->depth4_estimate : levels_of(m->n_tris) / 2u + 2u;
            return RTR_MAKE_REF(REF_NODE, m->node4_base);
        }
        // a 2-wide node folds the same way wherever it is reached from: the scene tree and the per-object entries of the
        // linear scan share the 4-wide subtrees below the nodes they have in common
        const size_t local2 = RTR_REF_INDEX(entry) - out.device_nodes;
        if (memo4_ref.size() < out.nodes.size() / 4) { memo4_ref.resize(out.nodes.size() / 4, REF_EMPTY_VALUE); memo4_depth.resize(out.nodes.size() / 4, 0u); }
        if (memo4_ref[local2] != REF_EMPTY_VALUE) { *depth = memo4_depth[local2]; return memo4_ref[local2]; }
        std::vector<Slot> s;
        slots_of(RTR_REF_INDEX(entry), s);
        for (;;) {
            int pick = -1;
            float best = -1.0f;
            for (size_t i = 0; i < s.size(); ++i) {
                if (RTR_REF_KIND(s[i].ref) != REF_NODE || RTR_REF_INDEX(s[i].ref) < out.device_nodes) continue;
                std::vector<Slot> kids;
                slots_of(RTR_REF_INDEX(s[i].ref), kids);
                if (kids.size() <= 1) { pick = (int)i; break; }  // a single-child pseudo parent costs no slot: look through it
                if (s.size() + 1 > 4) continue;
                const float area = half_area(s[i].box);
                if (area > best) { pick = (int)i; best = area; }  // else: the largest interior box that still fits
            }
            if (pick < 0) break;
            std::vector<Slot> kids;
            slots_of(RTR_REF_INDEX(s[pick].ref), kids);
            s.erase(s.begin() + pick);
            s.insert(s.begin() + pick, kids.begin(), kids.end());
        }
        const size_t local = out.nodes4.size() / 8;
        const uint32_t idx = out.device_nodes4 + (uint32_t)local;
        out.nodes4.resize(out.nodes4.size() + 8, F4{0, 0, 0, 0});
        uint32_t refs[4] = {REF_EMPTY_VALUE, REF_EMPTY_VALUE, REF_EMPTY_VALUE, REF_EMPTY_VALUE};
        Box boxes[4] = {};
        uint32_t deepest = 0;
        for (size_t i = 0; i < s.size(); ++i) {
            uint32_t d = 0;
            refs[i] = collapse4(s[i].ref, &d);  // recursion may grow nodes4: index, never hold pointers across it
            boxes[i] = s[i].box;
            deepest = std::max(deepest, d);
        }
        F4* q = &out.nodes4[local * 8];
        for (int i = 0; i < 4; ++i) {
            q[2 * i] = F4{boxes[i].min.x, boxes[i].min.y, boxes[i].min.z, 0.0f};
            q[2 * i + 1] = F4{boxes[i].max.x, boxes[i].max.y, boxes[i].max.z, 0.0f};
        }
        q[0].w = bits(refs[0]); q[1].w = bits(refs[1]); q[2].w = bits(refs[2]); q[3].w = bits(refs[3]);
        *depth = deepest + 1;
        memo4_ref[local2] = RTR_MAKE_REF(REF_NODE, idx); memo4_depth[local2] = deepest + 1;
        return RTR_MAKE_REF(REF_NODE, idx);
    }
    std::vector<uint32_t> memo4_ref, memo4_depth;  // by 2-wide node: its 4-wide node and depth, once folded

    // ---- leaves -------------------------------------------------------------
    Built emit_sphere(uint32_t sphere_index, uint32_t object) {
        const rtr_sphere& s = d.spheres[sphere_index];
        uint32_t slot = (uint32_t)out.spheres.size();
        out.spheres.push_back(F4{s.center.x, s.center.y, s.center.z, s.radius});
        out.sphere_info.push_back(PrimInfo{object, 0u, s.material, rank++});
        return Built{RTR_MAKE_REF(REF_SPHERE, slot), Box{}, false, 1};
    }

    struct MeshColl : Collection {
        Flattener& f;
        const rtr_mesh& m;
        uint32_t object;
        MeshColl(Flattener& fl, const rtr_mesh& mesh, uint32_t obj) : f(fl), m(mesh), object(obj) {}
        Box bounds(uint32_t h) const override {  // mesh.rs:201-213
            const uint32_t* t = &m.indices[3 * (size_t)h];
            V3 a = v3(&m.positions[3 * (size_t)t[0]]), b = v3(&m.positions[3 * (size_t)t[1]]), c = v3(&m.positions[3 * (size_t)t[2]]);
            return epsilon_expand(surround(Box{a, a}, surround(Box{b, b}, Box{c, c})), 0.01f);
        }
        Built leaf(uint32_t h) override {
            const uint32_t* t = &m.indices[3 * (size_t)h];
            V3 p0 = v3(&m.positions[3 * (size_t)t[0]]), p1 = v3(&m.positions[3 * (size_t)t[1]]), p2 = v3(&m.positions[3 * (size_t)t[2]]);
            // face normal as Mesh::from_buffers computes it (mesh.rs:85-90)
            V3 fn = normalize(cross(sub(p2, p1), sub(p0, p1)));
            // edges as Triangle::hit computes them (mesh.rs:160-161)
            V3 e1 = sub(p1, p0), e2 = sub(p2, p0);
            uint32_t slot = f.out.device_tri_slots + (uint32_t)f.out.tri_info.size();
            f.out.tris.push_back(F4{p0.x, p0.y, p0.z, fn.x});
            f.out.tris.push_back(F4{e1.x, e1.y, e1.z, fn.y});
            f.out.tris.push_back(F4{e2.x, e2.y, e2.z, fn.z});
            for (int k = 0; k < 3; ++k) {
                const float* nn = &m.normals[3 * (size_t)t[k]];
                f.out.tri_vn.push_back(F4{nn[0], nn[1], nn[2], 0.0f});
            }
            {
                const PrimInfo info{object, h, m.material, f.rank++};
                F4 bits;
                std::memcpy(&bits, &info, sizeof(bits));
                f.out.tris.push_back(bits);
                f.out.tri_info.push_back(info);
            }
            return Built{RTR_MAKE_REF(REF_TRI, slot), Box{}, false, 1};
        }
    };

    Built emit_mesh(uint32_t mesh_index, uint32_t object) {
        const rtr_mesh& m = d.meshes[mesh_index];
        for (DeviceMesh& dm : out.device_meshes) {
            if (dm.object != object) continue;
            // Built on the device: reserve the DFS ranks of its triangles and hand the enclosing collection the root
            // Split (node_base; both walks share the device-built nodes) with the box from_list gives it (bvh.rs:68-72).
            dm.rank_base = rank;
            rank += dm.n_tris;
            MeshColl coll(*this, m, object);
            Box root = coll.bounds(0);
            for (uint32_t h = 1; h < dm.n_tris; ++h) root = surround(root, coll.bounds(h));
            Built b(RTR_MAKE_REF(REF_NODE, dm.node_base), root, true, levels_of(dm.n_tris) + 1u);
            if (dm.hybrid) { b.fast_ref = RTR_MAKE_REF(REF_NODE, dm.sah_base); b.fast_depth = levels_of(dm.n_tris) + dm.slack; }  // upper bound
            return b;
        }
        MeshColl coll(*this, m, object);
        {   // the mesh's records are appended one triangle at a time: make room once
            const size_t nt = m.n_indices / 3;
            out.tris.reserve(out.tris.size() + 4 * nt); out.tri_vn.reserve(out.tri_vn.size() + 3 * nt);
            out.tri_info.reserve(out.tri_info.size() + nt); out.nodes.reserve(out.nodes.size() + 4 * (2 * nt + 4));
        }
        std::vector<uint32_t> handles(m.n_indices / 3);
        for (size_t i = 0; i < handles.size(); ++i) handles[i] = (uint32_t)i;
        return build_collection(handles, coll);  // mesh.build_bvh() (mesh.rs:119-121)
    }

    Box object_bounds(rtr_object_ref ref) const {
        if (ref.kind == RTR_OBJ_SPHERE) {  // objects.rs:53-60
            const rtr_sphere& s = d.spheres[ref.index];
            return Box{V3{s.center.x - s.radius, s.center.y - s.radius, s.center.z - s.radius},
                       V3{s.center.x + s.radius, s.center.y + s.radius, s.center.z + s.radius}};
        }
        if (ref.kind == RTR_OBJ_MESH) {  // mesh.rs:99-108: seeded with AABB::default() = origin
            const rtr_mesh& m = d.meshes[ref.index];
            Box b{{0, 0, 0}, {0, 0, 0}};
            for (uint32_t i = 0; i < m.n_vertices; ++i) { V3 p = v3(&m.positions[3 * (size_t)i]); b = surround(b, Box{p, p}); }
            return b;
        }
        return object_bounds(d.volumes[ref.index].boundary);  // volume.rs:89-93
    }

    struct SceneColl : Collection {
        Flattener& f;
        std::vector<Box> cached;
        std::vector<Built> built;  // per object: its leaf as emitted (for the linear-scan table)
        explicit SceneColl(Flattener& fl) : f(fl) {
            cached.resize(f.d.n_objects);
            built.resize(f.d.n_objects, Built{REF_EMPTY_VALUE, Box{}, false, 0});
            for (uint32_t i = 0; i < f.d.n_objects; ++i) cached[i] = f.object_bounds(f.d.objects[i]);
        }
        Box bounds(uint32_t h) const override { return cached[h]; }
        Built leaf(uint32_t h) override { return built[h] = f.emit_object(f.d.objects[h], h); }
    };

    Built emit_object(rtr_object_ref ref, uint32_t object) {
        if (ref.kind == RTR_OBJ_SPHERE) return emit_sphere(ref.index, object);
        if (ref.kind == RTR_OBJ_MESH) return emit_mesh(ref.index, object);
        // Volume: the leaf is the volume; its boundary is a separate tree queried twice per visit
        const rtr_volume& v = d.volumes[ref.index];
        VolumeRec rec;
        std::memset(&rec, 0, sizeof(rec));
        rec.rank = rank++;
        rtr_object_ref geom = v.boundary;  // the geometry at the end of the boundary chain: bounds() and material() forward to it
        while (geom.kind == RTR_OBJ_VOLUME) geom = d.volumes[geom.index].boundary;  // validated: terminates
        if (v.boundary.kind == RTR_OBJ_VOLUME) {
            // A Volume bounded by a Volume never hits.  Its first query is boundary.hit(ray, -inf, +inf) (volume.rs:36),
            // i.e. the inner Volume::hit with tmin = -inf: there the second boundary query starts at t1 + (-inf) = -inf
            // and returns the first one's hit again, so `hit1.t >= hit2.t` (volume.rs:47-49) ends it with None before any
            // random number is drawn — and the outer `?` propagates the None.  The object still takes part in the
            // BVH build through its bounds (volume.rs:89-93), so it is a leaf with an EMPTY boundary entry.
            rec.entry = rec.entry_reference = REF_EMPTY_VALUE;
        } else {
            uint32_t bdepth = 0;
            Built b = emit_object(v.boundary, object);
            rec.entry_reference = make_entry(b, &bdepth);
            rec.entry = traversal_entry(b);
            out.max_boundary_depth = std::max(out.max_boundary_depth, std::max(bdepth, b.fast_depth + 1));
        }
        rec.neg_inv_density = v.neg_inv_density;  // as stored by Volume::new (volume.rs:26-31)
        rec.material = geom.kind == RTR_OBJ_SPHERE ? d.spheres[geom.index].material : d.meshes[geom.index].material;
        rec.object = object;
        rec.index = ref.index;
        uint32_t vi = (uint32_t)out.volumes.size();
        out.volumes.push_back(rec);
        return Built{RTR_MAKE_REF(REF_VOLUME, vi), Box{}, false, 1};
    }
};

bool validate(const rtr_scene_desc& d, std::string& err) {
    auto bad = [&](const char* m) { err = m; return false; };
    if ((d.n_objects && !d.objects) || (d.n_spheres && !d.spheres) || (d.n_meshes && !d.meshes) ||
        (d.n_volumes && !d.volumes) || (d.n_materials && !d.materials))
        return bad("null array with non-zero count");
    auto ref_ok = [&](rtr_object_ref r, bool allow_volume) {
        if (r.kind == RTR_OBJ_SPHERE) return r.index < d.n_spheres;
        if (r.kind == RTR_OBJ_MESH) return r.index < d.n_meshes;
        if (r.kind == RTR_OBJ_VOLUME) return allow_volume && r.index < d.n_volumes;
        return false;
    };
    for (uint32_t i = 0; i < d.n_objects; ++i)
        if (!ref_ok(d.objects[i], true)) return bad("object reference out of range");
    for (uint32_t i = 0; i < d.n_volumes; ++i) {
        // boundary: Box<dyn Object> (volume.rs:20-23) — may again be a Volume; the index-based chain must end in geometry
        rtr_object_ref b = d.volumes[i].boundary;
        for (uint32_t steps = 0; b.kind == RTR_OBJ_VOLUME; ++steps) {
            if (b.index >= d.n_volumes || steps > d.n_volumes) return bad("volume boundary chain is cyclic or out of range");
            b = d.volumes[b.index].boundary;
        }
        if (!ref_ok(b, false)) return bad("volume boundary out of range");
    }
    for (uint32_t i = 0; i < d.n_spheres; ++i)
        if (d.spheres[i].material >= d.n_materials) return bad("sphere material handle out of range");
    for (uint32_t i = 0; i < d.n_meshes; ++i) {
        const rtr_mesh& m = d.meshes[i];
        if (m.material >= d.n_materials) return bad("mesh material handle out of range");
        if ((m.n_vertices && (!m.positions || !m.normals)) || (m.n_indices && !m.indices)) return bad("mesh with null buffers");
        for (uint32_t k = 0; k < m.n_indices; ++k)
            if (m.indices[k] >= m.n_vertices) return bad("mesh vertex index out of range");
    }
    for (uint32_t i = 0; i < d.n_materials; ++i) {
        const rtr_material& m = d.materials[i];
        if (m.kind > RTR_MAT_ISOTROPIC) return bad("unknown material kind");
        if (m.kind == RTR_MAT_MIX && (m.first >= d.n_materials || m.second >= d.n_materials)) return bad("MixMaterial child out of range");
    }
    // a MixMaterial cycle would recurse forever in the reference; reject
    for (uint32_t i = 0; i < d.n_materials; ++i) {
        if (d.materials[i].kind != RTR_MAT_MIX) continue;
        std::vector<uint32_t> stack{i};
        uint32_t steps = 0;
        while (!stack.empty()) {
            uint32_t k = stack.back(); stack.pop_back();
            if (++steps > 4 * d.n_materials + 8) return bad("MixMaterial cycle or nesting too deep");
            if (d.materials[k].kind == RTR_MAT_MIX) { stack.push_back(d.materials[k].first); stack.push_back(d.materials[k].second); }
        }
    }
    if (d.background.kind > RTR_BG_SKYMAP) return bad("unknown background kind");
    if (d.background.kind == RTR_BG_SKYMAP && (!d.background.texels || !d.background.width || !d.background.height))
        return bad("SkyMap without texels");
    return true;
}

}  // namespace

uint32_t count_nodes4(uint32_t n) {
    // bvh_build.cu k_flag4: the Splits at even depth holding three or more leaves, plus the root.  A median-split subtree
    // is determined by its size, and one level holds at most two distinct sizes: walk the levels with (size, count) pairs.
    if (n < 2) return 0;
    uint64_t total = 0;
    std::vector<std::pair<uint32_t, uint64_t>> level{{n, 1}}, next;
    for (uint32_t depth = 0; !level.empty(); ++depth) {
        next.clear();
        for (const auto& sc : level) {
            if (sc.first < 2) continue;
            if ((depth & 1u) == 0u && (sc.first >= 3 || depth == 0)) total += sc.second;
            const uint32_t parts[2] = {sc.first / 2, sc.first - sc.first / 2};
            for (uint32_t part : parts) {
                bool merged = false;
                for (auto& e : next) if (e.first == part) { e.second += sc.second; merged = true; break; }
                if (!merged) next.push_back({part, sc.second});
            }
        }
        level.swap(next);
    }
    return (uint32_t)total;
}

uint32_t count_gated(uint32_t n) {
    if (n == 0) return 0;
    uint64_t total = 0;
    std::vector<std::pair<uint32_t, uint64_t>> level{{n, 1}}, next;
    while (!level.empty()) {
        next.clear();
        for (const auto& sc : level) {
            if (sc.first <= 2) { total += sc.second; continue; }  // a single leaf, or a 2-leaf Split folded into one paired leaf
            const uint32_t parts[2] = {sc.first / 2, sc.first - sc.first / 2};
            for (uint32_t part : parts) {
                bool merged = false;
                for (auto& e : next) if (e.first == part) { e.second += sc.second; merged = true; break; }
                if (!merged) next.push_back({part, sc.second});
            }
        }
        level.swap(next);
    }
    return (uint32_t)total;
}

void build_traversal_over_gates(const F4* gated, uint32_t n_positions, uint32_t node_base, uint32_t slack, std::vector<F4>& nodes,
                                uint32_t* depth) {
    HostFlat tmp;
    tmp.device_nodes = node_base;  // the Flattener numbers its nodes from here
    rtr_scene_desc none;
    std::memset(&none, 0, sizeof(none));
    std::string err;
    Flattener f(tmp, none, err);
    f.slack = slack;
    std::vector<GatedLeaf> L;
    L.reserve(n_positions);
    for (uint32_t p = 0; p < n_positions; ++p) {
        uint32_t ref;
        std::memcpy(&ref, &gated[2 * (size_t)p].w, 4);
        if (ref == REF_EMPTY_VALUE) continue;
        const F4& mn = gated[2 * (size_t)p]; const F4& mx = gated[2 * (size_t)p + 1];
        L.push_back(GatedLeaf{ref, Box{V3{mn.x, mn.y, mn.z}, V3{mx.x, mx.y, mx.z}}, 1u});
    }
    uint32_t levels = 0;
    while ((size_t(1) << levels) < n_positions) ++levels;
    const Flattener::FastOut o = f.build_traversal(L, 0, L.size(), levels + slack);  // as build_collection does for a host-built mesh
    nodes.swap(tmp.nodes);
    *depth = o.depth;
}

void collapse_scene4(const HostFlat& flat, const std::vector<const std::vector<F4>*>& mesh_sah, std::vector<F4>& nodes4,
                     uint32_t* root_entry4, std::vector<VolumeRec>& volumes, std::vector<ObjectRec>& objects, uint32_t* max_depth4) {
    HostFlat tmp;  // every 2-wide traversal node under its own index: nothing is external to this fold
    tmp.nodes.assign(((size_t)flat.device_nodes) * 4 + flat.nodes.size(), F4{0, 0, 0, 0});
    std::memcpy(tmp.nodes.data() + (size_t)flat.device_nodes * 4, flat.nodes.data(), flat.nodes.size() * sizeof(F4));
    for (size_t k = 0; k < flat.device_meshes.size(); ++k)
        if (flat.device_meshes[k].hybrid && mesh_sah[k])
            std::memcpy(tmp.nodes.data() + (size_t)flat.device_meshes[k].sah_base * 4, mesh_sah[k]->data(), mesh_sah[k]->size() * sizeof(F4));
    rtr_scene_desc none;
    std::memset(&none, 0, sizeof(none));
    std::string err;
    Flattener f(tmp, none, err);
    volumes = flat.volumes; objects = flat.objects;
    uint32_t d4 = 0, v_depth4 = 0;
    *root_entry4 = f.collapse4(flat.root_entry, &d4);
    *max_depth4 = d4;
    for (VolumeRec& v : volumes) { v.entry4 = f.collapse4(v.entry, &d4); v_depth4 = std::max(v_depth4, d4); }
    for (ObjectRec& o : objects) { o.entry4 = f.collapse4(o.entry, &d4); *max_depth4 = std::max(*max_depth4, d4); }
    *max_depth4 += v_depth4;
    nodes4.swap(tmp.nodes4);
}

int flatten_scene(const rtr_scene_desc& d, HostFlat& out, std::string& err, uint32_t slack, uint32_t device_mesh_min_tris, bool hybrid) {
    if (!validate(d, err)) return err.find("unsupported") != std::string::npos ? RTR_UNSUPPORTED : RTR_INVALID_ARG;
    if (device_mesh_min_tris) {
        // every object that is a mesh, or a Volume bounded directly by one, and is big enough: its index ranges
        for (uint32_t i = 0; i < d.n_objects; ++i) {
            rtr_object_ref r = d.objects[i];
            if (r.kind == RTR_OBJ_VOLUME) r = d.volumes[r.index].boundary;  // a Volume bounded by a Volume emits no boundary
            if (r.kind != RTR_OBJ_MESH) continue;
            const rtr_mesh& m = d.meshes[r.index];
            const uint32_t n = m.n_indices / 3;
            if (n < (hybrid ? 3u : 2u) || n < device_mesh_min_tris) continue;
            DeviceMesh dm{r.index, i, m.material, n, out.device_tri_slots, out.device_nodes, out.device_nodes4, 0u, 0u, 0u, 0u, slack, 0u};
            out.device_tri_slots += n; out.device_nodes += n;
            if (hybrid) {
                dm.hybrid = 1u;
                dm.n_gated = count_gated(n);
                dm.sah_base = out.device_nodes;
                out.device_nodes += dm.n_gated - 1u;    // a binary tree over n_gated entries; the 4-wide tree of a hybrid
                                                        // scene is folded as a whole at upload time (collapse_scene4)
                uint32_t levels = 0;
                while ((1ull << levels) < n) ++levels;
                dm.depth4_estimate = (levels + slack) / 2u + 2u;  // checked against the tree actually built at upload time
            } else {
                out.device_nodes4 += count_nodes4(n);
            }
            out.device_meshes.push_back(dm);
        }
    }
    Flattener f(out, d, err);
    f.slack = slack;
    out.nodes.reserve(64);
    Flattener::SceneColl coll(f);
    std::vector<uint32_t> handles(d.n_objects);
    for (uint32_t i = 0; i < d.n_objects; ++i) handles[i] = i;
    Built root = f.build_collection(handles, coll);  // scene.build_bvh() (scene.rs:112-114)
    if (!err.empty()) return RTR_INVALID_ARG;
    uint32_t depth = 0;
    out.root_entry_reference = f.make_entry(root, &depth);
    out.root_entry = f.traversal_entry(root);
    out.max_depth = std::max(depth, root.fast_depth + 1);

    // linear-scan table (scene.rs:141-156): one entry per object, in add_object order.  A mesh
    // root that is a Split needs its own box tested on entry -> pseudo parent per mesh object.
    uint32_t v_depth4 = 0;
    out.objects.resize(d.n_objects);
    for (uint32_t i = 0; i < d.n_objects; ++i) {
        std::memset(&out.objects[i], 0, sizeof(ObjectRec));
        uint32_t dd = 0;
        out.objects[i].entry_reference = f.make_entry(coll.built[i], &dd);
        out.objects[i].entry = f.traversal_entry(coll.built[i]);
    }
    // 4-wide traversal trees of the scene, of every volume boundary and of every object (linear scan entries)
    {
        uint32_t d4 = 0;
        out.root_entry4 = f.collapse4(out.root_entry, &d4);
        out.max_depth4 = d4;
        for (VolumeRec& v : out.volumes) { v.entry4 = f.collapse4(v.entry, &d4); v_depth4 = std::max(v_depth4, d4); }
        for (ObjectRec& o : out.objects) { o.entry4 = f.collapse4(o.entry, &d4); out.max_depth4 = std::max(out.max_depth4, d4); }
        out.max_depth4 += v_depth4;  // a boundary walk is stacked on top of the suspended scene walk
    }
    out.materials.resize(d.n_materials);
    for (uint32_t i = 0; i < d.n_materials; ++i) {
        const rtr_material& m = d.materials[i];
        MaterialRec r;
        std::memset(&r, 0, sizeof(r));
        r.kind = m.kind; r.r = m.color.x; r.g = m.color.y; r.b = m.color.z;
        r.param = m.param; r.first = m.first; r.second = m.second;
        out.materials[i] = r;
    }
    std::memset(&out.background, 0, sizeof(out.background));
    out.background.kind = d.background.kind;
    out.background.r = d.background.color.x; out.background.g = d.background.color.y; out.background.b = d.background.color.z;
    out.background.br = d.background.bottom.x; out.background.bg = d.background.bottom.y; out.background.bb = d.background.bottom.z;
    if (d.background.kind == RTR_BG_SKYMAP) {
        out.background.width = d.background.width; out.background.height = d.background.height;
        // texels are expanded RGB -> float4 on the device (capi.cpp: expand_env), not here: the host loop over a
        // 4096x2048 map cost more than the whole H2D copy
    }
    return RTR_OK;
}

}  // namespace rtr
