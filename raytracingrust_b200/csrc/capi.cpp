// capi.cpp — implementation of the C ABI declared in include/rtr.h.
//
// No CPU fallback exists: every entry point that computes launches CUDA kernels (kernels.cu)
// and returns RTR_CUDA_ERROR when no device is usable.
#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <vector>

#include "../../include/rtr.h"
#include "bvh_build.h"
#include "context.h"
#include "host_flatten.h"
#include "kernels.h"

using namespace rtr;

static std::string g_last_error;
static std::mutex g_err_mu;

namespace rtr {
rtr_status fail(rtr_context* ctx, rtr_status st, const std::string& msg) {
    if (ctx) ctx->error = msg;
    std::lock_guard<std::mutex> lk(g_err_mu);
    g_last_error = msg;
    return st;
}
void set_global_error(const std::string& m) { fail(nullptr, RTR_OK, m); }
rtr_status cuda_fail(rtr_context* ctx, cudaError_t e, const char* what) {
    return fail(ctx, e == cudaErrorMemoryAllocation ? RTR_OOM : RTR_CUDA_ERROR,
                std::string(what) + ": " + cudaGetErrorString(e));
}
}  // namespace rtr

// ------------------------------------------------------------------ camera (camera.rs:37-56)
// host f32 arithmetic in the reference's order (this TU is built with -ffp-contract=off)
namespace {
inline v3 vsub(v3 a, v3 b) { return v3{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline v3 vadd(v3 a, v3 b) { return v3{a.x + b.x, a.y + b.y, a.z + b.z}; }
inline v3 vscale(float s, v3 a) { return v3{s * a.x, s * a.y, s * a.z}; }
inline v3 vdiv(v3 a, float s) { return v3{a.x / s, a.y / s, a.z / s}; }
inline float vdot(v3 a, v3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
inline v3 vcross(v3 a, v3 b) { return v3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
inline v3 vnormalize(v3 a) { float s = 1.0f / std::sqrt(vdot(a, a)); return v3{a.x * s, a.y * s, a.z * s}; }
inline v3 from(const rtr_vec3& v) { return v3{v.x, v.y, v.z}; }

CameraBasis ray_origin(const rtr_camera& cam) {
    const float PI = 3.14159265358979323846f;
    float theta = cam.vertical_fov * PI / 180.0f;  // degrees_to_radians (math.rs:32-34)
    float h = std::tan(theta / 2.0f);
    float viewport_height = 2.0f * h;
    float viewport_width = cam.aspect_ratio * viewport_height;
    v3 w = vnormalize(vsub(from(cam.lookfrom), from(cam.lookat)));
    v3 u = vnormalize(vcross(from(cam.vertical), w));
    v3 v = vcross(w, u);
    v3 horizontal = vscale(viewport_width, u);
    v3 vertical = vscale(viewport_height, v);
    CameraBasis b;
    b.origin = from(cam.lookfrom);
    b.horizontal = horizontal;
    b.vertical = vertical;
    b.lower_left_corner = vsub(vadd(vsub(from(cam.lookfrom), vdiv(horizontal, 2.0f)), vdiv(vertical, 2.0f)), w);
    return b;
}
}  // namespace

extern "C" {

uint32_t rtr_abi_version(void) { return RTR_ABI_VERSION; }

const char* rtr_last_error(const rtr_context* ctx) {
    if (ctx) return ctx->error.c_str();
    // a per-thread copy: another thread's failure may reallocate the global string while the caller reads this one
    static thread_local std::string copy;
    std::lock_guard<std::mutex> lk(g_err_mu);
    copy = g_last_error;
    return copy.c_str();
}

static rtr_status context_init(rtr_context* ctx, int device, void* stream) {
    ctx->device = device;
    CU(cudaSetDevice(device));
    if (stream) { ctx->stream = (cudaStream_t)stream; ctx->own_stream = false; }
    else { CU(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking)); ctx->own_stream = true; }
    CU(cudaEventCreate(&ctx->ev0));
    CU(cudaEventCreate(&ctx->ev1));
    for (cudaStream_t& ls : ctx->lane_stream) CU(cudaStreamCreateWithFlags(&ls, cudaStreamNonBlocking));
    for (cudaEvent_t& ev : ctx->wf_ev_wave) CU(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&ctx->wf_ev_fork, cudaEventDisableTiming));
    CU(cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device));
    {
        cudaMemPoolProps props;
        std::memset(&props, 0, sizeof(props));
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        CU(cudaMemPoolCreate(&ctx->pool, &props));
        unsigned long long keep = ~0ull;
        CU(cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep));
    }
    if (const char* e = std::getenv("RTR_REFILL_THRESHOLD")) {
        int v = std::atoi(e);
        ctx->refill_threshold = (uint32_t)(v < 1 ? 1 : v > 32 ? 32 : v);
    }
    if (const char* e = std::getenv("RTR_WF_REFILL_THRESHOLD")) {
        int v = std::atoi(e);
        ctx->wf_refill_threshold = (uint32_t)(v < 1 ? 1 : v > 32 ? 32 : v);
    }
    if (const char* e = std::getenv("RTR_WF_PATHS")) {
        long long v = std::atoll(e);
        if (v >= 1024) ctx->wf_target_paths = (uint64_t)v;
    }
    if (const char* e = std::getenv("RTR_WF_GRAPH")) ctx->wf_use_graph = std::atoi(e);
    if (const char* e = std::getenv("RTR_WF_ATOMICS")) ctx->wf_atomics_mode = (uint32_t)std::atoi(e) & 3u;
    if (const char* e = std::getenv("RTR_WF_BIN")) {  // 0 = off, 8 = octant, 24 = octant x dominant axis
        const int v = std::atoi(e);
        ctx->wf_bin_keys = v <= 0 ? 0u : v <= 8 ? 8u : 24u;
    }
    if (const char* e = std::getenv("RTR_WF_WIDE")) ctx->wf_wide4 = std::atoi(e) != 0;
    if (const char* e = std::getenv("RTR_WF_LANES")) ctx->wf_lanes = (uint32_t)std::min<int>(std::max(1, std::atoi(e)), (int)rtr_context::kMaxLanes);
    if (const char* e = std::getenv("RTR_WF_LANES_MIN_DEPTH")) ctx->wf_lanes_min_depth = (uint32_t)std::max(0, std::atoi(e));
    if (const char* e = std::getenv("RTR_WF_BIN_LAST")) ctx->wf_bin_last_bounce = (uint32_t)std::max(0, std::atoi(e));
    if (const char* e = std::getenv("RTR_SCAN_KERNEL")) ctx->use_scan_kernel = std::atoi(e);
    CU(cudaMalloc((void**)&ctx->work_counter, sizeof(uint32_t)));
    CU(cudaMalloc((void**)&ctx->wf_counters, rtr_context::kMaxLanes * WFC_WORDS * sizeof(uint32_t)));
    CU(cudaMalloc((void**)&ctx->stats, STAT_COUNT * sizeof(unsigned long long)));
    return RTR_OK;
}

rtr_status rtr_context_create(int device, void* stream, rtr_context** out) {
    if (!out) return fail(nullptr, RTR_INVALID_ARG, "rtr_context_create: out is NULL");
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(nullptr, RTR_CUDA_ERROR, std::string("no CUDA device: ") + cudaGetErrorString(e) +
                                                 " (this library has no CPU fallback)");
    if (device < 0 || device >= n) return fail(nullptr, RTR_INVALID_ARG, "rtr_context_create: bad device index");
    rtr_context* ctx = new (std::nothrow) rtr_context();
    if (!ctx) return fail(nullptr, RTR_OOM, "host allocation failed");
    const rtr_status st = context_init(ctx, device, stream);
    if (st != RTR_OK) {  // release whatever was created before the failing call; the message stays in the global slot
        (void)cudaGetLastError();
        rtr_context_destroy(ctx);
        return st;
    }
    *out = ctx;
    return RTR_OK;
}

void rtr_context_destroy(rtr_context* ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (cudaStream_t ls : ctx->lane_stream) if (ls) { cudaStreamSynchronize(ls); cudaStreamDestroy(ls); }
    for (cudaEvent_t ev : ctx->wf_ev_wave) if (ev) cudaEventDestroy(ev);
    if (ctx->wf_ev_fork) cudaEventDestroy(ctx->wf_ev_fork);
    cudaFree(ctx->wf_state); cudaFree(ctx->wf_counters);
    cudaFree(ctx->work_counter); cudaFree(ctx->stats); cudaFree(ctx->accum); cudaFree(ctx->partials);
    cudaFree(ctx->rgba); cudaFree(ctx->batch_in); cudaFree(ctx->batch_in2); cudaFree(ctx->batch_out);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

rtr_status rtr_synchronize(rtr_context* ctx) {
    if (!ctx) return fail(nullptr, RTR_INVALID_ARG, "null context");
    CU(cudaSetDevice(ctx->device));
    CU(cudaStreamSynchronize(ctx->stream));
    return RTR_OK;
}

}  // extern "C"

// ------------------------------------------------------------------ scene upload
template <typename T>
static cudaError_t upload(rtr_scene* s, const std::vector<T>& v, const T** dst, cudaStream_t st, size_t prefix = 0) {
    // `prefix` elements are left in front of the host data for records written on the device
    *dst = nullptr;
    size_t bytes = (prefix + v.size()) * sizeof(T);
    if (bytes == 0) bytes = sizeof(T);  // keep pointers valid
    void* p = nullptr;
    cudaError_t e = cudaMallocFromPoolAsync(&p, bytes, s->ctx->pool, st);
    if (e != cudaSuccess) return e;
    s->allocs.push_back(p);
    s->device_bytes += bytes;
    if (!v.empty()) {
        e = cudaMemcpyAsync((T*)p + prefix, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st);
        s->h2d_bytes += v.size() * sizeof(T);
    }
    *dst = (const T*)p;
    return e;
}

namespace rtr {

// Traversal-stack budget per thread (device_core.cuh RTR_STACK_CAP): scene levels + the deepest volume-boundary tree.
// The traversal tree is a depth-capped SAH tree (host_flatten.cpp: balanced depth + `slack` levels); when a scene
// would not fit the stack with the default slack it is flattened again with slack 0 — every collection then has
// its minimal depth ceil(log2 n) + 1, and with 29-bit primitive indices three nested collections (scene, mesh,
// volume boundary) always fit.  The reference recurses without a limit (bvh.rs:146-173).
static uint32_t stack_limit() {
    uint32_t lim = 96;  // == RTR_STACK_CAP
    if (const char* e = std::getenv("RTR_STACK_LIMIT")) {  // tests lower it to exercise the re-flatten path
        const int v = std::atoi(e);
        if (v >= 4 && v < 96) lim = (uint32_t)v;
    }
    return lim;
}

rtr_status prepare_scene(rtr_context* ctx, const rtr_scene_desc* desc, PreparedScene& out) {
    auto t0 = std::chrono::steady_clock::now();
    // Sphere scenes can have their BVH built on the device (bvh_build.cu): RTR_BVH_BUILD = host | device | auto
    // (auto: from 2048 objects up, where the level-by-level launches beat the host's recursive sort).
    {
        const char* mode = std::getenv("RTR_BVH_BUILD");
        const bool applies = device_build_applies(*desc);
        if (mode && std::strcmp(mode, "device") == 0) out.on_device = applies;
        else if (mode && std::strcmp(mode, "host") == 0) out.on_device = false;
        else out.on_device = applies && desc->n_objects >= 2048;
    }
    rtr_scene_desc host_desc = *desc;
    if (out.on_device) { host_desc.objects = nullptr; host_desc.n_objects = 0; }  // materials / background still flatten on the host
    // Triangle meshes: RTR_BVH_BUILD=device also builds every mesh's BVH on the device (bvh_build.cu build_mesh_on_device).
    // Not the default: the device builds the reference's median-split topology, the host a SAH tree over the same gates
    // that needs about half the node visits per ray — a faster upload against a slower render.
    // RTR_BVH_BUILD=hybrid: the device builds each mesh's reference tree (the O(n log^2 n) sorts, gates, ranks, records) and
    // the host builds the SAH traversal tree over the gates it reads back — the device's upload speed for the sorts, the
    // host tree's render speed.
    // auto: meshes of 65 536 triangles or more go the hybrid way (same tree, same render speed, upload 2-2.6x faster than
    // the threaded host build: profiles/r2/ab_hybrid_mesh_build.txt); smaller ones are built on the host in a few ms.
    uint32_t device_mesh_min_tris = 65536;
    bool hybrid = true;
    {
        const char* mode = std::getenv("RTR_BVH_BUILD");
        if (mode && std::strcmp(mode, "device") == 0) hybrid = false;
        if (mode && std::strcmp(mode, "device") == 0) device_mesh_min_tris = 2;
        if (mode && std::strcmp(mode, "hybrid") == 0) { device_mesh_min_tris = 3; hybrid = true; }
        if (const char* e = std::getenv("RTR_BVH_DEVICE_MESH_MIN_TRIS")) device_mesh_min_tris = (uint32_t)std::max(0, std::atoi(e));
        if (mode && std::strcmp(mode, "host") == 0) device_mesh_min_tris = 0;
    }
    for (int attempt = 0; attempt < 2; ++attempt) {
        out.flat = HostFlat();
        std::string err;
        int rc = RTR_OK;
        try {
            rc = flatten_scene(host_desc, out.flat, err, attempt == 0 ? 8u : 0u, device_mesh_min_tris, hybrid);
        } catch (const std::bad_alloc&) {
            return fail(ctx, RTR_OOM, "host allocation failed while flattening");
        }
        if (rc != RTR_OK) return fail(ctx, (rtr_status)rc, "rtr_scene_upload: " + err);
        const uint32_t need = out.flat.max_depth + out.flat.max_boundary_depth + 4;  // + the suspended outer hit and t1 of a volume query (wavefront.cu)
        if (need <= stack_limit()) break;
        if (attempt == 1) return fail(ctx, RTR_UNSUPPORTED, "BVH too deep for the traversal stack even when balanced");
    }
    // the 4-wide tree pushes up to three siblings per level; a scene whose worst case would not fit keeps the 2-wide walk
    if (3u * out.flat.max_depth4 + 4u > stack_limit()) { out.flat.nodes4.clear(); out.flat.device_nodes4 = 0; out.flat.root_entry4 = 0xFFFFFFFFu; }
    out.host_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    return RTR_OK;
}

rtr_status upload_prepared(rtr_context* ctx, const rtr_scene_desc* desc, const PreparedScene& prep, rtr_scene** out) {
    *out = nullptr;
    CU(cudaSetDevice(ctx->device));
    auto t0 = std::chrono::steady_clock::now();
    const HostFlat& flat = prep.flat;
    const bool on_device = prep.on_device;
    rtr_scene* s = new (std::nothrow) rtr_scene();
    if (!s) return fail(ctx, RTR_OOM, "host allocation failed");
    s->ctx = ctx;
    s->n_nodes = flat.device_nodes + (uint32_t)(flat.nodes.size() / 4);
    s->max_depth = flat.max_depth;
    SceneView& v = s->view;
    std::memset(&v, 0, sizeof(v));
    cudaError_t e = cudaSuccess;
    auto step = [&](cudaError_t r) { if (e == cudaSuccess) e = r; };
    // device-built meshes own the front of the node and triangle arrays (host_flatten.h DeviceMesh); host records follow
    const bool wide4 = ctx->wf_wide4 != 0u && flat.nodes4.size() + flat.device_nodes4 > 0;
    step(upload(s, flat.nodes, &v.nodes, ctx->stream, (size_t)flat.device_nodes * 4));
    if (wide4) step(upload(s, flat.nodes4, &v.nodes4, ctx->stream, (size_t)flat.device_nodes4 * 8));
    step(upload(s, flat.spheres, &v.spheres, ctx->stream));
    step(upload(s, flat.sphere_info, &v.sphere_info, ctx->stream));
    step(upload(s, flat.tris, &v.tris, ctx->stream, (size_t)flat.device_tri_slots * 4));
    step(upload(s, flat.tri_vn, &v.tri_vn, ctx->stream, (size_t)flat.device_tri_slots * 3));
    bool hybrid = false;
    std::vector<std::vector<F4>> mesh_sah(flat.device_meshes.size());  // hybrid: every mesh's host-built SAH tree, kept for the fold below
    for (size_t k = 0; k < flat.device_meshes.size(); ++k) {
        const DeviceMesh& dm = flat.device_meshes[k];
        if (e != cudaSuccess) break;
        bool nan_centroid = false, count_mismatch = false;
        std::vector<F4> gated;
        step(build_mesh_on_device(desc->meshes[dm.mesh_index], dm, ctx->stream, ctx->pool, const_cast<F4*>(v.nodes),
                                  dm.hybrid ? nullptr : const_cast<F4*>(v.nodes4), const_cast<F4*>(v.tris), const_cast<F4*>(v.tri_vn),
                                  &s->h2d_bytes, &nan_centroid, &count_mismatch, dm.hybrid ? &gated : nullptr));
        if (e == cudaSuccess && nan_centroid) { rtr_scene_destroy(s); return fail(ctx, RTR_INVALID_ARG, "rtr_scene_upload: NaN centroid (the reference panics at bvh.rs:99)"); }
        if (e == cudaSuccess && count_mismatch) { rtr_scene_destroy(s); return fail(ctx, RTR_CUDA_ERROR, "rtr_scene_upload: device mesh build emitted an unexpected number of 4-wide nodes"); }
        if (e == cudaSuccess && dm.hybrid) {  // the host's SAH tree over the device-built gates, into the range reserved for it
            hybrid = true;
            uint32_t depth = 0;
            try {
                build_traversal_over_gates(gated.data(), dm.n_tris, dm.sah_base, dm.slack, mesh_sah[k], &depth);
            } catch (const std::bad_alloc&) {
                rtr_scene_destroy(s);
                return fail(ctx, RTR_OOM, "host allocation failed while building the traversal tree");
            }
            if (mesh_sah[k].size() / 4 != (size_t)dm.n_gated - 1u) {
                rtr_scene_destroy(s);
                return fail(ctx, RTR_CUDA_ERROR, "rtr_scene_upload: the device-built gates do not match the reserved traversal-tree size");
            }
            step(cudaMemcpyAsync(const_cast<F4*>(v.nodes) + (size_t)dm.sah_base * 4, mesh_sah[k].data(), mesh_sah[k].size() * sizeof(F4), cudaMemcpyHostToDevice, ctx->stream));
            s->h2d_bytes += mesh_sah[k].size() * sizeof(F4);
        }
        s->built_on_device = true;
    }
    // hybrid: the 4-wide tree is folded now, over the scene level and the meshes' SAH trees together, exactly as a
    // host-built scene's is (a mesh's top folds into the scene's nodes); volumes / objects carry the new entry refs
    std::vector<F4> nodes4_hybrid;
    std::vector<VolumeRec> volumes_hybrid;
    std::vector<ObjectRec> objects_hybrid;
    uint32_t root_entry4_hybrid = 0xFFFFFFFFu;
    if (hybrid && e == cudaSuccess) {
        std::vector<const std::vector<F4>*> sah_of(mesh_sah.size(), nullptr);
        for (size_t k = 0; k < mesh_sah.size(); ++k) sah_of[k] = &mesh_sah[k];
        uint32_t depth4 = 0;
        try {
            collapse_scene4(flat, sah_of, nodes4_hybrid, &root_entry4_hybrid, volumes_hybrid, objects_hybrid, &depth4);
        } catch (const std::bad_alloc&) {
            rtr_scene_destroy(s);
            return fail(ctx, RTR_OOM, "host allocation failed while folding the traversal tree");
        }
        v.nodes4 = nullptr;
        if (ctx->wf_wide4 != 0u && !nodes4_hybrid.empty() && 3u * depth4 + 4u <= stack_limit()) step(upload(s, nodes4_hybrid, &v.nodes4, ctx->stream));
        step(cudaStreamSynchronize(ctx->stream));  // mesh_sah / nodes4_hybrid are read by the copies above
    }
    const std::vector<VolumeRec>& volumes_up = hybrid ? volumes_hybrid : flat.volumes;
    const std::vector<ObjectRec>& objects_up = hybrid ? objects_hybrid : flat.objects;
    step(upload(s, volumes_up, &v.volumes, ctx->stream));
    step(upload(s, objects_up, &v.objects, ctx->stream));
    step(upload(s, flat.materials, &v.materials, ctx->stream));
    {   // SkyMap texels: raw RGB goes up as it is, the float4 layout is produced on the device
        const size_t n_tex = desc->background.kind == RTR_BG_SKYMAP ? (size_t)desc->background.width * desc->background.height : 0;
        void* env = nullptr;
        float* raw = nullptr;
        step(cudaMallocFromPoolAsync(&env, (n_tex ? n_tex : 1) * sizeof(float4), ctx->pool, ctx->stream));
        if (env) { s->allocs.push_back(env); s->device_bytes += (n_tex ? n_tex : 1) * sizeof(float4); }
        v.env = (const F4*)env;
        if (n_tex && e == cudaSuccess) {
            step(cudaMallocFromPoolAsync((void**)&raw, n_tex * 3 * sizeof(float), ctx->pool, ctx->stream));
            if (raw) {
                step(cudaMemcpyAsync(raw, desc->background.texels, n_tex * 3 * sizeof(float), cudaMemcpyHostToDevice, ctx->stream));
                step(launch_expand_env(raw, (float4*)env, n_tex, ctx->stream));
                s->h2d_bytes += n_tex * 3 * sizeof(float);
                step(cudaStreamSynchronize(ctx->stream));  // the caller's texels may die at return
                cudaFreeAsync(raw, ctx->stream);
            }
        }
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);  // host vectors die at return
    if (e != cudaSuccess) {
        rtr_scene_destroy(s);
        return cuda_fail(ctx, e, "rtr_scene_upload");
    }
    v.background = flat.background;
    v.root_entry = flat.root_entry;
    v.root_entry4 = hybrid ? root_entry4_hybrid : flat.root_entry4;
    s->root_entry_reference = flat.root_entry_reference;
    v.n_objects = (uint32_t)flat.objects.size();
    v.n_materials = (uint32_t)flat.materials.size();
    v.n_volumes = (uint32_t)flat.volumes.size();
    v.n_nodes = flat.device_nodes + (uint32_t)(flat.nodes.size() / 4);
    v.n_sphere_slots = (uint32_t)flat.spheres.size();
    v.n_tri_slots = flat.device_tri_slots + (uint32_t)flat.tri_info.size();
    v.linear_needs_inv = 0;
    for (const ObjectRec& o : flat.objects)
        if (RTR_REF_KIND(o.entry) != REF_SPHERE && RTR_REF_KIND(o.entry) != REF_EMPTY) v.linear_needs_inv = 1;
    if (!on_device && v.n_objects > 0) {  // all-sphere scene: object-order sphere records for the BVH-off scan
        bool all_spheres = true;
        for (const ObjectRec& o : flat.objects) all_spheres = all_spheres && RTR_REF_KIND(o.entry) == REF_SPHERE;
        if (all_spheres) {
            std::vector<F4> lin(flat.objects.size());
            for (size_t i = 0; i < lin.size(); ++i) lin[i] = flat.spheres[RTR_REF_INDEX(flat.objects[i].entry)];
            cudaError_t le = upload(s, lin, &v.linear_spheres, ctx->stream);
            if (le == cudaSuccess) le = cudaStreamSynchronize(ctx->stream);
            if (le != cudaSuccess) { rtr_scene_destroy(s); return cuda_fail(ctx, le, "rtr_scene_upload"); }
        }
    }
    if (on_device) {
        DeviceBvh bvh;
        cudaError_t be = build_sphere_scene_on_device(*desc, ctx->stream, ctx->pool, &bvh, s->allocs, ctx->wf_wide4 != 0u);
        if (be != cudaSuccess) { rtr_scene_destroy(s); return cuda_fail(ctx, be, "device BVH build"); }
        if (bvh.nan_centroid) { rtr_scene_destroy(s); return fail(ctx, RTR_INVALID_ARG, "rtr_scene_upload: NaN centroid (the reference panics at bvh.rs:99)"); }
        if (bvh.max_depth + 2 > stack_limit()) { rtr_scene_destroy(s); return fail(ctx, RTR_UNSUPPORTED, "BVH too deep for the traversal stack"); }
        v.nodes = bvh.nodes; v.spheres = bvh.spheres; v.sphere_info = bvh.sphere_info; v.objects = bvh.objects;
        v.linear_spheres = bvh.linear_spheres;
        v.root_entry = bvh.root_entry; s->root_entry_reference = bvh.root_entry;
        if (bvh.nodes4 && 3u * bvh.max_depth4 + 4u <= stack_limit()) { v.nodes4 = bvh.nodes4; v.root_entry4 = bvh.root_entry4; }
        v.n_objects = desc->n_objects; v.n_nodes = bvh.n_nodes; v.n_sphere_slots = desc->n_objects;
        s->n_nodes = bvh.n_nodes; s->max_depth = bvh.max_depth;
        s->device_bytes += bvh.device_bytes; s->h2d_bytes += bvh.h2d_bytes;
        s->built_on_device = true;
    }
    s->build_ms = prep.host_ms + std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    *out = s;
    return RTR_OK;
}

}  // namespace rtr

extern "C" {

rtr_status rtr_scene_upload(rtr_context* ctx, const rtr_scene_desc* desc, rtr_scene** out) {
    if (!ctx || !desc || !out) return fail(ctx, RTR_INVALID_ARG, "rtr_scene_upload: null argument");
    *out = nullptr;
    PreparedScene prep;
    const rtr_status st = prepare_scene(ctx, desc, prep);
    if (st != RTR_OK) return st;
    return upload_prepared(ctx, desc, prep, out);
}

void rtr_scene_destroy(rtr_scene* s) {
    if (!s) return;
    if (s->ctx) { cudaSetDevice(s->ctx->device); cudaStreamSynchronize(s->ctx->stream); }
    for (void* p : s->allocs) {
        if (s->ctx) cudaFreeAsync(p, s->ctx->stream);  // back into the context's pool
        else cudaFree(p);
    }
    delete s;
}

rtr_status rtr_scene_info(const rtr_scene* s, uint64_t* device_bytes, uint64_t* h2d_bytes, double* build_ms,
                          uint32_t* n_nodes, uint32_t* max_depth) {
    if (!s) return fail(nullptr, RTR_INVALID_ARG, "null scene");
    if (device_bytes) *device_bytes = s->device_bytes;
    if (h2d_bytes) *h2d_bytes = s->h2d_bytes;
    if (build_ms) *build_ms = s->build_ms;
    if (n_nodes) *n_nodes = s->n_nodes;
    if (max_depth) *max_depth = s->max_depth;
    return RTR_OK;
}

}  // extern "C"

// the scene as a launch sees it: traversal tree by default, the reference topology on request
static SceneView view_for(const rtr_scene* scene, uint32_t flags) {
    SceneView v = scene->view;
    if (flags & RTR_FLAG_REFERENCE_TREE) { v.root_entry = scene->root_entry_reference; v.reference_tree = 1u; }
    return v;
}

// ------------------------------------------------------------------ render
rtr_status rtr::check_render_args(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p) {
    if (!ctx || !scene || !p) return fail(ctx, RTR_INVALID_ARG, "render: null argument");
    if (scene->ctx != ctx) return fail(ctx, RTR_INVALID_ARG, "render: scene belongs to another context");
    if (p->settings.max_ray_depth > 255) return fail(ctx, RTR_INVALID_ARG, "max_ray_depth is a u8 in the reference");
    if (p->settings.mode > RTR_MODE_NORMAL) return fail(ctx, RTR_INVALID_ARG, "unknown render mode");
    if (p->kernel > RTR_KERNEL_WAVEFRONT) return fail(ctx, RTR_INVALID_ARG, "unknown kernel selector");
    if ((uint64_t)p->width * p->height > 0x7FFFFFFFull) return fail(ctx, RTR_INVALID_ARG, "image too large");
    const uint32_t spp = p->settings.samples_per_pixel;
    const uint32_t cnt = p->sample_count ? p->sample_count : spp;
    if ((uint64_t)p->sample_begin + cnt > 0xFFFFFFFFull) return fail(ctx, RTR_INVALID_ARG, "sample range overflows u32");
    return RTR_OK;
}

// ------------------------------------------------------------------ wavefront variant (wavefront.cu)
// One wave = generate, max_ray_depth x [extend, shade+compact] (this function), then the ordered sample sum and the
// next-wave bookkeeping (launch_wf_end_wave, issued by the caller so that it can be ordered across wave sequences).
// Every queue size lives on the device, so the launch sequence of a wave is fixed: it is captured into a
// CUDA graph once per call and replayed for each wave (plain launches when capture or instantiation fails).
static cudaError_t wf_launch_wave(rtr_context* ctx, WfArgs a, const WfPaths buf[2], bool exhaustive, bool count_work,
                                  int gen_grid, int ext_grid, int shade_grid, int bin_grid, cudaStream_t st, uint32_t* launches) {
    cudaError_t e = cudaSuccess;
    auto step = [&](cudaError_t r) { if (e == cudaSuccess) e = r; ++*launches; };
    a.bounce = 0; a.in = buf[0]; a.out = buf[1];
    step(launch_wf_generate(a, gen_grid, st));
    uint32_t* const perm = a.perm_out;
    for (uint32_t b = 0; b < a.max_ray_depth; ++b) {
        a.bounce = b; a.in = buf[b & 1]; a.out = buf[(b + 1) & 1];
        a.perm = nullptr;
        // camera rays arrive in tile order and are coherent as they are; scattered rays are regrouped by direction
        if (perm && b >= 1 && b <= ctx->wf_bin_last_bounce) {
            step(launch_wf_bin(a, bin_grid, st));
            a.perm = perm;
        }
        step(launch_wf_extend(a, ext_grid, exhaustive, count_work, st));
        step(launch_wf_shade(a, shade_grid, count_work, st));
    }
    return e;
}

static rtr_status render_wavefront(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p,
                                   float4* accum_dev, bool timed, rtr_stats* stats_out, bool reset_stats) {
    const uint32_t W = p->width, H = p->height;
    const uint32_t count = p->sample_count ? p->sample_count : p->settings.samples_per_pixel;
    const bool exhaustive = (p->flags & RTR_FLAG_EXHAUSTIVE) != 0;
    const bool count_work = (stats_out != nullptr && (p->flags & RTR_FLAG_COUNT_WORK) != 0) || (p->flags & RTR_FLAG_REFERENCE_TREE) != 0;  // the reference-topology walk lives in the counting instantiations
    const uint64_t n_pixels = (uint64_t)W * H;
    WfArgs a;
    std::memset(&a, 0, sizeof(a));
    a.sc = view_for(scene, p->flags);
    a.cam = ray_origin(p->camera);
    a.width = W; a.height = H; a.n_pixels = (uint32_t)n_pixels;
    a.max_ray_depth = p->settings.max_ray_depth;
    a.clamp_indirect = p->settings.clamp_indirect;
    a.enable_bvh_tree = p->settings.enable_bvh_tree;
    a.mode = p->settings.mode;
    a.k0 = (uint32_t)p->seed; a.k1 = (uint32_t)(p->seed >> 32);
    a.sample_begin = p->sample_begin; a.sample_end = p->sample_begin + count;
    a.refill_threshold = ctx->wf_refill_threshold;
    a.atomics_mode = ctx->wf_atomics_mode;
    a.tmin = 0.00001f;  // render.rs:38
    a.accumulate = (p->flags & RTR_FLAG_ACCUMULATE) ? 1u : 0u;
    // samples of every pixel in flight per wave: enough paths to fill the chip, bounded scratch
    uint64_t ws = n_pixels ? ctx->wf_target_paths / n_pixels : 1;
    if (ws < 1) ws = 1;
    if (ws > count) ws = count ? count : 1;
    if (ws > 0xFFFFFFull) ws = 0xFFFFFFull;  // (sample - wave_begin) is packed into 24 bits
    // Two wave sequences (context.h wf_lanes): the waves alternate between two streams, each with its own path state and
    // queue counters, so that one wave's launch tails and thin late bounces overlap the other's full kernels.
    uint32_t lanes = 1;
    if (ctx->wf_lanes >= 2 && a.max_ray_depth >= ctx->wf_lanes_min_depth && count >= 2 && n_pixels > 0) {
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(ctx->stream, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone) lanes = std::min<uint32_t>(ctx->wf_lanes, count);
        else (void)cudaGetLastError();
    }
    // each sequence keeps the full wave size (2 x 2.1 GB of path state: bigger waves AND overlap measured best,
    // profiles/r2/ab_wave_sequences_sizes.txt); a render of less than two such waves is cut into two
    if (lanes > 1 && ws > (count + lanes - 1u) / lanes) ws = (count + lanes - 1u) / lanes;
    a.wave_samples = (uint32_t)ws;
    a.wave_step = lanes;
    const uint64_t n_paths = ws * n_pixels;
    if (n_paths > 0x7FFFFFFFull) return fail(ctx, RTR_INVALID_ARG, "wavefront: too many paths per wave");
    const uint32_t n_waves = count ? (uint32_t)((count + ws - 1) / ws) : 0;
    const bool use_lbuf = ws > 1 || lanes > 1;  // several sequences: each wave's samples wait in lbuf for their turn in the pixel sums
    // direction binning (wavefront.cu wf_bin_kernel): BVH walks only, and only when a bounce follows the first
    const bool use_bin = ctx->wf_bin_keys != 0u && a.enable_bvh_tree != 0u && a.max_ray_depth > 1u && n_paths >= RTR_WF_BIN_BLOCK;
    const size_t streams = 7 + (use_lbuf ? 1 : 0);
    const size_t perm_elems = use_bin ? ((size_t)n_paths + 3) / 4 : 0;  // n_paths uint32 in float4 units
    const size_t lane_elems = streams * (size_t)n_paths + perm_elems;
    CU(grow(&ctx->wf_state, &ctx->wf_state_elems, lanes * lane_elems));
    constexpr uint32_t ML = rtr_context::kMaxLanes;
    WfArgs la[ML];
    WfPaths buf[ML][2];
    cudaStream_t lst[ML] = {ctx->stream, ctx->lane_stream[0], ctx->lane_stream[1], ctx->lane_stream[2]};
    for (uint32_t l = 0; l < lanes; ++l) {
        float4* s0 = ctx->wf_state + l * lane_elems;
        buf[l][0].o = s0; buf[l][0].d = s0 + n_paths; buf[l][0].thr = s0 + 2 * n_paths;
        buf[l][1].o = s0 + 3 * n_paths; buf[l][1].d = s0 + 4 * n_paths; buf[l][1].thr = s0 + 5 * n_paths;
        la[l] = a;
        la[l].hit = s0 + 6 * n_paths;
        la[l].lbuf = use_lbuf ? s0 + 7 * n_paths : nullptr;
        la[l].perm = nullptr;
        la[l].perm_out = use_bin ? reinterpret_cast<uint32_t*>(s0 + streams * (size_t)n_paths) : nullptr;
        la[l].bin_keys = ctx->wf_bin_keys;
        la[l].accum = accum_dev;
        la[l].counters = ctx->wf_counters + l * WFC_WORDS;
        la[l].stats = ctx->stats;
        la[l].wave_first = l;
    }
    const int per_sm = wf_extend_max_blocks_per_sm(exhaustive, count_work, a.sc.n_volumes != 0u || a.enable_bvh_tree == 0u, false, wf_extend_wide4(a, exhaustive, count_work));
    if (per_sm <= 0) return fail(ctx, RTR_CUDA_ERROR, "wf_extend cannot be resident on this device");
    auto blocks_for = [&](uint64_t items, int cap_per_sm) {
        uint64_t b = (items + RTR_WF_THREADS - 1) / RTR_WF_THREADS;
        const uint64_t cap = (uint64_t)ctx->sm_count * cap_per_sm;
        return (int)(b < 1 ? 1 : b > cap ? cap : b);
    };
    const int ext_grid = blocks_for(n_paths * RTR_WF_THREADS / RTR_WF_EXTEND_THREADS, per_sm);
    const int gen_grid = blocks_for(n_paths + 32 * (uint64_t)ws * (W / 8 + H / 4 + 2), 16);
    const int shade_grid = blocks_for(n_paths * RTR_WF_THREADS / wf_shade_threads(), 16 * RTR_WF_THREADS / wf_shade_threads());
    const int sum_grid = blocks_for(n_pixels, 8);
    const int bin_grid = (int)std::min<uint64_t>((n_paths + RTR_WF_BIN_BLOCK - 1) / RTR_WF_BIN_BLOCK, (uint64_t)ctx->sm_count * 8);
    if (reset_stats) CU(cudaMemsetAsync(ctx->stats, 0, STAT_COUNT * sizeof(unsigned long long), ctx->stream));
    uint32_t launches = 0;
    if (timed) CU(cudaEventRecord(ctx->ev0, ctx->stream));
    CU(launch_wf_begin(la[0], ctx->stream));
    ++launches;
    if (lanes > 1) {  // the other sequences start once the image and the counters are clear
        for (uint32_t l = 1; l < lanes; ++l) { CU(launch_wf_begin_lane(la[l], ctx->stream)); ++launches; }
        CU(cudaEventRecord(ctx->wf_ev_fork, ctx->stream));
        for (uint32_t l = 1; l < lanes; ++l) CU(cudaStreamWaitEvent(lst[l], ctx->wf_ev_fork, 0));
    }
    // the launch sequence of one wave, per lane: a CUDA graph when there is something to replay
    cudaGraph_t graph[ML] = {nullptr, nullptr, nullptr, nullptr};
    cudaGraphExec_t exec[ML] = {nullptr, nullptr, nullptr, nullptr};
    uint32_t per_wave = 0;
    bool use_graph = ctx->wf_use_graph && n_waves > lanes;
    for (uint32_t l = 0; l < lanes && use_graph; ++l) {
        per_wave = 0;
        cudaError_t e = cudaStreamBeginCapture(lst[l], cudaStreamCaptureModeThreadLocal);
        if (e == cudaSuccess) {
            cudaError_t le = wf_launch_wave(ctx, la[l], buf[l], exhaustive, count_work, gen_grid, ext_grid, shade_grid, bin_grid, lst[l], &per_wave);
            e = cudaStreamEndCapture(lst[l], &graph[l]);  // always ends the capture, also after a failed launch
            if (e == cudaSuccess && le != cudaSuccess) e = le;
        }
        if (e == cudaSuccess) e = cudaGraphInstantiate(&exec[l], graph[l], 0);
        if (e != cudaSuccess) {
            // capture / instantiation is unavailable here (e.g. the caller's stream is itself being captured): nothing
            // of a wave has run yet, so clear the error and issue the same sequences as plain launches
            use_graph = false;
            (void)cudaGetLastError();
        }
    }
    cudaError_t we = cudaSuccess;
    for (uint32_t w = 0; w < n_waves && we == cudaSuccess; ++w) {
        const uint32_t l = w % lanes;
        if (use_graph) { we = cudaGraphLaunch(exec[l], lst[l]); launches += per_wave; }
        else we = wf_launch_wave(ctx, la[l], buf[l], exhaustive, count_work, gen_grid, ext_grid, shade_grid, bin_grid, lst[l], &launches);
        // pixel sums in wave order (render.rs:99 adds the samples of a pixel one after the other): this wave's sum waits
        // for the previous wave's, which ran on the other stream
        if (we == cudaSuccess && lanes > 1 && w > 0) we = cudaStreamWaitEvent(lst[l], ctx->wf_ev_wave[(w - 1) % ML], 0);
        if (we == cudaSuccess) we = launch_wf_end_wave(la[l], sum_grid, lst[l]);
        launches += la[l].lbuf ? 2u : 1u;
        if (we == cudaSuccess && lanes > 1) we = cudaEventRecord(ctx->wf_ev_wave[w % ML], lst[l]);
    }
    // join: the last wave's sum is ordered after every earlier one; if it ran on another stream, wait for it
    if (we == cudaSuccess && lanes > 1 && n_waves > 0 && ((n_waves - 1) % lanes) != 0) we = cudaStreamWaitEvent(ctx->stream, ctx->wf_ev_wave[(n_waves - 1) % ML], 0);
    for (uint32_t l = 0; l < ML; ++l) {
        if (exec[l]) cudaGraphExecDestroy(exec[l]);
        if (graph[l]) cudaGraphDestroy(graph[l]);
    }
    if (we != cudaSuccess) {
        for (uint32_t l = 1; l < lanes; ++l) cudaStreamSynchronize(lst[l]);
        return cuda_fail(ctx, we, "wavefront wave launch");
    }
    if (timed) CU(cudaEventRecord(ctx->ev1, ctx->stream));
    if (stats_out) {
        std::memset(stats_out, 0, sizeof(*stats_out));
        stats_out->kernel_launches = launches;
        stats_out->kernel_used = RTR_KERNEL_WAVEFRONT;
        stats_out->samples = n_pixels * count;
    }
    return RTR_OK;
}

// launches the render kernels for this call's sample slice into accum_dev; fills stats (async)
rtr_status rtr::render_into(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p,
                            float4* accum_dev, bool timed, rtr_stats* stats_out, bool reset_stats) {
    // RTR_KERNEL_AUTO (measured on B200, DESIGN.md "Megakernel or wavefront"): a BVH walk whose length varies
    // from ray to ray is where the wavefront's cheap lane refill pays for its path-state traffic; the linear
    // scan (BVH off), tiny trees and renders too small to fill a wave stay in the single-launch megakernel.
    uint32_t kernel = p->kernel;
    if (kernel == RTR_KERNEL_AUTO) {
        const uint64_t paths = (uint64_t)p->width * p->height * (p->sample_count ? p->sample_count : p->settings.samples_per_pixel);
        const bool wf = p->settings.enable_bvh_tree != 0 && scene->n_nodes >= 64 && paths >= (1ull << 20);
        kernel = wf ? RTR_KERNEL_WAVEFRONT : RTR_KERNEL_MEGAKERNEL;
    }
    if (kernel == RTR_KERNEL_WAVEFRONT) return render_wavefront(ctx, scene, p, accum_dev, timed, stats_out, reset_stats);
    const uint32_t W = p->width, H = p->height;
    const uint32_t count = p->sample_count ? p->sample_count : p->settings.samples_per_pixel;
    const bool exhaustive = (p->flags & RTR_FLAG_EXHAUSTIVE) != 0;
    const bool count_work = (stats_out != nullptr && (p->flags & RTR_FLAG_COUNT_WORK) != 0) || (p->flags & RTR_FLAG_REFERENCE_TREE) != 0;  // the reference-topology walk lives in the counting instantiations
    RenderArgs a;
    std::memset(&a, 0, sizeof(a));
    a.sc = view_for(scene, p->flags);
    a.cam = ray_origin(p->camera);
    a.width = W; a.height = H;
    a.max_ray_depth = p->settings.max_ray_depth;
    a.clamp_indirect = p->settings.clamp_indirect;
    a.enable_bvh_tree = p->settings.enable_bvh_tree;
    a.mode = p->settings.mode;
    a.k0 = (uint32_t)p->seed; a.k1 = (uint32_t)(p->seed >> 32);
    a.pk = make_philox_keys(a.k0, a.k1);
    a.sample_begin = p->sample_begin; a.sample_count = count;
    a.tiles_x = (W + 7) / 8;
    a.n_tiles = a.tiles_x * ((H + 3) / 4);
    // persistent grid: one block per resident slot
    // BVH off and every object a sphere (scenes/benchmark.json): the scan kernel (scan.cu), same results
    const bool scan = ctx->use_scan_kernel != 0 && p->settings.enable_bvh_tree == 0 && a.sc.linear_spheres != nullptr &&
                      p->settings.max_ray_depth >= 1;
    int per_sm = scan ? scan_kernel_max_blocks_per_sm(count_work)
                      : megakernel_max_blocks_per_sm(exhaustive, count_work, p->settings.enable_bvh_tree != 0, a.sc.n_volumes != 0u);
    if (per_sm <= 0) return fail(ctx, RTR_CUDA_ERROR, "megakernel cannot be resident on this device");
    const uint32_t resident_warps = (uint32_t)ctx->sm_count * per_sm * (RTR_MEGA_THREADS / 32);
    // small image + many samples: split the samples of a tile into chunks so the chip is full
    a.n_chunks = 1; a.chunk_samples = count ? count : 1;
    if (a.n_tiles < 4 * resident_warps && count > 1) {
        uint32_t want = (4 * resident_warps + a.n_tiles - 1) / (a.n_tiles ? a.n_tiles : 1);
        if (want > count) want = count;
        a.chunk_samples = (count + want - 1) / want;
        a.n_chunks = (count + a.chunk_samples - 1) / a.chunk_samples;
    }
    a.refill_threshold = ctx->refill_threshold;
    a.accumulate = (p->flags & RTR_FLAG_ACCUMULATE) ? 1u : 0u;
    a.accum = accum_dev;
    if (a.n_chunks > 1) {
        CU(grow(&ctx->partials, &ctx->partials_elems, (size_t)a.n_chunks * W * H));
        a.partials = ctx->partials;
    }
    a.work_counter = ctx->work_counter;
    a.stats = ctx->stats;
    CU(cudaMemsetAsync(ctx->work_counter, 0, sizeof(uint32_t), ctx->stream));
    if (reset_stats) CU(cudaMemsetAsync(ctx->stats, 0, STAT_COUNT * sizeof(unsigned long long), ctx->stream));
    uint32_t launches = 0;
    if (timed) CU(cudaEventRecord(ctx->ev0, ctx->stream));
    if (W && H) {
        uint32_t items = a.n_tiles * a.n_chunks;
        uint32_t blocks = (uint32_t)ctx->sm_count * per_sm;
        uint32_t need_blocks = (items + (RTR_MEGA_THREADS / 32) - 1) / (RTR_MEGA_THREADS / 32);
        if (blocks > need_blocks) blocks = need_blocks;
        if (scan) CU(launch_scan_kernel(a, (int)blocks, count_work, ctx->stream));
        else CU(launch_megakernel(a, (int)blocks, exhaustive, count_work, ctx->stream));
        ++launches;
        if (a.n_chunks > 1) {
            CU(launch_sum_chunks(a.partials, accum_dev, W * H, a.n_chunks, a.accumulate != 0, ctx->stream));
            ++launches;
        }
    }
    if (timed) CU(cudaEventRecord(ctx->ev1, ctx->stream));
    if (stats_out) {
        std::memset(stats_out, 0, sizeof(*stats_out));
        stats_out->kernel_launches = launches;
        stats_out->kernel_used = RTR_KERNEL_MEGAKERNEL;
        stats_out->samples = (uint64_t)W * H * count;
    }
    return RTR_OK;
}

rtr_status rtr::collect_stats(rtr_context* ctx, bool timed, rtr_stats* stats_out) {
    if (!stats_out) return RTR_OK;
    unsigned long long h[STAT_COUNT];
    CU(cudaMemcpyAsync(h, ctx->stats, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    stats_out->rays = h[STAT_RAYS]; stats_out->nodes_visited = h[STAT_NODES]; stats_out->tri_tests = h[STAT_TRIS];
    stats_out->sphere_tests = h[STAT_SPHERES]; stats_out->env_lookups = h[STAT_ENV];
    if (timed) {
        float ms = 0.0f;
        CU(cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        stats_out->device_ms = ms;
    }
    return RTR_OK;
}

extern "C" {

rtr_status rtr_render(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p, uint8_t* rgba8_out,
                      float* accum_out, rtr_stats* stats_out) {
    rtr_status st = check_render_args(ctx, scene, p);
    if (st != RTR_OK) return st;
    CU(cudaSetDevice(ctx->device));
    const size_t n = (size_t)p->width * p->height;
    if (n == 0) { if (stats_out) std::memset(stats_out, 0, sizeof(*stats_out)); return RTR_OK; }
    CU(grow(&ctx->accum, &ctx->accum_pixels, n));
    CU(grow(&ctx->rgba, &ctx->rgba_pixels, n));
    st = render_into(ctx, scene, p, ctx->accum, true, stats_out);
    if (st != RTR_OK) return st;
    const uint32_t count = p->sample_count ? p->sample_count : p->settings.samples_per_pixel;
    if (rgba8_out) {
        CU(launch_resolve(ctx->accum, ctx->rgba, (uint32_t)n, (float)count, ctx->stream));
        if (stats_out) stats_out->kernel_launches++;
        CU(cudaMemcpyAsync(rgba8_out, ctx->rgba, n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    }
    if (accum_out) CU(cudaMemcpyAsync(accum_out, ctx->accum, n * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return collect_stats(ctx, true, stats_out);
}

rtr_status rtr_render_accumulate_device(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p,
                                        float* accum_dev, rtr_stats* stats_out) {
    rtr_status st = check_render_args(ctx, scene, p);
    if (st != RTR_OK) return st;
    if (!accum_dev) return fail(ctx, RTR_INVALID_ARG, "accum_dev is NULL");
    CU(cudaSetDevice(ctx->device));
    st = render_into(ctx, scene, p, (float4*)accum_dev, stats_out != nullptr, stats_out);
    if (st != RTR_OK) return st;
    return collect_stats(ctx, true, stats_out);  // synchronises only when stats are requested
}

rtr_status rtr_resolve_device(rtr_context* ctx, const float* accum_dev, uint32_t width, uint32_t height,
                              uint32_t samples_per_pixel, uint8_t* rgba8_dev) {
    if (!ctx || !accum_dev || !rgba8_dev) return fail(ctx, RTR_INVALID_ARG, "rtr_resolve_device: null argument");
    CU(cudaSetDevice(ctx->device));
    if ((uint64_t)width * height == 0) return RTR_OK;
    CU(launch_resolve((const float4*)accum_dev, (uchar4*)rgba8_dev, width * height, (float)samples_per_pixel, ctx->stream));
    return RTR_OK;
}

// ------------------------------------------------------------------ batch entry points
rtr_status rtr_trace_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays, size_t n, float tmin,
                           uint32_t enable_bvh_tree, uint32_t flags, uint64_t seed, rtr_hit* hits_out,
                           rtr_stats* stats_out) {
    if (!ctx || !scene || (n && (!rays || !hits_out))) return fail(ctx, RTR_INVALID_ARG, "rtr_trace_batch: null argument");
    if (scene->ctx != ctx) return fail(ctx, RTR_INVALID_ARG, "scene belongs to another context");
    if (stats_out) std::memset(stats_out, 0, sizeof(*stats_out));
    if (n == 0) return RTR_OK;
    if (n > 0xFFFFFFFFull) return fail(ctx, RTR_INVALID_ARG, "batch too large (ray index is the u32 RNG counter)");
    CU(cudaSetDevice(ctx->device));
    // The walk is the render's own: wf_extend (persistent warps, majority-rule stepping, lane refill, shared-memory
    // short stack), fed from the queue of bounce 0.  Work counters only under RTR_FLAG_COUNT_WORK / RTR_FLAG_REFERENCE_TREE;
    // stats_out->device_ms always times the three kernels (events on the context stream, copies outside).
    const bool exhaustive = (flags & RTR_FLAG_EXHAUSTIVE) != 0;
    const bool count_work = (flags & (RTR_FLAG_COUNT_WORK | RTR_FLAG_REFERENCE_TREE)) != 0;  // the reference-topology walk lives in the counting instantiations
    const size_t piece = std::min<size_t>(n, (size_t)1 << 24);  // rays per pass: bounds the scratch (48 B per ray)
    CU(grow(&ctx->batch_in, &ctx->batch_in_bytes, n * sizeof(rtr_ray)));
    CU(grow(&ctx->batch_out, &ctx->batch_out_bytes, n * sizeof(rtr_hit)));
    CU(grow(&ctx->wf_state, &ctx->wf_state_elems, 3 * piece));
    WfArgs a;
    std::memset(&a, 0, sizeof(a));
    a.sc = view_for(scene, flags);
    a.max_ray_depth = 1;
    a.enable_bvh_tree = enable_bvh_tree;
    a.k0 = (uint32_t)seed; a.k1 = (uint32_t)(seed >> 32);
    a.sample_begin = 0; a.sample_end = 1; a.wave_samples = 1;
    a.refill_threshold = ctx->wf_refill_threshold;
    a.atomics_mode = ctx->wf_atomics_mode;
    a.tmin = tmin;
    a.in.o = ctx->wf_state; a.in.d = ctx->wf_state + piece; a.hit = ctx->wf_state + 2 * piece;
    a.counters = ctx->wf_counters;
    a.stats = ctx->stats;
    const int per_sm = wf_extend_max_blocks_per_sm(exhaustive, count_work, a.sc.n_volumes != 0u || enable_bvh_tree == 0u, tmin != 0.00001f, wf_extend_wide4(a, exhaustive, count_work));
    if (per_sm <= 0) return fail(ctx, RTR_CUDA_ERROR, "wf_extend cannot be resident on this device");
    CU(cudaMemcpyAsync(ctx->batch_in, rays, n * sizeof(rtr_ray), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemsetAsync(ctx->stats, 0, STAT_COUNT * sizeof(unsigned long long), ctx->stream));
    CU(cudaEventRecord(ctx->ev0, ctx->stream));
    uint32_t launches = 0;
    for (size_t first = 0; first < n; first += piece) {
        const uint32_t m = (uint32_t)std::min(piece, n - first);
        CU(cudaMemsetAsync(ctx->wf_counters, 0, WFC_WORDS * sizeof(uint32_t), ctx->stream));
        CU(launch_trace_unpack(a, (const rtr_ray*)ctx->batch_in + first, m, (uint32_t)first, ctx->stream));
        uint64_t blocks = ((uint64_t)m + RTR_WF_EXTEND_THREADS - 1) / RTR_WF_EXTEND_THREADS;
        const uint64_t cap = (uint64_t)ctx->sm_count * per_sm;
        CU(launch_wf_extend(a, (int)(blocks < 1 ? 1 : blocks > cap ? cap : blocks), exhaustive, count_work, ctx->stream));
        CU(launch_trace_finish(a, m, (rtr_hit*)ctx->batch_out + first, ctx->stream));
        launches += 3;
    }
    CU(cudaEventRecord(ctx->ev1, ctx->stream));
    CU(cudaMemcpyAsync(hits_out, ctx->batch_out, n * sizeof(rtr_hit), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    if (stats_out) {
        stats_out->kernel_launches = launches; stats_out->kernel_used = RTR_KERNEL_WAVEFRONT;
        rtr_status st = collect_stats(ctx, true, stats_out);
        stats_out->rays = n;
        return st;
    }
    return RTR_OK;
}

rtr_status rtr_scatter_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_scatter_in* in, size_t n,
                             uint32_t mode, uint64_t seed, rtr_scatter_out* out) {
    if (!ctx || !scene || (n && (!in || !out))) return fail(ctx, RTR_INVALID_ARG, "rtr_scatter_batch: null argument");
    if (scene->ctx != ctx) return fail(ctx, RTR_INVALID_ARG, "scene belongs to another context");
    if (mode != RTR_MODE_FULL && mode != RTR_MODE_CLAY) return fail(ctx, RTR_INVALID_ARG, "scatter mode must be FULL or CLAY");
    if (n == 0) return RTR_OK;
    if (mode == RTR_MODE_FULL)
        for (size_t i = 0; i < n; ++i)
            if (in[i].material >= scene->view.n_materials) return fail(ctx, RTR_INVALID_ARG, "material handle out of range");
    CU(cudaSetDevice(ctx->device));
    CU(grow(&ctx->batch_in, &ctx->batch_in_bytes, n * sizeof(rtr_scatter_in)));
    CU(grow(&ctx->batch_out, &ctx->batch_out_bytes, n * sizeof(rtr_scatter_out)));
    CU(cudaMemcpyAsync(ctx->batch_in, in, n * sizeof(rtr_scatter_in), cudaMemcpyHostToDevice, ctx->stream));
    CU(launch_scatter_batch(scene->view, (const rtr_scatter_in*)ctx->batch_in, n, mode, (uint32_t)seed,
                            (uint32_t)(seed >> 32), (rtr_scatter_out*)ctx->batch_out, ctx->stream));
    CU(cudaMemcpyAsync(out, ctx->batch_out, n * sizeof(rtr_scatter_out), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return RTR_OK;
}

rtr_status rtr_background_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays, size_t n,
                                float* rgb_out) {
    if (!ctx || !scene || (n && (!rays || !rgb_out))) return fail(ctx, RTR_INVALID_ARG, "rtr_background_batch: null argument");
    if (scene->ctx != ctx) return fail(ctx, RTR_INVALID_ARG, "scene belongs to another context");
    if (n == 0) return RTR_OK;
    CU(cudaSetDevice(ctx->device));
    CU(grow(&ctx->batch_in, &ctx->batch_in_bytes, n * sizeof(rtr_ray)));
    CU(grow(&ctx->batch_out, &ctx->batch_out_bytes, n * 3 * sizeof(float)));
    CU(cudaMemcpyAsync(ctx->batch_in, rays, n * sizeof(rtr_ray), cudaMemcpyHostToDevice, ctx->stream));
    CU(launch_background_batch(scene->view, (const rtr_ray*)ctx->batch_in, n, (float*)ctx->batch_out, ctx->stream));
    CU(cudaMemcpyAsync(rgb_out, ctx->batch_out, n * 3 * sizeof(float), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return RTR_OK;
}

rtr_status rtr_camera_rays(rtr_context* ctx, const rtr_render_params* p, const uint32_t* pixels,
                           const uint32_t* samples, size_t n, rtr_ray* rays_out) {
    if (!ctx || !p || (n && (!pixels || !samples || !rays_out))) return fail(ctx, RTR_INVALID_ARG, "rtr_camera_rays: null argument");
    if (p->width == 0 || p->height == 0) return fail(ctx, RTR_INVALID_ARG, "empty render target");
    if (n == 0) return RTR_OK;
    CU(cudaSetDevice(ctx->device));
    CU(grow(&ctx->batch_in, &ctx->batch_in_bytes, n * sizeof(uint32_t)));
    CU(grow(&ctx->batch_in2, &ctx->batch_in2_bytes, n * sizeof(uint32_t)));
    CU(grow(&ctx->batch_out, &ctx->batch_out_bytes, n * sizeof(rtr_ray)));
    CU(cudaMemcpyAsync(ctx->batch_in, pixels, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(cudaMemcpyAsync(ctx->batch_in2, samples, n * sizeof(uint32_t), cudaMemcpyHostToDevice, ctx->stream));
    CU(launch_camera_rays(ray_origin(p->camera), p->width, p->height, (uint32_t)p->seed, (uint32_t)(p->seed >> 32),
                          (const uint32_t*)ctx->batch_in, (const uint32_t*)ctx->batch_in2, n, (rtr_ray*)ctx->batch_out,
                          ctx->stream));
    CU(cudaMemcpyAsync(rays_out, ctx->batch_out, n * sizeof(rtr_ray), cudaMemcpyDeviceToHost, ctx->stream));
    CU(cudaStreamSynchronize(ctx->stream));
    return RTR_OK;
}

}  // extern "C"
