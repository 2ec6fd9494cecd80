// noalloc.cu — rtr_trace_batch_device_noalloc: world.hit(ray, tmin, +INF) for a batch of rays WITHOUT any allocation.
//
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It is the device-pointer
// variant of rtr_trace_batch (capi.cpp): the same three kernels (trace_unpack -> wf_extend -> trace_finish), but
//   * rays and hits are the CALLER's device buffers, and
//   * the scratch wf_extend needs — the unpacked ray streams, the hit stream and the queue counters that
//     rtr_trace_batch keeps in cudaMalloc'ed, grown-on-demand buffers — lives in the file-scope array below.
// The array costs every GPU that loads the library RTR_NOALLOC_SCRATCH_BYTES of device memory until the process exits.
// Every call on a device shares it and it keeps the last call's contents: the counter block is cleared at the start
// of each call (the streams are fully rewritten for the slots that are read).  Calls that overlap on the GPU — two
// streams or two contexts of one device — race on it; the caller serialises them.
//
// The name `g_noalloc_scratch` in HOST code is a host-side stand-in, not a device address: its device address is
// looked up once per device with cudaGetSymbolAddress (device current) and that pointer is what the kernels get.
#include <cstring>
#include <mutex>
#include <string>

#include "context.h"
#include "kernels.h"

namespace {

// layout: [0, HEADER) counters (WFC_WORDS u32) + stats (STAT_COUNT u64); then 3 float4 streams of MAX_RAYS slots
constexpr size_t kHeaderBytes = 4096;
constexpr size_t kMaxRays = RTR_NOALLOC_MAX_RAYS;
constexpr size_t kScratchBytes = kHeaderBytes + 3 * kMaxRays * sizeof(float4);
static_assert(rtr::WFC_WORDS * sizeof(uint32_t) + rtr::STAT_COUNT * sizeof(unsigned long long) <= kHeaderBytes, "header too small");
static_assert(kScratchBytes == RTR_NOALLOC_SCRATCH_BYTES, "rtr.h states the size of the scratch array");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

constexpr int kMaxDevices = 64;
std::mutex g_mu;
unsigned char* g_scratch_ptr[kMaxDevices] = {};  // device address of g_noalloc_scratch on device i (null: not looked up yet)

// once per device, with that device current
cudaError_t scratch_for_device(int device, unsigned char** out) {
    std::lock_guard<std::mutex> lk(g_mu);
    if (!g_scratch_ptr[device]) {
        unsigned char* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        g_scratch_ptr[device] = p;
    }
    *out = g_scratch_ptr[device];
    return cudaSuccess;
}

}  // namespace

using namespace rtr;

extern "C" rtr_status rtr_trace_batch_device_noalloc(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays_dev,
                                                     size_t n, float tmin, uint32_t enable_bvh_tree, uint32_t flags,
                                                     uint64_t seed, rtr_hit* hits_dev) {
    if (!ctx || !scene || (n && (!rays_dev || !hits_dev)))
        return fail(ctx, RTR_INVALID_ARG, "rtr_trace_batch_device_noalloc: null argument");
    if (scene->ctx != ctx) return fail(ctx, RTR_INVALID_ARG, "scene belongs to another context");
    if (n > kMaxRays)
        return fail(ctx, RTR_INVALID_ARG, "rtr_trace_batch_device_noalloc: " + std::to_string(n) + " rays exceed RTR_NOALLOC_MAX_RAYS (" +
                                              std::to_string(kMaxRays) + "): the static scratch array holds no more; split the batch or use rtr_trace_batch");
    if (flags & ~(uint32_t)(RTR_FLAG_EXHAUSTIVE | RTR_FLAG_REFERENCE_TREE))
        return fail(ctx, RTR_INVALID_ARG, "rtr_trace_batch_device_noalloc: only RTR_FLAG_EXHAUSTIVE / RTR_FLAG_REFERENCE_TREE are accepted");
    if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, RTR_INVALID_ARG, "device ordinal out of range");
    if (n == 0) return RTR_OK;
    CU(cudaSetDevice(ctx->device));
    unsigned char* base = nullptr;
    CU(scratch_for_device(ctx->device, &base));
    const bool exhaustive = (flags & RTR_FLAG_EXHAUSTIVE) != 0;
    const bool count_work = (flags & RTR_FLAG_REFERENCE_TREE) != 0;  // the reference-topology walk lives in the counting instantiations
    WfArgs a;
    std::memset(&a, 0, sizeof(a));
    a.sc = scene->view;
    if (flags & RTR_FLAG_REFERENCE_TREE) { a.sc.root_entry = scene->root_entry_reference; a.sc.reference_tree = 1u; }
    a.max_ray_depth = 1;
    a.enable_bvh_tree = enable_bvh_tree;
    a.k0 = (uint32_t)seed; a.k1 = (uint32_t)(seed >> 32);
    a.sample_begin = 0; a.sample_end = 1; a.wave_samples = 1;
    a.refill_threshold = ctx->wf_refill_threshold;
    a.tmin = tmin;
    a.counters = reinterpret_cast<uint32_t*>(base);
    a.stats = reinterpret_cast<unsigned long long*>(base + WFC_WORDS * sizeof(uint32_t));
    float4* streams = reinterpret_cast<float4*>(base + kHeaderBytes);
    a.in.o = streams; a.in.d = streams + kMaxRays; a.hit = streams + 2 * kMaxRays;
    const int per_sm = wf_extend_max_blocks_per_sm(exhaustive, count_work, a.sc.n_volumes != 0u || enable_bvh_tree == 0u, tmin != 0.00001f, wf_extend_wide4(a, exhaustive, count_work));
    if (per_sm <= 0) return fail(ctx, RTR_CUDA_ERROR, "wf_extend cannot be resident on this device");
    // the array keeps the previous call's queue size and fetch position: both must start at zero
    CU(cudaMemsetAsync(base, 0, kHeaderBytes, ctx->stream));
    CU(launch_trace_unpack(a, rays_dev, (uint32_t)n, 0u, ctx->stream));
    uint64_t blocks = ((uint64_t)n + RTR_WF_EXTEND_THREADS - 1) / RTR_WF_EXTEND_THREADS;
    const uint64_t cap = (uint64_t)ctx->sm_count * per_sm;
    CU(launch_wf_extend(a, (int)(blocks < 1 ? 1 : blocks > cap ? cap : blocks), exhaustive, count_work, ctx->stream));
    CU(launch_trace_finish(a, (uint32_t)n, hits_dev, ctx->stream));
    return RTR_OK;  // asynchronous on the context's stream: rtr_synchronize (or the caller's own stream) orders the result
}
