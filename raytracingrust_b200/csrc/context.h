// context.h — internals shared by capi.cpp (one device) and group.cpp (N devices of one process).
// Not part of the ABI: include/rtr.h only sees the opaque handles.
#ifndef RTR_CONTEXT_H_
#define RTR_CONTEXT_H_

#include <cuda_runtime_api.h>

#include <string>
#include <vector>

#include "../../include/rtr.h"
#include "host_flatten.h"
#include "kernels.h"

struct rtr_context {
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    int sm_count = 0;
    // scene arrays and build scratch come from this stream-ordered pool (release threshold: never): a scene upload is
    // ~30 allocations, and plain cudaMalloc / cudaFree stalled for 0.3-0.6 s at random while multi-GB buffers were live
    cudaMemPool_t pool = nullptr;
    uint32_t refill_threshold = 12;  // tuning knob (RTR_REFILL_THRESHOLD), see DESIGN.md
    uint32_t* work_counter = nullptr;
    unsigned long long* stats = nullptr;  // STAT_COUNT
    // scratch, grown on demand
    float4* accum = nullptr; size_t accum_pixels = 0;
    float4* partials = nullptr; size_t partials_elems = 0;
    uchar4* rgba = nullptr; size_t rgba_pixels = 0;
    float4* wf_state = nullptr; size_t wf_state_elems = 0;  // wavefront: 7 path streams (+ lbuf), see render_wavefront
    uint32_t* wf_counters = nullptr;    // kMaxLanes x WFC_WORDS: one block per wave sequence
    static constexpr uint32_t kMaxLanes = 4;
    cudaStream_t lane_stream[kMaxLanes - 1] = {nullptr, nullptr, nullptr};  // the further wave sequences of a wavefront render (RTR_WF_LANES)
    cudaEvent_t wf_ev_wave[kMaxLanes] = {nullptr, nullptr, nullptr, nullptr}, wf_ev_fork = nullptr;
    // RTR_WF_LANES: 2 = consecutive waves run as two launch sequences on two streams, so that the launch tails and the thin
    // late bounces of one wave (few paths alive, each launch waiting on its longest walk) overlap the full kernels of the
    // other; the per-pixel sums stay in wave order (events), so the image is bit-identical.  With 2^25-path waves: +3.0..+5.0 %
    // on C1 / C3 / C4 / C5, +36 % on the depth-50 mesh bench (profiles/r2/ab_wave_sequences*.txt; 3 and 4 sequences add little).  Used from max_ray_depth >= wf_lanes_min_depth (RTR_WF_LANES_MIN_DEPTH).
    uint32_t wf_lanes = 2, wf_lanes_min_depth = 1;
    uint32_t wf_refill_threshold = 16;  // RTR_WF_REFILL_THRESHOLD
    uint64_t wf_target_paths = 1ull << 25;  // RTR_WF_PATHS: paths in flight per wave and sequence (4.3 GB of path state each; swept 2^22..2^25 with one and with several sequences on B200)
    int wf_use_graph = 1;  // RTR_WF_GRAPH
    // Same-address atomics run at ~1 per ns on B200 (ncu, profiles/r2: wf_generate issued 24 % of its slots and used 19 % of
    // DRAM while doing 512 K atomicAdds on the queue counter in 0.48 ms).  RTR_WF_ATOMICS: bit 0 = atomic-free wf_generate
    // when the image tiles exactly, bit 1 = one compaction atomic per CTA in wf_shade (0 = one per warp, as in round 1).
    uint32_t wf_atomics_mode = 3;
    uint32_t wf_wide4 = 1;     // RTR_WF_WIDE (0 = off): scenes uploaded through this context also get the 4-wide traversal tree (flat.h) and wf_extend walks it
    uint32_t wf_bin_keys = 0;  // RTR_WF_BIN: direction-binning keys of bounces >= 1 (0 = off: measured -1..-7 %; 8 = octant, 24 = octant x major axis)
    uint32_t wf_bin_last_bounce = 255;  // RTR_WF_BIN_LAST: bin bounces 1..this
    int use_scan_kernel = 1;  // RTR_SCAN_KERNEL: 0 sends all-sphere BVH-off renders through the generic megakernel
    char* batch_in = nullptr; size_t batch_in_bytes = 0;
    char* batch_in2 = nullptr; size_t batch_in2_bytes = 0;
    char* batch_out = nullptr; size_t batch_out_bytes = 0;
    std::string error;
};

struct rtr_scene {
    rtr_context* ctx = nullptr;
    rtr::SceneView view;
    std::vector<void*> allocs;
    uint64_t device_bytes = 0, h2d_bytes = 0;
    double build_ms = 0.0;
    uint32_t n_nodes = 0, max_depth = 0;
    uint32_t root_entry_reference = 0xFFFFFFFFu;
    bool built_on_device = false;
};

namespace rtr {

rtr_status fail(rtr_context* ctx, rtr_status st, const std::string& msg);
rtr_status cuda_fail(rtr_context* ctx, cudaError_t e, const char* what);

#define CU(call)                                                  \
    do {                                                          \
        cudaError_t e_ = (call);                                  \
        if (e_ != cudaSuccess) return rtr::cuda_fail(ctx, e_, #call);  \
    } while (0)

template <typename T>
static cudaError_t grow(T** p, size_t* have, size_t want) {
    if (*have >= want && *p) return cudaSuccess;
    if (*p) cudaFree(*p);
    *p = nullptr; *have = 0;
    cudaError_t e = cudaMalloc((void**)p, want * sizeof(T));
    if (e == cudaSuccess) *have = want;
    return e;
}

// The host half of rtr_scene_upload: validate, rebuild the reference topology, flatten (host_flatten.cpp), decide
// whether the BVH is built on the device instead.  Done ONCE per upload, also when the scene goes to N devices.
struct PreparedScene {
    HostFlat flat;
    bool on_device = false;  // sphere scene whose BVH is built by bvh_build.cu
    double host_ms = 0.0;
};
rtr_status prepare_scene(rtr_context* err_ctx, const rtr_scene_desc* desc, PreparedScene& out);
// The device half: H2D of the flattened arrays (and the device BVH build) on ctx's stream.
rtr_status upload_prepared(rtr_context* ctx, const rtr_scene_desc* desc, const PreparedScene& prep, rtr_scene** out);

rtr_status check_render_args(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p);
// launches the render kernels for this call's sample slice into accum_dev (async on ctx->stream)
// reset_stats = false: keep adding to the device-side work counters of the previous call (a render issued in pieces)
rtr_status render_into(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* p, float4* accum_dev,
                       bool timed, rtr_stats* stats_out, bool reset_stats = true);
// synchronises ctx->stream and reads the device counters / event times into stats_out
rtr_status collect_stats(rtr_context* ctx, bool timed, rtr_stats* stats_out);

}  // namespace rtr
#endif
