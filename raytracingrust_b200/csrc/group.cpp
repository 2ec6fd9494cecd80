// group.cpp — one render() over N GPUs of this process (include/rtr.h, "rtr_group_*").
//
// The reference parallelises render() over pixels with rayon's work-stealing pool (lib/core/render.rs:114-119).  Across
// GPUs the unit is a SLICE OF THE SAMPLES of every pixel (render.rs:86-100): every pixel-sample is independent, the
// per-sample clamp happens before accumulation (render.rs:95-99), so per-device partial sums add linearly, and with
// Philox keyed by the global (pixel, sample) index the image does not depend on the device count except for fp32
// summation order.  One host thread per device issues that device's kernels (the wavefront path replays hundreds of
// graph launches per call; issued from one thread, a full launch queue on device 0 would stall the others).
//
// Cross-device step, once per render:
//   RTR_REDUCE_P2P   reduce_resolve_peers_kernel (kernels.cu) on every device: 1/N of the image each, summed over all
//                    peers' buffers by NVLink loads in rank order, accumulation + RGBA8 stored into the root's buffers.
//                    Ordering between devices is by CUDA events only (no kernel ever spins on another device).
//   RTR_REDUCE_NCCL  ncclReduce(sum, fp32, root 0) inside ncclGroupStart/End, then resolve_kernel on the root.
//                    NCCL is loaded with dlopen (libnccl.so.2): the library has no link-time dependency on it.
#include <dlfcn.h>

#include <atomic>
#include <chrono>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <new>
#include <string>
#include <thread>
#include <vector>

#include "context.h"

using namespace rtr;

namespace {

// ---- the five NCCL entry points used, resolved at run time (prototypes as in nccl.h 2.27)
struct NcclApi {
    void* handle = nullptr;
    int (*CommInitAll)(void** comms, int ndev, const int* devlist) = nullptr;
    int (*CommDestroy)(void* comm) = nullptr;
    int (*Reduce)(const void* send, void* recv, size_t count, int datatype, int op, int root, void* comm, cudaStream_t st) = nullptr;
    int (*GroupStart)() = nullptr;
    int (*GroupEnd)() = nullptr;
    const char* (*GetErrorString)(int) = nullptr;
    bool ok() const { return handle && CommInitAll && CommDestroy && Reduce && GroupStart && GroupEnd && GetErrorString; }
};
const int kNcclFloat = 7, kNcclSum = 0;  // ncclFloat32, ncclSum

NcclApi& nccl() {
    static NcclApi api;
    static std::once_flag once;
    std::call_once(once, [] {
        const char* names[] = {"libnccl.so.2", "libnccl.so"};
        for (const char* n : names) {
            api.handle = dlopen(n, RTLD_NOW | RTLD_LOCAL);
            if (api.handle) break;
        }
        if (!api.handle) return;
        api.CommInitAll = (decltype(api.CommInitAll))dlsym(api.handle, "ncclCommInitAll");
        api.CommDestroy = (decltype(api.CommDestroy))dlsym(api.handle, "ncclCommDestroy");
        api.Reduce = (decltype(api.Reduce))dlsym(api.handle, "ncclReduce");
        api.GroupStart = (decltype(api.GroupStart))dlsym(api.handle, "ncclGroupStart");
        api.GroupEnd = (decltype(api.GroupEnd))dlsym(api.handle, "ncclGroupEnd");
        api.GetErrorString = (decltype(api.GetErrorString))dlsym(api.handle, "ncclGetErrorString");
    });
    return api;
}

}  // namespace

struct rtr_group {
    uint32_t n = 0;
    std::vector<int> devices;
    std::vector<rtr_context*> ctx;   // one per member, each with its own stream
    std::vector<float4*> accum;      // per member: W*H float4 on its device
    size_t accum_pixels = 0;
    uchar4* rgba_root = nullptr;     // on devices[0]
    std::vector<cudaEvent_t> ev_started, ev_rendered, ev_reduced;  // per member, recorded on its stream
    std::vector<cudaEvent_t> ev_piece[RTR_MAX_GROUP];  // dynamic split: two pieces in flight per member
    cudaEvent_t ev_end = nullptr;    // root stream: end of the cross-device step
    uint32_t reduce = RTR_REDUCE_P2P;
    bool distinct = true;            // no device ordinal appears twice
    std::vector<void*> comms;        // ncclComm_t per member (NCCL mode)
    uint32_t split = RTR_SPLIT_STATIC, chunk = 0;
    std::vector<rtr_stats> last_stats;
    std::vector<uint32_t> last_begin, last_count;
    std::string error;
};

struct rtr_group_scene {
    rtr_group* group = nullptr;
    std::vector<rtr_scene*> scenes;  // per member
};

static rtr_status gfail(rtr_group* g, rtr_status st, const std::string& msg) {
    if (g) g->error = msg;
    return fail(nullptr, st, msg);
}

// contiguous, balanced partition of `count` samples over `n` members: (begin offset, count) of member r
static void slice_of(uint32_t count, uint32_t r, uint32_t n, uint32_t* begin, uint32_t* cnt) {
    const uint32_t base = count / n, rem = count % n;
    *begin = r * base + (r < rem ? r : rem);
    *cnt = base + (r < rem ? 1u : 0u);
}

extern "C" {

void rtr_group_destroy(rtr_group* g) {
    if (!g) return;
    for (uint32_t r = 0; r < g->ctx.size(); ++r) {
        if (!g->ctx[r]) continue;
        cudaSetDevice(g->devices[r]);
        cudaStreamSynchronize(g->ctx[r]->stream);
    }
    if (!g->comms.empty() && nccl().ok())
        for (void* c : g->comms) if (c) nccl().CommDestroy(c);
    for (uint32_t r = 0; r < g->ctx.size(); ++r) {
        if (!g->ctx[r]) continue;
        cudaSetDevice(g->devices[r]);
        if (r < g->accum.size() && g->accum[r]) cudaFree(g->accum[r]);
        if (r < g->ev_rendered.size() && g->ev_rendered[r]) cudaEventDestroy(g->ev_rendered[r]);
        if (r < g->ev_started.size() && g->ev_started[r]) cudaEventDestroy(g->ev_started[r]);
        for (cudaEvent_t e : g->ev_piece[r]) if (e) cudaEventDestroy(e);
        if (r < g->ev_reduced.size() && g->ev_reduced[r]) cudaEventDestroy(g->ev_reduced[r]);
        if (r == 0) {
            if (g->rgba_root) cudaFree(g->rgba_root);
            if (g->ev_end) cudaEventDestroy(g->ev_end);
        }
        rtr_context_destroy(g->ctx[r]);
    }
    delete g;
}

rtr_status rtr_group_create(const int* devices, uint32_t n_devices, uint32_t reduce, rtr_group** out) {
    if (!out) return gfail(nullptr, RTR_INVALID_ARG, "rtr_group_create: out is NULL");
    *out = nullptr;
    if (!devices || n_devices == 0 || n_devices > RTR_MAX_GROUP)
        return gfail(nullptr, RTR_INVALID_ARG, "rtr_group_create: need 1..16 devices");
    if (reduce > RTR_REDUCE_P2P) return gfail(nullptr, RTR_INVALID_ARG, "rtr_group_create: unknown reduce mode");
    rtr_group* g = new (std::nothrow) rtr_group();
    if (!g) return gfail(nullptr, RTR_OOM, "host allocation failed");
    g->n = n_devices;
    g->devices.assign(devices, devices + n_devices);
    g->ctx.assign(n_devices, nullptr);
    g->accum.assign(n_devices, nullptr);
    g->ev_started.assign(n_devices, nullptr);
    g->ev_rendered.assign(n_devices, nullptr);
    for (uint32_t r = 0; r < n_devices; ++r) g->ev_piece[r].assign(2, nullptr);
    g->ev_reduced.assign(n_devices, nullptr);
    g->last_stats.assign(n_devices, rtr_stats());
    g->last_begin.assign(n_devices, 0u);
    g->last_count.assign(n_devices, 0u);
    for (uint32_t a = 0; a < n_devices; ++a)
        for (uint32_t b = a + 1; b < n_devices; ++b)
            if (devices[a] == devices[b]) g->distinct = false;
    auto bail = [&](rtr_status st, const std::string& m) { rtr_group_destroy(g); return gfail(nullptr, st, m); };
    for (uint32_t r = 0; r < n_devices; ++r) {
        rtr_status st = rtr_context_create(devices[r], nullptr, &g->ctx[r]);
        if (st != RTR_OK) return bail(st, std::string("rtr_group_create: ") + rtr_last_error(nullptr));
        cudaError_t e = cudaEventCreate(&g->ev_rendered[r]);
        if (e == cudaSuccess) e = cudaEventCreate(&g->ev_started[r]);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&g->ev_reduced[r], cudaEventDisableTiming);
        for (int k = 0; k < 2 && e == cudaSuccess; ++k) e = cudaEventCreateWithFlags(&g->ev_piece[r][k], cudaEventDisableTiming);
        if (e == cudaSuccess && r == 0) e = cudaEventCreate(&g->ev_end);
        if (e != cudaSuccess) return bail(RTR_CUDA_ERROR, std::string("rtr_group_create: ") + cudaGetErrorString(e));
    }
    // peer access between every pair of distinct devices (NVSwitch: every peer at full bandwidth)
    bool all_peers = true;
    for (uint32_t a = 0; a < n_devices && all_peers; ++a) {
        for (uint32_t b = 0; b < n_devices; ++b) {
            if (devices[a] == devices[b]) continue;
            int can = 0;
            cudaSetDevice(devices[a]);
            if (cudaDeviceCanAccessPeer(&can, devices[a], devices[b]) != cudaSuccess || !can) { all_peers = false; break; }
            const cudaError_t e = cudaDeviceEnablePeerAccess(devices[b], 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) { all_peers = false; break; }
            (void)cudaGetLastError();
        }
    }
    (void)cudaGetLastError();
    if (reduce == RTR_REDUCE_AUTO) {
        const char* env = std::getenv("RTR_GROUP_REDUCE");
        if (env && std::strcmp(env, "nccl") == 0) reduce = RTR_REDUCE_NCCL;
        else if (env && std::strcmp(env, "p2p") == 0) reduce = RTR_REDUCE_P2P;
        else reduce = (all_peers || n_devices == 1) ? RTR_REDUCE_P2P : RTR_REDUCE_NCCL;
    }
    if (reduce == RTR_REDUCE_P2P && !all_peers && n_devices > 1)
        return bail(RTR_UNSUPPORTED, "rtr_group_create: RTR_REDUCE_P2P needs peer access between every pair of devices");
    if (reduce == RTR_REDUCE_NCCL && n_devices > 1) {
        if (!g->distinct) return bail(RTR_INVALID_ARG, "rtr_group_create: NCCL needs distinct devices");
        if (!nccl().ok()) return bail(RTR_NCCL_ERROR, "rtr_group_create: libnccl.so.2 could not be loaded");
        g->comms.assign(n_devices, nullptr);
        const int rc = nccl().CommInitAll(g->comms.data(), (int)n_devices, g->devices.data());
        if (rc != 0) return bail(RTR_NCCL_ERROR, std::string("ncclCommInitAll: ") + nccl().GetErrorString(rc));
    }
    g->reduce = reduce;
    *out = g;
    return RTR_OK;
}

uint32_t rtr_group_size(const rtr_group* g) { return g ? g->n : 0u; }
uint32_t rtr_group_reduce_mode(const rtr_group* g) { return g ? g->reduce : 0u; }
const char* rtr_group_last_error(const rtr_group* g) { return g ? g->error.c_str() : rtr_last_error(nullptr); }

rtr_status rtr_group_set_split(rtr_group* g, uint32_t split, uint32_t chunk) {
    if (!g || split > RTR_SPLIT_DYNAMIC) return gfail(g, RTR_INVALID_ARG, "rtr_group_set_split: bad argument");
    g->split = split; g->chunk = chunk;
    return RTR_OK;
}

void rtr_group_scene_destroy(rtr_group_scene* s) {
    if (!s) return;
    for (rtr_scene* sc : s->scenes) rtr_scene_destroy(sc);
    delete s;
}

rtr_status rtr_group_scene_upload(rtr_group* g, const rtr_scene_desc* desc, rtr_group_scene** out) {
    if (!g || !desc || !out) return gfail(g, RTR_INVALID_ARG, "rtr_group_scene_upload: null argument");
    *out = nullptr;
    PreparedScene prep;  // validate + rebuild the reference topology + flatten: once, whatever the device count
    rtr_status st = prepare_scene(g->ctx[0], desc, prep);
    if (st != RTR_OK) return gfail(g, st, g->ctx[0]->error);
    rtr_group_scene* s = new (std::nothrow) rtr_group_scene();
    if (!s) return gfail(g, RTR_OOM, "host allocation failed");
    s->group = g;
    s->scenes.assign(g->n, nullptr);
    std::vector<rtr_status> rc(g->n, RTR_OK);
    std::vector<std::thread> th;
    for (uint32_t r = 0; r < g->n; ++r)
        th.emplace_back([&, r] { rc[r] = upload_prepared(g->ctx[r], desc, prep, &s->scenes[r]); });
    for (auto& t : th) t.join();
    for (uint32_t r = 0; r < g->n; ++r)
        if (rc[r] != RTR_OK) {
            const std::string msg = g->ctx[r]->error;
            rtr_group_scene_destroy(s);
            return gfail(g, rc[r], msg);
        }
    *out = s;
    return RTR_OK;
}

rtr_status rtr_group_device_stats(const rtr_group* g, uint32_t i, rtr_stats* stats_out, uint32_t* sample_begin,
                                  uint32_t* sample_count) {
    if (!g || i >= g->n) return fail(nullptr, RTR_INVALID_ARG, "rtr_group_device_stats: bad argument");
    if (stats_out) *stats_out = g->last_stats[i];
    if (sample_begin) *sample_begin = g->last_begin[i];
    if (sample_count) *sample_count = g->last_count[i];
    return RTR_OK;
}

rtr_status rtr_group_render(rtr_group* g, const rtr_group_scene* scene, const rtr_render_params* p, uint8_t* rgba8_out,
                            float* accum_out, rtr_stats* stats_out) {
    if (!g || !scene || !p) return gfail(g, RTR_INVALID_ARG, "rtr_group_render: null argument");
    if (scene->group != g) return gfail(g, RTR_INVALID_ARG, "rtr_group_render: scene belongs to another group");
    const uint32_t N = g->n;
    for (uint32_t r = 0; r < N; ++r) {
        const rtr_status st = check_render_args(g->ctx[r], scene->scenes[r], p);
        if (st != RTR_OK) return gfail(g, st, g->ctx[r]->error);
    }
    if (stats_out) std::memset(stats_out, 0, sizeof(*stats_out));
    const size_t n_pix = (size_t)p->width * p->height;
    if (n_pix == 0) return RTR_OK;
    const uint32_t count = p->sample_count ? p->sample_count : p->settings.samples_per_pixel;
    // ---- buffers: one accumulation image per member, RGBA8 on the root
    if (g->accum_pixels < n_pix) {
        for (uint32_t r = 0; r < N; ++r) {
            cudaSetDevice(g->devices[r]);
            cudaStreamSynchronize(g->ctx[r]->stream);
            if (g->accum[r]) cudaFree(g->accum[r]);
            g->accum[r] = nullptr;
            cudaError_t e = cudaMalloc((void**)&g->accum[r], n_pix * sizeof(float4));
            if (e == cudaSuccess && r == 0) {
                if (g->rgba_root) cudaFree(g->rgba_root);
                g->rgba_root = nullptr;
                e = cudaMalloc((void**)&g->rgba_root, n_pix * sizeof(uchar4));
            }
            if (e != cudaSuccess) { g->accum_pixels = 0; return gfail(g, RTR_OOM, std::string("rtr_group_render: ") + cudaGetErrorString(e)); }
        }
        g->accum_pixels = n_pix;
    }
    // ---- render: one host thread per member
    std::vector<rtr_status> rc(N, RTR_OK);
    std::vector<rtr_stats> st(N);
    std::vector<uint32_t> launches(N, 0u);
    std::atomic<uint32_t> next_chunk{0};
    const bool dynamic = g->split == RTR_SPLIT_DYNAMIC && N > 1;
    uint32_t chunk = g->chunk ? g->chunk : (count + 8u * N - 1u) / (8u * N);  // default: ~8 pieces per device
    if (chunk == 0) chunk = 1;
    const uint32_t n_chunks = (count + chunk - 1) / chunk;
    auto worker = [&](uint32_t r) {
        rtr_context* ctx = g->ctx[r];
        std::memset(&st[r], 0, sizeof(rtr_stats));
        auto cu = [&](cudaError_t e, const char* what) { if (e != cudaSuccess && rc[r] == RTR_OK) rc[r] = cuda_fail(ctx, e, what); return e == cudaSuccess; };
        if (!cu(cudaSetDevice(g->devices[r]), "cudaSetDevice")) return;
        if (!cu(cudaEventRecord(g->ev_started[r], ctx->stream), "cudaEventRecord")) return;
        rtr_render_params q = *p;
        auto add = [&](const rtr_stats& one) {
            st[r].samples += one.samples; st[r].kernel_launches += one.kernel_launches; st[r].kernel_used = one.kernel_used;
        };
        if (!dynamic) {
            uint32_t off, cnt;
            slice_of(count, r, N, &off, &cnt);
            g->last_begin[r] = p->sample_begin + off; g->last_count[r] = cnt;
            if (cnt == 0) {  // more devices than samples: an all-zero partial image
                cu(cudaMemsetAsync(g->accum[r], 0, n_pix * sizeof(float4), ctx->stream), "cudaMemsetAsync");
                cu(cudaMemsetAsync(ctx->stats, 0, STAT_COUNT * sizeof(unsigned long long), ctx->stream), "cudaMemsetAsync");
            } else {
                q.sample_begin = p->sample_begin + off; q.sample_count = cnt;
                q.flags &= ~(uint32_t)RTR_FLAG_ACCUMULATE;
                rtr_stats one;
                rc[r] = render_into(ctx, scene->scenes[r], &q, g->accum[r], false, &one);
                if (rc[r] == RTR_OK) add(one);
            }
        } else {
            // dynamic: pull `chunk`-sample pieces from the shared counter (rayon's work stealing, render.rs:114-119, at
            // device granularity).  The first piece overwrites this device's buffer, the others accumulate into it; two
            // pieces are kept in flight so that the device never waits for the host.
            uint32_t issued = 0;
            g->last_begin[r] = 0; g->last_count[r] = 0;
            for (;;) {
                if (issued >= 2u && !cu(cudaEventSynchronize(g->ev_piece[r][issued & 1u]), "cudaEventSynchronize")) break;
                const uint32_t c = next_chunk.fetch_add(1u);
                if (c >= n_chunks) break;
                q.sample_begin = p->sample_begin + c * chunk;
                q.sample_count = (uint64_t)(c + 1u) * chunk <= count ? chunk : count - c * chunk;
                q.flags = (p->flags & ~(uint32_t)RTR_FLAG_ACCUMULATE) | (issued ? (uint32_t)RTR_FLAG_ACCUMULATE : 0u);
                rtr_stats one;
                const rtr_status s1 = render_into(ctx, scene->scenes[r], &q, g->accum[r], false, &one, /*reset_stats=*/issued == 0u);
                if (s1 != RTR_OK) { rc[r] = s1; break; }
                add(one);
                if (!cu(cudaEventRecord(g->ev_piece[r][issued & 1u], ctx->stream), "cudaEventRecord")) break;
                if (issued == 0u) g->last_begin[r] = q.sample_begin;
                g->last_count[r] += q.sample_count;
                ++issued;
            }
            if (issued == 0u && rc[r] == RTR_OK) {
                cu(cudaMemsetAsync(g->accum[r], 0, n_pix * sizeof(float4), ctx->stream), "cudaMemsetAsync");
                cu(cudaMemsetAsync(ctx->stats, 0, STAT_COUNT * sizeof(unsigned long long), ctx->stream), "cudaMemsetAsync");
            }
        }
        if (rc[r] == RTR_OK) cu(cudaEventRecord(g->ev_rendered[r], ctx->stream), "cudaEventRecord");
    };
    if (N == 1) worker(0);
    else {
        std::vector<std::thread> th;
        for (uint32_t r = 0; r < N; ++r) th.emplace_back(worker, r);
        for (auto& t : th) t.join();
    }
    for (uint32_t r = 0; r < N; ++r)
        if (rc[r] != RTR_OK) {
            for (uint32_t k = 0; k < N; ++k) { cudaSetDevice(g->devices[k]); cudaStreamSynchronize(g->ctx[k]->stream); }
            return gfail(g, rc[r], g->ctx[r]->error);
        }
    // ---- the one cross-device step
    rtr_context* ctx = g->ctx[0];  // CU() reports through the root context
    uint32_t extra_launches = 0;
    if (g->reduce == RTR_REDUCE_NCCL && N > 1) {
        int nrc = nccl().GroupStart();
        for (uint32_t r = 0; r < N && nrc == 0; ++r)
            nrc = nccl().Reduce(g->accum[r], g->accum[0], n_pix * 4, kNcclFloat, kNcclSum, 0, g->comms[r], g->ctx[r]->stream);
        const int erc = nccl().GroupEnd();
        if (nrc == 0) nrc = erc;
        if (nrc != 0) return gfail(g, RTR_NCCL_ERROR, std::string("ncclReduce: ") + nccl().GetErrorString(nrc));
        CU(cudaSetDevice(g->devices[0]));
        CU(launch_resolve(g->accum[0], g->rgba_root, (uint32_t)n_pix, (float)count, ctx->stream));
        extra_launches = 1;
    } else {
        PeerAccum peers;
        std::memset(&peers, 0, sizeof(peers));
        for (uint32_t r = 0; r < N; ++r) peers.accum[r] = g->accum[r];
        for (uint32_t d = 0; d < N; ++d) {
            CU(cudaSetDevice(g->devices[d]));
            for (uint32_t r = 0; r < N; ++r)
                if (r != d) CU(cudaStreamWaitEvent(g->ctx[d]->stream, g->ev_rendered[r], 0));
            const uint32_t lo = (uint32_t)((uint64_t)n_pix * d / N), hi = (uint32_t)((uint64_t)n_pix * (d + 1) / N);
            CU(launch_reduce_resolve_peers(peers, N, g->accum[0], g->rgba_root, lo, hi, (float)count, g->ctx[d]->stream));
            CU(cudaEventRecord(g->ev_reduced[d], g->ctx[d]->stream));
            ++extra_launches;
        }
        CU(cudaSetDevice(g->devices[0]));
        for (uint32_t d = 1; d < N; ++d) CU(cudaStreamWaitEvent(ctx->stream, g->ev_reduced[d], 0));
    }
    CU(cudaEventRecord(g->ev_end, ctx->stream));
    if (rgba8_out) CU(cudaMemcpyAsync(rgba8_out, g->rgba_root, n_pix * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (accum_out) CU(cudaMemcpyAsync(accum_out, g->accum[0], n_pix * sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    for (uint32_t r = 0; r < N; ++r) {  // nobody may overwrite a buffer a peer is still reading; device counters -> host
        CU(cudaSetDevice(g->devices[r]));
        const rtr_status cs = collect_stats(g->ctx[r], false, &st[r]);  // synchronises the member's stream
        if (cs != RTR_OK) return gfail(g, cs, g->ctx[r]->error);
        float ms = 0.0f;
        CU(cudaEventElapsedTime(&ms, g->ev_started[r], g->ev_rendered[r]));
        st[r].device_ms = ms;
    }
    CU(cudaSetDevice(g->devices[0]));
    float total_ms = 0.0f;
    CU(cudaEventElapsedTime(&total_ms, g->ev_started[0], g->ev_end));
    for (uint32_t r = 0; r < N; ++r) g->last_stats[r] = st[r];
    if (stats_out) {
        double max_ms = 0.0;
        for (uint32_t r = 0; r < N; ++r) {
            stats_out->rays += st[r].rays; stats_out->samples += st[r].samples;
            stats_out->nodes_visited += st[r].nodes_visited; stats_out->tri_tests += st[r].tri_tests;
            stats_out->sphere_tests += st[r].sphere_tests; stats_out->env_lookups += st[r].env_lookups;
            stats_out->kernel_launches += st[r].kernel_launches;
            if (st[r].kernel_used) stats_out->kernel_used = st[r].kernel_used;
            if (st[r].device_ms > max_ms) max_ms = st[r].device_ms;
        }
        stats_out->kernel_launches += extra_launches;
        stats_out->device_ms = total_ms > 0.0f ? (double)total_ms : max_ms;
        stats_out->upload_ms = max_ms;  // slowest member's render alone: device_ms - this = the cross-device step + skew
    }
    return RTR_OK;
}

}  // extern "C"
