// kernels.h — host-visible launch interface of kernels.cu (plain structs, no device code)
#ifndef RTR_KERNELS_H_
#define RTR_KERNELS_H_

#include <cuda_runtime_api.h>
#include <vector_types.h>
#include <stdint.h>

#include "../../include/rtr.h"
#include "flat.h"

namespace rtr {

#define RTR_MEGA_THREADS 128
#ifndef RTR_MEGA_MIN_BLOCKS
#define RTR_MEGA_MIN_BLOCKS 8  // resident CTAs per SM the register allocation must allow (swept on B200 4..8; 8 since Volume::hit is inlined)
#endif

#ifndef RTR_MEGA_MIN_BLOCKS_SCAN
#define RTR_MEGA_MIN_BLOCKS_SCAN 7  // the same for the BVH-off instantiations (linear scan, no traversal state)
#endif

#ifndef RTR_SCAN_MIN_BLOCKS
#define RTR_SCAN_MIN_BLOCKS 7  // resident CTAs per SM of the all-sphere scan kernel (scan.cu)
#endif

#ifndef RTR_SHORT_STACK
#define RTR_SHORT_STACK 8  // traversal-stack entries per thread kept in shared memory (deeper ones spill to local memory)
#endif

struct v3 { float x, y, z; };

// RayOrigin (lib/core/camera.rs:11-16), computed once per render on the host
struct CameraBasis { v3 origin, horizontal, vertical, lower_left_corner; };

// the ten Philox4x32-10 round keys (seed + i * Weyl constant), filled by the host once per launch
struct PhiloxKeys { uint32_t rk0[10], rk1[10]; };
inline PhiloxKeys make_philox_keys(uint32_t k0, uint32_t k1) {
    PhiloxKeys pk;
    for (uint32_t i = 0; i < 10; ++i) { pk.rk0[i] = k0 + i * 0x9E3779B9u; pk.rk1[i] = k1 + i * 0xBB67AE85u; }
    return pk;
}

enum { STAT_RAYS = 0, STAT_NODES = 1, STAT_TRIS = 2, STAT_SPHERES = 3, STAT_ENV = 4, STAT_COUNT = 8 };

struct RenderArgs {
    SceneView sc;
    CameraBasis cam;
    uint32_t width, height;
    uint32_t max_ray_depth;
    float clamp_indirect;
    uint32_t enable_bvh_tree;
    uint32_t mode;
    uint32_t k0, k1;             // Philox key = seed
    PhiloxKeys pk;               // round keys of (k0, k1): LOP3 reads them straight from the constant bank
    uint32_t sample_begin, sample_count;
    uint32_t tiles_x, n_tiles;   // 8x4-pixel warp tiles
    uint32_t n_chunks, chunk_samples;  // sample chunks per tile (1 unless the image is small)
    uint32_t refill_threshold;   // 1..32: refill finished lanes once fewer than this many are still walking
    uint32_t accumulate;         // RTR_FLAG_ACCUMULATE: add to accum instead of overwriting it
    float4* accum;               // W*H: (sum r, sum g, sum b, samples)
    float4* partials;            // n_chunks * W*H when n_chunks > 1
    uint32_t* work_counter;      // zeroed before launch
    unsigned long long* stats;   // STAT_COUNT counters
};

int megakernel_max_blocks_per_sm(bool exhaustive, bool count, bool bvh, bool volumes);
cudaError_t launch_megakernel(const RenderArgs& a, int grid, bool exhaustive, bool count, cudaStream_t st);
// scan.cu: BVH off + every object a sphere (scenes/benchmark.json); same RenderArgs, same results as the megakernel
int scan_kernel_max_blocks_per_sm(bool count);
cudaError_t launch_scan_kernel(const RenderArgs& a, int grid, bool count, cudaStream_t st);
cudaError_t launch_sum_chunks(const float4* partials, float4* accum, uint32_t n_pixels, uint32_t n_chunks, bool accumulate, cudaStream_t st);
cudaError_t launch_expand_env(const float* rgb, float4* out, size_t n, cudaStream_t st);
cudaError_t launch_resolve(const float4* accum, uchar4* rgba, uint32_t n_pixels, float spp, cudaStream_t st);
// Multi-device reduce fused with the resolve (group.cpp): for the pixels [lo, hi) this device owns, sum the N devices'
// accumulation buffers IN RANK ORDER (peer loads over NVLink), store the sum in the root's accumulation buffer and
// sqrt(sum/spp) -> RGBA8 in the root's image (peer stores).  One launch per device, no other collective.
#define RTR_MAX_GROUP 16
struct PeerAccum { const float4* accum[RTR_MAX_GROUP]; };
cudaError_t launch_reduce_resolve_peers(const PeerAccum& peers, uint32_t n_ranks, float4* accum_root, uchar4* rgba_root,
                                        uint32_t lo, uint32_t hi, float spp, cudaStream_t st);
cudaError_t launch_scatter_batch(const SceneView& sc, const rtr_scatter_in* in, size_t n, uint32_t mode, uint32_t k0,
                                 uint32_t k1, rtr_scatter_out* out, cudaStream_t st);
cudaError_t launch_background_batch(const SceneView& sc, const rtr_ray* rays, size_t n, float* rgb, cudaStream_t st);
cudaError_t launch_camera_rays(const CameraBasis& cam, uint32_t W, uint32_t H, uint32_t k0, uint32_t k1,
                               const uint32_t* pixels, const uint32_t* samples, size_t n, rtr_ray* out,
                               cudaStream_t st);

// ------------------------------------------------------------------ wavefront variant (wavefront.cu)
// One WAVE = `wave_samples` consecutive global samples of every pixel, traced to completion:
//   wf_generate -> [ wf_extend -> wf_shade(+compact) ] x max_ray_depth -> (wf_sum_samples)
// Path state lives in HBM as SoA float4 streams, double-buffered so that wf_shade writes the
// surviving paths DENSELY (warp ballot + popc + one atomic per warp) into the other buffer.
#define RTR_WF_THREADS 128
#ifndef RTR_WF_EXTEND_THREADS
#define RTR_WF_EXTEND_THREADS 128  // CTA size of wf_extend (swept 64 / 128 / 256 on B200)
#endif
#ifndef RTR_WF_EXTEND_MIN_BLOCKS
#define RTR_WF_EXTEND_MIN_BLOCKS 8  // wf_extend with Volume::hit compiled in (64 registers)
#endif
#ifndef RTR_WF_EXTEND_MIN_BLOCKS_W4
#define RTR_WF_EXTEND_MIN_BLOCKS_W4 8        // the 4-wide instantiations (swept 6..8 with Volume::hit, 7..9 without: profiles/r2/ab_bvh4_variants.txt)
#endif
#ifndef RTR_WF_EXTEND_MIN_BLOCKS_NOVOL_W4
#define RTR_WF_EXTEND_MIN_BLOCKS_NOVOL_W4 9  // 56 registers, no spills; under two wave sequences +1..+2.5 % over 8 (profiles/r2/ab_extend_occupancy_two_sequences.txt)
#endif
#ifndef RTR_WF_EXTEND_MIN_BLOCKS_NOVOL
#define RTR_WF_EXTEND_MIN_BLOCKS_NOVOL 9  // without it: 56 registers, swept 6..10 on B200 (9: +2..5 % over 8)
#endif
#define RTR_WF_MAX_DEPTH 256      // max_ray_depth is a u8 in the reference
enum { WFC_WAVE = 0, WFC_QCOUNT = 8, WFC_FETCH = WFC_QCOUNT + RTR_WF_MAX_DEPTH + 8, WFC_WORDS = WFC_FETCH + RTR_WF_MAX_DEPTH + 8 };

struct WfPaths {      // one path per slot
    float4* o;        // ray origin.xyz, w = bits(pixel index)
    float4* d;        // ray direction.xyz, w = bits((sample - wave_begin) << 8 | depth)
    float4* thr;      // throughput T.xyz (product of attenuations so far), w unused
};

struct WfArgs {
    SceneView sc;
    CameraBasis cam;
    uint32_t width, height, n_pixels;
    uint32_t max_ray_depth;
    float clamp_indirect;
    uint32_t enable_bvh_tree;
    uint32_t mode;
    uint32_t k0, k1;
    uint32_t sample_begin, sample_end;  // this call's global sample range
    uint32_t wave_samples;              // samples of every pixel in flight per wave
    uint32_t wave_first, wave_step;     // this launch sequence renders waves wave_first, wave_first + wave_step, ... (two sequences overlap: capi.cpp)
    uint32_t bounce;                    // which [extend, shade] pair this launch is
    uint32_t refill_threshold;          // extend refills idle lanes when fewer lanes than this are walking
    uint32_t atomics_mode;              // bit 0: wf_generate hands out slots without atomics when the image tiles exactly;
                                        // bit 1: wf_shade aggregates its compaction atomics per CTA
    float tmin;                         // world.hit(ray, tmin, +INF): 0.00001 in render() (render.rs:38); the caller's in rtr_trace_batch
    uint32_t accumulate;                // RTR_FLAG_ACCUMULATE: keep what accum already holds
    WfPaths in, out;
    float4* hit;                        // per slot of `in`: (t, bits(ref), u, v)
    float4* accum;                      // W*H (sum r, sum g, sum b, samples)
    float4* lbuf;                       // wave_samples * n_pixels clamped sample radiances (wave_samples > 1)
    uint32_t* counters;                 // WFC_WORDS words, see enum
    unsigned long long* stats;
    // direction binning of bounces >= 1 (wf_bin_kernel): perm[i] = slot of `in` the i-th fetch of wf_extend takes
    const uint32_t* perm;               // null: slots are fetched in stream order
    uint32_t* perm_out;                 // what wf_bin writes (the same buffer)
    uint32_t bin_keys;                  // 8: direction octant; 24: octant x dominant axis
};

int wf_extend_max_blocks_per_sm(bool exhaustive, bool count, bool volumes, bool tmin_arg = false, bool wide4 = false);
bool wf_extend_wide4(const WfArgs& a, bool exhaustive, bool count);  // will launch_wf_extend walk the 4-wide tree for these arguments?
cudaError_t launch_wf_begin(const WfArgs& a, cudaStream_t st);       // zero accum + wave counter
cudaError_t launch_wf_generate(const WfArgs& a, int grid, cudaStream_t st);
cudaError_t launch_wf_extend(const WfArgs& a, int grid, bool exhaustive, bool count, cudaStream_t st);
cudaError_t launch_wf_shade(const WfArgs& a, int grid, bool count, cudaStream_t st);
int wf_shade_threads();  // CTA size of wf_shade (grid sizing)
cudaError_t launch_wf_begin_lane(const WfArgs& a, cudaStream_t st);  // the counters of a second wave sequence only
cudaError_t launch_wf_end_wave(const WfArgs& a, int grid, cudaStream_t st);  // ordered sum of lbuf + advance the wave
// rtr_trace_batch = unpack (AoS rays -> the wavefront's ray streams, queue size n) -> wf_extend (the render's own
// persistent, majority-rule, short-stack walk) -> finish (HitRecord of every winner -> AoS rtr_hit)
cudaError_t launch_trace_unpack(const WfArgs& a, const rtr_ray* rays, uint32_t n, uint32_t first_index, cudaStream_t st);
cudaError_t launch_trace_finish(const WfArgs& a, uint32_t n, rtr_hit* out, cudaStream_t st);
// stable counting sort of every RTR_WF_BIN_BLOCK consecutive queued paths by ray-direction key -> perm_out
#define RTR_WF_BIN_BLOCK 1024
cudaError_t launch_wf_bin(const WfArgs& a, int grid, cudaStream_t st);

}  // namespace rtr
#endif
