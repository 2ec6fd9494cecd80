// wavefront.cu — the wavefront variant of render() (lib/core/render.rs:61-127) for sm_100a.
//
// One WAVE traces `wave_samples` consecutive samples of every pixel to completion:
//
//   wf_generate   per-pixel sample generation (render.rs:88-92): camera rays -> path stream, bounce 0
//   wf_extend     world.hit(ray, 1e-5, +INF) (render.rs:38) for every live path.  Persistent warps pull
//                 rays from the stream with one atomic per refill; a lane whose walk has finished stores
//                 its hit and takes the next ray while its neighbours keep walking, so the node/leaf
//                 steps run with most lanes busy (the megakernel can only refill by running the whole
//                 shading routine divergently).
//   wf_shade      background / HitRecord / Material::scatter for every live path (render.rs:39-57),
//                 then COMPACT: surviving paths are written densely into the other path buffer
//                 (warp ballot + popc prefix + one atomicAdd per warp); finished paths add their clamped
//                 radiance to the pixel (render.rs:95-99).
//   wf_sum_samples / wf_next_wave   ordered per-pixel sum of the wave's samples, counters for the next wave
//
// All queue sizes live on the device (counters[]), so a wave is a fixed launch sequence with no host
// synchronisation; capi.cpp captures it once into a CUDA graph and replays it per wave.
// Results are bit-identical to the megakernel: the same per-path arithmetic and Philox counters, and each
// pixel's samples are summed in sample order.
#include <cstdlib>

#include "device_core.cuh"
#include "kernels.h"

#ifndef RTR_WF_SHADE_THREADS
#define RTR_WF_SHADE_THREADS 256  // CTA size of wf_shade = paths per compaction atomic in the per-CTA mode (A/B 128..1024: profiles/r2/ab_shade_cta.txt)
#endif
#ifndef RTR_WF_SHADE_MIN_BLOCKS
#define RTR_WF_SHADE_MIN_BLOCKS 4  // wf_shade waits on memory (ncu: 3.8 long-scoreboard stalls per issue); swept 9 / 12 / 16
#endif
#ifndef RTR_WF_SHADE_PREFETCH
#define RTR_WF_SHADE_PREFETCH 1  // +0.2..+0.7 % (profiles/r2/ab_shade_prefetch_occupancy.txt)
#endif
#ifndef RTR_WF_NODE_BURST
#define RTR_WF_NODE_BURST 0  // 1: two node steps per scheduling decision
#endif

namespace rtr {

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

struct WaveInfo { uint32_t begin, samples, n_paths; };
__device__ __forceinline__ WaveInfo wave_info(const WfArgs& a) {
    WaveInfo w;
    const uint32_t wave = a.counters[WFC_WAVE];
    w.begin = a.sample_begin + wave * a.wave_samples;
    const uint32_t left = a.sample_end > w.begin ? a.sample_end - w.begin : 0u;
    w.samples = min(a.wave_samples, left);
    w.n_paths = w.samples * a.n_pixels;
    return w;
}

// clamp + accumulate one finished sample (render.rs:95-99)
__device__ __forceinline__ void finish_sample(const WfArgs& a, uint32_t pixel, uint32_t rel, v3 L) {
    L.x = f32_clamp(L.x, 0.0f, a.clamp_indirect);
    L.y = f32_clamp(L.y, 0.0f, a.clamp_indirect);
    L.z = f32_clamp(L.z, 0.0f, a.clamp_indirect);
    if (a.lbuf) {
        a.lbuf[(size_t)rel * a.n_pixels + pixel] = make_float4(L.x, L.y, L.z, 1.0f);
    } else {  // one path per pixel per wave: this thread is the only writer of the pixel in this launch
        float4 c = a.accum[pixel];
        c.x = c.x + L.x; c.y = c.y + L.y; c.z = c.z + L.z; c.w = c.w + 1.0f;
        a.accum[pixel] = c;
    }
}

__global__ void wf_begin_kernel(const WfArgs a) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < a.n_pixels && !a.accumulate) a.accum[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (i < WFC_WORDS) a.counters[i] = i == (uint32_t)WFC_WAVE ? a.wave_first : 0u;
}

// ------------------------------------------------------------------ generate
// Slots are handed out in 8x4-pixel tile order so that a warp of wf_extend starts with coherent rays.
__global__ void __launch_bounds__(RTR_WF_THREADS) wf_generate_kernel(const WfArgs a) {
    const WaveInfo w = wave_info(a);
    const uint32_t tiles_x = (a.width + 7u) / 8u, tiles_y = (a.height + 3u) / 4u;
    const uint32_t per_sample = tiles_x * tiles_y * 32u;
    const uint64_t total = (uint64_t)per_sample * w.samples;
    const uint32_t lane = threadIdx.x & 31u;
    // One atomicAdd per warp on the queue counter made this kernel atomic-bound (~1 same-address atomic per ns: 512 K of
    // them for a 16 M-path wave = 0.48 ms at 24 % issue, 19 % DRAM).  When the image is a whole number of 8x4 tiles no
    // compaction is needed at all: slot = tile-order index, the queue size is written once.
    const bool direct = (a.atomics_mode & 1u) != 0u && (a.width & 7u) == 0u && (a.height & 3u) == 0u;
    if (direct && blockIdx.x == 0 && threadIdx.x == 0) a.counters[WFC_QCOUNT] = a.max_ray_depth != 0u ? (uint32_t)total : 0u;
    for (uint64_t base = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) - lane; base < total;
         base += (uint64_t)gridDim.x * blockDim.x) {
        const uint64_t i = base + lane;
        const uint32_t rel = (uint32_t)(i / per_sample), p = (uint32_t)(i - (uint64_t)rel * per_sample);
        const uint32_t tile = p >> 5, tx = tile % tiles_x, ty = tile / tiles_x;
        const uint32_t x = tx * 8u + (lane & 7u), y = ty * 4u + (lane >> 3);
        const bool valid = i < total && x < a.width && y < a.height;
        const uint32_t pixel = y * a.width + x;
        if (valid && a.max_ray_depth == 0u) finish_sample(a, pixel, rel, mk(1.0f, 1.0f, 1.0f) * mk(0.0f, 0.0f, 0.0f));  // render.rs:34-36
        const bool live = valid && a.max_ray_depth != 0u;
        const uint32_t m = __ballot_sync(0xffffffffu, live);
        if (m == 0u) continue;
        uint32_t slot = 0;
        if (direct) {
            slot = (uint32_t)i;  // every lane of every tile is a pixel: the tile-order index IS the dense slot
        } else {
            if (lane == (uint32_t)(__ffs(m) - 1)) slot = atomicAdd(&a.counters[WFC_QCOUNT], (uint32_t)__popc(m));
            slot = __shfl_sync(0xffffffffu, slot, __ffs(m) - 1) + __popc(m & lanemask_lt());
        }
        if (live) {
            RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = w.begin + rel; key.bounce = 0;
            const Ray r = camera_sample_ray(a.cam, a.width, a.height, x, y, key);  // render.rs:88-92
            a.in.o[slot] = make_float4(r.o.x, r.o.y, r.o.z, __uint_as_float(pixel));
            a.in.d[slot] = make_float4(r.d.x, r.d.y, r.d.z, __uint_as_float(rel << 8));
            a.in.thr[slot] = make_float4(1.0f, 1.0f, 1.0f, 0.0f);
        }
    }
}

// ------------------------------------------------------------------ direction binning (bounces >= 1)
// Compaction keeps the paths in stream order, i.e. in the 8x4-tile order of wf_generate: after a bounce, neighbouring
// slots still START close together (neighbouring pixels hit neighbouring surface points) but their scattered
// DIRECTIONS are independent, so the 32 rays a warp of wf_extend fetches fan out into all octants at once: every
// lane sits in its own node (ncu, round 1: 14-16 of 32 lanes per instruction and the L1 data pipe at 71-79 % of
// peak on bounces 1-4, against 20 lanes / 54 % on the coherent camera rays).  This kernel regroups each block of
// RTR_WF_BIN_BLOCK consecutive slots by a direction key with a STABLE counting sort (stream order, hence origin
// locality, survives inside a group) and writes the permutation; wf_extend then fetches slot perm[i] instead of i.
// Path data does not move (the permutation is 4 B per path against 112 B of state), wf_shade is unchanged, and every
// path's arithmetic is what it was: results stay bit-identical.
__device__ __forceinline__ uint32_t direction_key(const float4& d, uint32_t bin_keys) {
    const uint32_t oct = (__float_as_uint(d.x) >> 31) | ((__float_as_uint(d.y) >> 31) << 1) | ((__float_as_uint(d.z) >> 31) << 2);
    if (bin_keys <= 8u) return oct;
    const float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    const uint32_t major = (ax >= ay && ax >= az) ? 0u : (ay >= az ? 1u : 2u);
    return major * 8u + oct;  // < 24
}

__global__ void __launch_bounds__(256) wf_bin_kernel(const WfArgs a) {
    constexpr uint32_t KEYS = 32u, CHUNKS = RTR_WF_BIN_BLOCK / 32u;  // a chunk = 32 consecutive slots (one warp pass)
    __shared__ uint32_t table[KEYS * CHUNKS];                     // table[key][chunk] = how many of the chunk hold key
    __shared__ uint32_t warp_sums[8];
    const uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    const uint32_t n = a.counters[WFC_QCOUNT + a.bounce];
    const uint32_t n_blocks = (n + RTR_WF_BIN_BLOCK - 1u) / RTR_WF_BIN_BLOCK;
    for (uint32_t blk = blockIdx.x; blk < n_blocks; blk += gridDim.x) {
        const uint32_t base = blk * RTR_WF_BIN_BLOCK;
        if (base + RTR_WF_BIN_BLOCK > n) {  // the ragged tail keeps stream order
            for (uint32_t i = base + threadIdx.x; i < n; i += blockDim.x) a.perm_out[i] = i;
            continue;
        }
        for (uint32_t i = threadIdx.x; i < KEYS * CHUNKS; i += blockDim.x) table[i] = 0u;
        __syncthreads();
        uint32_t key[CHUNKS / 8u], rank[CHUNKS / 8u];
#pragma unroll
        for (uint32_t k = 0; k < CHUNKS / 8u; ++k) {  // chunk = k * 8 + warp
            const uint32_t chunk = k * 8u + warp;
            key[k] = direction_key(a.in.d[base + chunk * 32u + lane], a.bin_keys);
            const uint32_t peers = __match_any_sync(FULL, key[k]);
            rank[k] = (uint32_t)__popc(peers & lanemask_lt());
            if (rank[k] == 0u) table[key[k] * CHUNKS + chunk] = (uint32_t)__popc(peers);
        }
        __syncthreads();
        {   // exclusive prefix over the table in (key, chunk) order: 4 entries per thread + a block scan
            const uint32_t i0 = threadIdx.x * 4u;
            const uint32_t c0 = table[i0], c1 = table[i0 + 1u], c2 = table[i0 + 2u], c3 = table[i0 + 3u];
            const uint32_t mine = c0 + c1 + c2 + c3;
            uint32_t incl = mine;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t v = __shfl_up_sync(FULL, incl, o);
                if ((int)lane >= o) incl += v;
            }
            if (lane == 31u) warp_sums[warp] = incl;
            __syncthreads();
            uint32_t before = 0u;
            for (uint32_t w = 0; w < warp; ++w) before += warp_sums[w];
            const uint32_t excl = before + incl - mine;
            table[i0] = excl; table[i0 + 1u] = excl + c0; table[i0 + 2u] = excl + c0 + c1; table[i0 + 3u] = excl + c0 + c1 + c2;
        }
        __syncthreads();
#pragma unroll
        for (uint32_t k = 0; k < CHUNKS / 8u; ++k) {
            const uint32_t chunk = k * 8u + warp;
            a.perm_out[base + table[key[k] * CHUNKS + chunk] + rank[k]] = base + chunk * 32u + lane;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------ extend
// VOLUMES = false: the scene holds no Volume and the BVH is on (every shipped scene but C5).  Volume::hit is a
// non-inlined call with two nested walks; compiling it out of the walk removes the call's register save/restore
// and the per-ray RNG key from the traversal state (ncu: local-memory traffic was 44 % of wf_extend's L1 wavefronts).
// TMIN_ARG = true: world.hit's tmin comes from the arguments (rtr_trace_batch with a tmin other than render()'s 1e-5);
// the render path keeps the literal, which costs the volume instantiation 30 bytes of spills less.
// WIDE4 = true: interior steps walk the 4-wide tree (flat.h: nodes4, node4_step); culled, render-tmin walks of the traversal tree only.
template <bool EXHAUSTIVE, bool COUNT, bool VOLUMES, bool TMIN_ARG = false, bool WIDE4 = false>
__global__ void __launch_bounds__(RTR_WF_EXTEND_THREADS, WIDE4 ? (VOLUMES ? RTR_WF_EXTEND_MIN_BLOCKS_W4 : RTR_WF_EXTEND_MIN_BLOCKS_NOVOL_W4)
                                                               : (VOLUMES ? RTR_WF_EXTEND_MIN_BLOCKS : RTR_WF_EXTEND_MIN_BLOCKS_NOVOL)) wf_extend_kernel(const WfArgs a) {
    const float tmin0 = TMIN_ARG ? a.tmin : 0.00001f;  // render.rs:38
    StackEntry stack[RTR_STACK_CAP];  // spill area of the short stack + scratch of nested (volume boundary) walks
    __shared__ uint2 short_stack[RTR_SHORT_STACK][RTR_WF_EXTEND_THREADS];
    ShortStack<RTR_SHORT_STACK, RTR_WF_EXTEND_THREADS> wstack;
    wstack.sh = &short_stack[0][threadIdx.x];
    wstack.loc = stack;
    const uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = a.counters[WFC_QCOUNT + a.bounce];
    const WaveInfo w = wave_info(a);
    uint32_t* fetch = &a.counters[WFC_FETCH + a.bounce];
    Counters cnt; cnt.nodes = cnt.tris = cnt.spheres = cnt.env = 0;

    if (VOLUMES && a.enable_bvh_tree == 0u) {  // Hittable for Scene: linear scan (scene.rs:141-156), uniform work per ray
        for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
            const float4 o = a.in.o[i], d = a.in.d[i];
            Ray r; r.o = xyz(o); r.d = xyz(d);
            RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = __float_as_uint(o.w);
            key.sample = w.begin + (__float_as_uint(d.w) >> 8); key.bounce = __float_as_uint(d.w) & 0xFFu;
            const LiteHit h = world_hit<EXHAUSTIVE, COUNT>(a.sc, false, r, tmin0, key, stack, &cnt);
            a.hit[i] = make_float4(h.t, __uint_as_float(h.ref), h.u, h.v);
        }
    } else {
        bool has = false, drained = false;
        uint32_t idx = 0, cur = RTR_REF_DONE;
        int sp = 0;
        Ray r; r.o = r.d = mk(0.0f, 0.0f, 0.0f);
        RayInv inv; inv.ix = inv.iy = inv.iz = 0.0f;
        RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = key.sample = key.bounce = 0;
        LiteHit best; best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
        // Volume::hit as part of the lane's resumable walk (VOLUMES only).  A lane that reaches a Volume leaf does not
        // run the two boundary queries of volume.rs:36-37 as nested loops while the rest of the warp waits (ncu on C5:
        // 3.4 of 32 lanes per instruction); it SUSPENDS its outer walk — the outer best hit goes onto its traversal
        // stack, `floor` marks the stack level the nested walk must not pop below — and walks the boundary with the
        // same node / leaf steps the other lanes are taking: phase 1 over (-inf, +inf), phase 2 from t1 + tmin.  When
        // phase 2 ends the medium is sampled (volume_sample), the outer walk is restored and offered the hit.
        uint32_t vstate = 0u;       // 0: outer walk; else phase (1 | 2) << 30 | index in volumes[]
        int floor = 0;              // the nested walk's stack bottom
        float tmin_cur = tmin0;  // tmin of the walk in progress (render.rs:38; volume.rs:36-37)
        for (;;) {
            // ---- refill idle lanes from the stream: one atomic per warp  [bucket:wf_fetch]
            const uint32_t idle = __ballot_sync(FULL, !has);
            if (!drained && idle != 0u) {
                uint32_t base = 0;
                if (lane == 0) base = atomicAdd(fetch, (uint32_t)__popc(idle));
                base = __shfl_sync(FULL, base, 0);
                if (!has) {
                    idx = base + __popc(idle & lanemask_lt());
                    if (idx < n) {
                        if (a.perm) idx = a.perm[idx];  // direction-binned order (wf_bin_kernel); idx < n still holds
                        const float4 o = a.in.o[idx], d = a.in.d[idx];
                        r.o = xyz(o); r.d = xyz(d);
                        if (VOLUMES) {  // only Volume::hit draws inside world.hit (volume.rs:57)
                            key.pixel = __float_as_uint(o.w);
                            key.sample = w.begin + (__float_as_uint(d.w) >> 8);
                            key.bounce = __float_as_uint(d.w) & 0xFFu;
                        }
                        inv.ix = 1.0f / r.d.x; inv.iy = 1.0f / r.d.y; inv.iz = 1.0f / r.d.z;
                        best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
                        cur = walk_entry(WIDE4 ? a.sc.root_entry4 : a.sc.root_entry);
                        sp = 0;
                        if (VOLUMES) { vstate = 0u; floor = 0; tmin_cur = tmin0; }
                        has = true;
                    }
                }
                drained = base + (uint32_t)__popc(idle) >= n;
            }
            const uint32_t busy = __ballot_sync(FULL, has);
            if (busy == 0u) break;
            const uint32_t keep = drained ? 1u : min(a.refill_threshold, (uint32_t)__popc(busy));
            // ---- majority-rule step scheduling, as in the megakernel  [bucket:schedule]
            do {
                for (;;) {
                    const bool at_node = RTR_REF_KIND(cur) == REF_NODE;
                    const uint32_t mn = __ballot_sync(FULL, at_node);
                    const uint32_t ml = __ballot_sync(FULL, !at_node && cur != RTR_REF_DONE);
                    if (mn == 0u || __popc(mn) < __popc(ml)) break;
                    if (at_node && WIDE4) {
                        node4_step<COUNT>(a.sc, cur, sp, best, r, inv, VOLUMES ? tmin_cur : tmin0, RTR_INF, wstack, &cnt, VOLUMES ? floor : 0);
                    } else if (at_node) {
                        node_step<EXHAUSTIVE, COUNT>(a.sc, cur, sp, best, r, inv, VOLUMES ? tmin_cur : tmin0, RTR_INF, wstack, &cnt, VOLUMES ? floor : 0);
#if RTR_WF_NODE_BURST
                        if (RTR_REF_KIND(cur) == REF_NODE) node_step<EXHAUSTIVE, COUNT>(a.sc, cur, sp, best, r, inv, VOLUMES ? tmin_cur : tmin0, RTR_INF, wstack, &cnt, VOLUMES ? floor : 0);
#endif
                    }
                }
                if (RTR_REF_KIND(cur) != REF_NODE && cur != RTR_REF_DONE) {
                    if (VOLUMES && RTR_REF_KIND(cur) == REF_VOLUME && vstate == 0u) {  // [bucket:volume_fsm]
                        const uint32_t vidx = RTR_REF_INDEX(cur);
                        const VolumeRec& vol = a.sc.volumes[vidx];
                        const uint32_t entry = walk_entry(WIDE4 ? vol.entry4 : a.sc.reference_tree ? vol.entry_reference : vol.entry);
                        if (entry == RTR_REF_DONE) {  // empty boundary: Volume::hit is None
                            cur = pop_next<EXHAUSTIVE>(wstack, sp, best.t, 0);
                        } else {  // suspend the outer walk, start boundary.hit(ray, -inf, +inf)  (volume.rs:36)
                            wstack.put(sp, best.ref, best.t); ++sp;
                            wstack.put(sp, __float_as_uint(best.u), best.v); ++sp;
                            floor = sp;
                            vstate = (1u << 30) | vidx;
                            tmin_cur = -RTR_INF;
                            best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
                            cur = entry;
                        }
                    } else {
                        // nested walks never meet a Volume leaf (flat.h: a boundary is a sphere or a mesh)
                        leaf_step<EXHAUSTIVE, COUNT, false, false>(a.sc, cur, sp, best, r, inv, VOLUMES ? tmin_cur : tmin0, key, wstack, &cnt, VOLUMES ? floor : 0);
                    }
                }
                if (VOLUMES && vstate != 0u && cur == RTR_REF_DONE) {  // a boundary query has finished  [bucket:volume_fsm]
                    const uint32_t vidx = vstate & 0x3FFFFFFFu;
                    const VolumeRec& vol = a.sc.volumes[vidx];
                    const bool second = (vstate >> 30) == 2u;
                    if (!second && best.ref != REF_EMPTY_VALUE) {  // boundary.hit(ray, t1 + tmin, +inf)  (volume.rs:37)
                        const float t1 = best.t;
                        wstack.put(sp, 0u, t1); ++sp;
                        floor = sp;
                        vstate = (2u << 30) | vidx;
                        tmin_cur = t1 + tmin0;
                        best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
                        cur = walk_entry(WIDE4 ? vol.entry4 : a.sc.reference_tree ? vol.entry_reference : vol.entry);
                    } else {
                        float t_hit = __int_as_float(0x7fc00000);  // None
                        int base = floor;                          // stack level of the saved outer hit + 2
                        if (second) {
                            base = floor - 1;
                            if (best.ref != REF_EMPTY_VALUE)
                                t_hit = volume_sample(vol, r, wstack.get(floor - 1).enter, best.t, tmin0, RTR_INF, key);
                        }
                        const StackEntry uv = wstack.get(base - 1), tr = wstack.get(base - 2);
                        sp = base - 2;
                        best.t = tr.enter; best.ref = tr.ref; best.u = __uint_as_float(uv.ref); best.v = uv.enter;
                        if (t_hit == t_hit) offer(a.sc, best, t_hit, RTR_MAKE_REF(REF_VOLUME, vidx), 0.0f, 0.0f);
                        vstate = 0u; floor = 0; tmin_cur = tmin0;
                        cur = pop_next<EXHAUSTIVE>(wstack, sp, best.t, 0);
                    }
                }
            } while ((uint32_t)__popc(__ballot_sync(FULL, cur != RTR_REF_DONE)) >= keep);
            // [bucket:wf_store]
            if (has && cur == RTR_REF_DONE) {
                a.hit[idx] = make_float4(best.t, __uint_as_float(best.ref), best.u, best.v);
                has = false;
            }
        }
    }
    if (COUNT) {
        unsigned long long n_nodes = cnt.nodes, n_tris = cnt.tris, n_spheres = cnt.spheres;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            n_nodes += __shfl_xor_sync(FULL, n_nodes, o);
            n_tris += __shfl_xor_sync(FULL, n_tris, o);
            n_spheres += __shfl_xor_sync(FULL, n_spheres, o);
        }
        if (lane == 0) {
            atomicAdd(&a.stats[STAT_NODES], n_nodes);
            atomicAdd(&a.stats[STAT_TRIS], n_tris);
            atomicAdd(&a.stats[STAT_SPHERES], n_spheres);
        }
    }
}

// ------------------------------------------------------------------ shade + compact
template <bool COUNT>
__global__ void __launch_bounds__(RTR_WF_SHADE_THREADS, RTR_WF_SHADE_MIN_BLOCKS) wf_shade_kernel(const WfArgs a) {
    const uint32_t FULL = 0xffffffffu;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t n = a.counters[WFC_QCOUNT + a.bounce];
    const WaveInfo w = wave_info(a);
    uint32_t* n_out = &a.counters[WFC_QCOUNT + a.bounce + 1];
    Counters cnt; cnt.nodes = cnt.tris = cnt.spheres = cnt.env = 0;
    // compaction bookkeeping of the per-CTA mode (double buffered: a warp may start trip k+1 while another still reads trip k)
    constexpr uint32_t NW = RTR_WF_SHADE_THREADS / 32;
    __shared__ uint32_t s_cnt[2][NW];  // survivors per warp, then (after the scan) the warp's first output slot
    const bool per_cta = (a.atomics_mode & 2u) != 0u;
    const uint32_t warp = threadIdx.x >> 5;
    uint32_t trip = 0;
    // the loop is uniform per CTA (block base), so __syncthreads inside it is safe
    for (uint32_t bb = blockIdx.x * blockDim.x; bb < n; bb += gridDim.x * blockDim.x, ++trip) {
        const uint32_t i = bb + threadIdx.x;
        bool cont = false;
        Ray ray; ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
        v3 T = mk(0.0f, 0.0f, 0.0f);
        uint32_t pixel = 0, rel = 0, depth = 0;
#if RTR_WF_SHADE_PREFETCH
        {   // the next trip's four stream records towards L2 while this trip's dependent fetches (hit -> primitive -> material) run
            const uint64_t nx = (uint64_t)i + (uint64_t)gridDim.x * blockDim.x;
            if (nx < n && (threadIdx.x & 7u) == 0u) {  // one hint per 128-byte line
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.in.o + nx));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.in.d + nx));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.in.thr + nx));
                asm volatile("prefetch.global.L2 [%0];" ::"l"(a.hit + nx));
            }
        }
#endif
        if (i < n) {
            const float4 o = a.in.o[i], d = a.in.d[i], t = a.in.thr[i], h = a.hit[i];
            ray.o = xyz(o); ray.d = xyz(d);
            T = xyz(t);
            pixel = __float_as_uint(o.w);
            rel = __float_as_uint(d.w) >> 8;
            depth = __float_as_uint(d.w) & 0xFFu;
            RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = w.begin + rel; key.bounce = depth;
            LiteHit lh; lh.t = h.x; lh.ref = __float_as_uint(h.y); lh.u = h.z; lh.v = h.w;
            v3 L = mk(0.0f, 0.0f, 0.0f);
            if (lh.ref == REF_EMPTY_VALUE) {
                L = T * background_sample(a.sc, ray, COUNT ? &cnt : nullptr);  // render.rs:57
            } else {
                const Hit hit = finish_hit(a.sc, ray, lh);
                if (a.mode == RTR_MODE_FULL || a.mode == RTR_MODE_CLAY) {
                    const Scatter sc = scatter_event(a.sc, a.mode == RTR_MODE_CLAY, hit.material, ray, hit, key);  // :40-41
                    if (sc.some) {  // attenuation * trace_ray(scattered, depth + 1)  (:51-52)
                        T = T * sc.attenuation;
                        ray = sc.ray;
                        ++depth;
                        if (depth >= a.max_ray_depth) L = T * mk(0.0f, 0.0f, 0.0f);  // :34-36
                        else cont = true;
                    } else {
                        L = T * sc.attenuation;  // :54
                    }
                } else if (a.mode == RTR_MODE_NORMAL) {  // :42-45
                    L = T * (0.5f * (normalize(hit.normal) + mk(1.0f, 1.0f, 1.0f)));
                } else {
                    L = T * mk(0.0f, 0.0f, 0.0f);  // Random (:46-48)
                }
            }
            if (!cont) finish_sample(a, pixel, rel, L);
        }
        // ---- compaction: survivors go densely into the other buffer  [bucket:wf_compact]
        const uint32_t m = __ballot_sync(FULL, cont);
        if (per_cta) {
            // one atomicAdd on the queue counter per CTA and trip instead of one per warp (the counter is a single address:
            // ~1 atomic per ns, and a 16 M-path bounce did 512 K of them)
            const uint32_t buf = trip & 1u;
            if (lane == 0) s_cnt[buf][warp] = (uint32_t)__popc(m);
            __syncthreads();
            if (warp == 0) {  // exclusive scan of the warps' counts, one atomicAdd for the CTA's total
                const uint32_t c = lane < NW ? s_cnt[buf][lane] : 0u;
                uint32_t inc = c;
#pragma unroll
                for (uint32_t o = 1; o < NW; o <<= 1) {
                    const uint32_t v = __shfl_up_sync(FULL, inc, o);
                    if (lane >= o) inc += v;
                }
                const uint32_t total = __shfl_sync(FULL, inc, NW - 1);
                uint32_t cta_base = 0;
                if (lane == 0 && total != 0u) cta_base = atomicAdd(n_out, total);
                cta_base = __shfl_sync(FULL, cta_base, 0);
                if (lane < NW) s_cnt[buf][lane] = cta_base + inc - c;
            }
            __syncthreads();
            if (cont) {
                const uint32_t slot = s_cnt[buf][warp] + (uint32_t)__popc(m & lanemask_lt());
                a.out.o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(pixel));
                a.out.d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float((rel << 8) | depth));
                a.out.thr[slot] = make_float4(T.x, T.y, T.z, 0.0f);
            }
        } else if (m != 0u) {
            const int leader = __ffs(m) - 1;
            uint32_t slot = 0;
            if ((int)lane == leader) slot = atomicAdd(n_out, (uint32_t)__popc(m));
            slot = __shfl_sync(FULL, slot, leader) + __popc(m & lanemask_lt());
            if (cont) {
                a.out.o[slot] = make_float4(ray.o.x, ray.o.y, ray.o.z, __uint_as_float(pixel));
                a.out.d[slot] = make_float4(ray.d.x, ray.d.y, ray.d.z, __uint_as_float((rel << 8) | depth));
                a.out.thr[slot] = make_float4(T.x, T.y, T.z, 0.0f);
            }
        }
    }
    if (COUNT) {
        unsigned long long n_env = cnt.env;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n_env += __shfl_xor_sync(FULL, n_env, o);
        if (lane == 0 && n_env) atomicAdd(&a.stats[STAT_ENV], n_env);
    }
}

// ------------------------------------------------------------------ end of wave
// accum[p] += lbuf[0][p] + lbuf[1][p] + ... one sample at a time, i.e. the reference's sequential sum (render.rs:99)
__global__ void wf_sum_samples_kernel(const WfArgs a) {
    const WaveInfo w = wave_info(a);
    for (uint32_t p = blockIdx.x * blockDim.x + threadIdx.x; p < a.n_pixels; p += gridDim.x * blockDim.x) {
        float4 c = a.accum[p];
        for (uint32_t rel = 0; rel < w.samples; ++rel) {
            const float4 l = a.lbuf[(size_t)rel * a.n_pixels + p];
            c.x = c.x + l.x; c.y = c.y + l.y; c.z = c.z + l.z; c.w = c.w + l.w;
        }
        a.accum[p] = c;
    }
}

// one block: count the wave's rays (every queued path-bounce is one world.hit query), reset the queues, next wave
__global__ void wf_next_wave_kernel(const WfArgs a) {
    unsigned long long rays = 0;
    for (uint32_t b = threadIdx.x; b < RTR_WF_MAX_DEPTH; b += blockDim.x)
        if (b < a.max_ray_depth) rays += a.counters[WFC_QCOUNT + b];
    if (rays) atomicAdd(&a.stats[STAT_RAYS], rays);
    __syncthreads();
    for (uint32_t i = WFC_QCOUNT + threadIdx.x; i < WFC_WORDS; i += blockDim.x) a.counters[i] = 0u;
    if (threadIdx.x == 0) a.counters[WFC_WAVE] += a.wave_step;
}

// ------------------------------------------------------------------ rtr_trace_batch on the wavefront walk
// world.hit(ray, tmin, +INF) for caller-supplied rays (include/rtr.h): the rays are unpacked into the queue of bounce 0,
// wf_extend walks them exactly as it walks a render's rays, and the winners' HitRecords are rebuilt for the caller.
// Ray i draws its volume random numbers from counter (pixel = i, sample 0, bounce 0).
__global__ void trace_unpack_kernel(const WfArgs a, const rtr_ray* rays, uint32_t n, uint32_t first_index) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) a.counters[WFC_QCOUNT] = n;
    if (i >= n) return;
    const rtr_ray r = rays[i];
    a.in.o[i] = make_float4(r.origin.x, r.origin.y, r.origin.z, __uint_as_float(first_index + i));
    a.in.d[i] = make_float4(r.direction.x, r.direction.y, r.direction.z, __uint_as_float(0u));
}

__global__ void trace_finish_kernel(const WfArgs a, uint32_t n, rtr_hit* out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 o = a.in.o[i], d = a.in.d[i], hq = a.hit[i];
    Ray r; r.o = xyz(o); r.d = xyz(d);
    LiteHit lh; lh.t = hq.x; lh.ref = __float_as_uint(hq.y); lh.u = hq.z; lh.v = hq.w;
    rtr_hit h;
    if (lh.ref == REF_EMPTY_VALUE) {
        h.object = RTR_NO_HIT; h.primitive = 0; h.t = 0.0f;
        h.point.x = h.point.y = h.point.z = 0.0f; h.normal.x = h.normal.y = h.normal.z = 0.0f;
        h.front_face = 0; h.material = 0;
    } else {
        const Hit f = finish_hit(a.sc, r, lh);
        h.object = f.object; h.primitive = f.primitive; h.t = f.t;
        h.point.x = f.point.x; h.point.y = f.point.y; h.point.z = f.point.z;
        h.normal.x = f.normal.x; h.normal.y = f.normal.y; h.normal.z = f.normal.z;
        h.front_face = f.front_face ? 1u : 0u; h.material = f.material;
    }
    out[i] = h;
}

cudaError_t launch_trace_unpack(const WfArgs& a, const rtr_ray* rays, uint32_t n, uint32_t first_index, cudaStream_t st) {
    trace_unpack_kernel<<<(n + 255u) / 256u, 256, 0, st>>>(a, rays, n, first_index);
    return cudaGetLastError();
}

cudaError_t launch_trace_finish(const WfArgs& a, uint32_t n, rtr_hit* out, cudaStream_t st) {
    trace_finish_kernel<<<(n + 127u) / 128u, 128, 0, st>>>(a, n, out);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ launchers
template <typename F>
static auto pick_wf_extend(bool exhaustive, bool count, bool volumes, bool tmin_arg, bool wide4, F f) {
    if (wide4 && !exhaustive && !tmin_arg) {
        if (count) { if (volumes) return f(wf_extend_kernel<false, true, true, false, true>); return f(wf_extend_kernel<false, true, false, false, true>); }
        if (volumes) return f(wf_extend_kernel<false, false, true, false, true>);
        return f(wf_extend_kernel<false, false, false, false, true>);
    }
#define RTR_PICK(E, C, V) do { if (tmin_arg) return f(wf_extend_kernel<E, C, V, true>); return f(wf_extend_kernel<E, C, V, false>); } while (0)
    if (exhaustive) {
        if (count) { if (volumes) RTR_PICK(true, true, true); else RTR_PICK(true, true, false); }
        else { if (volumes) RTR_PICK(true, false, true); else RTR_PICK(true, false, false); }
    } else {
        if (count) { if (volumes) RTR_PICK(false, true, true); else RTR_PICK(false, true, false); }
        else { if (volumes) RTR_PICK(false, false, true); else RTR_PICK(false, false, false); }
    }
#undef RTR_PICK
}

bool wf_extend_wide4(const WfArgs& a, bool exhaustive, bool count) {
    (void)count;  // the counting instantiations walk it too (nodes_visited = boxes tested), unless they walk the reference topology
    return a.sc.nodes4 != nullptr && a.sc.reference_tree == 0u && a.enable_bvh_tree != 0u && !exhaustive && a.tmin == 0.00001f;
}

int wf_extend_max_blocks_per_sm(bool exhaustive, bool count, bool volumes, bool tmin_arg, bool wide4) {
    int n = 0;
    // RTR_WF_CARVEOUT (percent of the SM's L1/shared array given to shared memory; unset = the driver's choice): wf_extend
    // keeps 8 KB of short stacks per CTA, the rest of the array serves node and triangle fetches as L1
    static const int carveout = [] { const char* e = std::getenv("RTR_WF_CARVEOUT"); return e ? std::atoi(e) : -1; }();
    cudaError_t e = pick_wf_extend(exhaustive, count, volumes, tmin_arg, wide4, [&](void (*k)(const WfArgs)) {
        if (carveout >= 0) (void)cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, carveout);
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k, RTR_WF_EXTEND_THREADS, 0);
    });
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_wf_begin(const WfArgs& a, cudaStream_t st) {
    const uint32_t n = a.n_pixels > (uint32_t)WFC_WORDS ? a.n_pixels : (uint32_t)WFC_WORDS;
    wf_begin_kernel<<<(n + 255) / 256, 256, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_wf_begin_lane(const WfArgs& a, cudaStream_t st) {
    WfArgs b = a;
    b.n_pixels = 0;  // counters only: the image is cleared by the first sequence's wf_begin
    wf_begin_kernel<<<(WFC_WORDS + 255) / 256, 256, 0, st>>>(b);
    return cudaGetLastError();
}

cudaError_t launch_wf_generate(const WfArgs& a, int grid, cudaStream_t st) {
    wf_generate_kernel<<<grid, RTR_WF_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_wf_extend(const WfArgs& a, int grid, bool exhaustive, bool count, cudaStream_t st) {
    // the generic instantiation also carries the BVH-off linear scan
    const bool volumes = a.sc.n_volumes != 0u || a.enable_bvh_tree == 0u;
    const bool tmin_arg = a.tmin != 0.00001f;
    pick_wf_extend(exhaustive, count, volumes, tmin_arg, wf_extend_wide4(a, exhaustive, count), [&](void (*k)(const WfArgs)) {
        k<<<grid, RTR_WF_EXTEND_THREADS, 0, st>>>(a);
        return cudaSuccess;
    });
    return cudaGetLastError();
}

int wf_shade_threads() { return RTR_WF_SHADE_THREADS; }

cudaError_t launch_wf_shade(const WfArgs& a, int grid, bool count, cudaStream_t st) {
    if (count) wf_shade_kernel<true><<<grid, RTR_WF_SHADE_THREADS, 0, st>>>(a);
    else wf_shade_kernel<false><<<grid, RTR_WF_SHADE_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_wf_bin(const WfArgs& a, int grid, cudaStream_t st) {
    wf_bin_kernel<<<grid, 256, 0, st>>>(a);
    return cudaGetLastError();
}

cudaError_t launch_wf_end_wave(const WfArgs& a, int grid, cudaStream_t st) {
    if (a.lbuf) wf_sum_samples_kernel<<<grid, 256, 0, st>>>(a);
    wf_next_wave_kernel<<<1, 256, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace rtr
