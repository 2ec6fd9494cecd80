// kernels.cu — sm_100a kernels of the render hot path.
//
//   render_megakernel   render() + trace_ray() (lib/core/render.rs:32-127): persistent warps pull
//                       8x4-pixel tiles (x sample chunks) from an atomic counter; every lane runs
//                       a flat path state machine (regenerates its next sample as soon as its
//                       path ends, no per-sample warp sync); stack-based ordered BVH walk.
//   sum_chunks          deterministic in-order sum of sample-chunk partials (small images only)
//   resolve             sqrt(sum/spp) + Color::into_raw (render.rs:103-110, color.rs:34-41)
//   trace/scatter/background/camera batch kernels: the parity entry points of include/rtr.h
#include "device_core.cuh"
#include "kernels.h"

namespace rtr {

// ------------------------------------------------------------------ megakernel
// Lane = pixel (samples of a pixel are traced and summed in order, render.rs:86-100, so the image is
// bit-reproducible); a warp = an 8x4 pixel tile.  Each lane runs a small path state machine
//     FRESH (camera ray) -> LAUNCH (start world.hit) -> [walk rounds] -> SHADE (scatter / terminate) -> ...
// The BVH walk is resumable (walk_round), which lets the warp REFILL lanes: when fewer than
// `refill_threshold` lanes are still walking, the lanes whose walk has finished shade their hit and
// start their next ray while the others keep their traversal state, instead of idling until the
// longest walk of the warp ends (ncu before this: 4.3 of 32 lanes active in the node loop on the mesh scene).
enum { ST_FRESH = 0, ST_LAUNCH = 1, ST_SHADE = 2, ST_DEAD = 3 };

// VOLUMES = false (BVH walks of scenes without a Volume): Volume::hit, a non-inlined call, is compiled out of the walk.
template <bool EXHAUSTIVE, bool COUNT, bool BVH, bool VOLUMES>
__global__ void __launch_bounds__(RTR_MEGA_THREADS, BVH ? RTR_MEGA_MIN_BLOCKS : RTR_MEGA_MIN_BLOCKS_SCAN) render_megakernel(const RenderArgs a) {
    StackEntry stack[RTR_STACK_CAP];  // spill area of the short stack + scratch of nested (volume boundary) walks
    __shared__ uint2 short_stack[BVH ? RTR_SHORT_STACK : 1][RTR_MEGA_THREADS];
    ShortStack<RTR_SHORT_STACK, RTR_MEGA_THREADS> wstack;
    wstack.sh = &short_stack[0][threadIdx.x];
    wstack.loc = stack;
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t FULL = 0xffffffffu;
    unsigned long long n_rays = 0, n_nodes = 0, n_tris = 0, n_spheres = 0, n_env = 0;
    const bool bvh = BVH;  // settings.enable_bvh_tree (render.rs:68-72); false = linear scan, no resumable walk

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(a.work_counter, 1u);
        item = __shfl_sync(FULL, item, 0);
        if (item >= a.n_tiles * a.n_chunks) break;
        const uint32_t tile = item / a.n_chunks, chunk = item - tile * a.n_chunks;
        const uint32_t tx = tile % a.tiles_x, ty = tile / a.tiles_x;
        const uint32_t x = tx * 8u + (lane & 7u), y = ty * 4u + (lane >> 3);
        const bool valid = x < a.width && y < a.height;
        const uint32_t pixel = y * a.width + x;  // index = y*W + x (render.rs:76-77)
        uint32_t s = a.sample_begin + chunk * a.chunk_samples;
        const uint32_t s_first = s;
        const uint32_t s_end = min(s + a.chunk_samples, a.sample_begin + a.sample_count);
        Counters cnt; cnt.nodes = cnt.tris = cnt.spheres = cnt.env = 0;
        uint32_t rays = 0;
        v3 sum = mk(0.0f, 0.0f, 0.0f);
        RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = s; key.bounce = 0;
        Ray ray; ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
        RayInv inv; inv.ix = inv.iy = inv.iz = 0.0f;
        v3 T = mk(1.0f, 1.0f, 1.0f);
        uint32_t depth = 0;
        uint32_t state = valid ? ST_FRESH : ST_DEAD;
        bool walking = false;
        uint32_t cur = RTR_REF_DONE;
        int sp = 0;
        LiteHit best; best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;

        for (;;) {
            // ---- lanes without a walk in flight: shade what they found, start their next ray  [bucket:path_fsm]
            while (state != ST_DEAD && !walking) {
                v3 L = mk(0.0f, 0.0f, 0.0f);
                bool finished = false;  // this sample's radiance L is final
                if (state == ST_SHADE) {  // the rest of one trace_ray frame (render.rs:39-57)
                    if (best.ref == REF_EMPTY_VALUE) {
                        L = T * background_sample(a.sc, ray, COUNT ? &cnt : nullptr);  // :57
                        finished = true;
                    } else {
                        const Hit hit = finish_hit(a.sc, ray, best);
                        if (a.mode == RTR_MODE_FULL || a.mode == RTR_MODE_CLAY) {
                            const Scatter sc = scatter_event(a.sc, a.mode == RTR_MODE_CLAY, hit.material, ray, hit, key);  // :40-41
                            if (sc.some) {  // attenuation * trace_ray(scattered, depth + 1)  (:51-52)
                                T = T * sc.attenuation;
                                ray = sc.ray;
                                ++depth;
                                state = ST_LAUNCH;
                            } else {
                                L = T * sc.attenuation;  // :54
                                finished = true;
                            }
                        } else if (a.mode == RTR_MODE_NORMAL) {  // :42-45
                            L = T * (0.5f * (normalize(hit.normal) + mk(1.0f, 1.0f, 1.0f)));
                            finished = true;
                        } else {
                            L = T * mk(0.0f, 0.0f, 0.0f);  // Random (:46-48)
                            finished = true;
                        }
                    }
                } else if (state == ST_LAUNCH) {  // only reached with depth >= max_ray_depth (:34-36)
                    L = T * mk(0.0f, 0.0f, 0.0f);
                    finished = true;
                }
                if (finished) {
                    L.x = f32_clamp(L.x, 0.0f, a.clamp_indirect);  // render.rs:95-97
                    L.y = f32_clamp(L.y, 0.0f, a.clamp_indirect);
                    L.z = f32_clamp(L.z, 0.0f, a.clamp_indirect);
                    sum = sum + L;  // :99
                    ++s;
                    state = ST_FRESH;
                }
                if (state == ST_FRESH) {
                    if (s >= s_end) { state = ST_DEAD; break; }
                    key.sample = s; key.bounce = 0;
                    ray = camera_sample_ray(a.cam, a.width, a.height, x, y, key);  // render.rs:88-92
                    T = mk(1.0f, 1.0f, 1.0f);
                    depth = 0;
                    state = ST_LAUNCH;
                }
                // state == ST_LAUNCH: world.hit(ray, 0.00001, INF) (:38) unless the depth limit ends the path
                if (depth >= a.max_ray_depth) continue;
                key.bounce = depth;
                ++rays;
                state = ST_SHADE;
                if (bvh) {
                    inv.ix = 1.0f / ray.d.x; inv.iy = 1.0f / ray.d.y; inv.iz = 1.0f / ray.d.z;
                    best.t = RTR_INF; best.ref = REF_EMPTY_VALUE; best.u = best.v = 0.0f;
                    cur = walk_entry(a.sc.root_entry);
                    sp = 0;
                    walking = cur != RTR_REF_DONE;
                } else {
                    best = world_hit<EXHAUSTIVE, COUNT>(a.sc, false, ray, 0.00001f, key, stack, &cnt);
                }
            }
            const uint32_t alive = __ballot_sync(FULL, state != ST_DEAD);
            if (!BVH || alive == 0u) break;
            const uint32_t keep = min(a.refill_threshold, (uint32_t)__popc(alive));
            // ---- walk until enough lanes have finished to make a refill worth its divergence.  Each turn  [bucket:schedule]
            // runs whichever step has more lanes waiting (majority rule): node steps while at least as
            // many lanes sit at interior nodes as at primitives, then one primitive step for the rest.
            do {
                for (;;) {
                    const bool at_node = RTR_REF_KIND(cur) == REF_NODE;
                    const uint32_t mn = __ballot_sync(FULL, at_node);
                    const uint32_t ml = __ballot_sync(FULL, !at_node && cur != RTR_REF_DONE);
                    if (mn == 0u || __popc(mn) < __popc(ml)) break;
                    if (at_node) node_step<EXHAUSTIVE, COUNT>(a.sc, cur, sp, best, ray, inv, 0.00001f, RTR_INF, wstack, &cnt);
                }
                if (RTR_REF_KIND(cur) != REF_NODE && cur != RTR_REF_DONE)
                    leaf_step<EXHAUSTIVE, COUNT, false, VOLUMES>(a.sc, cur, sp, best, ray, inv, 0.00001f, key, wstack, &cnt);
            } while ((uint32_t)__popc(__ballot_sync(FULL, cur != RTR_REF_DONE)) >= keep);
            walking = cur != RTR_REF_DONE;
        }
        // [bucket:other]
        if (valid) {
            const float n = s_first < s_end ? (float)(s_end - s_first) : 0.0f;
            float4* dst = a.n_chunks == 1 ? a.accum : a.partials + (size_t)chunk * a.width * a.height;
            float4 o = make_float4(sum.x, sum.y, sum.z, n);
            if (a.accumulate && a.n_chunks == 1) { const float4 p = dst[pixel]; o.x = p.x + o.x; o.y = p.y + o.y; o.z = p.z + o.z; o.w = p.w + o.w; }
            dst[pixel] = o;
        }
        n_rays += rays;
        if (COUNT) { n_nodes += cnt.nodes; n_tris += cnt.tris; n_spheres += cnt.spheres; n_env += cnt.env; }
    }
    // per-warp reduction of the counters, one atomic per warp and counter
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        n_rays += __shfl_xor_sync(0xffffffffu, n_rays, o);
        if (COUNT) {
            n_nodes += __shfl_xor_sync(0xffffffffu, n_nodes, o);
            n_tris += __shfl_xor_sync(0xffffffffu, n_tris, o);
            n_spheres += __shfl_xor_sync(0xffffffffu, n_spheres, o);
            n_env += __shfl_xor_sync(0xffffffffu, n_env, o);
        }
    }
    if (lane == 0) {
        atomicAdd(&a.stats[STAT_RAYS], n_rays);
        if (COUNT) {
            atomicAdd(&a.stats[STAT_NODES], n_nodes);
            atomicAdd(&a.stats[STAT_TRIS], n_tris);
            atomicAdd(&a.stats[STAT_SPHERES], n_spheres);
            atomicAdd(&a.stats[STAT_ENV], n_env);
        }
    }
}

__global__ void sum_chunks_kernel(const float4* partials, float4* accum, uint32_t n_pixels, uint32_t n_chunks, bool accumulate) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    float4 s = partials[i];
    if (accumulate) { const float4 p = accum[i]; s.x = p.x + s.x; s.y = p.y + s.y; s.z = p.z + s.z; s.w = p.w + s.w; }
    for (uint32_t c = 1; c < n_chunks; ++c) {  // chunk order == sample order
        const float4 p = partials[(size_t)c * n_pixels + i];
        s.x = s.x + p.x; s.y = s.y + p.y; s.z = s.z + p.z; s.w = s.w + p.w;
    }
    accum[i] = s;
}

// render.rs:103-110 + color.rs:34-41
__global__ void resolve_kernel(const float4* accum, uchar4* rgba, uint32_t n_pixels, float spp) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const float4 c = accum[i];
    const float r = sqrtf(c.x / spp), g = sqrtf(c.y / spp), b = sqrtf(c.z / spp);
    rgba[i] = make_uchar4(f32_as_u8(r * 255.0f), f32_as_u8(g * 255.0f), f32_as_u8(b * 255.0f), 255);
}

// The cross-device step of a multi-GPU render, fused with the resolve: this device's share of the image is
// reduced over the N accumulation buffers by 128-bit PEER loads (NVLink / NVSwitch; the local buffer is one of them),
// summed in rank order so the result does not depend on timing, and leaves as one float4 + one uchar4 peer store into
// the root's buffers.  Per pixel N x 16 B cross the links inbound and 20 B outbound; every device moves 1/N of the image,
// so all NVLink ports are busy at once (a root-only gather would pull N x the image through one port).
__global__ void __launch_bounds__(256) reduce_resolve_peers_kernel(const PeerAccum peers, uint32_t n_ranks, float4* accum_root,
                                                                   uchar4* rgba_root, uint32_t lo, uint32_t hi, float spp) {
    for (uint32_t p = lo + blockIdx.x * blockDim.x + threadIdx.x; p < hi; p += gridDim.x * blockDim.x) {
        float4 v[RTR_MAX_GROUP];
#pragma unroll
        for (uint32_t r = 0; r < RTR_MAX_GROUP; ++r)  // all loads in flight before the first add
            if (r < n_ranks) v[r] = __ldcv(peers.accum[r] + p);
        float4 s = v[0];
#pragma unroll
        for (uint32_t r = 1; r < RTR_MAX_GROUP; ++r)
            if (r < n_ranks) { s.x = s.x + v[r].x; s.y = s.y + v[r].y; s.z = s.z + v[r].z; s.w = s.w + v[r].w; }
        accum_root[p] = s;
        const float r8 = sqrtf(s.x / spp), g8 = sqrtf(s.y / spp), b8 = sqrtf(s.z / spp);  // render.rs:103-110
        rgba_root[p] = make_uchar4(f32_as_u8(r8 * 255.0f), f32_as_u8(g8 * 255.0f), f32_as_u8(b8 * 255.0f), 255);
    }
}
cudaError_t launch_reduce_resolve_peers(const PeerAccum& peers, uint32_t n_ranks, float4* accum_root, uchar4* rgba_root,
                                        uint32_t lo, uint32_t hi, float spp, cudaStream_t st) {
    if (hi <= lo) return cudaSuccess;
    const uint32_t n = hi - lo;
    uint32_t blocks = (n + 255u) / 256u;
    if (blocks > 148u * 8u) blocks = 148u * 8u;
    reduce_resolve_peers_kernel<<<blocks, 256, 0, st>>>(peers, n_ranks, accum_root, rgba_root, lo, hi, spp);
    return cudaGetLastError();
}

// SkyMap.image (backgrounds.rs:59-81: RGB f32, alpha dropped) -> float4 texels, one 128-bit fetch per lookup
__global__ void expand_env_kernel(const float* rgb, float4* out, size_t n) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        out[i] = make_float4(rgb[3 * i], rgb[3 * i + 1], rgb[3 * i + 2], 0.0f);
}
cudaError_t launch_expand_env(const float* rgb, float4* out, size_t n, cudaStream_t st) {
    expand_env_kernel<<<1184, 256, 0, st>>>(rgb, out, n);
    return cudaGetLastError();
}

// ------------------------------------------------------------------ batch (parity) kernels
// (rtr_trace_batch runs on the wavefront's wf_extend: trace_unpack / trace_finish in wavefront.cu)
__global__ void scatter_batch_kernel(const SceneView sc, const rtr_scatter_in* in, size_t n, uint32_t mode,
                                     uint32_t k0, uint32_t k1, rtr_scatter_out* out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const rtr_scatter_in q = in[i];
    Ray r; r.o = mk(q.ray.origin.x, q.ray.origin.y, q.ray.origin.z);
    r.d = mk(q.ray.direction.x, q.ray.direction.y, q.ray.direction.z);
    Hit h;
    h.point = mk(q.point.x, q.point.y, q.point.z);
    h.normal = mk(q.normal.x, q.normal.y, q.normal.z);
    h.t = q.t; h.front_face = q.front_face != 0; h.material = q.material; h.object = 0; h.primitive = 0;
    RngKey key; key.k0 = k0; key.k1 = k1; key.pixel = q.pixel; key.sample = q.sample; key.bounce = q.bounce;
    const Scatter s = scatter_event(sc, mode == RTR_MODE_CLAY, h.material, r, h, key);
    rtr_scatter_out o;
    o.attenuation.x = s.attenuation.x; o.attenuation.y = s.attenuation.y; o.attenuation.z = s.attenuation.z;
    o.scattered = s.some ? 1u : 0u;
    o.ray.origin.x = s.ray.o.x; o.ray.origin.y = s.ray.o.y; o.ray.origin.z = s.ray.o.z;
    o.ray.direction.x = s.some ? s.ray.d.x : 0.0f;
    o.ray.direction.y = s.some ? s.ray.d.y : 0.0f;
    o.ray.direction.z = s.some ? s.ray.d.z : 0.0f;
    out[i] = o;
}

__global__ void background_batch_kernel(const SceneView sc, const rtr_ray* rays, size_t n, float* rgb) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r; r.o = mk(rays[i].origin.x, rays[i].origin.y, rays[i].origin.z);
    r.d = mk(rays[i].direction.x, rays[i].direction.y, rays[i].direction.z);
    const v3 c = background_sample(sc, r, nullptr);
    rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
}

__global__ void camera_rays_kernel(const CameraBasis cam, uint32_t W, uint32_t H, uint32_t k0, uint32_t k1,
                                   const uint32_t* pixels, const uint32_t* samples, size_t n, rtr_ray* out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    RngKey key; key.k0 = k0; key.k1 = k1; key.pixel = pixels[i]; key.sample = samples[i]; key.bounce = 0;
    const Ray r = camera_sample_ray(cam, W, H, pixels[i] % W, pixels[i] / W, key);
    out[i].origin.x = r.o.x; out[i].origin.y = r.o.y; out[i].origin.z = r.o.z;
    out[i].direction.x = r.d.x; out[i].direction.y = r.d.y; out[i].direction.z = r.d.z;
}

// ------------------------------------------------------------------ launchers (called from capi.cpp)
template <typename F>
static auto pick_megakernel(bool exhaustive, bool count, bool bvh, bool volumes, F f) {
#define RTR_PICK(E, C) \
    do { if (!bvh) return f(render_megakernel<E, C, false, true>); \
         if (volumes) return f(render_megakernel<E, C, true, true>); \
         return f(render_megakernel<E, C, true, false>); } while (0)
    if (exhaustive) { if (count) RTR_PICK(true, true); else RTR_PICK(true, false); }
    else { if (count) RTR_PICK(false, true); else RTR_PICK(false, false); }
#undef RTR_PICK
}

int megakernel_max_blocks_per_sm(bool exhaustive, bool count, bool bvh, bool volumes) {
    int n = 0;
    cudaError_t e = pick_megakernel(exhaustive, count, bvh, volumes, [&](void (*k)(const RenderArgs)) {
        return cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k, RTR_MEGA_THREADS, 0);
    });
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_megakernel(const RenderArgs& a, int grid, bool exhaustive, bool count, cudaStream_t st) {
    pick_megakernel(exhaustive, count, a.enable_bvh_tree != 0, a.sc.n_volumes != 0u, [&](void (*k)(const RenderArgs)) {
        k<<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
        return cudaSuccess;
    });
    return cudaGetLastError();
}

cudaError_t launch_sum_chunks(const float4* partials, float4* accum, uint32_t n_pixels, uint32_t n_chunks, bool accumulate, cudaStream_t st) {
    sum_chunks_kernel<<<(n_pixels + 255) / 256, 256, 0, st>>>(partials, accum, n_pixels, n_chunks, accumulate);
    return cudaGetLastError();
}

cudaError_t launch_resolve(const float4* accum, uchar4* rgba, uint32_t n_pixels, float spp, cudaStream_t st) {
    resolve_kernel<<<(n_pixels + 255) / 256, 256, 0, st>>>(accum, rgba, n_pixels, spp);
    return cudaGetLastError();
}

cudaError_t launch_scatter_batch(const SceneView& sc, const rtr_scatter_in* in, size_t n, uint32_t mode, uint32_t k0,
                                 uint32_t k1, rtr_scatter_out* out, cudaStream_t st) {
    scatter_batch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(sc, in, n, mode, k0, k1, out);
    return cudaGetLastError();
}

cudaError_t launch_background_batch(const SceneView& sc, const rtr_ray* rays, size_t n, float* rgb, cudaStream_t st) {
    background_batch_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(sc, rays, n, rgb);
    return cudaGetLastError();
}

cudaError_t launch_camera_rays(const CameraBasis& cam, uint32_t W, uint32_t H, uint32_t k0, uint32_t k1,
                               const uint32_t* pixels, const uint32_t* samples, size_t n, rtr_ray* out,
                               cudaStream_t st) {
    camera_rays_kernel<<<(unsigned)((n + 127) / 128), 128, 0, st>>>(cam, W, H, k0, k1, pixels, samples, n, out);
    return cudaGetLastError();
}

}  // namespace rtr
