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
template <bool EXHAUSTIVE, bool COUNT>
__global__ void __launch_bounds__(RTR_MEGA_THREADS, RTR_MEGA_MIN_BLOCKS) render_megakernel(const RenderArgs a) {
    StackEntry stack[RTR_STACK_CAP];
    const uint32_t lane = threadIdx.x & 31u;
    unsigned long long n_rays = 0, n_nodes = 0, n_tris = 0, n_spheres = 0, n_env = 0;

    for (;;) {
        uint32_t item = 0;
        if (lane == 0) item = atomicAdd(a.work_counter, 1u);
        item = __shfl_sync(0xffffffffu, item, 0);
        if (item >= a.n_tiles * a.n_chunks) break;
        const uint32_t tile = item / a.n_chunks, chunk = item - tile * a.n_chunks;
        const uint32_t tx = tile % a.tiles_x, ty = tile / a.tiles_x;
        const uint32_t x = tx * 8u + (lane & 7u), y = ty * 4u + (lane >> 3);
        if (x < a.width && y < a.height) {
            const uint32_t pixel = y * a.width + x;  // index = y*W + x (render.rs:76-77)
            uint32_t s = a.sample_begin + chunk * a.chunk_samples;
            const uint32_t s_last = a.sample_begin + a.sample_count;
            const uint32_t s_end = min(s + a.chunk_samples, s_last);
            Counters cnt; cnt.nodes = cnt.tris = cnt.spheres = cnt.env = 0;
            uint32_t rays = 0;
            v3 sum = mk(0.0f, 0.0f, 0.0f);
            RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = s; key.bounce = 0;
            Ray ray; ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
            v3 T = mk(1.0f, 1.0f, 1.0f);
            uint32_t depth = 0;
            bool fresh = true;
            for (;;) {
                if (fresh) {
                    if (s >= s_end) break;
                    key.sample = s; key.bounce = 0;
                    ray = camera_sample_ray(a.cam, a.width, a.height, x, y, key);  // render.rs:88-92
                    T = mk(1.0f, 1.0f, 1.0f);
                    depth = 0;
                    fresh = false;
                }
                // one trace_ray frame (render.rs:32-59); L = radiance when the path ends here
                v3 L;
                bool done = true;
                if (depth >= a.max_ray_depth) {
                    L = T * mk(0.0f, 0.0f, 0.0f);
                } else {
                    key.bounce = depth;
                    ++rays;
                    const LiteHit lh = world_hit<EXHAUSTIVE, COUNT>(a.sc, a.enable_bvh_tree != 0, ray, 0.00001f, key,
                                                                    stack, &cnt);
                    if (lh.ref == REF_EMPTY_VALUE) {
                        L = T * background_sample(a.sc, ray, COUNT ? &cnt : nullptr);  // :57
                    } else {
                        const Hit hit = finish_hit(a.sc, ray, lh);
                        if (a.mode == RTR_MODE_FULL || a.mode == RTR_MODE_CLAY) {
                            const Scatter sc = scatter_event(a.sc, a.mode == RTR_MODE_CLAY, hit.material, ray, hit, key);  // :40-41
                            if (sc.some) {  // attenuation * trace_ray(scattered, depth + 1)  (:51-52)
                                T = T * sc.attenuation;
                                ray = sc.ray;
                                ++depth;
                                done = false;
                            } else {
                                L = T * sc.attenuation;  // :54
                            }
                        } else if (a.mode == RTR_MODE_NORMAL) {  // :42-45
                            L = T * (0.5f * (normalize(hit.normal) + mk(1.0f, 1.0f, 1.0f)));
                        } else {
                            L = T * mk(0.0f, 0.0f, 0.0f);  // Random (:46-48)
                        }
                    }
                }
                if (done) {
                    L.x = f32_clamp(L.x, 0.0f, a.clamp_indirect);  // render.rs:95-97
                    L.y = f32_clamp(L.y, 0.0f, a.clamp_indirect);
                    L.z = f32_clamp(L.z, 0.0f, a.clamp_indirect);
                    sum = sum + L;  // :99
                    ++s;
                    fresh = true;
                }
            }
            const uint32_t first = a.sample_begin + chunk * a.chunk_samples;
            const float n = first < s_end ? (float)(s_end - first) : 0.0f;
            float4* dst = a.n_chunks == 1 ? a.accum : a.partials + (size_t)chunk * a.width * a.height;
            dst[pixel] = make_float4(sum.x, sum.y, sum.z, n);
            n_rays += rays;
            if (COUNT) { n_nodes += cnt.nodes; n_tris += cnt.tris; n_spheres += cnt.spheres; n_env += cnt.env; }
        }
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

__global__ void sum_chunks_kernel(const float4* partials, float4* accum, uint32_t n_pixels, uint32_t n_chunks) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    float4 s = partials[i];
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

// ------------------------------------------------------------------ batch (parity) kernels
template <bool EXHAUSTIVE>
__global__ void __launch_bounds__(128) trace_batch_kernel(const SceneView sc, const rtr_ray* rays, size_t n, float tmin,
                                                          uint32_t bvh, uint32_t k0, uint32_t k1, rtr_hit* out,
                                                          unsigned long long* stats) {
    StackEntry stack[RTR_STACK_CAP];
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    Ray r;
    r.o = mk(rays[i].origin.x, rays[i].origin.y, rays[i].origin.z);
    r.d = mk(rays[i].direction.x, rays[i].direction.y, rays[i].direction.z);
    RngKey key; key.k0 = k0; key.k1 = k1; key.pixel = (uint32_t)i; key.sample = 0; key.bounce = 0;
    Counters cnt; cnt.nodes = cnt.tris = cnt.spheres = cnt.env = 0;
    const LiteHit lh = world_hit<EXHAUSTIVE, true>(sc, bvh != 0, r, tmin, key, stack, &cnt);
    rtr_hit h;
#ifdef RTR_DEBUG_BOUNDS
    if (lh.t == -12345.0f || lh.t == -54321.0f) {
        h.object = 0xDEADBEEFu; h.primitive = lh.ref; h.t = lh.t; h.point.x = lh.u; h.point.y = h.point.z = 0.0f;
        h.normal.x = h.normal.y = h.normal.z = 0.0f; h.front_face = 0; h.material = 0;
        out[i] = h;
        return;
    }
#endif
    if (lh.ref == REF_EMPTY_VALUE) {
        h.object = RTR_NO_HIT; h.primitive = 0; h.t = 0.0f;
        h.point.x = h.point.y = h.point.z = 0.0f; h.normal.x = h.normal.y = h.normal.z = 0.0f;
        h.front_face = 0; h.material = 0;
    } else {
        const Hit f = finish_hit(sc, r, lh);
        h.object = f.object; h.primitive = f.primitive; h.t = f.t;
        h.point.x = f.point.x; h.point.y = f.point.y; h.point.z = f.point.z;
        h.normal.x = f.normal.x; h.normal.y = f.normal.y; h.normal.z = f.normal.z;
        h.front_face = f.front_face ? 1u : 0u; h.material = f.material;
    }
    out[i] = h;
    atomicAdd(&stats[STAT_NODES], (unsigned long long)cnt.nodes);
    atomicAdd(&stats[STAT_TRIS], (unsigned long long)cnt.tris);
    atomicAdd(&stats[STAT_SPHERES], (unsigned long long)cnt.spheres);
}

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
int megakernel_max_blocks_per_sm(bool exhaustive, bool count) {
    int n = 0;
    cudaError_t e;
    if (exhaustive) e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_megakernel<true, true>, RTR_MEGA_THREADS, 0)
                              : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_megakernel<true, false>, RTR_MEGA_THREADS, 0);
    else e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_megakernel<false, true>, RTR_MEGA_THREADS, 0)
                   : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_megakernel<false, false>, RTR_MEGA_THREADS, 0);
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_megakernel(const RenderArgs& a, int grid, bool exhaustive, bool count, cudaStream_t st) {
    if (exhaustive) {
        if (count) render_megakernel<true, true><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
        else render_megakernel<true, false><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
    } else {
        if (count) render_megakernel<false, true><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
        else render_megakernel<false, false><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
    }
    return cudaGetLastError();
}

cudaError_t launch_sum_chunks(const float4* partials, float4* accum, uint32_t n_pixels, uint32_t n_chunks, cudaStream_t st) {
    sum_chunks_kernel<<<(n_pixels + 255) / 256, 256, 0, st>>>(partials, accum, n_pixels, n_chunks);
    return cudaGetLastError();
}

cudaError_t launch_resolve(const float4* accum, uchar4* rgba, uint32_t n_pixels, float spp, cudaStream_t st) {
    resolve_kernel<<<(n_pixels + 255) / 256, 256, 0, st>>>(accum, rgba, n_pixels, spp);
    return cudaGetLastError();
}

cudaError_t launch_trace_batch(const SceneView& sc, const rtr_ray* rays, size_t n, float tmin, uint32_t bvh,
                               bool exhaustive, uint32_t k0, uint32_t k1, rtr_hit* out, unsigned long long* stats,
                               cudaStream_t st) {
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (exhaustive) trace_batch_kernel<true><<<grid, 128, 0, st>>>(sc, rays, n, tmin, bvh, k0, k1, out, stats);
    else trace_batch_kernel<false><<<grid, 128, 0, st>>>(sc, rays, n, tmin, bvh, k0, k1, out, stats);
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
