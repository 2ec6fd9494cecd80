// scan.cu — render() + trace_ray() (lib/core/render.rs:32-127) for the reference's BVH-off path
// (`Hittable for Scene`, lib/core/scene.rs:141-156) when EVERY object is a sphere: scenes/benchmark.json as shipped.
//
// The generic megakernel (kernels.cu) runs this case at 13 of 32 lanes per instruction (ncu,
// profiles/r1_benchmark_megakernel.txt): the linear scan itself is uniform work, but the per-lane path state
// machine executes its camera ray (one Philox call, ~7 lanes), its scatter (the unit-ball rejection loop, ~8 lanes:
// the loop runs until the unluckiest lane is accepted) and its background one after the other.  This kernel keeps
// the same lane = pixel, in-order sample sum (bit-reproducible, render.rs:99) and the same Philox counters — so it
// is bit-identical to the generic kernel and to the oracle — but arranges one loop trip as phases every live lane
// takes part in:
//
//   classify   finish the lane's last world.hit: background (render.rs:57) or HitRecord + material record;
//              paths that end here (miss, Emission, depth limit) are clamped and summed at once (render.rs:95-99)
//   draw       ONE Philox site per trip, executed by all lanes: a lane that starts a new sample draws its pixel
//              jitter (render.rs:88-89), a lane that scatters draws candidate 0 of its unit-ball rejection loop
//              (sample.rs:44-52) or its Dielectric uniform (materials.rs:127)
//   tail       WARP-COOPERATIVE rejection loop: the lanes whose candidate 0 was rejected (47.6 % of them) are
//              served by all 32 lanes, each evaluating candidate k of some failing lane's counter stream
//              (counter-based RNG: any lane can compute any lane's k-th draw); the first accepted k wins, exactly
//              as in the sequential loop.  Two or three full-width rounds replace five or six thin ones.
//   epilogue   camera ray (camera.rs:19-24) or the material's few arithmetic instructions (materials.rs)
//   scan       world.hit: the linear scan over the spheres, all live lanes
#include "device_core.cuh"
#include "kernels.h"

namespace rtr {

enum { W_NONE = 0, W_JITTER = 1, W_BALL = 2, W_UNIFORM = 3 };

template <bool COUNT>
__global__ void __launch_bounds__(RTR_MEGA_THREADS, RTR_SCAN_MIN_BLOCKS) render_scan_kernel(const RenderArgs a) {
    __shared__ uint4 serve_tab[RTR_MEGA_THREADS / 32][32];  // coop_ball_tail's per-warp table
    const uint32_t lane = threadIdx.x & 31u;
    uint4* tab = serve_tab[threadIdx.x >> 5];
    const uint32_t FULL = 0xffffffffu;
    const uint32_t NONE = 0xFFFFFFFFu;
    unsigned long long n_rays = 0, n_env = 0;
    const SceneView& sc = a.sc;
    const uint32_t n_obj = sc.n_objects;
    const float4* spheres = reinterpret_cast<const float4*>(sc.linear_spheres);

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
        uint32_t rays = 0, envs = 0;  // rays: warp total (every lane holds the same count)
        v3 sum = mk(0.0f, 0.0f, 0.0f);
        Ray ray; ray.o = ray.d = mk(0.0f, 0.0f, 0.0f);
        v3 T = mk(1.0f, 1.0f, 1.0f);
        uint32_t depth = 0;
        bool alive = valid && s < s_end;
        bool traced = false;     // the lane holds the result of a world.hit that has not been shaded yet
        float hit_t = RTR_INF;
        uint32_t winner = NONE;  // Scene.objects index of the closest sphere

        for (;;) {
            // ---------------------------------------------------------------- classify  [bucket:path_fsm]
            uint32_t want = W_NONE;
            bool fin = false;
            v3 L = mk(0.0f, 0.0f, 0.0f);
            uint32_t kind = RTR_MAT_EMISSION;
            v3 color = mk(0.0f, 0.0f, 0.0f), normal = mk(0.0f, 0.0f, 0.0f), point = mk(0.0f, 0.0f, 0.0f), unit = mk(0.0f, 0.0f, 0.0f);
            float param = 0.0f, ratio = 1.0f, cos_theta = 0.0f;
            bool cannot_refract = false;
            if (alive && traced) {
                if (winner == NONE) {  // render.rs:57
                    Counters c; c.env = 0;
                    L = T * background_sample(sc, ray, COUNT ? &c : nullptr);
                    if (COUNT) envs += c.env;
                    fin = true;
                } else {
                    // HitRecord of the winner: objects.rs:45-50, ray.rs:39-54  [bucket:finish_hit]
                    const float4 sp = spheres[winner];
                    const uint32_t slot = RTR_REF_INDEX(sc.objects[winner].entry);
                    const uint32_t material = sc.sphere_info[slot].z;
                    point = ray_at(ray, hit_t);
                    const v3 n = (point - xyz(sp)) / sp.w;
                    const bool front_face = dot(ray.d, n) < 0.0f;
                    normal = front_face ? n : -n;
                    if (a.mode == RTR_MODE_FULL || a.mode == RTR_MODE_CLAY) {  // render.rs:40-41  [bucket:scatter]
                        kind = RTR_MAT_LAMBERTIAN; color = mk(0.8f, 0.8f, 0.8f);
                        if (a.mode == RTR_MODE_FULL) {
                            MaterialRec m = sc.materials[material];
                            if (m.kind == RTR_MAT_MIX) {  // MixMaterial (materials.rs:150-156)
                                RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = s; key.bounce = depth;
                                for (uint32_t level = 0; m.kind == RTR_MAT_MIX; ++level) {
                                    const bool first = gen_range_01(draw(key, P_MIX, level).x) >= m.param;
                                    m = sc.materials[first ? m.first : m.second];
                                }
                            }
                            kind = m.kind; color = mk(m.r, m.g, m.b); param = m.param;
                        }
                        if (kind == RTR_MAT_EMISSION) {  // materials.rs:88-90, render.rs:54
                            L = T * color;
                            fin = true;
                        } else if (depth + 1u >= a.max_ray_depth && kind != RTR_MAT_METAL) {
                            // the scattered ray would meet the depth limit (render.rs:34-36): attenuation * 0.  The
                            // attenuation of these materials does not depend on the draw, so the draw is skipped.
                            const v3 att = kind == RTR_MAT_DIELECTRIC ? mk(1.0f, 1.0f, 1.0f) : color;
                            L = (T * att) * mk(0.0f, 0.0f, 0.0f);
                            fin = true;
                        } else if (kind == RTR_MAT_DIELECTRIC || kind == RTR_MAT_METAL) {
                            // one normalize site for both: the Dielectric's unit direction (materials.rs:120) and the
                            // Metal's unit reflection (:59); a normalize is ~25 instructions with IEEE sqrt and division
                            unit = normalize(kind == RTR_MAT_METAL ? reflect(ray.d, normal) : ray.d);
                            want = W_BALL;
                            if (kind == RTR_MAT_DIELECTRIC) {  // prologue, materials.rs:114-124
                                ratio = front_face ? 1.0f / param : param;
                                cos_theta = fminf(dot(normal, -unit), 1.0f);
                                const float sin_theta = sqrtf(1.0f - cos_theta * cos_theta);
                                cannot_refract = (ratio * sin_theta) > 1.0f;
                                want = cannot_refract ? W_NONE : W_UNIFORM;  // no draw under total internal reflection (Q11)
                            }
                        } else {
                            want = W_BALL;  // Lambertian / Metal / Isotropic
                        }
                    } else if (a.mode == RTR_MODE_NORMAL) {  // render.rs:42-45
                        L = T * (0.5f * (normalize(normal) + mk(1.0f, 1.0f, 1.0f)));
                        fin = true;
                    } else {  // Random (render.rs:46-48)
                        L = T * mk(0.0f, 0.0f, 0.0f);
                        fin = true;
                    }
                }
            }
            bool fresh = alive && !traced;
            if (fin) {  // render.rs:95-99  [bucket:path_fsm]
                L.x = f32_clamp(L.x, 0.0f, a.clamp_indirect);
                L.y = f32_clamp(L.y, 0.0f, a.clamp_indirect);
                L.z = f32_clamp(L.z, 0.0f, a.clamp_indirect);
                sum = sum + L;
                ++s;
                fresh = true;
            }
            if (fresh) {
                traced = false;
                depth = 0;
                if (s >= s_end) alive = false;
                else want = W_JITTER;
            }
            if (__ballot_sync(FULL, alive) == 0u) break;

            // ---------------------------------------------------------------- the one draw site  [bucket:rng]
            const uint32_t purpose = want == W_JITTER ? P_JITTER : want == W_BALL ? P_SPHERE : P_DIELECTRIC;
            const uint4 r0 = philox_rk(pixel, s, (depth & 0xFFu) | (purpose << 8), 0u, a.pk);
            v3 ball = mk(gen_range_m11(r0.x), gen_range_m11(r0.y), gen_range_m11(r0.z));

            // ---------------------------------------------------------------- cooperative rejection tail (device_core.cuh)
            coop_ball_tail(want == W_BALL && !(dot(ball, ball) <= 1.0f), pixel, s, depth, tab, a.pk, ball);

            // ---------------------------------------------------------------- epilogue
            bool late = false;  // the path ended after its draw (absorbed Metal, or Metal at the depth limit)
            if (want == W_JITTER) {  // render.rs:88-92, camera.rs:19-24  [bucket:camera]
                const float u = ((float)x + gen_range_01_inclusive(r0.x)) / (float)(a.width - 1);
                const float v = ((float)y + gen_range_01_inclusive(r0.y)) / (float)(a.height - 1);
                ray.o = a.cam.origin;
                ray.d = a.cam.lower_left_corner + a.cam.horizontal * u - a.cam.vertical * v - a.cam.origin;
                T = mk(1.0f, 1.0f, 1.0f);
            } else if (alive) {  // Material::scatter, second half  [bucket:scatter]
                bool some = true;
                v3 att = color, dir;
                v3 nball = ball;  // unit_sphere_surface() of Lambertian and Metal (sample.rs:66-70), one site for both
                if (kind == RTR_MAT_LAMBERTIAN || kind == RTR_MAT_METAL) nball = normalize(ball);
                if (kind == RTR_MAT_LAMBERTIAN) {  // materials.rs:31-41
                    dir = normal + nball;
                    if (near_zero(dir)) dir = normal;
                } else if (kind == RTR_MAT_METAL) {  // :58-70
                    dir = unit + param * nball;
                    some = dot(dir, normal) > 0.0f;
                    if (!some) att = mk(0.0f, 0.0f, 0.0f);
                } else if (kind == RTR_MAT_DIELECTRIC) {  // :125-138
                    const bool refl = cannot_refract || reflectance(cos_theta, ratio) > gen_range_01(r0.x);
                    dir = refl ? reflect(unit, normal) : refract(unit, normal, ratio);
                    att = mk(1.0f, 1.0f, 1.0f);
                } else {  // Isotropic (volume.rs:82-87)
                    dir = ball;
                }
                if (some) {  // attenuation * trace_ray(scattered, depth + 1)  (render.rs:51-52)
                    T = T * att;
                    ray.o = point; ray.d = dir;
                    ++depth;
                    if (depth >= a.max_ray_depth) { L = T * mk(0.0f, 0.0f, 0.0f); late = true; }  // :34-36
                } else {
                    L = T * att;  // :54
                    late = true;
                }
            }
            if (late) {  // rare: clamp, sum and start the next sample with a draw of its own
                L.x = f32_clamp(L.x, 0.0f, a.clamp_indirect);
                L.y = f32_clamp(L.y, 0.0f, a.clamp_indirect);
                L.z = f32_clamp(L.z, 0.0f, a.clamp_indirect);
                sum = sum + L;
                ++s;
                depth = 0;
                if (s >= s_end) {
                    alive = false;
                } else {
                    RngKey key; key.k0 = a.k0; key.k1 = a.k1; key.pixel = pixel; key.sample = s; key.bounce = 0;
                    ray = camera_sample_ray(a.cam, a.width, a.height, x, y, key);
                    T = mk(1.0f, 1.0f, 1.0f);
                }
            }

            // ---------------------------------------------------------------- world.hit (scene.rs:141-156)  [bucket:primitive_tests]
            traced = alive;
            rays += (uint32_t)__popc(__ballot_sync(FULL, alive));
            if (alive) {
                float closest = RTR_INF;  // closest_so_far (scene.rs:144)
                winner = NONE;
                for (uint32_t i = 0; i < n_obj; ++i) {
                    const float4 sp = __ldg(spheres + i);
                    float t;
                    if (sphere_t<true>(sp, ray, 0.00001f, closest, t)) { closest = t; winner = i; }
                }
                hit_t = closest;
            }
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
        if (COUNT) n_env += envs;
    }
    if (COUNT) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) n_env += __shfl_xor_sync(0xffffffffu, n_env, o);
    }
    if (lane == 0) {
        atomicAdd(&a.stats[STAT_RAYS], n_rays);
        if (COUNT) {
            atomicAdd(&a.stats[STAT_SPHERES], n_rays * n_obj);  // every ray tests every sphere (scene.rs:146)
            atomicAdd(&a.stats[STAT_ENV], n_env);
        }
    }
}

int scan_kernel_max_blocks_per_sm(bool count) {
    int n = 0;
    cudaError_t e = count ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_scan_kernel<true>, RTR_MEGA_THREADS, 0)
                          : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, render_scan_kernel<false>, RTR_MEGA_THREADS, 0);
    return e == cudaSuccess ? n : 0;
}

cudaError_t launch_scan_kernel(const RenderArgs& a, int grid, bool count, cudaStream_t st) {
    if (count) render_scan_kernel<true><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
    else render_scan_kernel<false><<<grid, RTR_MEGA_THREADS, 0, st>>>(a);
    return cudaGetLastError();
}

}  // namespace rtr
