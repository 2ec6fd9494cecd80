/*
 * rtr.h — C ABI of the B200 render hot path (librtr_b200.so).
 *
 * This is the drop-in boundary for the reference crate's
 *     pub fn render(target: &mut RenderTarget, scene: &Scene)      (lib/core/render.rs:61)
 * and for everything that call consumes.  The reference has no FFI of its own; the
 * Rust-side `lower()` pass walks the crate's `Scene` into the plain
 * structs below and calls these entry points.  No torch / C++ types appear here.
 *
 * Each struct cites the reference type it replaces (paths relative to the reference
 * checkout).  All floating point is IEEE f32 (lib/utils/types.rs:4).
 *
 * Ownership: every pointer passed IN is borrowed for the duration of the call only
 * (the library copies what it needs); every handle passed OUT is owned by the
 * library until the matching *_destroy.  Output buffers are caller-allocated.
 * Errors: never unwinds; returns rtr_status, message via rtr_last_error().
 * Threading: a context is not re-entrant; distinct contexts are independent.
 */
#ifndef RTR_H_
#define RTR_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTR_ABI_VERSION 2u
#define RTR_NO_HIT 0xFFFFFFFFu

typedef enum rtr_status {
    RTR_OK = 0,
    RTR_INVALID_ARG = 1,  /* the reference would panic (unwrap/expect/index) */
    RTR_CUDA_ERROR = 2,
    RTR_OOM = 3,
    RTR_UNSUPPORTED = 4,
    RTR_IO_ERROR = 5,
    RTR_NCCL_ERROR = 6
} rtr_status;

/* cgmath::Vector3<f32> (lib/utils/types.rs:8) and Color (lib/utils/color.rs:6-11) */
typedef struct rtr_vec3 { float x, y, z; } rtr_vec3;

/* Camera (lib/core/camera.rs:27-34) */
typedef struct rtr_camera {
    rtr_vec3 lookfrom, lookat, vertical;
    float vertical_fov;  /* degrees */
    float aspect_ratio;  /* independent of W/H, as in the reference */
} rtr_camera;

/* RenderMode (lib/core/scene.rs:15-20), serde names "Full","Clay","Random","Normal" */
enum { RTR_MODE_FULL = 0, RTR_MODE_CLAY = 1, RTR_MODE_RANDOM = 2, RTR_MODE_NORMAL = 3 };

/* RenderSettings (lib/core/scene.rs:23-30); defaults 5/6/10.0/true/true/Full (:52-59) */
typedef struct rtr_settings {
    uint32_t samples_per_pixel;
    uint32_t max_ray_depth;          /* u8 in the reference: 0..255 */
    float clamp_indirect;
    uint32_t enable_multithreading;  /* accepted, ignored on device */
    uint32_t enable_bvh_tree;        /* 0 = Hittable for Scene linear scan (scene.rs:141-156) */
    uint32_t mode;                   /* RTR_MODE_* */
} rtr_settings;

/* Box<dyn Object> (lib/core/traits.rs:17-20): tagged reference into the typed arrays */
enum { RTR_OBJ_SPHERE = 0, RTR_OBJ_MESH = 1, RTR_OBJ_VOLUME = 2 };
typedef struct rtr_object_ref { uint32_t kind, index; } rtr_object_ref;

/* Sphere (lib/objects.rs:15-20) */
typedef struct rtr_sphere { rtr_vec3 center; float radius; uint32_t material; } rtr_sphere;

/* Mesh after Mesh::from_file/from_buffers + build_bvh() (lib/core/mesh.rs:31-38,63-121).
 * positions/normals: 3 floats per vertex; indices: 3 per triangle (u32). */
typedef struct rtr_mesh {
    const float* positions;
    const float* normals;
    const uint32_t* indices;
    uint32_t n_vertices;
    uint32_t n_indices;
    uint32_t material;
    uint32_t _pad;
} rtr_mesh;

/* Volume (lib/volume.rs:19-23): boundary is a SPHERE, MESH or VOLUME ref (Box<dyn Object>; a chain of volumes must
 * end in geometry, a cyclic chain is RTR_INVALID_ARG); the boundary object is owned by the volume (it need not appear
 * in `objects`).  A volume bounded by a volume never hits in the reference (its inner query runs with tmin = -inf,
 * volume.rs:36-49) but still takes part in the BVH through its bounds; that is reproduced.  neg_inv_density is the
 * struct's stored field, i.e. -1.0/density as Volume::new computes it (volume.rs:26-31). */
typedef struct rtr_volume { rtr_object_ref boundary; float neg_inv_density; } rtr_volume;

/* Box<dyn Material> (lib/materials.rs, lib/volume.rs:75-87) */
enum {
    RTR_MAT_LAMBERTIAN = 0,  /* color = albedo */
    RTR_MAT_METAL = 1,       /* color = albedo, param = fuzz */
    RTR_MAT_EMISSION = 2,    /* color = color*strength, already multiplied (materials.rs:79-83) */
    RTR_MAT_DIELECTRIC = 3,  /* param = ir */
    RTR_MAT_MIX = 4,         /* param = factor, first/second = indices into materials[] */
    RTR_MAT_ISOTROPIC = 5    /* color */
};
typedef struct rtr_material {
    uint32_t kind;
    rtr_vec3 color;
    float param;
    uint32_t first, second;
} rtr_material;

/* Box<dyn Background> (lib/backgrounds.rs) */
enum { RTR_BG_UNIFORM = 0, RTR_BG_GRADIENT = 1, RTR_BG_SKYMAP = 2 };
typedef struct rtr_background {
    uint32_t kind;
    rtr_vec3 color;       /* uniform colour | gradient top */
    rtr_vec3 bottom;      /* gradient bottom */
    const float* texels;  /* SkyMap.image: RGB f32, row-major, image[x + y*width] */
    uint32_t width, height;
} rtr_background;

/* Scene minus camera/settings (lib/core/scene.rs:33-40).  MaterialHandle(i) == materials[i];
 * children of a MixMaterial are further entries of materials[]. */
typedef struct rtr_scene_desc {
    const rtr_object_ref* objects;  /* Scene.objects, in add_object order */
    const rtr_sphere* spheres;
    const rtr_mesh* meshes;
    const rtr_volume* volumes;
    const rtr_material* materials;
    uint32_t n_objects, n_spheres, n_meshes, n_volumes, n_materials;
    uint32_t _pad;
    rtr_background background;
} rtr_scene_desc;

/* ---- render ------------------------------------------------------------------- */

enum { RTR_KERNEL_AUTO = 0, RTR_KERNEL_MEGAKERNEL = 1, RTR_KERNEL_WAVEFRONT = 2 };
enum {
    RTR_FLAG_EXHAUSTIVE = 1u, /* traverse like the reference: both children, no t-culling */
    RTR_FLAG_COUNT_WORK = 2u, /* also count nodes / primitive tests / env lookups into rtr_stats */
    RTR_FLAG_ACCUMULATE = 8u, /* rtr_render_accumulate_device ADDS this sample slice to accum_dev instead of
                                 overwriting it: progressive rendering (the GUI's continuous mode, src/gui.rs:333-337) */
    RTR_FLAG_REFERENCE_TREE = 4u /* walk the reference's own BVH topology (bvh.rs:59-144) instead of the
                                    traversal tree built over the same gate boxes; same results, used to count
                                    work on the reference topology and to validate the traversal tree */
};

typedef struct rtr_render_params {
    uint32_t width, height;  /* RenderTarget.width/height (render.rs:14-19) */
    rtr_camera camera;       /* Scene.camera */
    rtr_settings settings;   /* Scene.settings */
    uint64_t seed;           /* Philox key; the reference is OS-seeded */
    uint32_t sample_begin;   /* this call renders global samples [begin, begin+count) */
    uint32_t sample_count;   /* 0 = all of settings.samples_per_pixel */
    uint32_t kernel;         /* RTR_KERNEL_* */
    uint32_t flags;          /* RTR_FLAG_* */
} rtr_render_params;

typedef struct rtr_stats {
    double device_ms;        /* CUDA events around the render kernels only */
    double upload_ms;        /* host flatten + H2D, filled by rtr_scene_upload queries */
    uint64_t rays;           /* world.hit queries (render.rs:38) */
    uint64_t samples;        /* iterations of render.rs:86 */
    uint64_t nodes_visited;  /* interior-node boxes slab-tested (32 B each) */
    uint64_t tri_tests;
    uint64_t sphere_tests;
    uint64_t env_lookups;    /* SkyMap texel fetches */
    uint32_t kernel_launches;
    uint32_t kernel_used;    /* RTR_KERNEL_* actually run */
} rtr_stats;

/* ---- batch parity entry points -------------------------------------------------- */

typedef struct rtr_ray { rtr_vec3 origin, direction; } rtr_ray;

/* HitRecord<MaterialHandle> (lib/utils/ray.rs:30-36) + which primitive produced it */
typedef struct rtr_hit {
    uint32_t object;     /* index into Scene.objects, RTR_NO_HIT on miss */
    uint32_t primitive;  /* triangle index inside the mesh (original order); 0 otherwise */
    float t;
    rtr_vec3 point;
    rtr_vec3 normal;     /* already front-face flipped */
    uint32_t front_face;
    uint32_t material;
} rtr_hit;

typedef struct rtr_scatter_in {
    rtr_ray ray;
    rtr_vec3 point, normal;
    float t;
    uint32_t front_face;
    uint32_t material;
    uint32_t pixel, sample, bounce;  /* RNG counter for this event */
} rtr_scatter_in;

typedef struct rtr_scatter_out {
    rtr_vec3 attenuation;
    uint32_t scattered;  /* Option<Ray>::is_some */
    rtr_ray ray;
} rtr_scatter_out;

typedef struct rtr_context rtr_context;
typedef struct rtr_scene rtr_scene;

uint32_t rtr_abi_version(void);

/* One context per GPU / per process rank.  stream: a cudaStream_t to launch on, or NULL
 * for a library-owned stream. */
rtr_status rtr_context_create(int device, void* stream, rtr_context** out);
void rtr_context_destroy(rtr_context* ctx);
const char* rtr_last_error(const rtr_context* ctx); /* ctx may be NULL: last global error */

/* Replaces scene.build_bvh() at render.rs:67 + mesh.build_bvh() (mesh.rs:119-121):
 * rebuilds the reference BVH topology (bvh.rs:59-144), flattens, uploads once. */
rtr_status rtr_scene_upload(rtr_context* ctx, const rtr_scene_desc* desc, rtr_scene** out);
void rtr_scene_destroy(rtr_scene* scene);
/* bytes resident on device / H2D bytes moved by the upload, host build ms */
rtr_status rtr_scene_info(const rtr_scene* scene, uint64_t* device_bytes, uint64_t* h2d_bytes,
                          double* build_ms, uint32_t* n_nodes, uint32_t* max_depth);

/* render() (render.rs:61-127).  rgba8_out: host, W*H*4 (may be NULL). accum_out: host,
 * W*H*4 f32 = per-pixel sum of clamped samples (rgb) + sample count (a) (may be NULL). */
rtr_status rtr_render(rtr_context* ctx, const rtr_scene* scene, const rtr_render_params* params,
                      uint8_t* rgba8_out, float* accum_out, rtr_stats* stats_out);

/* Multi-GPU building blocks (device pointers; async on the context stream):
 * accumulate this rank's sample slice into accum_dev (W*H float4, overwritten), then the
 * caller reduces accum_dev across ranks (one NCCL reduce) and resolves on the root. */
rtr_status rtr_render_accumulate_device(rtr_context* ctx, const rtr_scene* scene,
                                        const rtr_render_params* params, float* accum_dev,
                                        rtr_stats* stats_out);
/* sqrt(sum/spp) then Color::into_raw (render.rs:103-110, color.rs:34-41) */
rtr_status rtr_resolve_device(rtr_context* ctx, const float* accum_dev, uint32_t width,
                              uint32_t height, uint32_t samples_per_pixel, uint8_t* rgba8_dev);
rtr_status rtr_synchronize(rtr_context* ctx);

/* ---- one render() over N GPUs of this process -------------------------------------------------
 * render() (lib/core/render.rs:61-127) splits its pixels over host threads with rayon (render.rs:114-119); here the
 * SAMPLES of every pixel (render.rs:86-100) are split over the devices of a group: the scene is replicated, device r
 * renders a contiguous slice of the call's global samples into its own W*H float4 buffer (Philox is keyed by the
 * global (pixel, sample), so the image does not depend on N except for fp32 summation order), and ONE cross-device
 * step follows: either ncclReduce(sum, fp32, root 0) + resolve on the root (RTR_REDUCE_NCCL), or the library's own
 * kernel in which every device sums its 1/N of the image from all peers over NVLink in rank order and stores
 * accumulation + RGBA8 straight into the root's buffers (RTR_REDUCE_P2P, deterministic; needs peer access between all
 * pairs).  A Rust caller binds rtr_group_* exactly like the single-device calls. */
typedef struct rtr_group rtr_group;
typedef struct rtr_group_scene rtr_group_scene;
enum { RTR_REDUCE_AUTO = 0, RTR_REDUCE_NCCL = 1, RTR_REDUCE_P2P = 2 };
enum { RTR_SPLIT_STATIC = 0, RTR_SPLIT_DYNAMIC = 1 };

/* devices: CUDA device ordinals, root first (the image is produced on devices[0]); n_devices <= 16.  reduce:
 * RTR_REDUCE_*; AUTO = P2P when every pair of devices has peer access, else NCCL (env RTR_GROUP_REDUCE=nccl|p2p
 * overrides AUTO).  The same ordinal may appear more than once (P2P only): the slices then time-share that device,
 * which is how the N-way split is tested on a one-GPU box. */
rtr_status rtr_group_create(const int* devices, uint32_t n_devices, uint32_t reduce, rtr_group** out);
void rtr_group_destroy(rtr_group* group);
uint32_t rtr_group_size(const rtr_group* group);
uint32_t rtr_group_reduce_mode(const rtr_group* group); /* what AUTO resolved to */
const char* rtr_group_last_error(const rtr_group* group);
/* flattens ONCE on the host, then uploads to every device of the group concurrently */
rtr_status rtr_group_scene_upload(rtr_group* group, const rtr_scene_desc* desc, rtr_group_scene** out);
void rtr_group_scene_destroy(rtr_group_scene* scene);
/* Same contract as rtr_render: host buffers, synchronous.  params->sample_begin/sample_count name the call's global
 * samples; they are partitioned over the devices.  stats_out: rays/samples/work summed over the devices, device_ms =
 * root-stream time from the start of its render to the end of the cross-device step, kernel_launches summed. */
rtr_status rtr_group_render(rtr_group* group, const rtr_group_scene* scene, const rtr_render_params* params,
                            uint8_t* rgba8_out, float* accum_out, rtr_stats* stats_out);
/* per-device view of the last rtr_group_render: stats (device_ms = that device's render kernels), its sample slice */
rtr_status rtr_group_device_stats(const rtr_group* group, uint32_t index, rtr_stats* stats_out,
                                  uint32_t* sample_begin, uint32_t* sample_count);
/* RTR_SPLIT_STATIC (default): balanced contiguous slices.  RTR_SPLIT_DYNAMIC: the call's samples are cut into
 * `chunk` -sample pieces (0 = library default) that the devices' host threads pull from a shared counter, for scenes
 * whose cost per sample differs between devices' slices (rayon's work stealing, render.rs:114-119). */
rtr_status rtr_group_set_split(rtr_group* group, uint32_t split, uint32_t chunk);

/* world.hit(ray, tmin, +INF) for explicit rays (render.rs:38).  Ray i draws any
 * volume random numbers from counter (pixel=i, sample=0, bounce=0).  Runs the render's own traversal kernel
 * (wf_extend) on the batch.  stats_out (may be NULL): rays, device_ms of the kernels (copies excluded); the work
 * counters are filled only under RTR_FLAG_COUNT_WORK or RTR_FLAG_REFERENCE_TREE. */
rtr_status rtr_trace_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays,
                           size_t n, float tmin, uint32_t enable_bvh_tree, uint32_t flags,
                           uint64_t seed, rtr_hit* hits_out, rtr_stats* stats_out);
/* rtr_trace_batch for callers that cannot allocate.  NO COUNTERPART IN THE REFERENCE: added
 * for code that must not call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  Same result as rtr_trace_batch, but rays_dev and hits_dev are DEVICE buffers of the context's device supplied
 * by the caller, and the kernels' scratch (ray / hit streams, queue counters) is a static device array inside the
 * library: RTR_NOALLOC_SCRATCH_BYTES of device memory held by every GPU that loads the library until the process
 * exits.  The call allocates nothing (no cudaMalloc / cudaFree / pool allocation), copies nothing to or from the host
 * and does not synchronise: it enqueues one memset + three kernels on the context's stream; call rtr_synchronize (or
 * order your own stream after it) before reading hits_dev.  n > RTR_NOALLOC_MAX_RAYS -> RTR_INVALID_ARG.
 * flags: RTR_FLAG_EXHAUSTIVE and RTR_FLAG_REFERENCE_TREE only.  All calls on one device share the one scratch array:
 * calls that overlap on the GPU (two streams or two contexts of a device) race on it — serialise them.  The first
 * call on a device looks the array's address up (cudaGetSymbolAddress); make it before starting a graph capture. */
#define RTR_NOALLOC_MAX_RAYS 262144u
#define RTR_NOALLOC_SCRATCH_BYTES 12587008u /* 4096 + 3 streams x 16 B x RTR_NOALLOC_MAX_RAYS */
rtr_status rtr_trace_batch_device_noalloc(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays_dev,
                                          size_t n, float tmin, uint32_t enable_bvh_tree, uint32_t flags,
                                          uint64_t seed, rtr_hit* hits_dev);
/* Material::scatter (materials.rs, volume.rs:82-87); mode FULL or CLAY (render.rs:40-41) */
rtr_status rtr_scatter_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_scatter_in* in,
                             size_t n, uint32_t mode, uint64_t seed, rtr_scatter_out* out);
/* Background::sample (backgrounds.rs:27-29,46-48,85-92): 3 floats per ray */
rtr_status rtr_background_batch(rtr_context* ctx, const rtr_scene* scene, const rtr_ray* rays,
                                size_t n, float* rgb_out);
/* RayOrigin::get_ray for the jittered sample (pixel, sample) (render.rs:88-92) */
rtr_status rtr_camera_rays(rtr_context* ctx, const rtr_render_params* params,
                           const uint32_t* pixels, const uint32_t* samples, size_t n,
                           rtr_ray* rays_out);

/* ---- host loaders (Scene::from_file scene.rs:74-76, Mesh::from_file mesh.rs:63-76) ---- */

typedef struct rtr_host_scene rtr_host_scene;
/* Parses the typetag/serde JSON the crate writes; clamp_indirect defaults to 10.0 */
rtr_status rtr_scene_load_json(const char* path, rtr_host_scene** out);
const rtr_scene_desc* rtr_host_scene_desc(const rtr_host_scene* hs);
const rtr_camera* rtr_host_scene_camera(const rtr_host_scene* hs);
const rtr_settings* rtr_host_scene_settings(const rtr_host_scene* hs);
void rtr_host_scene_destroy(rtr_host_scene* hs);

typedef struct rtr_host_mesh rtr_host_mesh;
/* Triangulated OBJ with v//vn or v/vt/vn faces; (position,normal) pairs de-duplicated */
rtr_status rtr_mesh_load_obj(const char* path, uint32_t material, rtr_host_mesh** out);
const rtr_mesh* rtr_host_mesh_get(const rtr_host_mesh* hm);
void rtr_host_mesh_destroy(rtr_host_mesh* hm);

/* ---- headless output (replaces the window blit of RenderTarget.data, src/main.rs:92-97) ---- */
/* 8-bit RGBA PNG of a host RGBA8 buffer as rtr_render fills it */
rtr_status rtr_write_png(const char* path, const uint8_t* rgba8, uint32_t width, uint32_t height);
/* fp32 image (the accumulation buffer divided by spp, or any RGB/RGBA float image) as a scan-line OpenEXR file with
 * float channels and ZIP blocks; channels = 3 (RGB) or 4 (RGBA), interleaved, row 0 = top */
rtr_status rtr_write_exr(const char* path, const float* pixels, uint32_t width, uint32_t height,
                         uint32_t channels);

/* ---- environment maps: SkyMap::new (backgrounds.rs:58-80) ---- */
typedef struct rtr_host_image rtr_host_image;
/* First RGB(A) layer of a single-part scan-line OpenEXR file (half / float / uint channels; NONE, RLE, ZIPS, ZIP,
 * PIZ or PXR24 blocks), or a Radiance RGBE .hdr file.  Texels come back as the reference stores them: RGB f32,
 * image[x + y*width], row 0 = first row of the data window; pass them as rtr_background.texels.
 * RTR_UNSUPPORTED for tiled / multi-part / deep files and B44 / DWA blocks, message via rtr_last_error(NULL). */
rtr_status rtr_image_load(const char* path, rtr_host_image** out);
uint32_t rtr_host_image_width(const rtr_host_image* img);
uint32_t rtr_host_image_height(const rtr_host_image* img);
const float* rtr_host_image_rgb(const rtr_host_image* img);
void rtr_host_image_destroy(rtr_host_image* img);

#ifdef __cplusplus
}
#endif
#endif /* RTR_H_ */
