"""ctypes mirror of include/rtr.h (plain structs only; no compute).

Field order and types must match the header exactly; tests/test_abi.py checks sizeof/offsets
against a C probe compiled from the header.
"""
import ctypes as C

RTR_ABI_VERSION = 2
RTR_NO_HIT = 0xFFFFFFFF

RTR_OK, RTR_INVALID_ARG, RTR_CUDA_ERROR, RTR_OOM, RTR_UNSUPPORTED, RTR_IO_ERROR, RTR_NCCL_ERROR = range(7)
STATUS_NAMES = ["RTR_OK", "RTR_INVALID_ARG", "RTR_CUDA_ERROR", "RTR_OOM", "RTR_UNSUPPORTED", "RTR_IO_ERROR",
                "RTR_NCCL_ERROR"]

MODE_FULL, MODE_CLAY, MODE_RANDOM, MODE_NORMAL = range(4)
OBJ_SPHERE, OBJ_MESH, OBJ_VOLUME = range(3)
MAT_LAMBERTIAN, MAT_METAL, MAT_EMISSION, MAT_DIELECTRIC, MAT_MIX, MAT_ISOTROPIC = range(6)
BG_UNIFORM, BG_GRADIENT, BG_SKYMAP = range(3)
KERNEL_AUTO, KERNEL_MEGAKERNEL, KERNEL_WAVEFRONT = range(3)
FLAG_EXHAUSTIVE = 1
FLAG_COUNT_WORK = 2
FLAG_REFERENCE_TREE = 4
FLAG_ACCUMULATE = 8
REDUCE_AUTO, REDUCE_NCCL, REDUCE_P2P = range(3)
SPLIT_STATIC, SPLIT_DYNAMIC = range(2)
NOALLOC_MAX_RAYS = 262144

f32, u32, u64, f64 = C.c_float, C.c_uint32, C.c_uint64, C.c_double


class Vec3(C.Structure):
    _fields_ = [("x", f32), ("y", f32), ("z", f32)]

    def __init__(self, x=0.0, y=0.0, z=0.0):
        super().__init__(float(x), float(y), float(z))

    def tuple(self):
        return (self.x, self.y, self.z)

    def __repr__(self):
        return f"Vec3({self.x}, {self.y}, {self.z})"


class CameraS(C.Structure):
    _fields_ = [("lookfrom", Vec3), ("lookat", Vec3), ("vertical", Vec3),
                ("vertical_fov", f32), ("aspect_ratio", f32)]


class SettingsS(C.Structure):
    _fields_ = [("samples_per_pixel", u32), ("max_ray_depth", u32), ("clamp_indirect", f32),
                ("enable_multithreading", u32), ("enable_bvh_tree", u32), ("mode", u32)]


class ObjectRef(C.Structure):
    _fields_ = [("kind", u32), ("index", u32)]


class SphereS(C.Structure):
    _fields_ = [("center", Vec3), ("radius", f32), ("material", u32)]


class MeshS(C.Structure):
    _fields_ = [("positions", C.POINTER(f32)), ("normals", C.POINTER(f32)), ("indices", C.POINTER(u32)),
                ("n_vertices", u32), ("n_indices", u32), ("material", u32), ("_pad", u32)]


class VolumeS(C.Structure):
    _fields_ = [("boundary", ObjectRef), ("neg_inv_density", f32)]


class MaterialS(C.Structure):
    _fields_ = [("kind", u32), ("color", Vec3), ("param", f32), ("first", u32), ("second", u32)]


class BackgroundS(C.Structure):
    _fields_ = [("kind", u32), ("color", Vec3), ("bottom", Vec3), ("texels", C.POINTER(f32)),
                ("width", u32), ("height", u32)]


class SceneDesc(C.Structure):
    _fields_ = [("objects", C.POINTER(ObjectRef)), ("spheres", C.POINTER(SphereS)),
                ("meshes", C.POINTER(MeshS)), ("volumes", C.POINTER(VolumeS)),
                ("materials", C.POINTER(MaterialS)),
                ("n_objects", u32), ("n_spheres", u32), ("n_meshes", u32), ("n_volumes", u32),
                ("n_materials", u32), ("_pad", u32),
                ("background", BackgroundS)]


class RenderParams(C.Structure):
    _fields_ = [("width", u32), ("height", u32), ("camera", CameraS), ("settings", SettingsS),
                ("seed", u64), ("sample_begin", u32), ("sample_count", u32), ("kernel", u32), ("flags", u32)]


class Stats(C.Structure):
    _fields_ = [("device_ms", f64), ("upload_ms", f64), ("rays", u64), ("samples", u64),
                ("nodes_visited", u64), ("tri_tests", u64), ("sphere_tests", u64), ("env_lookups", u64),
                ("kernel_launches", u32), ("kernel_used", u32)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class RayS(C.Structure):
    _fields_ = [("origin", Vec3), ("direction", Vec3)]


class HitS(C.Structure):
    _fields_ = [("object", u32), ("primitive", u32), ("t", f32), ("point", Vec3), ("normal", Vec3),
                ("front_face", u32), ("material", u32)]


class ScatterIn(C.Structure):
    _fields_ = [("ray", RayS), ("point", Vec3), ("normal", Vec3), ("t", f32), ("front_face", u32),
                ("material", u32), ("pixel", u32), ("sample", u32), ("bounce", u32)]


class ScatterOut(C.Structure):
    _fields_ = [("attenuation", Vec3), ("scattered", u32), ("ray", RayS)]


# numpy views of the flat records (all fields are 4-byte, so plain structured dtypes match)
import numpy as np  # noqa: E402

RAY_DTYPE = np.dtype([("origin", np.float32, 3), ("direction", np.float32, 3)])
HIT_DTYPE = np.dtype([("object", np.uint32), ("primitive", np.uint32), ("t", np.float32),
                      ("point", np.float32, 3), ("normal", np.float32, 3),
                      ("front_face", np.uint32), ("material", np.uint32)])
SCATTER_IN_DTYPE = np.dtype([("origin", np.float32, 3), ("direction", np.float32, 3),
                             ("point", np.float32, 3), ("normal", np.float32, 3), ("t", np.float32),
                             ("front_face", np.uint32), ("material", np.uint32),
                             ("pixel", np.uint32), ("sample", np.uint32), ("bounce", np.uint32)])
SCATTER_OUT_DTYPE = np.dtype([("attenuation", np.float32, 3), ("scattered", np.uint32),
                              ("origin", np.float32, 3), ("direction", np.float32, 3)])
assert RAY_DTYPE.itemsize == C.sizeof(RayS)
assert HIT_DTYPE.itemsize == C.sizeof(HitS)
assert SCATTER_IN_DTYPE.itemsize == C.sizeof(ScatterIn)
assert SCATTER_OUT_DTYPE.itemsize == C.sizeof(ScatterOut)

# every symbol include/rtr.h declares (tests check the built library exports all of them)
EXPORTED_SYMBOLS = [
    "rtr_abi_version", "rtr_context_create", "rtr_context_destroy", "rtr_last_error",
    "rtr_scene_upload", "rtr_scene_destroy", "rtr_scene_info", "rtr_render",
    "rtr_render_accumulate_device", "rtr_resolve_device", "rtr_synchronize",
    "rtr_trace_batch", "rtr_trace_batch_device_noalloc", "rtr_scatter_batch", "rtr_background_batch", "rtr_camera_rays",
    "rtr_group_create", "rtr_group_destroy", "rtr_group_size", "rtr_group_reduce_mode", "rtr_group_last_error",
    "rtr_group_scene_upload", "rtr_group_scene_destroy", "rtr_group_render", "rtr_group_device_stats",
    "rtr_group_set_split",
    "rtr_scene_load_json", "rtr_host_scene_desc", "rtr_host_scene_camera", "rtr_host_scene_settings",
    "rtr_host_scene_destroy", "rtr_mesh_load_obj", "rtr_host_mesh_get", "rtr_host_mesh_destroy",
    "rtr_write_png", "rtr_write_exr", "rtr_image_load", "rtr_host_image_width", "rtr_host_image_height",
    "rtr_host_image_rgb", "rtr_host_image_destroy",
]
