"""ctypes binding of librtr_b200.so + the reference-facing `render(target, scene)` call.

`render` mirrors `pub fn render(target: &mut RenderTarget, scene: &Scene)` (lib/core/render.rs:61):
it lowers the scene (the reference rebuilds its BVH on every call, render.rs:67), renders on the
GPU and fills `target.data` with RGBA8.  There is NO CPU fallback: if the CUDA library is missing
or no device is usable this raises.
"""
import ctypes as C
import os

import numpy as np

from . import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
# RTR_LIBRARY: load another build of the same ABI (same-box A/B of kernel variants, tools/ab.py)
LIB_PATH = os.environ.get("RTR_LIBRARY") or os.path.join(_HERE, "librtr_b200.so")
_LIB = None


class RtrError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"{A.STATUS_NAMES[status] if 0 <= status < len(A.STATUS_NAMES) else status}: {message}")
        self.status = status


def lib():
    """Load the CUDA library (built in-tree by `make -C raytracingrust_b200/csrc` / __graft_entry__.build)."""
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `make -C {_HERE}/csrc` "
                          "(there is no CPU fallback for the render path)")
    L = C.CDLL(LIB_PATH)
    vp, u32, u64, sz = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t
    L.rtr_abi_version.restype = u32
    L.rtr_last_error.restype = C.c_char_p
    L.rtr_last_error.argtypes = [vp]
    L.rtr_context_create.argtypes = [C.c_int, vp, C.POINTER(vp)]
    L.rtr_context_destroy.argtypes = [vp]
    L.rtr_context_destroy.restype = None
    L.rtr_scene_upload.argtypes = [vp, C.POINTER(A.SceneDesc), C.POINTER(vp)]
    L.rtr_scene_destroy.argtypes = [vp]
    L.rtr_scene_destroy.restype = None
    L.rtr_scene_info.argtypes = [vp, C.POINTER(u64), C.POINTER(u64), C.POINTER(C.c_double), C.POINTER(u32), C.POINTER(u32)]
    L.rtr_render.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp, C.POINTER(A.Stats)]
    L.rtr_render_accumulate_device.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, C.POINTER(A.Stats)]
    L.rtr_resolve_device.argtypes = [vp, vp, u32, u32, u32, vp]
    L.rtr_synchronize.argtypes = [vp]
    L.rtr_trace_batch.argtypes = [vp, vp, vp, sz, C.c_float, u32, u32, u64, vp, C.POINTER(A.Stats)]
    L.rtr_trace_batch_device_noalloc.argtypes = [vp, vp, vp, sz, C.c_float, u32, u32, u64, vp]
    L.rtr_scatter_batch.argtypes = [vp, vp, vp, sz, u32, u64, vp]
    L.rtr_background_batch.argtypes = [vp, vp, vp, sz, vp]
    L.rtr_camera_rays.argtypes = [vp, C.POINTER(A.RenderParams), vp, vp, sz, vp]
    L.rtr_group_create.argtypes = [C.POINTER(C.c_int), u32, u32, C.POINTER(vp)]
    L.rtr_group_destroy.argtypes = [vp]
    L.rtr_group_destroy.restype = None
    L.rtr_group_size.argtypes = [vp]
    L.rtr_group_size.restype = u32
    L.rtr_group_reduce_mode.argtypes = [vp]
    L.rtr_group_reduce_mode.restype = u32
    L.rtr_group_last_error.argtypes = [vp]
    L.rtr_group_last_error.restype = C.c_char_p
    L.rtr_group_scene_upload.argtypes = [vp, C.POINTER(A.SceneDesc), C.POINTER(vp)]
    L.rtr_group_scene_destroy.argtypes = [vp]
    L.rtr_group_scene_destroy.restype = None
    L.rtr_group_render.argtypes = [vp, vp, C.POINTER(A.RenderParams), vp, vp, C.POINTER(A.Stats)]
    L.rtr_group_device_stats.argtypes = [vp, u32, C.POINTER(A.Stats), C.POINTER(u32), C.POINTER(u32)]
    L.rtr_group_set_split.argtypes = [vp, u32, u32]
    L.rtr_scene_load_json.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.rtr_host_scene_desc.argtypes = [vp]
    L.rtr_host_scene_desc.restype = C.POINTER(A.SceneDesc)
    L.rtr_host_scene_camera.argtypes = [vp]
    L.rtr_host_scene_camera.restype = C.POINTER(A.CameraS)
    L.rtr_host_scene_settings.argtypes = [vp]
    L.rtr_host_scene_settings.restype = C.POINTER(A.SettingsS)
    L.rtr_host_scene_destroy.argtypes = [vp]
    L.rtr_host_scene_destroy.restype = None
    L.rtr_mesh_load_obj.argtypes = [C.c_char_p, u32, C.POINTER(vp)]
    L.rtr_host_mesh_get.argtypes = [vp]
    L.rtr_host_mesh_get.restype = C.POINTER(A.MeshS)
    L.rtr_host_mesh_destroy.argtypes = [vp]
    L.rtr_host_mesh_destroy.restype = None
    L.rtr_write_png.argtypes = [C.c_char_p, vp, u32, u32]
    L.rtr_write_exr.argtypes = [C.c_char_p, vp, u32, u32, u32]
    L.rtr_image_load.argtypes = [C.c_char_p, C.POINTER(vp)]
    L.rtr_host_image_width.argtypes = [vp]
    L.rtr_host_image_width.restype = u32
    L.rtr_host_image_height.argtypes = [vp]
    L.rtr_host_image_height.restype = u32
    L.rtr_host_image_rgb.argtypes = [vp]
    L.rtr_host_image_rgb.restype = C.POINTER(C.c_float)
    L.rtr_host_image_destroy.argtypes = [vp]
    L.rtr_host_image_destroy.restype = None
    if L.rtr_abi_version() != A.RTR_ABI_VERSION:
        raise ImportError("librtr_b200.so ABI version mismatch")
    _LIB = L
    return L


def _check(rc, ctx=None):
    if rc != A.RTR_OK:
        msg = lib().rtr_last_error(ctx)
        raise RtrError(rc, msg.decode() if msg else "")


def make_params(scene, width, height, seed=0, sample_begin=0, sample_count=0, kernel=A.KERNEL_AUTO, flags=0,
                settings=None, camera=None):
    p = A.RenderParams()
    p.width, p.height = width, height
    p.camera = (camera or scene.camera).to_c()
    p.settings = (settings or scene.settings).to_c()
    p.seed, p.sample_begin, p.sample_count, p.kernel, p.flags = seed, sample_begin, sample_count, kernel, flags
    return p


# ---------------------------------------------------------------------- host loaders
def load_obj(path):
    """Mesh::from_file's loader (mesh.rs:63-76): returns (positions Nx3, normals Nx3, indices)."""
    h = C.c_void_p()
    _check(lib().rtr_mesh_load_obj(os.fsencode(path), 0, C.byref(h)))
    try:
        m = lib().rtr_host_mesh_get(h).contents
        nv, ni = m.n_vertices, m.n_indices
        pos = np.ctypeslib.as_array(m.positions, (nv, 3)).copy() if nv else np.zeros((0, 3), np.float32)
        nrm = np.ctypeslib.as_array(m.normals, (nv, 3)).copy() if nv else np.zeros((0, 3), np.float32)
        idx = np.ctypeslib.as_array(m.indices, (ni,)).copy() if ni else np.zeros(0, np.uint32)
    finally:
        lib().rtr_host_mesh_destroy(h)
    return pos, nrm, idx


class _LoadedScene:
    """A Scene whose description lives in the library's JSON loader (Scene::from_file)."""

    def __init__(self, path):
        from .scene import Camera, RenderSettings
        self._h = C.c_void_p()
        _check(lib().rtr_scene_load_json(os.fsencode(path), C.byref(self._h)))
        self.camera = Camera.from_c(lib().rtr_host_scene_camera(self._h).contents)
        self.settings = RenderSettings.from_c(lib().rtr_host_scene_settings(self._h).contents)

    def lower(self):
        from .scene import LoweredScene
        out = LoweredScene()
        out.desc = lib().rtr_host_scene_desc(self._h).contents
        out._keep.append(self)
        return out

    def __del__(self):
        if getattr(self, "_h", None):
            lib().rtr_host_scene_destroy(self._h)
            self._h = None


def load_scene_json(path):
    return _LoadedScene(path)


def write_png(path, rgba):
    """Headless stand-in for the window blit of RenderTarget.data (src/main.rs:92-97): RGBA8 (H, W, 4) -> PNG file."""
    rgba = np.ascontiguousarray(rgba, np.uint8)
    if rgba.ndim != 3 or rgba.shape[2] != 4:
        raise ValueError("expected an (H, W, 4) uint8 image")
    _check(lib().rtr_write_png(os.fsencode(path), rgba.ctypes.data, rgba.shape[1], rgba.shape[0]))


def write_exr(path, image):
    """fp32 (H, W, 3|4) image -> scan-line OpenEXR, float channels, ZIP blocks (rtr_write_exr)."""
    image = np.ascontiguousarray(image, np.float32)
    if image.ndim != 3 or image.shape[2] not in (3, 4):
        raise ValueError("expected an (H, W, 3) or (H, W, 4) float image")
    _check(lib().rtr_write_exr(os.fsencode(path), image.ctypes.data, image.shape[1], image.shape[0], image.shape[2]))


def load_image(path):
    """SkyMap::new's file read (backgrounds.rs:59-73): OpenEXR or Radiance HDR -> (H, W, 3) float32, row 0 = top."""
    h = C.c_void_p()
    _check(lib().rtr_image_load(os.fsencode(path), C.byref(h)))
    try:
        L = lib()
        W, H = L.rtr_host_image_width(h), L.rtr_host_image_height(h)
        return np.ctypeslib.as_array(L.rtr_host_image_rgb(h), shape=(H, W, 3)).copy()
    finally:
        lib().rtr_host_image_destroy(h)


# ---------------------------------------------------------------------- device objects
class Context:
    """One per GPU / process rank (rtr_context)."""

    def __init__(self, device=0, stream=None):
        self._h = C.c_void_p()
        _check(lib().rtr_context_create(device, stream, C.byref(self._h)))
        self.device = device
        # the caller's cudaStream_t as an integer (None: the library owns a private non-blocking stream)
        self.stream_handle = None if stream is None else (stream.value if hasattr(stream, "value") else int(stream))

    def close(self):
        if getattr(self, "_h", None):
            lib().rtr_context_destroy(self._h)
            self._h = None

    __del__ = close

    def upload(self, scene):
        return DeviceScene(self, scene)

    def synchronize(self):
        _check(lib().rtr_synchronize(self._h), self._h)

    def camera_rays(self, params, pixels, samples):
        pixels = np.ascontiguousarray(pixels, np.uint32)
        samples = np.ascontiguousarray(samples, np.uint32)
        rays = np.zeros(len(pixels), A.RAY_DTYPE)
        _check(lib().rtr_camera_rays(self._h, C.byref(params), pixels.ctypes.data, samples.ctypes.data, len(pixels),
                                     rays.ctypes.data), self._h)
        return rays

    def resolve_device(self, accum_ptr, width, height, spp, rgba_ptr):
        _check(lib().rtr_resolve_device(self._h, accum_ptr, width, height, spp, rgba_ptr), self._h)


class DeviceScene:
    """Flattened scene resident in HBM (rtr_scene)."""

    def __init__(self, ctx, scene):
        self.ctx = ctx
        lowered = scene.lower() if hasattr(scene, "lower") else scene
        self._h = C.c_void_p()
        _check(lib().rtr_scene_upload(ctx._h, C.byref(lowered.desc), C.byref(self._h)), ctx._h)

    def close(self):
        if getattr(self, "_h", None) and getattr(self.ctx, "_h", None):
            lib().rtr_scene_destroy(self._h)
        self._h = None

    __del__ = close

    def info(self):
        db, hb, ms, nn, md = C.c_uint64(), C.c_uint64(), C.c_double(), C.c_uint32(), C.c_uint32()
        _check(lib().rtr_scene_info(self._h, C.byref(db), C.byref(hb), C.byref(ms), C.byref(nn), C.byref(md)))
        return dict(device_bytes=db.value, h2d_bytes=hb.value, build_ms=ms.value, n_nodes=nn.value, max_depth=md.value)

    def render(self, params, want_rgba=True, want_accum=False):
        W, H = params.width, params.height
        rgba = np.zeros((H, W, 4), np.uint8) if want_rgba else None
        acc = np.zeros((H, W, 4), np.float32) if want_accum else None
        st = A.Stats()
        _check(lib().rtr_render(self.ctx._h, self._h, C.byref(params), rgba.ctypes.data if want_rgba else None,
                                acc.ctypes.data if want_accum else None, C.byref(st)), self.ctx._h)
        return rgba, acc, st.as_dict()

    def render_into(self, params, rgba_out, accum_out=None):
        """Same, into caller-owned host buffers (e.g. pinned torch tensors' data_ptr)."""
        st = A.Stats()
        _check(lib().rtr_render(self.ctx._h, self._h, C.byref(params), rgba_out, accum_out, C.byref(st)), self.ctx._h)
        return st.as_dict()

    def render_accumulate_device(self, params, accum_ptr, want_stats=True):
        st = A.Stats()
        _check(lib().rtr_render_accumulate_device(self.ctx._h, self._h, C.byref(params), accum_ptr,
                                                  C.byref(st) if want_stats else None), self.ctx._h)
        return st.as_dict() if want_stats else None

    def trace_batch(self, rays, tmin=1e-5, enable_bvh_tree=True, flags=0, seed=0, want_stats=False):
        rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
        hits = np.zeros(len(rays), A.HIT_DTYPE)
        st = A.Stats()
        _check(lib().rtr_trace_batch(self.ctx._h, self._h, rays.ctypes.data, len(rays), tmin, int(enable_bvh_tree),
                                     flags, seed, hits.ctypes.data, C.byref(st)), self.ctx._h)
        return (hits, st.as_dict()) if want_stats else hits

    def trace_batch_into(self, rays_ptr, n, hits_ptr, tmin=1e-5, enable_bvh_tree=True, flags=0, seed=0):
        """Same, with caller-owned HOST buffers (e.g. pinned): rays_ptr -> n rtr_ray, hits_ptr -> n rtr_hit.  Returns stats."""
        st = A.Stats()
        _check(lib().rtr_trace_batch(self.ctx._h, self._h, rays_ptr, n, tmin, int(enable_bvh_tree), flags, seed, hits_ptr,
                                     C.byref(st)), self.ctx._h)
        return st.as_dict()

    def scatter_batch(self, inp, mode=A.MODE_FULL, seed=0):
        inp = np.ascontiguousarray(inp, A.SCATTER_IN_DTYPE)
        out = np.zeros(len(inp), A.SCATTER_OUT_DTYPE)
        _check(lib().rtr_scatter_batch(self.ctx._h, self._h, inp.ctypes.data, len(inp), mode, seed, out.ctypes.data),
               self.ctx._h)
        return out

    def background_batch(self, rays):
        rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
        out = np.zeros((len(rays), 3), np.float32)
        _check(lib().rtr_background_batch(self.ctx._h, self._h, rays.ctypes.data, len(rays), out.ctypes.data), self.ctx._h)
        return out


class Group:
    """N GPUs of this process rendering ONE image (rtr_group): the scene is replicated, the samples of every pixel are
    split over the devices, one cross-device reduce (the library's NVLink peer kernel, or ncclReduce) + resolve."""

    def __init__(self, devices, reduce=A.REDUCE_AUTO):
        devices = list(devices)
        arr = (C.c_int * len(devices))(*devices)
        self._h = C.c_void_p()
        _check(lib().rtr_group_create(arr, len(devices), reduce, C.byref(self._h)))
        self.devices = devices

    def close(self):
        if getattr(self, "_h", None):
            lib().rtr_group_destroy(self._h)
            self._h = None

    __del__ = close

    @property
    def reduce_mode(self):
        return {A.REDUCE_NCCL: "nccl", A.REDUCE_P2P: "p2p"}[lib().rtr_group_reduce_mode(self._h)]

    def set_split(self, split, chunk=0):
        _check(lib().rtr_group_set_split(self._h, split, chunk))

    def upload(self, scene):
        return GroupScene(self, scene)

    def device_stats(self, i):
        st, b, c = A.Stats(), C.c_uint32(), C.c_uint32()
        _check(lib().rtr_group_device_stats(self._h, i, C.byref(st), C.byref(b), C.byref(c)))
        d = st.as_dict()
        d.update(sample_begin=b.value, sample_count=c.value)
        return d


class GroupScene:
    """The scene flattened once and resident on every device of a Group (rtr_group_scene)."""

    def __init__(self, group, scene):
        self.group = group
        lowered = scene.lower() if hasattr(scene, "lower") else scene
        self._h = C.c_void_p()
        rc = lib().rtr_group_scene_upload(group._h, C.byref(lowered.desc), C.byref(self._h))
        if rc != A.RTR_OK:
            raise RtrError(rc, (lib().rtr_group_last_error(group._h) or b"").decode())

    def close(self):
        if getattr(self, "_h", None) and getattr(self.group, "_h", None):
            lib().rtr_group_scene_destroy(self._h)
        self._h = None

    __del__ = close

    def _render(self, params, rgba_ptr, accum_ptr):
        st = A.Stats()
        rc = lib().rtr_group_render(self.group._h, self._h, C.byref(params), rgba_ptr, accum_ptr, C.byref(st))
        if rc != A.RTR_OK:
            raise RtrError(rc, (lib().rtr_group_last_error(self.group._h) or b"").decode())
        return st.as_dict()

    def render(self, params, want_rgba=True, want_accum=False):
        W, H = params.width, params.height
        rgba = np.zeros((H, W, 4), np.uint8) if want_rgba else None
        acc = np.zeros((H, W, 4), np.float32) if want_accum else None
        st = self._render(params, rgba.ctypes.data if want_rgba else None, acc.ctypes.data if want_accum else None)
        return rgba, acc, st

    def render_into(self, params, rgba_out, accum_out=None):
        """Same, into caller-owned host buffers (e.g. pinned torch tensors' data_ptr)."""
        return self._render(params, rgba_out, accum_out)


_DEFAULT_CTX = {}
_DEFAULT_GROUP = {}


def default_context(device=0):
    if device not in _DEFAULT_CTX:
        _DEFAULT_CTX[device] = Context(device)
    return _DEFAULT_CTX[device]


def render(target, scene, seed=0, device=0, kernel=A.KERNEL_AUTO, want_accum=False, devices=None):
    """render(&mut target, &scene) (lib/core/render.rs:61-127) on the GPU; devices=[...] splits the samples of the
    one image over several GPUs of this process (rtr_group_render)."""
    target.request_redraw = True  # render.rs:62
    if devices is not None and len(devices) > 1:
        key = tuple(devices)
        if key not in _DEFAULT_GROUP:
            _DEFAULT_GROUP[key] = Group(devices)
        gs = _DEFAULT_GROUP[key].upload(scene)  # the reference rebuilds the BVH on every call (render.rs:67)
        try:
            params = make_params(scene, target.width, target.height, seed=seed, kernel=kernel)
            rgba, acc, stats = gs.render(params, want_rgba=True, want_accum=want_accum)
        finally:
            gs.close()
        target.data[:] = rgba.reshape(-1)
        target.accum = acc
        target.stats = stats
        return stats
    ctx = default_context(device if devices is None else devices[0])
    dev_scene = ctx.upload(scene)  # the reference rebuilds the BVH on every call (render.rs:67)
    try:
        params = make_params(scene, target.width, target.height, seed=seed, kernel=kernel)
        rgba, acc, stats = dev_scene.render(params, want_rgba=True, want_accum=want_accum)
    finally:
        dev_scene.close()
    target.data[:] = rgba.reshape(-1)
    target.accum = acc
    target.stats = stats
    return stats
