"""ctypes wrapper of oracle/liboracle.so — TEST INFRASTRUCTURE ONLY (see oracle.cpp header).

Imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg.
Never imported by the raytracingrust_b200 package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

from raytracingrust_b200 import _abi as A

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    so = os.path.join(_HERE, "liboracle.so")
    src = os.path.join(_HERE, "oracle.cpp")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "liboracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.orc_build_info.restype = C.c_char_p
        L.orc_scene_create.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
        L.orc_scene_destroy.argtypes = [C.c_void_p]
        L.orc_trace_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_uint32, C.c_uint64, C.c_void_p]
        L.orc_count_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_uint64, C.c_void_p, C.c_void_p]
        L.orc_scatter_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_uint32, C.c_uint64, C.c_void_p]
        L.orc_background_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_camera_rays.argtypes = [C.POINTER(A.RenderParams), C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.orc_render.argtypes = [C.c_void_p, C.POINTER(A.RenderParams), C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                 C.POINTER(A.Stats)]
        L.orc_philox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_aabb_hit_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float, C.c_void_p]
        _LIB = L
    return _LIB


def philox(ctr, key):
    c = np.asarray(ctr, np.uint32)
    k = np.asarray(key, np.uint32)
    out = np.zeros(4, np.uint32)
    lib().orc_philox(c.ctypes.data, k.ctypes.data, out.ctypes.data)
    return out


def _check(rc, what):
    if rc != A.RTR_OK:
        raise RuntimeError(f"oracle {what} failed: {A.STATUS_NAMES[rc]}")


class OracleScene:
    """CPU scene built from the same rtr_scene_desc the product consumes."""

    def __init__(self, scene):
        """scene: raytracingrust_b200.scene.Scene (lowered here) or a LoweredScene."""
        lowered = scene.lower() if hasattr(scene, "lower") else scene
        self._lowered = lowered
        h = C.c_void_p()
        err = C.create_string_buffer(256)
        rc = lib().orc_scene_create(C.byref(lowered.desc), C.byref(h), err, 256)
        if rc != A.RTR_OK:
            raise ValueError(f"oracle scene: {err.value.decode()}")
        self._h = h

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_scene_destroy(self._h)
            self._h = None

    def trace_batch(self, rays, tmin=1e-5, enable_bvh_tree=True, seed=0):
        rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
        hits = np.zeros(len(rays), A.HIT_DTYPE)
        _check(lib().orc_trace_batch(self._h, rays.ctypes.data, len(rays), tmin, int(enable_bvh_tree), seed,
                                     hits.ctypes.data), "trace_batch")
        return hits

    def count_batch(self, rays, tmin=1e-5, seed=0):
        """Ordered, t-culled walk counts: dict(rays,nodes,tris,spheres) and closest t per ray."""
        rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
        counts = np.zeros(4, np.uint64)
        t = np.zeros(len(rays), np.float32)
        _check(lib().orc_count_batch(self._h, rays.ctypes.data, len(rays), tmin, seed, counts.ctypes.data,
                                     t.ctypes.data), "count_batch")
        return dict(rays=int(counts[0]), nodes=int(counts[1]), tris=int(counts[2]), spheres=int(counts[3])), t

    def scatter_batch(self, inp, mode=A.MODE_FULL, seed=0):
        inp = np.ascontiguousarray(inp, A.SCATTER_IN_DTYPE)
        out = np.zeros(len(inp), A.SCATTER_OUT_DTYPE)
        _check(lib().orc_scatter_batch(self._h, inp.ctypes.data, len(inp), mode, seed, out.ctypes.data), "scatter_batch")
        return out

    def background_batch(self, rays):
        rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
        out = np.zeros((len(rays), 3), np.float32)
        _check(lib().orc_background_batch(self._h, rays.ctypes.data, len(rays), out.ctypes.data), "background_batch")
        return out

    def render(self, params, threads=0, count_work=False, want_rgba=True, want_accum=True):
        """params: A.RenderParams.  Returns (rgba8 HxWx4 | None, accum HxWx4 | None, stats dict)."""
        W, H = params.width, params.height
        rgba = np.zeros((H, W, 4), np.uint8) if want_rgba else None
        acc = np.zeros((H, W, 4), np.float32) if want_accum else None
        st = A.Stats()
        _check(lib().orc_render(self._h, C.byref(params), threads, int(count_work),
                                rgba.ctypes.data if want_rgba else None,
                                acc.ctypes.data if want_accum else None, C.byref(st)), "render")
        return rgba, acc, st.as_dict()


def aabb_hit_batch(boxes, rays, tmin, tmax):
    """AABB::hit (aabb.rs:81-137) for n (box, ray) pairs; boxes (n, 6) = min.xyz, max.xyz."""
    boxes = np.ascontiguousarray(boxes, np.float32)
    rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
    out = np.zeros(len(rays), np.uint8)
    _check(lib().orc_aabb_hit_batch(boxes.ctypes.data, rays.ctypes.data, len(rays), tmin, tmax, out.ctypes.data), "aabb_hit_batch")
    return out.astype(bool)


def camera_rays(params, pixels, samples):
    pixels = np.ascontiguousarray(pixels, np.uint32)
    samples = np.ascontiguousarray(samples, np.uint32)
    rays = np.zeros(len(pixels), A.RAY_DTYPE)
    _check(lib().orc_camera_rays(C.byref(params), pixels.ctypes.data, samples.ctypes.data, len(pixels),
                                 rays.ctypes.data), "camera_rays")
    return rays


def make_params(scene, width, height, seed=0, sample_begin=0, sample_count=0, kernel=A.KERNEL_AUTO, flags=0,
                settings=None, camera=None):
    p = A.RenderParams()
    p.width, p.height = width, height
    p.camera = (camera or scene.camera).to_c()
    p.settings = (settings or scene.settings).to_c()
    p.seed, p.sample_begin, p.sample_count, p.kernel, p.flags = seed, sample_begin, sample_count, kernel, flags
    return p
