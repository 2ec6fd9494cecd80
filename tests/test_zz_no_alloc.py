"""rtr_trace_batch_device_noalloc (include/rtr.h): the allocation-free, device-pointer variant of rtr_trace_batch.

Collected last on purpose.  Checks, through the C ABI, that the new entry point
returns exactly what rtr_trace_batch returns (the hit records are bit-identical: same kernels, same arithmetic — the
tolerance the level-1 parity tests use for device-vs-device comparisons is zero) on two calls in a row that share the
library's static scratch array, that device memory does not shrink across a call, and that oversize batches are refused.
Device buffers for the rays and hits are the TEST's (torch); the entry point itself must not allocate."""
import ctypes as C

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
import _scenes

pytestmark = pytest.mark.gpu


def noalloc_trace(ctx, ds, rays, tmin=1e-5, enable_bvh_tree=True, flags=0, seed=0):
    """Rays up (test-owned device buffers), the entry point, synchronise, hit records back."""
    import torch
    rays = np.ascontiguousarray(rays, A.RAY_DTYPE)
    n = len(rays)
    dev = torch.device("cuda", ctx.device)
    rays_dev = torch.from_numpy(rays.view(np.uint8).reshape(-1).copy()).to(dev)
    hits_dev = torch.full((n * A.HIT_DTYPE.itemsize,), 0xAB, dtype=torch.uint8, device=dev)  # poison: every record must be written
    torch.cuda.synchronize(dev)  # the buffers are settled before the library's own stream touches them
    free_before = torch.cuda.mem_get_info(dev)[0]
    rc = rt.lib().rtr_trace_batch_device_noalloc(ctx._h, ds._h, rays_dev.data_ptr(), n, tmin, int(enable_bvh_tree), flags, seed,
                                                 hits_dev.data_ptr())
    assert rc == A.RTR_OK, rt.lib().rtr_last_error(ctx._h)
    ctx.synchronize()
    free_after = torch.cuda.mem_get_info(dev)[0]
    hits = hits_dev.cpu().numpy().view(A.HIT_DTYPE)
    return hits, free_before, free_after


def test_noalloc_trace_matches_trace_batch_two_calls_in_a_row():
    ctx = rt.Context(0)
    try:
        scene = _scenes.mesh_hdri(64, 32)
        ds = ctx.upload(scene)
        rays1 = _scenes.random_rays(60_000, seed=1, scale=4.0)
        rays2 = _scenes.random_rays(20_001, seed=2, scale=6.0)  # fewer rays: stale slots / counters of call 1 must not matter
        want1, want2 = ds.trace_batch(rays1), ds.trace_batch(rays2)
        got1, _, _ = noalloc_trace(ctx, ds, rays1)          # first call on this device: looks the scratch address up
        got2, before, after = noalloc_trace(ctx, ds, rays2)  # second call: shares the array, allocates nothing
        for got, want, what in ((got1, want1, "call 1"), (got2, want2, "call 2")):
            assert (want["object"] != A.RTR_NO_HIT).mean() > 0.05, what
            assert got.tobytes() == want.tobytes(), f"{what}: hit records differ from rtr_trace_batch"
        assert after >= before, f"device memory shrank across the call: {before} -> {after}"
    finally:
        ctx.close()


def test_noalloc_trace_flags_tmin_and_linear_scan():
    ctx = rt.Context(0)
    try:
        scene = _scenes.volumes(64, 32)  # volumes: the generic wf_extend instantiation, nested boundary walks
        ds = ctx.upload(scene)
        rays = _scenes.random_rays(30_000, seed=3, scale=3.0)
        for kw in (dict(), dict(flags=A.FLAG_EXHAUSTIVE), dict(flags=A.FLAG_REFERENCE_TREE), dict(tmin=1e-3),
                   dict(enable_bvh_tree=False), dict(seed=7)):
            got, _, _ = noalloc_trace(ctx, ds, rays, **kw)
            assert got.tobytes() == ds.trace_batch(rays, **kw).tobytes(), kw
    finally:
        ctx.close()


def test_noalloc_trace_rejects_oversize_and_bad_flags():
    ctx = rt.Context(0)
    try:
        ds = ctx.upload(_scenes.simple())
        L = rt.lib()
        fake = C.c_void_p(256)  # never dereferenced: the size check comes first
        rc = L.rtr_trace_batch_device_noalloc(ctx._h, ds._h, fake, A.NOALLOC_MAX_RAYS + 1, 1e-5, 1, 0, 0, fake)
        assert rc == A.RTR_INVALID_ARG and b"RTR_NOALLOC_MAX_RAYS" in L.rtr_last_error(ctx._h)
        rc = L.rtr_trace_batch_device_noalloc(ctx._h, ds._h, fake, 16, 1e-5, 1, A.FLAG_COUNT_WORK, 0, fake)
        assert rc == A.RTR_INVALID_ARG
        assert L.rtr_trace_batch_device_noalloc(ctx._h, ds._h, None, 0, 1e-5, 1, 0, 0, None) == A.RTR_OK  # empty batch
        # the largest batch the array holds
        rays = _scenes.random_rays(A.NOALLOC_MAX_RAYS, seed=4, scale=2.0)
        got, _, _ = noalloc_trace(ctx, ds, rays)
        assert got.tobytes() == ds.trace_batch(rays).tobytes()
    finally:
        ctx.close()
