"""The CUDA path against the committed second-opinion vectors (tests/golden/vectors/*.npz, written by the independent
numpy restatement tests/golden/make_vectors.py): level-1 hits and level-2 shading routines through the C ABI, bit for
bit except where libm enters (tolerances in tests/_vectors.py).  AABB::hit has no entry point of its own on the device;
its folded form is exercised against the oracle by the trace tests."""
import pytest

import _vectors as V

pytestmark = pytest.mark.gpu


class DeviceImpl:
    def __init__(self, ctx):
        self.ctx = ctx

    def trace(self, scene, rays, tmin, seed):
        return self.ctx.upload(scene).trace_batch(rays, tmin=tmin, enable_bvh_tree=True, seed=seed)

    def scatter(self, scene, inp, seed):
        return self.ctx.upload(scene).scatter_batch(inp, seed=seed)

    def background(self, scene, rays):
        return self.ctx.upload(scene).background_batch(rays)

    def camera(self, params, pixels, samples):
        return self.ctx.camera_rays(params, pixels, samples)


@pytest.mark.parametrize("what", ["sphere", "triangle", "volume", "scatter", "background", "camera"])
def test_device_matches_second_opinion(ctx, what):
    n = getattr(V, "run_" + what)(DeviceImpl(ctx))
    assert n > 1000, (what, n)
