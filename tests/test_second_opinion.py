"""The oracle against a SECOND, independent statement of the reference's arithmetic: tests/golden/vectors/*.npz hold
inputs, Philox words and outputs computed by tests/golden/make_vectors.py, a float32 numpy restatement written from
the .rs files (objects.rs:23-50, mesh.rs:147-196, aabb.rs:81-137, volume.rs:35-72, materials.rs:31-156,
backgrounds.rs:27-92, camera.rs:19-56) without reference to oracle/oracle.cpp.  Bit for bit, except where libm enters
(ln in Volume::hit, acos/atan2 in SkyMap, tan in the camera), where the tolerance is stated in tests/_vectors.py.
The reference itself ships no vectors and cannot be built here, so parity stays "unpinned"; this removes the
single-transcription risk."""
import os
import sys

import numpy as np
import pytest

import raytracingrust_b200 as rt
from oracle import orc
import _vectors as V

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden"))


class OracleImpl:
    def trace(self, scene, rays, tmin, seed):
        return orc.OracleScene(scene).trace_batch(rays, tmin=tmin, enable_bvh_tree=True, seed=seed)

    def scatter(self, scene, inp, seed):
        return orc.OracleScene(scene).scatter_batch(inp, seed=seed)

    def background(self, scene, rays):
        return orc.OracleScene(scene).background_batch(rays)

    def camera(self, params, pixels, samples):
        return orc.camera_rays(params, pixels, samples)

    def aabb(self, boxes, rays, tmin, tmax):
        return orc.aabb_hit_batch(boxes, rays, tmin, tmax)


@pytest.mark.parametrize("what", ["aabb", "sphere", "triangle", "volume", "scatter", "background", "camera"])
def test_oracle_matches_second_opinion(what):
    n = getattr(V, "run_" + what)(OracleImpl())
    assert n > 1000, (what, n)  # the vectors exercise the routine, not just its early-outs


def test_second_opinion_philox_known_answers():
    """The numpy Philox4x32-10 (Random123 known-answer vectors) and its agreement with the oracle's generator."""
    import make_vectors as M
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0), (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        assert tuple(int(x) for x in M.philox4x32_10(np.array([ctr], np.uint32), key)[0]) == want
    rng = np.random.default_rng(0)
    for _ in range(64):
        ctr, key = rng.integers(0, 1 << 32, 4, dtype=np.uint64), rng.integers(0, 1 << 32, 2, dtype=np.uint64)
        assert np.array_equal(M.philox4x32_10(ctr[None, :].astype(np.uint32), tuple(int(k) for k in key))[0], orc.philox(ctr, key))


def test_vectors_are_reproducible():
    """The committed files are what the committed generator writes (spot check: the cheapest generator)."""
    import make_vectors as M
    fresh = M.gen_camera(np.random.default_rng(sum(map(ord, "camera"))))
    z = V.load("camera")
    for k, v in fresh.items():
        assert np.array_equal(np.asarray(v), z[k]), k
