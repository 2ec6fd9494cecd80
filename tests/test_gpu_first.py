"""First GPU parity checks: device closest-hit / render vs the CPU oracle, through the C ABI."""
import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes

pytestmark = pytest.mark.gpu


def compare_hits(dev, ora, what):
    assert np.array_equal(dev["object"], ora["object"]), f"{what}: object ids differ at {np.flatnonzero(dev['object'] != ora['object'])[:10]}"
    assert np.array_equal(dev["primitive"], ora["primitive"]), f"{what}: primitive ids differ"
    hit = ora["object"] != A.RTR_NO_HIT
    for f in ("t", "point", "normal"):
        a, b = dev[f][hit], ora[f][hit]
        assert np.array_equal(a, b), f"{what}: {f} not bit-identical, max abs diff {np.abs(a - b).max()}"
    assert np.array_equal(dev["front_face"][hit], ora["front_face"][hit])
    assert np.array_equal(dev["material"][hit], ora["material"][hit])


@pytest.mark.parametrize("name", ["simple", "benchmark", "mesh", "grid"])
def test_trace_batch_matches_oracle(ctx, name):
    scene = {"simple": _scenes.simple, "benchmark": _scenes.benchmark, "mesh": _scenes.mesh_bench,
             "grid": lambda: _scenes.grid_spheres(5)}[name]()
    rays = np.concatenate([_scenes.camera_rays_np(scene, 96, 64), _scenes.random_rays(4096, 3)])
    ds, os_ = ctx.upload(scene), orc.OracleScene(scene)
    for bvh in (True, False):
        if name == "grid" and not bvh:
            continue
        ora = os_.trace_batch(rays, enable_bvh_tree=bvh)
        compare_hits(ds.trace_batch(rays, enable_bvh_tree=bvh), ora, f"{name} bvh={bvh} culled")
        compare_hits(ds.trace_batch(rays, enable_bvh_tree=bvh, flags=A.FLAG_EXHAUSTIVE), ora, f"{name} bvh={bvh} exhaustive")


@pytest.mark.parametrize("name", ["simple", "benchmark"])
def test_render_matches_oracle(ctx, name):
    scene = getattr(_scenes, name)()
    p = rt.make_params(scene, 100, 100, seed=1)
    rgba_o, acc_o, st_o = orc.OracleScene(scene).render(p)
    rgba_d, acc_d, st_d = ctx.upload(scene).render(p, want_accum=True)
    assert st_d["rays"] == st_o["rays"]
    # iterative throughput product vs the reference's recursive product: few-ulp differences
    np.testing.assert_allclose(acc_d, acc_o, rtol=1e-5, atol=1e-6)
    assert np.abs(rgba_d.astype(int) - rgba_o.astype(int)).max() <= 1
