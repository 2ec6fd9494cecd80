"""The wavefront variant (generate / extend / shade+compact, wavefront.cu) against the megakernel and the oracle.

Both kernels run the same per-path arithmetic with the same Philox counters and sum each pixel's samples in sample
order, so wherever the megakernel does not chunk its samples the two accumulation buffers must be BIT-identical."""
import os

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes
from _parity import assert_images_close
from test_gpu_render import CASES

pytestmark = pytest.mark.gpu


def wf(scene, W, H, **kw):
    return rt.make_params(scene, W, H, kernel=A.KERNEL_WAVEFRONT, **kw)


@pytest.mark.parametrize("name", sorted(CASES))
def test_wavefront_matches_oracle_same_stream(ctx, name):
    factory, W, H, spp, off = CASES[name]
    scene = factory()
    scene.settings.samples_per_pixel = spp
    rgba_o, acc_o, st_o = orc.OracleScene(scene).render(rt.make_params(scene, W, H, seed=1))
    rgba_d, acc_d, st_d = ctx.upload(scene).render(wf(scene, W, H, seed=1), want_accum=True)
    assert st_d["kernel_used"] == A.KERNEL_WAVEFRONT and st_d["samples"] == W * H * spp
    assert_images_close(rgba_d, acc_d, rgba_o, acc_o, name, frac_pixels_off=off)
    if off == 0.0:
        assert st_d["rays"] == st_o["rays"]


@pytest.mark.parametrize("name,W,H,spp", [("simple", 1024, 768, 3), ("mesh_hdri", 1024, 768, 2), ("volumes", 1024, 768, 2),
                                          ("benchmark", 1024, 768, 5)])
def test_wavefront_bit_identical_to_megakernel(ctx, name, W, H, spp):
    scene = _scenes.mesh_hdri(256, 128) if name == "mesh_hdri" else _scenes.volumes(128, 64) if name == "volumes" \
        else getattr(_scenes, name)()
    scene.settings.samples_per_pixel = spp
    ds = ctx.upload(scene)
    rgba_m, acc_m, st_m = ds.render(rt.make_params(scene, W, H, seed=9, kernel=A.KERNEL_MEGAKERNEL, flags=A.FLAG_COUNT_WORK),
                                    want_accum=True)
    rgba_w, acc_w, st_w = ds.render(wf(scene, W, H, seed=9, flags=A.FLAG_COUNT_WORK), want_accum=True)
    assert st_m["kernel_launches"] == 2, "the megakernel chunked its samples: pick a larger image for this test"
    assert np.array_equal(acc_m, acc_w) and np.array_equal(rgba_m, rgba_w)
    # the wavefront walks the 4-wide tree (other visiting order, other culling): the per-ray events match, the walk's work need not
    for k in ("rays", "env_lookups"):
        assert st_m[k] == st_w[k], k
    assert st_w["nodes_visited"] > 0 or name == "benchmark"
    # on the same 2-wide tree the two kernels do exactly the same work
    narrow = _context_with({"RTR_WF_WIDE": "0"})
    try:
        rgba_n, acc_n, st_n = narrow.upload(scene).render(wf(scene, W, H, seed=9, flags=A.FLAG_COUNT_WORK), want_accum=True)
    finally:
        narrow.close()
    assert np.array_equal(acc_m, acc_n)
    for k in ("rays", "nodes_visited", "tri_tests", "sphere_tests", "env_lookups"):
        assert st_m[k] == st_n[k], k


def test_wavefront_wave_sizes_and_graph_replay_do_not_change_the_image():
    """One sample per wave, many samples per wave (per-sample buffer + ordered sum), a ragged last wave, plain launches
    instead of the CUDA graph, and one to four overlapped wave sequences: all bit-identical."""
    scene = _scenes.simple()
    scene.settings.samples_per_pixel = 7
    W, H = 97, 53  # not multiples of the 8x4 tile
    images = []
    for env in ({"RTR_WF_PATHS": "1024"}, {"RTR_WF_PATHS": str(97 * 53 * 3)}, {"RTR_WF_PATHS": str(1 << 21)},
                {"RTR_WF_PATHS": str(97 * 53 * 2), "RTR_WF_GRAPH": "0"}, {"RTR_WF_REFILL_THRESHOLD": "1"},
                {"RTR_WF_REFILL_THRESHOLD": "32"},
                # one, three and four wave sequences instead of two (capi.cpp render_wavefront): the pixel sums stay in sample order
                {"RTR_WF_LANES": "1"}, {"RTR_WF_LANES": "3", "RTR_WF_PATHS": "1024"}, {"RTR_WF_LANES": "4", "RTR_WF_PATHS": str(97 * 53)},
                {"RTR_WF_LANES": "2", "RTR_WF_PATHS": str(97 * 53), "RTR_WF_GRAPH": "0"}):
        old = {k: os.environ.get(k) for k in env}
        os.environ.update(env)
        try:
            c = rt.Context(0)  # the knobs are read when the context is created
        finally:
            for k, v in old.items():
                os.environ.pop(k) if v is None else os.environ.__setitem__(k, v)
        rgba, acc, st = c.upload(scene).render(wf(scene, W, H, seed=2), want_accum=True)
        images.append((acc, st["rays"]))
        c.close()
    rgba_o, acc_o, st_o = orc.OracleScene(scene).render(rt.make_params(scene, W, H, seed=2))
    for acc, rays in images:
        assert np.array_equal(acc, images[0][0]) and rays == st_o["rays"]
    np.testing.assert_allclose(images[0][0], acc_o, rtol=1e-5, atol=1e-6)


def test_wavefront_sample_slices_modes_and_depth_limits(ctx):
    scene = _scenes.benchmark()
    scene.settings.samples_per_pixel = 12
    ds = ctx.upload(scene)
    full = ds.render(wf(scene, 80, 45, seed=8), want_accum=True)[1]
    parts = [ds.render(wf(scene, 80, 45, seed=8, sample_begin=b, sample_count=c), want_accum=True)[1]
             for b, c in ((0, 5), (5, 4), (9, 3))]
    np.testing.assert_allclose(sum(parts), full, rtol=2e-6, atol=1e-6)
    assert np.all(full[..., 3] == 12) and np.all(parts[1][..., 3] == 4)
    for depth in (0, 1, 2):
        for mode in (rt.RenderMode.Full, rt.RenderMode.Clay, rt.RenderMode.Normal, rt.RenderMode.Random):
            for bvh in (True, False):
                st = rt.RenderSettings(samples_per_pixel=3, max_ray_depth=depth, mode=mode, enable_bvh_tree=bvh)
                p = wf(scene, 64, 36, seed=4, settings=st)
                rgba_d, acc_d, sd = ds.render(p, want_accum=True)
                rgba_o, acc_o, so = orc.OracleScene(scene).render(p)
                assert_images_close(rgba_d, acc_d, rgba_o, acc_o, f"depth {depth} mode {mode} bvh {bvh}")
                assert sd["rays"] == so["rays"]


def test_wavefront_zero_sized_and_empty(ctx):
    s = rt.Scene.default()
    ds = ctx.upload(s)
    rgba, acc, st = ds.render(wf(s, 33, 17), want_accum=True)
    assert st["rays"] == 33 * 17 * 5 and np.all(acc[..., 3] == 5)
    for W, H in ((0, 10), (10, 0)):
        rgba, acc, st = ds.render(wf(s, W, H))
        assert rgba.size == 0 and st["rays"] == 0


def _context_with(env):
    old = {k: os.environ.get(k) for k in env}
    os.environ.update(env)
    try:
        return rt.Context(0)  # the knobs are read when the context is created
    finally:
        for k, v in old.items():
            os.environ.pop(k) if v is None else os.environ.__setitem__(k, v)


@pytest.mark.parametrize("name,W,H,spp", [("mesh_hdri", 1280, 720, 2), ("volumes", 1280, 720, 2), ("nested_volumes", 1024, 1024, 2),
                                          ("spheres_host_built", 1280, 720, 2), ("spheres_device_built", 1280, 720, 2),
                                          ("grid_depth50", 1024, 768, 2)])
def test_four_wide_walk_is_bit_identical_to_the_two_wide_walk(name, W, H, spp):
    """wf_extend walks the 4-wide traversal tree by default (flat.h nodes4; RTR_WF_WIDE=0 keeps the 2-wide tree).  Both
    test the same gate boxes, so images, ray counts and closest-hit batches must not differ by a bit."""
    from test_gpu_edge import nested_volume_scene
    env = {}
    if name == "mesh_hdri": scene = _scenes.mesh_hdri(256, 128)
    elif name == "volumes": scene = _scenes.volumes(128, 64)
    elif name == "nested_volumes": scene = nested_volume_scene()
    elif name == "grid_depth50": scene = _scenes.grid_spheres(10)
    else:
        scene = _scenes.random_spheres(30_000)
        env = {"RTR_BVH_BUILD": "host" if name == "spheres_host_built" else "device"}
    scene.settings.samples_per_pixel = spp
    rays = _scenes.camera_rays_np(scene, 320, 200, seed=3)
    out = []
    for wide in ("1", "0"):
        c = _context_with(dict(env, RTR_WF_WIDE=wide))
        try:
            ds = c.upload(scene)
            rgba, acc, st = ds.render(wf(scene, W, H, seed=5), want_accum=True)
            assert st["kernel_used"] == A.KERNEL_WAVEFRONT
            out.append((rgba, acc, st["rays"], ds.trace_batch(rays, seed=5)))
        finally:
            c.close()
    (rgba4, acc4, rays4, hits4), (rgba2, acc2, rays2, hits2) = out
    assert rays4 == rays2 and np.array_equal(acc4, acc2) and np.array_equal(rgba4, rgba2)
    assert hits4.tobytes() == hits2.tobytes()
    assert (hits4["object"] != A.RTR_NO_HIT).mean() > 0.05


def test_auto_kernel_choice(ctx):
    """RTR_KERNEL_AUTO: wavefront for BVH walks big enough to fill a wave, megakernel otherwise (capi.cpp)."""
    mesh = _scenes.mesh_hdri(64, 32)
    mesh.settings.samples_per_pixel = 2
    ds = ctx.upload(mesh)
    assert ds.render(rt.make_params(mesh, 1024, 768, seed=1))[2]["kernel_used"] == A.KERNEL_WAVEFRONT
    assert ds.render(rt.make_params(mesh, 64, 64, seed=1))[2]["kernel_used"] == A.KERNEL_MEGAKERNEL
    flat = _scenes.benchmark()  # ships enable_bvh_tree = false
    flat.settings.samples_per_pixel = 2
    assert ctx.upload(flat).render(rt.make_params(flat, 1024, 768, seed=1))[2]["kernel_used"] == A.KERNEL_MEGAKERNEL


def test_progressive_accumulation(ctx):
    """RTR_FLAG_ACCUMULATE: sample slices added into one device buffer (the GUI's continuous mode, src/gui.rs:333-337).
    The wavefront adds sample by sample, so slices reproduce the one-shot render bit for bit."""
    import torch
    scene = _scenes.simple()
    scene.settings.samples_per_pixel = 9
    W, H = 1024, 768
    ds = ctx.upload(scene)
    acc = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()  # the library launches on its own stream
    for kernel, exact in ((A.KERNEL_WAVEFRONT, True), (A.KERNEL_MEGAKERNEL, False)):
        full = ds.render(rt.make_params(scene, W, H, seed=3, kernel=kernel), want_accum=True)[1]
        for i, (b, c) in enumerate(((0, 4), (4, 3), (7, 2))):
            p = rt.make_params(scene, W, H, seed=3, kernel=kernel, sample_begin=b, sample_count=c,
                               flags=A.FLAG_ACCUMULATE if i else 0)
            ds.render_accumulate_device(p, acc.data_ptr(), want_stats=False)
        ctx.synchronize()
        got = acc.cpu().numpy()
        assert np.all(got[..., 3] == 9)
        if exact:
            assert np.array_equal(got, full)
        else:
            np.testing.assert_allclose(got, full, rtol=2e-6, atol=1e-6)
