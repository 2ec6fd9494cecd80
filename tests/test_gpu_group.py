"""One render() over N devices of a process (rtr_group_*): the N-way image equals the 1-device image up to fp32
summation order, through both cross-device steps (the library's NVLink peer kernel, ncclReduce).

On a one-GPU box the split is exercised with the same ordinal repeated (the slices time-share the device; P2P mode
only — NCCL refuses duplicate devices).  On a multi-GPU box the distinct-device cases run as well."""
import os
import sys

import numpy as np
import pytest

import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
from oracle import orc
import _scenes

pytestmark = pytest.mark.gpu

SCENES = {
    # name: (factory, W, H, spp, kernel)
    "simple": (_scenes.simple, 96, 64, 11, A.KERNEL_AUTO),
    "benchmark_scan": (_scenes.benchmark, 96, 54, 9, A.KERNEL_AUTO),
    "mesh_hdri": (lambda: _scenes.mesh_hdri(256, 128), 96, 54, 8, A.KERNEL_WAVEFRONT),
    "volumes": (lambda: _scenes.volumes(128, 64), 64, 36, 8, A.KERNEL_WAVEFRONT),
}


def n_devices():
    import torch
    return torch.cuda.device_count()


def check_against_single(ctx, scene, W, H, kernel, group, seed=3, rtol=2e-6):
    p = rt.make_params(scene, W, H, seed=seed, kernel=kernel)
    rgba1, acc1, st1 = ctx.upload(scene).render(p, want_accum=True)
    rgbaN, accN, stN = group.upload(scene).render(p, want_accum=True)
    spp = scene.settings.samples_per_pixel
    assert np.all(accN[..., 3] == spp)
    np.testing.assert_allclose(accN, acc1, rtol=rtol, atol=1e-6)  # fp32 summation order only
    assert np.abs(rgbaN.astype(int) - rgba1.astype(int)).max() <= 1
    assert stN["rays"] == st1["rays"] and stN["samples"] == W * H * spp
    slices = [group.device_stats(i) for i in range(len(group.devices))]
    assert sum(s["sample_count"] for s in slices) == spp
    assert sum(s["rays"] for s in slices) == stN["rays"]
    return stN, slices


@pytest.mark.parametrize("name", sorted(SCENES))
@pytest.mark.parametrize("members", [2, 3])
def test_group_on_one_device_matches_single_render(ctx, name, members):
    factory, W, H, spp, kernel = SCENES[name]
    scene = factory()
    scene.settings.samples_per_pixel = spp
    g = rt.Group([0] * members, reduce=A.REDUCE_P2P)
    try:
        st, slices = check_against_single(ctx, scene, W, H, kernel, g)
        begins = [s["sample_begin"] for s in slices]
        assert begins == sorted(begins) and begins[0] == 0  # contiguous slices in rank order
    finally:
        g.close()


def test_group_is_deterministic_and_matches_the_oracle(ctx):
    scene = _scenes.simple()
    scene.settings.samples_per_pixel = 6
    g = rt.Group([0, 0], reduce=A.REDUCE_P2P)
    try:
        gs = g.upload(scene)
        p = rt.make_params(scene, 80, 60, seed=5)
        a = gs.render(p, want_accum=True)
        b = gs.render(p, want_accum=True)
        assert np.array_equal(a[1], b[1]) and np.array_equal(a[0], b[0])  # rank-order sum: no run-to-run noise
        rgba_o, acc_o, st_o = orc.OracleScene(scene).render(p)
        assert a[2]["rays"] == st_o["rays"]
        np.testing.assert_allclose(a[1], acc_o, rtol=1e-5, atol=1e-6)
        assert np.abs(a[0].astype(int) - rgba_o.astype(int)).max() <= 1
    finally:
        g.close()


def test_more_members_than_samples_and_empty_image(ctx):
    scene = _scenes.benchmark()
    scene.settings.samples_per_pixel = 2
    g = rt.Group([0, 0, 0, 0], reduce=A.REDUCE_P2P)
    try:
        st, slices = check_against_single(ctx, scene, 48, 27, A.KERNEL_AUTO, g)
        assert [s["sample_count"] for s in slices] == [1, 1, 0, 0]
        rgba, acc, st0 = g.upload(scene).render(rt.make_params(scene, 0, 0), want_accum=True)
        assert rgba.size == 0 and st0["rays"] == 0
    finally:
        g.close()


def test_dynamic_split_matches_static(ctx):
    """RTR_SPLIT_DYNAMIC: devices pull sample chunks from a shared counter; same image up to summation order."""
    scene = _scenes.simple()
    scene.settings.samples_per_pixel = 24
    g = rt.Group([0, 0, 0], reduce=A.REDUCE_P2P)
    try:
        g.set_split(A.SPLIT_DYNAMIC, 5)  # 5 pieces: 5+5+5+5+4
        st, slices = check_against_single(ctx, scene, 64, 48, A.KERNEL_AUTO, g, rtol=1e-5)
        assert st["kernel_launches"] >= 5
    finally:
        g.close()


def test_group_rejects_bad_arguments():
    with pytest.raises(rt.RtrError):
        rt.Group([])
    with pytest.raises(rt.RtrError):
        rt.Group([0, 0], reduce=A.REDUCE_NCCL)  # NCCL needs distinct devices
    with pytest.raises(rt.RtrError):
        rt.Group([9999])


def test_reference_facing_render_over_devices():
    """render(&mut target, &scene) with devices=[...] (render.rs:61)."""
    scene = _scenes.simple()
    t1, t2 = rt.RenderTarget(64, 64), rt.RenderTarget(64, 64)
    rt.render(t1, scene, seed=2)
    rt.render(t2, scene, seed=2, devices=[0, 0])
    assert t2.request_redraw and np.abs(t1.image().astype(int) - t2.image().astype(int)).max() <= 1


# ---------------------------------------------------------------- distinct devices (multi-GPU box only)
@pytest.mark.parametrize("reduce", [A.REDUCE_P2P, A.REDUCE_NCCL])
@pytest.mark.parametrize("name", ["simple", "mesh_hdri", "volumes"])
def test_group_over_distinct_devices(ctx, name, reduce):
    n = n_devices()
    if n < 2:
        pytest.skip("needs >= 2 GPUs")
    factory, W, H, spp, kernel = SCENES[name]
    scene = factory()
    scene.settings.samples_per_pixel = spp
    g = rt.Group(list(range(min(n, 8))), reduce=reduce)
    try:
        assert g.reduce_mode == ("p2p" if reduce == A.REDUCE_P2P else "nccl")
        check_against_single(ctx, scene, W, H, kernel, g, rtol=2e-6 if reduce == A.REDUCE_P2P else 1e-5)
    finally:
        g.close()


def _nccl_worker(rank, world, port, out_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(root, "tests"))
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    import raytracingrust_b200 as rt
    from raytracingrust_b200.distributed import render_distributed_gpu
    import _scenes
    scene = _scenes.simple()
    scene.settings.samples_per_pixel = 9
    ctx = rt.Context(rank)  # a DEFAULT context: its own non-blocking stream, not torch's
    ds = ctx.upload(scene)
    rgba, acc = render_distributed_gpu(ctx, ds, scene, 96, 64, seed=4)
    if rank == 0:
        full = ds.render(rt.make_params(scene, 96, 64, seed=4), want_accum=True)
        np.savez(out_path, rgba=rgba.cpu().numpy(), acc=acc.cpu().numpy(), full_rgba=full[0], full_acc=full[1])
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(600)
def test_torch_distributed_nccl_ranks_with_default_contexts(tmp_path):
    """One process per GPU (the torchrun layout): library launches on the context's own stream, NCCL reduce on
    torch's — render_distributed_gpu must order the three steps."""
    if n_devices() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch.multiprocessing as mp
    out = str(tmp_path / "out.npz")
    mp.spawn(_nccl_worker, args=(2, 29500 + os.getpid() % 2000, out), nprocs=2, join=True)
    z = np.load(out)
    assert np.all(z["acc"][..., 3] == 9)
    np.testing.assert_allclose(z["acc"], z["full_acc"], rtol=2e-6, atol=1e-6)
    assert np.abs(z["rgba"].astype(int) - z["full_rgba"].astype(int)).max() <= 1
