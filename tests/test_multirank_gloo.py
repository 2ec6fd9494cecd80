"""N>1 path on CPU: world_size-2 gloo, each rank accumulates its sample slice (with the oracle as
the renderer), ONE reduce, resolve on rank 0 — the same render_distributed() the GPU path uses."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, out_path):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import raytracingrust_b200 as rt
    from raytracingrust_b200.distributed import render_distributed
    from oracle import orc
    import _scenes
    scene = _scenes.benchmark()
    scene.settings.samples_per_pixel = 9  # odd: uneven slices
    W, H, seed = 40, 24, 5
    o = orc.OracleScene(scene)

    def accumulate(begin, count):
        p = rt.make_params(scene, W, H, seed=seed, sample_begin=begin, sample_count=count)
        return torch.from_numpy(o.render(p, threads=1, want_rgba=False)[1].copy())

    def resolve(acc, spp):  # render.rs:103-110
        c = np.sqrt(acc.numpy()[..., :3] / np.float32(spp)) * np.float32(255.0)
        return np.clip(np.nan_to_num(c), 0, 255).astype(np.uint8)

    rgba, acc = render_distributed(accumulate, resolve, W, H, scene.settings.samples_per_pixel)
    if rank == 0:
        full_rgba, full_acc, _ = o.render(rt.make_params(scene, W, H, seed=seed), threads=1)
        np.savez(out_path, rgba=rgba, acc=acc.numpy(), full_rgba=full_rgba, full_acc=full_acc)
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_reduce_matches_single_render(tmp_path):
    port = 29500 + os.getpid() % 2000
    out = str(tmp_path / "out.npz")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    z = np.load(out)
    assert np.all(z["acc"][..., 3] == 9)
    np.testing.assert_allclose(z["acc"], z["full_acc"], rtol=2e-6, atol=1e-6)  # fp32 summation order only
    assert np.abs(z["rgba"].astype(int) - z["full_rgba"][..., :3].astype(int)).max() <= 1
