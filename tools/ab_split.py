#!/usr/bin/env python
"""SURVEY 8f-4 A/B on one process over all visible GPUs: static sample slices vs the dynamic chunk-pulling split of
rtr_group (and NCCL vs the library's P2P reduce), one image per call.  usage: python tools/ab_split.py workload:spp ...
Prints wall ms per call, the library's device_ms (root stream, render start -> end of the cross-device step), the
slowest member's render ms and every member's render ms / samples taken."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import raytracingrust_b200 as rt  # noqa: E402
from raytracingrust_b200 import _abi as A  # noqa: E402
import bench  # noqa: E402


def main():
    n = torch.cuda.device_count()
    groups = {"p2p": rt.Group(list(range(n)), reduce=A.REDUCE_P2P)}
    if n > 1:
        groups["nccl"] = rt.Group(list(range(n)), reduce=A.REDUCE_NCCL)
    for wl in sys.argv[1:]:
        name, spp = wl.split(":")
        spp = int(spp)
        w = bench.WORKLOADS[name]
        scene = bench.make_scene(name)
        scene.settings.samples_per_pixel = spp
        p = rt.make_params(scene, w["W"], w["H"], seed=0)
        host = torch.empty((w["H"], w["W"], 4), dtype=torch.uint8).pin_memory()
        ref = None
        for gname, g in groups.items():
            gs = g.upload(scene)
            for mode, chunk in (("static", 0), ("dynamic", max(1, spp // (4 * n))), ("dynamic", max(1, spp // (16 * n)))):
                if gname == "nccl" and mode == "dynamic":
                    continue
                g.set_split(A.SPLIT_DYNAMIC if mode == "dynamic" else A.SPLIT_STATIC, chunk)
                gs.render_into(p, host.data_ptr())
                wall, dev, slow = [], [], []
                for _ in range(3):
                    t0 = time.perf_counter()
                    st = gs.render_into(p, host.data_ptr())
                    wall.append((time.perf_counter() - t0) * 1e3); dev.append(st["device_ms"]); slow.append(st["upload_ms"])
                img = host.numpy().astype(np.int32).copy()
                if ref is None:
                    ref = img
                per = [g.device_stats(i) for i in range(n)]
                print(f"{name:13s} spp {spp:4d} N={n} {gname:4s} {mode:7s} chunk {chunk:3d}: wall {np.mean(wall):8.2f} ms  device {np.mean(dev):8.2f} ms  "
                      f"slowest member {np.mean(slow):8.2f} ms  Mrays/s {st['rays'] / np.mean(dev) / 1e3:9.0f}  max|rgba - first| {np.abs(img - ref).max()}  "
                      f"member ms {[round(x['device_ms'], 1) for x in per]} spp {[x['sample_count'] for x in per]}", flush=True)
            gs.close()


if __name__ == "__main__":
    main()
