#!/usr/bin/env python
"""Same-box A/B of library tuning knobs (environment variables read at context creation).
usage: python tools/ab.py "NAME:ENV=V,ENV=V" ... -- workload:spp ...
Each variant gets its own rtr_context; per workload the scene is built once; 1 warm-up + 3 timed renders, device time
from the library's own CUDA events (rtr_stats.device_ms).  Prints Mrays/s per (variant, workload)."""
import os
import sys
import json

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import raytracingrust_b200 as rt  # noqa: E402
import bench  # noqa: E402


def main():
    args = sys.argv[1:]
    sep = args.index("--")
    variants, workloads = args[:sep], args[sep + 1:]
    results = {}
    for wl in workloads:
        name, spp = wl.split(":")
        spp = int(spp)
        w = bench.WORKLOADS[name]
        scene = bench.make_scene(name)
        scene.settings.samples_per_pixel = spp
        for v in variants:
            vname, _, envs = v.partition(":")
            saved = {}
            for kv in filter(None, envs.split(",")):
                k, _, val = kv.partition("=")
                saved[k] = os.environ.get(k)
                os.environ[k] = val
            ctx = rt.Context(0)
            ds = ctx.upload(scene)
            p = rt.make_params(scene, w["W"], w["H"], seed=0)
            acc = None
            ms = []
            for i in range(4):
                rgba, _, st = ds.render(p, want_rgba=True, want_accum=False)
                if i:
                    ms.append(st["device_ms"])
            mr = st["rays"] / (np.mean(ms) * 1e-3) / 1e6
            results[(vname, name)] = mr
            print(f"{vname:24s} {name:14s} spp {spp:4d} kernel {st['kernel_used']} {mr:9.0f} Mrays/s  {np.mean(ms):9.2f} ms (min {min(ms):.2f})  crc {int(rgba.astype(np.uint64).sum())}", flush=True)
            ds.close(); ctx.close()
            for k, old in saved.items():
                if old is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = old
    base = variants[0].partition(":")[0]
    print("\nrelative to", base)
    for v in variants[1:]:
        vname = v.partition(":")[0]
        print(f"{vname:24s} " + "  ".join(f"{wl.split(':')[0]} {results[(vname, wl.split(':')[0])] / results[(base, wl.split(':')[0])] - 1:+.1%}" for wl in workloads))


if __name__ == "__main__":
    main()
