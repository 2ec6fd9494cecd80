#!/usr/bin/env python
"""One render of a bench workload inside a cudaProfilerStart/Stop range (ncu --profile-from-start off): a warm-up
render, then the profiled one.  usage: python tools/profile_one.py WORKLOAD SPP [KERNEL]
Prints a JSON line with the rays and samples of the profiled render (profiles/make_counters.py divides by them)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import raytracingrust_b200 as rt  # noqa: E402
import bench  # noqa: E402


def main():
    name, spp = sys.argv[1], int(sys.argv[2])
    kernel = int(sys.argv[3]) if len(sys.argv) > 3 else 0
    w = bench.WORKLOADS[name]
    scene = bench.make_scene(name)
    scene.settings.samples_per_pixel = spp
    ctx = rt.Context(0)
    ds = ctx.upload(scene)
    p = rt.make_params(scene, w["W"], w["H"], seed=0, kernel=kernel)
    ds.render(p, want_rgba=True)
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    _, _, st = ds.render(p, want_rgba=True)
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(json.dumps({"workload": name, "spp": spp, "rays": st["rays"], "samples": st["samples"], "device_ms": st["device_ms"],
                      "kernel_used": st["kernel_used"], "launches": st["kernel_launches"]}))


if __name__ == "__main__":
    main()
