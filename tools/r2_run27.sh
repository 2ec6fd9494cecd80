#!/bin/bash
# hybrid mesh build (device reference tree + host SAH over its gates): tests, upload time, render speed
mkdir -p gpurun_out/r2
timeout 900 python -m pytest tests/test_gpu_bvh_build.py -m gpu -x -q 2>&1 | tail -15 | tee gpurun_out/r2/run27_pytest_bvh_build.txt
{
timeout 900 python tools/ab.py "host_sah" "hybrid:RTR_BVH_BUILD=hybrid" "device_median:RTR_BVH_BUILD=device" -- mesh_hdri:32 volumes:16 mesh_depth50:16 2>&1 | grep -v "^$"
python - <<'PY'
import os, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import raytracingrust_b200 as rt
import _scenes, bench
from test_gpu_bvh_build import random_triangle_mesh
ctx = rt.Context(0)
def big(n):
    s = rt.Scene.default(); lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4))); s.add_object(random_triangle_mesh(n, 7, lam)); return s
cases = [("mesh_hdri (C4 scene)", bench.make_scene("mesh_hdri")), ("volumes (C5 scene)", bench.make_scene("volumes")), ("monkey", _scenes.mesh_bench()),
         ("28316 random tris (cool_cube.obj size)", big(28316)), ("200k random tris", big(200_000)), ("1M random tris", big(1_000_000))]
for name, scene in cases:
    row = []
    for mode in ("host", "hybrid", "device"):
        os.environ["RTR_BVH_BUILD"] = mode
        ms = []
        for i in range(3):
            t0 = time.perf_counter(); ds = ctx.upload(scene); ms.append((time.perf_counter() - t0) * 1e3); info = ds.info(); ds.close()
        row.append(f"{mode}: {min(ms):8.1f} ms ({info['n_nodes']} nodes)")
    print(f"{name:42s} rtr_scene_upload wall  " + "   ".join(row), flush=True)
PY
} > gpurun_out/r2/ab_hybrid_mesh_build.txt 2>&1
cat gpurun_out/r2/ab_hybrid_mesh_build.txt
