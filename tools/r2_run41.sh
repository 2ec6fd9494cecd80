#!/bin/bash
mkdir -p gpurun_out/r2
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 | tee gpurun_out/r2/run41_pytest.txt
python - <<'PY' | tee gpurun_out/r2/upload_times_threaded_host.txt
import sys, time, os
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import raytracingrust_b200 as rt
import bench, _scenes
from test_gpu_bvh_build import random_triangle_mesh
ctx = rt.Context(0)
def big(n):
    s = rt.Scene.default(); lam = s.add_material(rt.Lambertian(rt.Color(0.6, 0.5, 0.4))); s.add_object(random_triangle_mesh(n, 7, lam)); return s
cases = [("mesh_hdri (C4 scene)", bench.make_scene("mesh_hdri")), ("volumes (C5 scene)", bench.make_scene("volumes")),
         ("28316 random tris (cool_cube.obj size)", big(28316)), ("200k random tris", big(200_000)), ("1M random tris", big(1_000_000))]
for name, scene in cases:
    row = []
    for mode, thr in (("host", "1"), ("host", None), ("hybrid", None), ("device", None)):
        os.environ["RTR_BVH_BUILD"] = mode
        if thr: os.environ["RTR_HOST_THREADS"] = thr
        else: os.environ.pop("RTR_HOST_THREADS", None)
        ms = []
        for i in range(3):
            t0 = time.perf_counter(); ds = ctx.upload(scene); ms.append((time.perf_counter() - t0) * 1e3); ds.close()
        row.append(f"{mode}{' (1 thread)' if thr else ''}: {min(ms):7.1f} ms")
    print(f"{name:40s} rtr_scene_upload wall  " + "   ".join(row), flush=True)
PY
