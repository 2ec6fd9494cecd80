import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np, torch
import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
import bench
name = sys.argv[1]; spp = int(sys.argv[2])
scene = bench.make_scene(name); scene.settings.samples_per_pixel = spp
ctx = rt.Context(0)
host = torch.empty((1080, 1920, 4), dtype=torch.uint8).pin_memory()
for kernel in (1, 2, 1, 2):
    t0 = time.perf_counter(); low = scene.lower(); t1 = time.perf_counter()
    ds = ctx.upload(low); t2 = time.perf_counter()
    st = ds.render_into(rt.make_params(scene, 1920, 1080, seed=0, kernel=kernel), host.data_ptr()); t3 = time.perf_counter()
    ds.close(); t4 = time.perf_counter()
    print(f"kernel {kernel}: lower {1e3*(t1-t0):.1f} ms, upload {1e3*(t2-t1):.1f} (build {ds and 0}), render+D2H {1e3*(t3-t2):.1f} (device {st['device_ms']:.1f}), close {1e3*(t4-t3):.1f}")
