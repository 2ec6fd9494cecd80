"""Where an end-to-end step goes: lower() (Python), rtr_scene_upload (host BVH build + H2D), render + D2H.
usage (GPU box): python tools/e2e_phases.py WORKLOAD SPP"""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import raytracingrust_b200 as rt
import bench
name = sys.argv[1]; spp = int(sys.argv[2])
scene = bench.make_scene(name); scene.settings.samples_per_pixel = spp
W, H = bench.WORKLOADS[name][1:3]
ctx = rt.Context(0)
host = torch.empty((H, W, 4), dtype=torch.uint8).pin_memory()
for it in range(int(sys.argv[3]) if len(sys.argv) > 3 else 3):
    t0 = time.perf_counter(); low = scene.lower(); t1 = time.perf_counter()
    ds = ctx.upload(low); t2 = time.perf_counter()
    info = ds.info()
    st = ds.render_into(rt.make_params(scene, W, H, seed=it), host.data_ptr()); t3 = time.perf_counter()
    ds.close(); t4 = time.perf_counter()
    print(f"{name}: lower {1e3*(t1-t0):.1f} ms, upload {1e3*(t2-t1):.1f} ms (of which BVH build {info.get('build_ms', -1):.1f}, "
          f"h2d {info.get('h2d_bytes', 0)/1e6:.1f} MB), render+D2H {1e3*(t3-t2):.1f} (device {st['device_ms']:.1f}), close {1e3*(t4-t3):.1f}")
