"""Work per ray on the traversal tree vs the reference topology (device counters).  usage: count_work.py workload..."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import raytracingrust_b200 as rt
from raytracingrust_b200 import _abi as A
import bench
ctx = rt.Context(0)
for name in sys.argv[1:]:
    scene = bench.make_scene(bench.WORKLOADS[name][0])
    scene.settings.samples_per_pixel = 2
    ds = ctx.upload(scene)
    for label, fl in (("traversal", 0), ("reference", A.FLAG_REFERENCE_TREE)):
        st = ds.render(rt.make_params(scene, 960, 540, seed=1, flags=A.FLAG_COUNT_WORK | fl, kernel=A.KERNEL_MEGAKERNEL))[2]
        r = st["rays"]
        print(f"{name:10s} {label:9s} boxes/ray {st['nodes_visited']/r:7.2f} tris/ray {st['tri_tests']/r:6.2f} spheres/ray {st['sphere_tests']/r:6.2f} device_ms {st['device_ms']:.2f} info {ds.info()['build_ms']:.0f} ms build")
