import sys; sys.path.insert(0,'/root/repo'); sys.path.insert(0,'/root/repo/tests')
import numpy as np, raytracingrust_b200 as rt, _scenes
from raytracingrust_b200 import _abi as A
s=_scenes.grid_spheres(10)
ctx=rt.Context(0); ds=ctx.upload(s); print(ds.info())
rays=np.concatenate([_scenes.camera_rays_np(s,96,64), _scenes.random_rays(6000,3)])
for fl in (0,1):
    h=ds.trace_batch(rays, flags=fl)
    bad=h[h['object']==0xDEADBEEF]
    print('flags',fl,'bad',len(bad)); 
    for b in bad[:5]: print(hex(b['primitive']), b['t'], b['point'][0])
