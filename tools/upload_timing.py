import sys, time, os
sys.path.insert(0, "/root/repo")
import numpy as np
from cuda_pathtracer_b200 import api, abi, scenes
from cuda_pathtracer_b200.scene import Scene, material
v = scenes.heightfield_soup()
for dev in (0,):
    s = Scene().set_build_device(dev); s.addMaterial(material())
    t0=time.perf_counter(); s.addModelSoup(v); print("addModelSoup wall %.1f ms (build %.1f)" % ((time.perf_counter()-t0)*1e3, s.last_build_ms()))
    s.addObject(0); s.finalize()
    t0=time.perf_counter(); arrays = s.arrays(); print("arrays() %.1f ms" % ((time.perf_counter()-t0)*1e3))
    for flags in (0, 4, 0):
        t0=time.perf_counter(); pt = api.Pathtracer(arrays, abi.default_config(64, 64, max_ray_depth=2, interactive_depth=1, flags=flags)); t1=time.perf_counter(); pt.Init(); print("flags %d: ctor %.1f ms Init %.1f ms" % (flags, (t1-t0)*1e3, (time.perf_counter()-t1)*1e3))
        pt.close()
