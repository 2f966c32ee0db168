"""One device build of the atrium under ncu's launch list: which kernels the build time goes to.
  ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file out.csv python tools/ncu_builder.py [sah|ploc|device]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gpupathtracer_b200 import Renderer, Scene
how = sys.argv[1] if len(sys.argv) > 1 else "sah"
wl = bench.WORKLOADS[sys.argv[2] if len(sys.argv) > 2 else "atrium"]
s = Scene(bench.build_scene(wl).data)
r = Renderer().Initialize(64, 64, s, bvh=how)
print(how, "build_ms", r.bvh_stats.build_ms, "depth", r.bvh_stats.max_depth, "sah", r.bvh_stats.sah_cost)
r.CleanUp()
