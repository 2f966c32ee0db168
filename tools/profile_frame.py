"""Two sequential frames of a workload as full-grid, non-overlapping launches (pt_set_stage_timing: the launch set bench.py's
roofline.avg_launch_ms is measured on) -- the target of the ncu captures under profiles/.
  python tools/profile_frame.py [atrium|bathroom|4k] [frames]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gpupathtracer_b200 import Renderer

name = sys.argv[1] if len(sys.argv) > 1 else "atrium"
frames = int(sys.argv[2]) if len(sys.argv) > 2 else 2
wl = bench.WORKLOADS[name]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
r = Renderer().Initialize(W, H, scene, max_depth=wl["depth"], microfacet_model=wl["model"], seed=0xB200, bvh="build")
cam = bench.make_camera(scene, W, H)
r.SetStageTiming(True)
r.RenderFrame(cam, frames); r.Synchronize()
print({k: v / frames for k, v in r.StageMs().items()}, r.Counters())
r.CleanUp()
