"""Where does the RenderFrame + Present loop spend its time? Host enqueue time vs wall time per frame, lazy join on / off."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch
import bench
from gpupathtracer_b200 import Renderer
wl = bench.WORKLOADS["atrium"]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
r = Renderer().Initialize(W, H, scene, max_depth=5, seed=1, bvh="build")
cam = bench.make_camera(scene, W, H)
pinned = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
for present in (False, True):
    for _ in range(4):
        r.RenderFrame(cam, 1)
    r.Synchronize()
    n = 64
    t0 = time.perf_counter()
    for i in range(n):
        cam.GenerateImagePlane()
        r.RenderFrame(cam, 1)
        if present:
            r.PresentAsync(pinned[i & 1].data_ptr())
    t1 = time.perf_counter()
    r.Synchronize()
    t2 = time.perf_counter()
    print(f"PT_LAZY_JOIN={os.environ.get('PT_LAZY_JOIN', '1')} present={present}: host enqueue {1e3 * (t1 - t0) / n:.3f} ms/frame, wall {1e3 * (t2 - t0) / n:.3f} ms/frame")
t0 = time.perf_counter(); r.RenderFrame(cam, 64); t1 = time.perf_counter(); r.Synchronize(); t2 = time.perf_counter()
print(f"batch of 64: host enqueue {1e3 * (t1 - t0) / 64:.3f} ms/frame, wall {1e3 * (t2 - t0) / 64:.3f} ms/frame")
r.CleanUp()
