"""Experiment: do S independent wavefronts (contexts/streams) on one GPU overlap each other's kernel tails?"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from gpupathtracer_b200 import Renderer
import torch
wl = bench.WORKLOADS["atrium"]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
cam = bench.make_camera(scene, W, H)
for S in (1, 2, 3, 4):
    rs = [Renderer().Initialize(W, H, scene, max_depth=5, seed=100 + i, bvh="build" if scene.nodes is None else "given") for i in range(S)]
    for r in rs: r.RenderFrame(cam, 4)
    for r in rs: r.Synchronize()
    c0 = [r.Counters() for r in rs]
    K = 48 // S
    t0 = time.perf_counter()
    for r in rs: r.RenderFrame(cam, K)
    for r in rs: r.Synchronize()
    dt = time.perf_counter() - t0
    c1 = [r.Counters() for r in rs]
    rays = sum(b["closest_rays"] + b["shadow_rays"] - a["closest_rays"] - a["shadow_rays"] for a, b in zip(c0, c1))
    print(f"S={S}: {rays / dt / 1e6:.1f} Mrays/s wall, {dt / (K * S) * 1e3:.3f} ms/frame", flush=True)
    for r in rs: r.CleanUp()
