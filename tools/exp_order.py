"""Experiment: ray ordering (launch_ray_order) on an L2-resident scene -- 4 M random rays vs the 262k-triangle atrium,
natural vs binned order, reference and wide layouts. PT_RAY_ORDER is read at pt_create."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import bench
from gpupathtracer_b200 import Renderer, scenes
from gpupathtracer_b200.api import PT_BVH_WIDE, PT_BVH_REFERENCE

wl = bench.WORKLOADS["atrium"]
scene = bench.build_scene(wl)
lo, hi = scene.bounds()
n = 1 << 22
rays = scenes.random_rays(n, lo, hi, seed=5)
anyr = scenes.random_rays(n, lo, hi, seed=6, any_hit=True)
d = torch.from_numpy(rays).cuda(); dh = torch.empty((n, 4), device="cuda")
da = torch.from_numpy(anyr).cuda(); df = torch.empty((n,), dtype=torch.uint8, device="cuda")
for name, layout in (("ref", PT_BVH_REFERENCE), ("wide", PT_BVH_WIDE)):
    for mode in ("-1", "1"):
        os.environ["PT_RAY_ORDER"] = mode
        r = Renderer().Initialize(64, 64, scene, max_depth=5, seed=7, bvh="build" if scene.nodes is None else "given", bvh_layout=layout)
        out = []
        for ptr, res, fn in ((d, dh, r.TraceClosestDevice), (da, df, r.TraceAnyDevice)):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(2):
                fn(ptr.data_ptr(), n, res.data_ptr())
            r.Synchronize()
            import time
            t0 = time.perf_counter()
            for _ in range(5):
                fn(ptr.data_ptr(), n, res.data_ptr())
            r.Synchronize()
            out.append(5 * n / (time.perf_counter() - t0) / 1e6)
        print(f"{name} order {mode}: closest {out[0]:.0f} Mrays/s, any {out[1]:.0f} Mrays/s (binning included)", flush=True)
        r.CleanUp()
