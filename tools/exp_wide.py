"""Experiment: PT_BVH_WIDE vs PT_BVH_REFERENCE -- id agreement and speed on fixed ray batches + frames."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from gpupathtracer_b200 import Renderer, scenes, Scene
from gpupathtracer_b200.api import PT_BVH_WIDE, PT_BVH_REFERENCE, PT_BVH_WIDE4

wl = bench.WORKLOADS["atrium"]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
cam = bench.make_camera(scene, W, H)
lo, hi = scene.bounds()
n = 1 << 22
rays = scenes.random_rays(n, lo, hi, seed=5)
anyr = scenes.random_rays(n, lo, hi, seed=6, any_hit=True)
res = {}
for name, layout in (("ref", PT_BVH_REFERENCE), ("wide", PT_BVH_WIDE), ("wide4", PT_BVH_WIDE4)):
    r = Renderer().Initialize(W, H, scene, max_depth=5, seed=7, bvh="build" if scene.nodes is None else "given", bvh_layout=layout)
    hits = r.TraceClosest(rays); flags = r.TraceAny(anyr)
    import torch
    d = torch.from_numpy(rays).cuda(); dh = torch.empty((n, 4), device="cuda"); torch.cuda.synchronize()
    for _ in range(3):
        r.TraceClosestDevice(d.data_ptr(), n, dh.data_ptr())
    ms = r.LastDeviceMs()
    da = torch.from_numpy(anyr).cuda(); df = torch.empty((n,), dtype=torch.uint8, device="cuda"); torch.cuda.synchronize()
    for _ in range(3):
        r.TraceAnyDevice(da.data_ptr(), n, df.data_ptr())
    msa = r.LastDeviceMs()
    r.RenderFrame(cam, 4); r.Synchronize()
    c0 = r.Counters(); t0 = time.perf_counter()
    r.RenderFrame(cam, 32); r.Synchronize()
    dt = time.perf_counter() - t0; c1 = r.Counters()
    mr = (c1["closest_rays"] + c1["shadow_rays"] - c0["closest_rays"] - c0["shadow_rays"]) / dt / 1e6
    acc = r.DownloadAccum()
    res[name] = (hits, flags, acc)
    print(f"{name}: closest {n / ms / 1e3:.1f} Mrays/s, any {n / msa / 1e3:.1f} Mrays/s, frames {mr:.1f} Mrays/s", flush=True)
    r.CleanUp()
for other in ("wide", "wide4"):
  a, b = res["ref"], res[other]
  print("==", other)
  ida, idb = a[0][:, 3].view(np.int32), b[0][:, 3].view(np.int32)
  diff = ida != idb
  print("closest id mismatches:", int(diff.sum()), "of", n, "| t differs where ids equal:", int((a[0][~diff, 0] != b[0][~diff, 0]).sum()))
  if diff.any():
      k = np.nonzero(diff)[0][:8]
      print(" examples ref(id,t) wide(id,t):", [(int(ida[i]), float(a[0][i, 0]), int(idb[i]), float(b[0][i, 0])) for i in k])
  print("any-hit flag mismatches:", int((a[1] != b[1]).sum()), "of", n)
  fin = np.isfinite(a[2]).all(axis=2) & np.isfinite(b[2]).all(axis=2)
  print("frame pixels bit-identical:", float((a[2][fin] == b[2][fin]).all(axis=1).mean()))
