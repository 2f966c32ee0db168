import os, sys
sys.path.insert(0, "/root/repo" if os.path.isdir("/root/repo") else ".")
import numpy as np, torch
import bench
from gpupathtracer_b200 import Renderer, Scene, scenes
wl = bench.WORKLOADS["bunny"]; W, H = wl["width"], wl["height"]
base = Scene(scenes.bunny_like(seed=1))
cam = bench.make_camera(base, W, H)
n = W * H
cases = [("host SBVH", "build", None, None), ("device LBVH", "device", None, None), ("device SAH", "sah", None, None)] + [(f"PLOC r={r}", "ploc", None, r) for r in (16, 32)] + \
        [(f"split {g} + PLOC", "split", g, 16) for g in (32,)] + [(f"split {g} + LBVH", "split_lbvh", g, 16) for g in ()]
if os.environ.get('EXP_ONLY'):
    cases = [c for c in cases if c[1] in os.environ['EXP_ONLY'].split(',')]
for label, how, g, rad in cases:
    if g: os.environ["PT_SPLIT_GRID"] = str(g)
    if rad: os.environ["PT_PLOC_RADIUS"] = str(rad)
    s = Scene(base.data)
    r = Renderer().Initialize(W, H, s, max_depth=1, seed=1, bvh=how)
    if os.environ.get("EXP_OPT"):
        it, fr = os.environ["EXP_OPT"].split(","); o = r.OptimizeBVH(int(it), float(fr))
        label = label + f" + opt {it}x{fr}"; r.bvh_stats.sah_cost = o.sah_cost; r.bvh_stats.max_depth = o.max_depth
    r.lib.pt_set_camera(r.ctx, cam.plane)
    pr = r.RaygenProbe(0, n)
    rays = np.zeros((n, 8), np.float32); rays[:, 0:3] = pr[:, 2:5]; rays[:, 4:7] = pr[:, 5:8]; rays[:, 7] = 1e20
    d = torch.from_numpy(rays).cuda(); h = torch.empty((n, 4), dtype=torch.float32, device="cuda"); torch.cuda.synchronize()
    ms = []
    for _ in range(10):
        r.TraceClosestDevice(d.data_ptr(), n, h.data_ptr()); ms.append(r.LastDeviceMs())
    st = r.bvh_stats
    print(f"{label:20s} build {st.build_ms:7.2f} ms refs {st.n_refs:7d} depth {st.max_depth:3d} SAH {st.sah_cost:6.1f} primary rays warm L2 {n / min(ms) / 1e3:.0f} Mrays/s", flush=True)
    r.CleanUp()
