"""How much would coherence buy on secondary rays? Proxy rays for bounce 1 of the atrium (cosine-distributed directions about the
normals of the primary hit points; shadow rays from those points to area-sampled emitter points) traced as device batches in
queue order (primary-pixel order = what k_shade's compaction produces), in random order, and binned by origin cell + direction
octant (launch_ray_order), with the binning time included and separated."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from gpupathtracer_b200 import Renderer

wl = bench.WORKLOADS[sys.argv[1] if len(sys.argv) > 1 else "atrium"]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]


def renderer(order):
    os.environ["PT_RAY_ORDER"] = str(order)
    return Renderer().Initialize(W, H, scene, max_depth=5, seed=1, bvh="build" if scene.nodes is None else "given")


r = renderer(-1)
cam = bench.make_camera(scene, W, H)
r.lib.pt_set_camera(r.ctx, cam.plane)
pr = r.RaygenProbe(0, W * H)
rays = np.zeros((W * H, 8), np.float32); rays[:, 0:3] = pr[:, 2:5]; rays[:, 4:7] = pr[:, 5:8]; rays[:, 7] = 1e20
hits = r.TraceClosest(rays)
ok = hits[:, 3].view(np.int32) >= 0
tri = hits[ok, 3].view(np.int32)
p = rays[ok, 0:3] + rays[ok, 4:7] * hits[ok, 0:1]
n = scene.tris[tri][:, 15:18]
g = np.random.default_rng(1)
m = p.shape[0]
# cosine-distributed bounce directions
u1, u2 = g.uniform(size=m), g.uniform(size=m)
rr, phi = np.sqrt(u1), 2 * np.pi * u2
a = np.where(np.abs(n[:, 1:2]) > 0.99, [[1.0, 0, 0]], [[0, 1.0, 0]])
t = np.cross(a, n); t /= np.linalg.norm(t, axis=1, keepdims=True); b = np.cross(t, n)
d = t * (rr * np.sin(phi))[:, None] + b * (rr * np.cos(phi))[:, None] + n * np.sqrt(1 - u1)[:, None]
ext = np.zeros((m, 8), np.float32); ext[:, 0:3] = p + 0.003 * n; ext[:, 4:7] = d; ext[:, 7] = 1e20
# shadow rays to area-sampled emitter points
cum, ltri = scene.lights[:, 0], scene.lights[:, 1].view(np.uint32)
li = ltri[np.searchsorted(cum, g.uniform(size=m) * scene.total_light_area).clip(0, len(cum) - 1)]
lt = scene.tris[li]; p0 = lt[:, 0:3]; e1 = lt[:, 3:6]; e2 = lt[:, 6:9]
s1 = np.sqrt(g.uniform(size=m)); v = g.uniform(size=m) * s1; uu = 1 - s1
lp = p0 + e1 * v[:, None] + e2 * (1 - uu - v)[:, None]
o = p + 0.001 * n; dd = lp - o; dist = np.linalg.norm(dd, axis=1); dd /= dist[:, None]
keep = (dd * n).sum(axis=1) > 0
sh = np.zeros((int(keep.sum()), 8), np.float32); sh[:, 0:3] = o[keep]; sh[:, 4:7] = dd[keep]; sh[:, 7] = dist[keep] - 0.005
r.CleanUp()
print(f"{m} bounce rays, {sh.shape[0]} shadow rays (of {W * H} pixels)")


def timeit(rr, fn, d_rays, nrays, d_out, reps=8):
    ms = []
    for _ in range(reps):
        fn(d_rays.data_ptr(), nrays, d_out.data_ptr()); ms.append(rr.LastDeviceMs())
    return float(np.median(ms[2:]))


for name, batch, any_hit in (("extend (bounce 1 proxy)", ext, False), ("shadow (bounce 0 proxy)", sh, True)):
    nrays = batch.shape[0]
    variants = {"queue order": batch, "random order": batch[g.permutation(nrays)]}
    for vname, arr in variants.items():
        d_rays = torch.from_numpy(np.ascontiguousarray(arr)).cuda()
        d_out = torch.empty((nrays, 4), dtype=torch.float32, device="cuda")
        for order, bits in ((-1, 0), (1, 3), (1, 4), (1, 5), (1, 6)):
            os.environ["PT_RAY_ORDER_BITS"] = str(bits or 6); os.environ["PT_RAY_ORDER_PER_BIN"] = "1"
            rr_ = renderer(order)
            fn = rr_.TraceAnyDevice if any_hit else rr_.TraceClosestDevice
            ms = timeit(rr_, fn, d_rays, nrays, d_out)
            print(f"{name:26s} {vname:13s} {'natural' if order < 0 else f'binned {bits} bits/axis'}: {ms:.3f} ms = {nrays / ms / 1e6:.2f} Grays/s")
            rr_.CleanUp()
