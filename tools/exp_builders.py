"""Device builders against the host SBVH on a path-tracing workload: build time, references, SAH cost, frame time.
  python tools/exp_builders.py [atrium|bathroom] [grids...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import bench
from gpupathtracer_b200 import Renderer, Scene

name = sys.argv[1] if len(sys.argv) > 1 else "atrium"
grids = [int(g) for g in sys.argv[2:]] or [32, 64, 96, 128]
wl = bench.WORKLOADS[name]
base = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
cam = bench.make_camera(base, W, H)
cases = [("host SBVH", "build", None), ("device LBVH", "device", None), ("device PLOC", "ploc", None), ("device SAH", "sah", None)]
for g in grids:
    cases += [(f"split {g} + PLOC", "split", g), (f"split {g} + LBVH", "split_lbvh", g)]
if os.environ.get('EXP_ONLY'):
    cases = [c for c in cases if c[1] in os.environ['EXP_ONLY'].split(',')]
for label, how, g in cases:
    if g:
        os.environ["PT_SPLIT_GRID"] = str(g)
    s = Scene(base.data)
    r = Renderer().Initialize(W, H, s, max_depth=wl["depth"], microfacet_model=wl["model"], seed=1, bvh=how)
    st = r.bvh_stats
    if os.environ.get("EXP_OPT"):
        it, fr = os.environ["EXP_OPT"].split(","); o = r.OptimizeBVH(int(it), float(fr))
        label = label + f" + opt {it}x{fr}"; st.sah_cost = o.sah_cost; st.max_depth = o.max_depth; st.total_ms += o.total_ms
    for _ in range(3):
        r.RenderFrame(cam, 8); r.Synchronize()
    ms = []
    for _ in range(5):
        r.RenderFrame(cam, 32); r.Synchronize(); ms.append(r.LastDeviceMs() / 32)
    c0 = r.Counters(); r.RenderFrame(cam, 4); r.Synchronize(); c1 = r.Counters()
    rays = (c1["closest_rays"] + c1["shadow_rays"] - c0["closest_rays"] - c0["shadow_rays"]) / 4
    ms = sorted(ms)[2]
    print(f"{label:22s} build {st.build_ms:8.2f} ms (call {st.total_ms:8.1f}) refs {st.n_refs:8d} ({st.n_refs / s.tris.shape[0]:.3f}x) depth {st.max_depth:3d} "
          f"SAH {st.sah_cost:7.1f}  frame {ms:.3f} ms = {rays / ms / 1e3:.0f} Mrays/s", flush=True)
    r.CleanUp()
