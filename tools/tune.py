"""Tuning sweep (GPU box): frame time of the headline workload for warp-scheduling thresholds / grid sizes."""
import itertools, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from gpupathtracer_b200 import Renderer

scene = bench.build_scene()
cam = bench.make_camera(scene, bench.WIDTH, bench.HEIGHT)
grid = {"PT_LEAF_THRESH": sys.argv[1].split(","), "PT_FETCH_THRESH": sys.argv[2].split(","), "PT_BLOCKS_PER_SM": sys.argv[3].split(",")}
first = True
for vals in itertools.product(*grid.values()):
    for k, v in zip(grid.keys(), vals):
        os.environ[k] = v
    r = Renderer().Initialize(bench.WIDTH, bench.HEIGHT, scene, max_depth=bench.MAX_DEPTH, seed=1, bvh="build" if first else "given")
    first = False
    r.RenderFrame(cam, 3); r.Synchronize()
    r.SetStageTiming(True)
    r.RenderFrame(cam, 8); r.Synchronize()
    ms = r.LastDeviceMs() / 8
    st = {k: round(v / 8, 4) for k, v in r.StageMs().items()}
    c = r.Counters()
    print(json.dumps({"cfg": dict(zip(grid.keys(), vals)), "ms_per_frame": round(ms, 4), "stage": st}), flush=True)
    r.CleanUp()
