"""N > 1: where does the one-frame-batch + reduce + present loop lose against plain chained frames? Run under torchrun.
  python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/exp_batch_e2e.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
import bench
from gpupathtracer_b200 import Renderer
from gpupathtracer_b200.distributed import init_comm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("gloo")
wl = bench.WORKLOADS["atrium"]
scene = bench.build_scene(wl)
W, H = wl["width"], wl["height"]
r = Renderer().Initialize(W, H, scene, max_depth=5, seed=1 + rank, bvh="build", device=local)
init_comm(r, dist)
cam = bench.make_camera(scene, W, H)
pinned = [torch.empty((H, W, 3), dtype=torch.uint8).pin_memory() for _ in range(2)]
n = 128
def run(label, step):
    for i in range(8): step(i)
    r.CommSynchronize(); dist.barrier()
    t0 = time.perf_counter()
    for i in range(n): step(i)
    t1 = time.perf_counter()
    r.CommSynchronize(); dist.barrier()
    t2 = time.perf_counter()
    if rank == 0:
        print(f"{label:52s} host enqueue {1e3 * (t1 - t0) / n:.3f} ms/step, wall {1e3 * (t2 - t0) / n:.3f} ms/step", flush=True)
def frames(i): cam.GenerateImagePlane(); r.RenderFrame(cam, 1)
def frames_present(i): cam.GenerateImagePlane(); r.RenderFrame(cam, 1); r.PresentAsync(pinned[i & 1].data_ptr())
def batch(i): cam.GenerateImagePlane(); r.RenderBatch(cam, 1)
def batch_present(i): cam.GenerateImagePlane(); r.RenderBatch(cam, 1); r.CommPresentAsync(pinned[i & 1].data_ptr())
def batch4_present(i): cam.GenerateImagePlane(); r.RenderBatch(cam, 4); r.CommPresentAsync(pinned[i & 1].data_ptr())
run("chained frames, no comm", frames)
run("chained frames + per-slice present, no comm", frames_present)
run("1-frame batch + reduce", batch)
run("1-frame batch + reduce + present of the total", batch_present)
n = 32
run("4-frame batch + reduce + present (per 4 frames)", batch4_present)
r.CleanUp()
dist.destroy_process_group()
