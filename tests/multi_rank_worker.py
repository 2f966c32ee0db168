"""Worker of tests/test_multi_gpu.py::test_one_process_per_gpu: launched by torch.distributed.run, one process per GPU.
torch.distributed (gloo) only carries the NCCL id; the communicator, the BVH broadcast and the reduce are the library's."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE)); sys.path.insert(0, HERE)


def main(out_dir):
    import torch.distributed as dist
    from gpupathtracer_b200 import Renderer, Scene, scenes
    from gpupathtracer_b200.distributed import env_rank_world, init_comm
    from common import make_camera, mixed_rays
    rank, world, local = env_rank_world()
    dist.init_process_group("gloo")
    s = Scene(scenes.atrium(target_tris=20000, seed=2))
    w, h = 96, 64
    r = Renderer().Initialize(w, h, s, max_depth=5, device=local, bvh="build" if rank == 0 else "none")
    init_comm(r, dist)
    assert (r.lib.pt_comm_rank(r.ctx), r.lib.pt_comm_world(r.ctx)) == (rank, world)
    r.CommBroadcastBVH(0)                              # the tree is built once, on rank 0
    r.CommSeedRNG(0xB200)
    rays = mixed_rays(s, 50000, seed=3)
    nodes, refs = r.GetBVH()                           # non-root ranks: host mirror fetched lazily from the broadcast device copy
    hits = r.TraceClosest(rays)
    cam = make_camera(s.data, w, h)
    batches = []
    for _ in range(3):
        r.RenderBatch(cam, 4)
        batches.append(r.CommDownloadBatch())          # this rank's own sum of the batch it just rendered
    r.CommSynchronize()
    np.savez(os.path.join(out_dir, f"rank{rank}.npz"), hits=hits, nodes=nodes, refs=refs, rng=r.GetRNG(), batches=np.stack(batches),
             total=r.CommDownloadTotal(), samples=r.CommNumSamples(), reduce_ms=r.CommLastReduceMs())
    dist.barrier()
    r.CleanUp()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1])
