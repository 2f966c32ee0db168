"""One process per GPU (torchrun): the glue between torch.distributed's rendezvous and the library's own NCCL communicator.

The sample split itself -- per-rank RNG seeds, double-buffered batch accumulators, the ncclReduce per batch on its own
stream, the BVH broadcast -- lives in the library (csrc/pt_comm.cu, include/b200pt.h "several GPUs"). torch.distributed is
only the launcher-side channel that carries the 128-byte NCCL id from rank 0 to the others; MPI or a file would do as well.
"""
import os

from .api import Renderer, load_library


def env_rank_world():
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("LOCAL_RANK", "0"))


def rank_seed(seed, rank):
    """pt_rank_seed: the seed rank `rank` gives pt_seed_rng (rank 0 keeps `seed`, so a 1-GPU run is unchanged)."""
    return int(load_library().pt_rank_seed(seed, rank))


def split_samples(total_spp, world, rank):
    """pt_split_samples: (first, count) of the samples per pixel that belong to `rank` when `total_spp` is fixed."""
    import ctypes as C
    f, n = C.c_uint32(), C.c_uint32()
    load_library().pt_split_samples(total_spp, world, rank, C.byref(f), C.byref(n))
    return int(f.value), int(n.value)


def exchange_unique_id(dist, rank, make_id=None):
    """Rank 0 creates the id (pt_comm_unique_id), everybody receives it through torch.distributed (any backend)."""
    box = [(make_id or Renderer.CommUniqueId)() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def init_comm(renderer, dist=None):
    """Attach the library's communicator to `renderer` (this process's context). `dist`: an initialised torch.distributed
    module; None reads RANK / WORLD_SIZE and does nothing for a single process."""
    rank, world, _ = env_rank_world()
    if world == 1:
        return rank, world
    if dist is None:
        import torch.distributed as dist
    renderer.CommInit(exchange_unique_id(dist, rank), rank, world)
    return rank, world
