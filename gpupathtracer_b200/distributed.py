"""Multi-GPU sample split: every rank renders its share of the samples of EVERY pixel with its own
RNG table; the accumulation buffers (running sums) are added with ONE reduce per batch. Backend-agnostic
(`nccl` on GPUs over NVLink, `gloo` in the CPU tests); there is no data-path collective besides that reduce."""
import torch
import torch.distributed as dist


def split_samples(total_spp, world_size):
    """spp per rank: as even as possible, remainder to the low ranks; sums to total_spp."""
    base, rem = divmod(int(total_spp), int(world_size))
    return [base + (1 if r < rem else 0) for r in range(world_size)]


def rank_seed(seed, rank):
    """Decorrelated per-rank RNG-table seed (seed ^ golden-ratio hash of the rank)."""
    return (int(seed) ^ ((int(rank) * 0x9E3779B97F4A7C15) & 0xFFFFFFFFFFFFFFFF)) & 0xFFFFFFFFFFFFFFFF


def reduce_accum(accum, dst=0, group=None):
    """Sum the per-rank accumulation buffers into rank `dst` (in place there). accum: (H, W, 4) float32 running sums;
    alpha (forced to 1 by every store, Iterative.comp:263) is restored to 1 on the destination."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return accum
    dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    if dist.get_rank(group) == dst:
        accum[..., 3] = 1.0
    return accum


def mean_image(accum_sum, total_spp):
    """Present.frag:30 normalisation of the reduced sum."""
    return accum_sum[..., :3] / float(total_spp)
