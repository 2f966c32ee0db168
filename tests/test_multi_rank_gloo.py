"""CPU, world_size 2 over gloo: the host side of the multi-GPU path -- the library's sample split and per-rank
seeds (pt_split_samples, pt_rank_seed), the NCCL id exchange through torch.distributed (gpupathtracer_b200.distributed), and
the arithmetic of one reduce per batch. The per-rank 'renderer' here is the CPU oracle (allowed in tests); the reduce itself
runs inside the library over NCCL on the GPU box (tests/test_multi_gpu.py)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpupathtracer_b200.distributed import split_samples, rank_seed, exchange_unique_id


def test_split_and_seeds():
    assert [split_samples(4096, 8, r) for r in range(8)] == [(512 * r, 512) for r in range(8)]
    assert [split_samples(10, 4, r)[1] for r in range(4)] == [3, 3, 2, 2]
    parts = [split_samples(1001, 8, r) for r in range(8)]
    assert sum(n for _, n in parts) == 1001 and all(parts[r][0] + parts[r][1] == parts[r + 1][0] for r in range(7))
    seeds = {rank_seed(0xB200, r) for r in range(8)}
    assert len(seeds) == 8 and rank_seed(5, 0) == 5


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, total_spp, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import oracle
    from common import get_scene, make_camera
    from gpupathtracer_b200 import make_rng_table
    s = get_scene("cornell_small")
    w = h = 32
    cam = make_camera(s.data, w, h).as_array()
    # the 128-byte NCCL id travels from rank 0 to the others through torch.distributed, whatever the backend
    uid = exchange_unique_id(dist, rank)
    assert isinstance(uid, bytes) and len(uid) == 128
    with open(os.path.join(out_dir, f"uid{rank}.bin"), "wb") as f:
        f.write(uid)
    spp = split_samples(total_spp, world, rank)[1]
    rng = make_rng_table(w * h, rank_seed(7, rank) & 0x7FFFFFFF)
    acc = np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam, w, h, rng, acc, n_frames=spp, max_depth=5, n_threads=1)
    t = torch.from_numpy(acc)
    dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)          # what ncclReduce(sum, root 0) does per batch inside the library
    if rank == 0:
        np.save(os.path.join(out_dir, "reduced.npy"), t.numpy())
    np.save(os.path.join(out_dir, f"rank{rank}.npy"), acc if rank else np.zeros(1))
    dist.destroy_process_group()


def test_two_ranks_reduce_equals_sum_of_independent_renders(tmp_path):
    import oracle
    from common import get_scene, make_camera
    from gpupathtracer_b200 import make_rng_table
    world, total_spp = 2, 5
    mp.spawn(_worker, args=(world, _free_port(), total_spp, str(tmp_path)), nprocs=world, join=True)
    reduced = np.load(tmp_path / "reduced.npy")
    # recompute both ranks' shares in this process and add them in rank order
    s = get_scene("cornell_small")
    w = h = 32
    cam = make_camera(s.data, w, h).as_array()
    want = np.zeros((h, w, 4), np.float32)
    for r in range(world):
        acc = np.zeros((h, w, 4), np.float32)
        oracle.Oracle(s).render(cam, w, h, make_rng_table(w * h, rank_seed(7, r) & 0x7FFFFFFF), acc,
                                n_frames=split_samples(total_spp, world, r)[1], max_depth=5, n_threads=1)
        want += acc
    np.testing.assert_allclose(reduced[..., :3], want[..., :3], rtol=1e-6, atol=1e-7)
    assert float((reduced[..., :3] / total_spp).mean()) > 0
    assert (tmp_path / "uid0.bin").read_bytes() == (tmp_path / "uid1.bin").read_bytes()
