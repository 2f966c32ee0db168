"""CPU, world_size 2 over gloo: the sample-split + single-reduce logic of the multi-GPU path.
The per-rank 'renderer' here is the CPU oracle (allowed in tests); the code under test is gpupathtracer_b200.distributed."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from gpupathtracer_b200.distributed import split_samples, rank_seed, reduce_accum, mean_image


def test_split_and_seeds():
    assert split_samples(4096, 8) == [512] * 8
    assert split_samples(10, 4) == [3, 3, 2, 2] and sum(split_samples(1001, 8)) == 1001
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
    spp = split_samples(total_spp, world)[rank]
    rng = make_rng_table(w * h, rank_seed(7, rank) & 0x7FFFFFFF)
    acc = np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam, w, h, rng, acc, n_frames=spp, max_depth=5, n_threads=1)
    t = torch.from_numpy(acc)
    reduce_accum(t, dst=0)
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
                                n_frames=split_samples(total_spp, world)[r], max_depth=5, n_threads=1)
        want += acc
    want[..., 3] = 1.0
    np.testing.assert_allclose(reduced, want, rtol=1e-6, atol=1e-7)
    img = mean_image(torch.from_numpy(reduced), total_spp)
    assert img.shape == (h, w, 3) and float(img.mean()) > 0
