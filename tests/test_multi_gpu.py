"""-m gpu, >= 2 devices: the sample split across GPUs inside the library (csrc/pt_comm.cu) -- one process driving n GPUs
(pt_create_multi) and one process per GPU (pt_comm_init). Checked: the reduced total is exactly the sum of the ranks' own
batch sums, ranks use different random numbers, the combined image agrees with a 1-GPU render at equal total spp within the
noise floor, the tree is built once and arrives intact. Skipped on a 1-GPU box (the driver's N > 1 runs exercise the same
code through bench.py)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from gpupathtracer_b200 import MultiRenderer, Renderer, load_library
from common import get_scene, make_camera, rel_mse

pytestmark = pytest.mark.gpu


def _n_devices():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_n_devices() < 2, reason="needs 2 GPUs")


def test_group_of_one_is_a_plain_renderer():
    s = get_scene("cornell_small")
    w, h = 64, 48
    cam = make_camera(s.data, w, h)
    m = MultiRenderer().Initialize(w, h, s, 1, seed=5, max_depth=5)
    m.RenderFrame(cam, 3); m.Synchronize()
    r = Renderer().Initialize(w, h, s, max_depth=5, seed=5)
    r.RenderFrame(cam, 3)
    assert m.GetNumSamples() == 3
    assert np.array_equal(np.nan_to_num(m.DownloadAccum(), nan=-1.0), np.nan_to_num(r.DownloadAccum(), nan=-1.0))
    m.CleanUp(); r.CleanUp()


@needs2
def test_two_gpus_one_process_total_is_the_sum_of_the_ranks():
    s = get_scene("atrium_small")
    w, h, frames = 96, 64, 8
    cam = make_camera(s.data, w, h)
    m = MultiRenderer().Initialize(w, h, s, 2, seed=0xB200, max_depth=5)
    want = np.zeros((h, w, 4), np.float32)
    for b in range(3):
        m.RenderFrame(cam, frames)
        parts = [m.rank(i).CommDownloadBatch() for i in range(2)]
        want[..., :3] = want[..., :3] + (parts[0][..., :3] + parts[1][..., :3])     # ncclReduce(sum) of 2 ranks, then total += batch
        assert not np.array_equal(parts[0], parts[1])                                # different random numbers per rank
    m.Synchronize()
    total = m.DownloadAccum()
    assert m.GetNumSamples() == 3 * frames * 2
    assert np.array_equal(np.nan_to_num(total[..., :3], nan=-1.0), np.nan_to_num(want[..., :3], nan=-1.0))
    assert not np.array_equal(m.rank(0).GetRNG(), m.rank(1).GetRNG())
    # against 1-GPU renders at the same total spp: within 2 x the seed-to-seed noise floor
    spp = m.GetNumSamples()
    singles = []
    for seed in (1, 2):
        r = Renderer().Initialize(w, h, s, max_depth=5, seed=seed)
        r.RenderFrame(cam, spp)
        singles.append(r.DownloadAccum()[..., :3] / spp)
        r.CleanUp()
    floor = rel_mse(singles[0], singles[1])
    err = 0.5 * (rel_mse(total[..., :3] / spp, singles[0]) + rel_mse(total[..., :3] / spp, singles[1]))
    print(f"2 GPUs vs 1 GPU at {spp} spp: relMSE {err:.3e}, noise floor {floor:.3e}")
    assert err <= 2.0 * floor
    # Present of the combined image, then ResetSamples
    out = np.zeros((h, w, 3), np.uint8)
    m.PresentAsync(out.ctypes.data); m.Synchronize()
    assert out.mean() > 5
    m.ResetSamples(); m.Synchronize()
    assert m.GetNumSamples() == 0 and not m.DownloadAccum()[..., :3].any()
    m.CleanUp()


@needs2
def test_one_process_per_gpu(tmp_path):
    """pt_comm_init under torch.distributed.run: NCCL id through the launcher's channel, BVH built on rank 0 only and
    broadcast, one reduce per batch; rank 0's total == the sum over batches of the sum over ranks."""
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    worker = os.path.join(os.path.dirname(os.path.abspath(__file__)), "multi_rank_worker.py")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), worker, str(tmp_path)]
    p = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-4000:]
    z = [np.load(tmp_path / f"rank{r}.npz") for r in range(2)]
    assert np.array_equal(z[0]["hits"].view(np.uint32), z[1]["hits"].view(np.uint32))       # the broadcast tree traces like the built one
    assert np.array_equal(z[0]["nodes"].view(np.uint32), z[1]["nodes"].view(np.uint32)) and np.array_equal(z[0]["refs"], z[1]["refs"])   # link words are ints: compare bits
    assert not np.array_equal(z[0]["rng"], z[1]["rng"])
    want = np.zeros_like(z[0]["total"][..., :3])
    for b in range(3):
        want = want + (z[0]["batches"][b][..., :3] + z[1]["batches"][b][..., :3])
    assert np.array_equal(np.nan_to_num(z[0]["total"][..., :3], nan=-1.0), np.nan_to_num(want, nan=-1.0))
    assert int(z[0]["samples"]) == 3 * 4 * 2 and float(z[0]["reduce_ms"]) >= 0.0
    assert not z[1]["total"].any()                                                              # only rank 0 holds the total
