"""-m gpu, collected last on purpose: pt_trace_closest_device_noalloc, the binned closest-hit trace
whose scratch lives in a fixed __device__ array inside the library (csrc/pt_noalloc.cu; no counterpart in the reference).
Called through the C ABI on device buffers, two calls in a row, against the existing
pt_trace_closest_device: ray traversal does not depend on the processing order, so the bar is the one the traversal tests use --
identical bits in (t, u, v, triangle id)."""
import ctypes as C

import numpy as np
import pytest

from gpupathtracer_b200 import Renderer, load_library
from common import get_scene, mixed_rays

pytestmark = pytest.mark.gpu


def _device_pair(torch, rays):
    d_rays = torch.from_numpy(np.ascontiguousarray(rays)).cuda()
    d_hits = torch.full((rays.shape[0], 4), -7.0, dtype=torch.float32, device="cuda")
    return d_rays, d_hits


def test_no_alloc_trace_matches_the_allocating_entry_point_twice_in_a_row():
    import torch
    s = get_scene("atrium_small")
    r = Renderer().Initialize(64, 64, s)
    lib = load_library()
    assert lib.pt_noalloc_max_rays() == 1 << 20
    # two different batches, the second one SMALLER: whatever the first call left in the shared scratch must not matter
    batches = [mixed_rays(s, 300000, seed=61), mixed_rays(s, 70001, seed=62)]
    results = []
    free_before = None
    for k, rays in enumerate(batches):
        d_rays, d_hits = _device_pair(torch, rays)
        torch.cuda.synchronize()
        if k == 1:
            free_before = torch.cuda.mem_get_info()[0]            # the first call already loaded the kernels
        r.TraceClosestDeviceNoAlloc(d_rays.data_ptr(), rays.shape[0], d_hits.data_ptr())
        r.Synchronize()
        if k == 1:
            assert torch.cuda.mem_get_info()[0] == free_before, "device memory in use changed during the no-alloc call"
        results.append(d_hits.cpu().numpy())
    for rays, got in zip(batches, results):
        d_rays, d_want = _device_pair(torch, rays)
        r.TraceClosestDevice(d_rays.data_ptr(), rays.shape[0], d_want.data_ptr())
        r.Synchronize()
        want = d_want.cpu().numpy()
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
        assert (want[:, 3].view(np.int32) >= 0).mean() > 0.2 and not (got == -7.0).all(axis=1).any()
    # and once more with the first batch: same bits as the first time
    d_rays, d_hits = _device_pair(torch, batches[0])
    r.TraceClosestDeviceNoAlloc(d_rays.data_ptr(), batches[0].shape[0], d_hits.data_ptr())
    r.Synchronize()
    assert np.array_equal(d_hits.cpu().numpy().view(np.uint32), results[0].view(np.uint32))
    r.CleanUp()


def test_no_alloc_trace_refuses_what_it_cannot_hold():
    import torch
    s = get_scene("cornell_small")
    r = Renderer().Initialize(64, 64, s)
    lib = load_library()
    d = torch.zeros(64, dtype=torch.float32, device="cuda")
    rc = lib.pt_trace_closest_device_noalloc(r.ctx, C.c_void_p(d.data_ptr()), (1 << 20) + 1, C.c_void_p(d.data_ptr()))
    assert rc == -6 and b"PT_NOALLOC_MAX_RAYS" in lib.pt_last_error()          # PT_ERR_LIMIT, nothing launched
    assert lib.pt_trace_closest_device_noalloc(r.ctx, None, 8, C.c_void_p(d.data_ptr())) == -1
    assert lib.pt_trace_closest_device_noalloc(r.ctx, C.c_void_p(d.data_ptr()), 0, C.c_void_p(d.data_ptr())) == 0   # empty batch
    # a ragged little batch still works after the refusals
    rays = mixed_rays(s, 33, seed=9)
    d_rays, d_hits = _device_pair(torch, rays)
    r.TraceClosestDeviceNoAlloc(d_rays.data_ptr(), 33, d_hits.data_ptr()); r.Synchronize()
    assert np.array_equal(d_hits.cpu().numpy().view(np.uint32), r.TraceClosest(rays).view(np.uint32))
    r.CleanUp()
