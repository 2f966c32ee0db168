"""-m gpu: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs."""
import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import Renderer, make_rng_table, REFERENCE_MAX_DEPTH
from common import get_scene, make_camera, mixed_rays, rel_mse

pytestmark = pytest.mark.gpu


def _renderer(scene, w=64, h=64, **kw):
    return Renderer().Initialize(w, h, scene, **kw)


@pytest.mark.parametrize("name,n", [("cornell", 200000), ("soup", 200000), ("bunny", 300000), ("atrium_small", 200000)])
def test_closest_hit_bit_exact(name, n):
    s = get_scene(name)
    rays = mixed_rays(s, n)
    want, _ = oracle.Oracle(s).trace_closest(rays)
    r = _renderer(s)
    got = r.TraceClosest(rays)
    r.CleanUp()
    assert np.array_equal(got[:, 3].view(np.int32), want[:, 3].view(np.int32)), "hit triangle ids differ"
    # same arithmetic, same order: t,u,v are expected to be BIT-identical (the stated bar is 1e-5 relative)
    assert np.array_equal(got[:, :3].view(np.uint32), want[:, :3].view(np.uint32))
    assert (want[:, 3].view(np.int32) >= 0).mean() > 0.2


@pytest.mark.parametrize("name,n", [("cornell", 200000), ("soup", 200000), ("bunny", 300000), ("atrium_small", 200000)])
def test_any_hit_bit_exact(name, n):
    s = get_scene(name)
    rays = mixed_rays(s, n, any_hit=True)
    want, _ = oracle.Oracle(s).trace_any(rays)
    r = _renderer(s)
    got = r.TraceAny(rays)
    r.CleanUp()
    assert np.array_equal(got, want)
    assert 0.05 < want.mean() < 0.95


@pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built (it travels with the snapshot when it was)")
@pytest.mark.parametrize("name,n", [("cornell", 200000), ("soup", 200000), ("bunny", 300000), ("atrium_small", 200000)])
def test_cuda_traversal_equals_reference_shader_source(name, n):
    """The CUDA kernels against the reference's OWN traversal shader source compiled as C++ (oracle/ref_glsl.cpp: IfIfClosestHit /
    IfIfAnyHit and their callees extracted from common/BVH.glsl + common/Geometry.glsl), no restatement in between: same
    triangle index, same (t, u, v) bits, same occlusion flag for every ray."""
    s = get_scene(name)
    g = oracle.RefGlsl(s.nodes, s.refs, s.tris)
    rays = mixed_rays(s, n, seed=41)
    anyr = mixed_rays(s, n, seed=42, any_hit=True)
    want = g.trace_closest(rays)
    wflag = g.trace_any(anyr)
    r = _renderer(s)
    got = r.TraceClosest(rays)
    gflag = r.TraceAny(anyr)
    r.CleanUp()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(gflag, wflag)


def test_empty_and_ragged_batches():
    s = get_scene("cornell_small")
    r = _renderer(s)
    assert r.TraceClosest(np.zeros((0, 8), np.float32)).shape == (0, 4)
    rays = mixed_rays(s, 33)                       # not a multiple of the 32-ray packet
    want, _ = oracle.Oracle(s).trace_closest(rays)
    assert np.array_equal(r.TraceClosest(rays).view(np.uint32), want.view(np.uint32))
    r.CleanUp()


def test_raygen_matches_oracle():
    s = get_scene("cornell_small")
    w, h = 64, 40
    cam = make_camera(s.data, w, h, aperture=0.3)
    rng = make_rng_table(w * h, 5)
    r = _renderer(s, w, h, rng_state=rng)
    r.lib.pt_set_camera(r.ctx, cam.plane)
    got = r.RaygenProbe(0, w * h)
    rng_after = r.GetRNG()
    r.CleanUp()
    rng_o = rng.copy()
    want = oracle.Oracle(s).raygen(cam.as_array(), w, h, rng_o, 0, w * h)
    assert np.array_equal(got[:, :2].view(np.int32), want[:, :2].view(np.int32))      # Morton pixel order
    assert np.array_equal(rng_after, rng_o)                                            # RNG advanced identically
    # sin/cos differ by <= 2 ulp between CUDA libm and glibc: 1e-6 absolute on unit directions / lens offsets
    np.testing.assert_allclose(got[:, 2:8], want[:, 2:8], rtol=0, atol=2e-6)


@pytest.mark.parametrize("name,depth,model", [("cornell", REFERENCE_MAX_DEPTH, 0), ("cornell", 5, 0), ("soup_small", 5, 1),
                                              ("atrium_small", 5, 0)])
def test_frames_match_oracle(name, depth, model):
    s = get_scene(name)
    w, h, frames = 64, 64, 4
    cam = make_camera(s.data, w, h)
    rng = make_rng_table(w * h, 9)
    r = _renderer(s, w, h, max_depth=depth, microfacet_model=model, rng_state=rng)
    r.RenderFrame(cam, frames)
    got = r.DownloadAccum()
    rng_got = r.GetRNG()
    cnt = r.Counters()
    r.CleanUp()
    rng_o = rng.copy()
    want = np.zeros((h, w, 4), np.float32)
    c = oracle.Oracle(s).render(cam.as_array(), w, h, rng_o, want, n_frames=frames, max_depth=depth, model=model)
    # Per-pixel: identical control flow unless a 1-2 ulp libm difference (sin/cos/log/exp) flips a decision.
    fin = np.isfinite(want[..., :3]).all(axis=2) & np.isfinite(got[..., :3]).all(axis=2)
    close = np.isclose(got[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2)
    frac = float((close & fin).sum()) / float(fin.sum())
    assert frac > 0.995, f"only {frac:.5f} of pixels agree to 1e-4"
    same_rng = float((rng_got == rng_o).all(axis=1).mean())
    assert same_rng > 0.995, f"only {same_rng:.5f} of per-pixel RNG states agree"
    assert np.array_equal(np.isnan(got[..., 0]), np.isnan(want[..., 0])) or frac > 0.999
    # ray counters (device-side) agree with the oracle's to the same tolerance
    assert abs(cnt["closest_rays"] - c["closest_rays"]) <= 0.005 * c["closest_rays"]
    # the oracle, like the reference, traces every NEE shadow ray; the CUDA path skips those whose radiance is exactly 0
    assert abs(cnt["shadow_rays"] + cnt["null_shadow_rays"] - c["shadow_rays"]) <= 0.005 * c["shadow_rays"]
    assert abs(cnt["null_shadow_rays"] - c["null_shadow_rays"]) <= 0.005 * c["shadow_rays"]
    assert rel_mse(got[..., :3], want[..., :3]) < 1e-3


@pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built (it travels with the snapshot when it was)")
@pytest.mark.parametrize("name,textured", [("cornell", False), ("cornell", True), ("atrium_small", True)])
def test_cuda_frames_equal_reference_main(name, textured):
    """RenderFrame against the reference's OWN main() (Iterative.comp:214-307 compiled as C++, oracle/ref_glsl.cpp), no
    restatement in between, at the shader's own path length: the same tolerance as test_frames_match_oracle (a 1-2 ulp libm
    difference in sin / cos / log / exp can flip a decision in a few pixels; everywhere else image and RNG words agree)."""
    from common import with_textures_and_sky
    s = get_scene(name)
    if textured:
        s = with_textures_and_sky(s)
    w, h, frames = 64, 64, 2
    cam = make_camera(s.data, w, h, aperture=0.1)
    rng = make_rng_table(w * h, 23)
    r = _renderer(s, w, h, max_depth=REFERENCE_MAX_DEPTH, microfacet_model=0, rng_state=rng)
    r.RenderFrame(cam, frames)
    got = r.DownloadAccum()
    rng_got = r.GetRNG()
    r.CleanUp()
    rng_ref = rng.copy()
    want = np.zeros((h, w, 4), np.float32)
    o = oracle.Oracle(s)
    for _ in range(frames):
        o.ref_glsl_render(cam.as_array(), w, h, rng_ref, want, 0)
    assert np.isfinite(want).all()
    close = np.isclose(got[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2)
    assert close.mean() > 0.995, f"only {close.mean():.5f} of pixels agree to 1e-4"
    assert float((rng_got == rng_ref).all(axis=1).mean()) > 0.995
    assert rel_mse(got[..., :3], want[..., :3]) < 1e-3


def test_reset_and_accumulate():
    s = get_scene("cornell_small")
    w, h = 32, 32
    cam = make_camera(s.data, w, h)
    r = _renderer(s, w, h, max_depth=5, seed=3)
    r.RenderFrame(cam, 2)
    assert r.GetNumSamples() == 2
    a = r.DownloadAccum()
    assert np.all(a[..., 3] == 1.0) and np.nansum(a[..., :3]) > 0
    r.ResetSamples()
    assert r.GetNumSamples() == 0
    assert np.all(r.DownloadAccum() == 0)
    r.RenderFrame(cam, 1)
    img = r.Present()
    assert img.shape == (h, w, 3) and np.isfinite(img).mean() > 0.99
    r.CleanUp()


def test_present_matches_present_frag():
    """k_resolve against Present.frag (compiled as C++ when oracle/_ref travels, else the oracle's restatement, which is held to
    it bit for bit): the float image to a few ulp (CUDA's powf against glibc's), the 8-bit image to one step at rounding
    boundaries. Everything before the pow is IEEE arithmetic without FMA on both sides."""
    s = get_scene("cornell")
    w, h = 64, 48
    cam = make_camera(s.data, w, h)
    r = _renderer(s, w, h, max_depth=5, seed=5)
    r.RenderFrame(cam, 3)
    a = r.DownloadAccum()
    n = r.GetNumSamples()
    got = r.Present()
    got8 = r.Present(as_u8=True)
    r.CleanUp()
    assert n == 3 and np.isfinite(a).all() and (a[..., :3] > 0).any()
    want = oracle.ref_glsl_present(a, n) if oracle.have_glsl() else oracle.tonemap(a[..., :3], n).reshape(h, w, 3)
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-7)
    want8 = np.floor(np.clip(want, 0.0, 1.0) * np.float32(255.0) + np.float32(0.5)).astype(np.int32)
    d = np.abs(got8.astype(np.int32) - want8)
    assert d.max() <= 1 and (d == 0).mean() > 0.999


def test_slices_do_not_change_the_image():
    """pt_config.n_slices only changes scheduling (pixel-disjoint slices on concurrent streams): the accumulation
    buffer and the RNG table must be BIT-identical for 1, 3 and 4 slices, and ragged slice sizes must work."""
    s = get_scene("cornell")
    w, h, frames = 72, 40, 3                       # 2880 px: 3 slices of 960, 4 slices of 768 (last one ragged: 576)
    cam = make_camera(s.data, w, h)
    rng = make_rng_table(w * h, 21)
    outs = []
    for n_slices in (1, 3, 4):
        r = _renderer(s, w, h, max_depth=5, rng_state=rng, n_slices=n_slices)
        r.RenderFrame(cam, frames)
        outs.append((r.DownloadAccum(), r.GetRNG(), r.Counters()))
        r.CleanUp()
    for a, g, c in outs[1:]:
        assert np.array_equal(a.view(np.uint32), outs[0][0].view(np.uint32))
        assert np.array_equal(g, outs[0][1])
        assert c["closest_rays"] == outs[0][2]["closest_rays"] and c["shadow_rays"] == outs[0][2]["shadow_rays"]
        assert c["paths"] == frames * w * h


def test_present_async_matches_present():
    """pt_present_async (resolve on the render stream, D2H on the copy stream, double-buffered) must deliver the same
    8-bit image as the blocking Present, frame after frame, while later frames are already queued."""
    import ctypes as C
    s = get_scene("cornell_small")
    w, h = 64, 48
    cam = make_camera(s.data, w, h)
    r = _renderer(s, w, h, max_depth=5, seed=5)
    bufs = [np.zeros((h, w, 3), np.uint8) for _ in range(3)]
    for i in range(3):
        r.RenderFrame(cam, 1)
        r.PresentAsync(bufs[i].ctypes.data)
    r.Synchronize()
    r2 = _renderer(s, w, h, max_depth=5, seed=5)
    for i in range(3):
        r2.RenderFrame(cam, 1)
        assert np.array_equal(r2.Present(as_u8=True), bufs[i]), f"frame {i}"
    r.CleanUp(); r2.CleanUp()


def test_checkpoint_resume_is_bit_identical(tmp_path):
    """3 frames + save + 5 frames == 3 frames, restart in a fresh context, load, 5 frames."""
    s = get_scene("cornell_small")
    w, h = 64, 48
    cam = make_camera(s.data, w, h)
    ck = str(tmp_path / "render.ckpt")
    r = _renderer(s, w, h, max_depth=5, seed=11)
    r.RenderFrame(cam, 3)
    r.SaveCheckpoint(ck)
    r.RenderFrame(cam, 5)
    want, want_rng = r.DownloadAccum(), r.GetRNG()
    r.CleanUp()
    assert not (tmp_path / "render.ckpt.part").exists()

    r2 = _renderer(s, w, h, max_depth=5, seed=999)         # different seed: everything must come from the file
    r2.LoadCheckpoint(ck)
    assert r2.GetNumSamples() == 3
    r2.RenderFrame(cam, 5)
    assert r2.GetNumSamples() == 8
    got, got_rng = r2.DownloadAccum(), r2.GetRNG()
    assert np.array_equal(got_rng, want_rng)
    assert np.array_equal(np.nan_to_num(got, nan=-1.0), np.nan_to_num(want, nan=-1.0))
    # wrong size and damaged files are refused, the context stays usable
    r3 = _renderer(s, 32, 32, max_depth=5, seed=1)
    with pytest.raises(Exception):
        r3.LoadCheckpoint(ck)
    data = open(ck, "rb").read()
    bad = tmp_path / "bad.ckpt"
    bad.write_bytes(data[:-7])
    with pytest.raises(Exception):
        r2.LoadCheckpoint(str(bad))
    flipped = bytearray(data); flipped[len(data) // 2] ^= 0x40
    bad.write_bytes(bytes(flipped))
    with pytest.raises(Exception):
        r2.LoadCheckpoint(str(bad))
    r2.RenderFrame(cam, 1)
    assert r2.GetNumSamples() == 9
    r2.CleanUp(); r3.CleanUp()


def test_null_radiance_shadow_samples_need_no_ray():
    """An NEE sample whose radiance is exactly 0 (light below the horizon) cannot change the pixel: by default it is settled
    without a shadow ray. With trace_null_shadow the reference's ray population is traced; image and RNG are the same bits."""
    s = get_scene("atrium_small")
    w, h = 64, 64
    cam = make_camera(s.data, w, h)
    rng = make_rng_table(w * h, 17)
    res = []
    for trace_all in (False, True):
        r = _renderer(s, w, h, max_depth=5, rng_state=rng, trace_null_shadow=trace_all)
        r.RenderFrame(cam, 4)
        res.append((r.DownloadAccum(), r.GetRNG(), r.Counters()))
        r.CleanUp()
    (a0, g0, c0), (a1, g1, c1) = res
    assert np.array_equal(np.nan_to_num(a0, nan=-1.0), np.nan_to_num(a1, nan=-1.0)) and np.array_equal(g0, g1)
    assert c1["null_shadow_rays"] == 0 and c0["null_shadow_rays"] > 0.1 * c1["shadow_rays"]
    assert c0["shadow_rays"] + c0["null_shadow_rays"] == c1["shadow_rays"] and c0["closest_rays"] == c1["closest_rays"]
