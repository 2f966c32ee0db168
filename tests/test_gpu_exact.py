"""-m gpu: EQUALITY with the reference, at test sizes and at the BASELINE.json sizes.

With pt_config.parity_flags = PT_PARITY_LIBM the CUDA path evaluates GLSL's sin / cos / log / exp / pow through
csrc/pt_parity_libm.h, the header the reference's shader source is compiled with here (oracle/ref_glsl.cpp, model codes 4 / 6), so
the accumulation image and the RNG table must be the SAME BITS as the reference's own main() produces -- not "99.5 % of pixels".
The default build (CUDA's fast libm) is held to a stated mismatch fraction next to it."""
import functools
import os

import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import Renderer, Scene, scenes, make_rng_table, REFERENCE_MAX_DEPTH
from common import get_scene, make_camera, mixed_rays, rel_mse, with_textures_and_sky

pytestmark = pytest.mark.gpu
needs_glsl = pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built (it travels with the snapshot when it was)")


def same_bits(a, b):
    """Equal bit patterns; a NaN equals a NaN (x86 and the GPU sign their default NaN differently)."""
    a = np.ascontiguousarray(a); b = np.ascontiguousarray(b)
    eq = a.view(np.uint32) == b.view(np.uint32)
    if a.dtype == np.float32:
        eq |= np.isnan(a) & np.isnan(b)
    return eq


def test_parity_libm_device_equals_host():
    """pt_parity_libm.h compiled by nvcc for sm_100a and by g++ for the host gives the same bits on every argument."""
    from test_parity_libm import libm_cases
    s = get_scene("cornell_small")
    r = Renderer().Initialize(64, 64, s)
    for fn, c in libm_cases(n=300000, seed=5).items():
        x, y = c if fn == "pow" else (c, None)
        got = r.ParityLibmProbe(fn, x, y)
        want = oracle.parity_libm_probe(fn, x, y)
        assert same_bits(got, want).all(), fn
    r.CleanUp()


def _gpu_frames(s, w, h, cam, rng, frames, **kw):
    r = Renderer().Initialize(w, h, s, rng_state=rng, **kw)
    r.RenderFrame(cam, frames)
    out = r.DownloadAccum(), r.GetRNG(), r.Counters()
    r.CleanUp()
    return out


@needs_glsl
@pytest.mark.parametrize("name,model,textured,aperture", [("cornell", 0, False, 0.0), ("cornell", 0, True, 0.3), ("atrium_small", 0, True, 0.1),
                                                          ("cornell", 1, True, 0.3), ("atrium_small", 1, False, 0.25), ("soup", 1, True, 0.0)])
def test_frames_equal_reference_main_bit_for_bit(name, model, textured, aperture):
    """RenderFrame x 3 under PT_PARITY_LIBM == three dispatches of the reference's main() (Iterative.comp:214-307 compiled as C++
    with the same libm header; Beckmann through the IEEE maxNum build) at the shader's own path length: every
    pixel of the accumulation image and every RNG word."""
    s = get_scene(name)
    if textured:
        s = with_textures_and_sky(s)
    w, h, frames = 64, 64, 3
    cam = make_camera(s.data, w, h, aperture=aperture)
    rng = make_rng_table(w * h, 31)
    got, rng_got, _ = _gpu_frames(s, w, h, cam, rng, frames, max_depth=REFERENCE_MAX_DEPTH, microfacet_model=model, parity_libm=True)
    rng_ref, want = rng.copy(), np.zeros((h, w, 4), np.float32)
    o = oracle.Oracle(s)
    for _ in range(frames):
        o.ref_glsl_render(cam.as_array(), w, h, rng_ref, want, 6 if model else 4)
    assert same_bits(got, want).all(), f"{(~same_bits(got, want).all(axis=2)).sum()} pixels differ"
    assert np.array_equal(rng_got, rng_ref)
    assert np.isfinite(want[..., :3]).mean() > 0.99 and (want[..., :3] > 0).any()


@pytest.mark.parametrize("name,model,depth,sun", [("cornell", 0, 5, False), ("soup_small", 1, 5, False), ("cornell", 1, 7, True)])
def test_frames_equal_oracle_bit_for_bit(name, model, depth, sun):
    """The same against the restated oracle (pinned to the reference's source, tests/test_parity_libm.py) where the compiled
    shader has no build: BASELINE's depth-5 path length and the SUN_LIGHT branch."""
    s = get_scene(name)
    w, h, frames = 64, 48, 4
    cam = make_camera(s.data, w, h, aperture=0.2)
    rng = make_rng_table(w * h, 77)
    r = Renderer().Initialize(w, h, s, rng_state=rng, max_depth=depth, microfacet_model=model, parity_libm=True, sun_light=sun)
    sun_arg = None
    if sun:
        d = np.array([0.3, 0.9, 0.2], np.float32); d /= np.linalg.norm(d)
        r.SetSun(d, 0.08); sun_arg = (d, 0.08)
    r.RenderFrame(cam, frames)
    got, rng_got = r.DownloadAccum(), r.GetRNG()
    r.CleanUp()
    rng_o, want = rng.copy(), np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam.as_array(), w, h, rng_o, want, n_frames=frames, max_depth=depth, model=model, sun=sun_arg, parity_libm=True)
    assert same_bits(got, want).all() and np.array_equal(rng_got, rng_o)


@needs_glsl
def test_default_libm_differs_on_few_pixels_only():
    """The default build keeps CUDA's sincosf / logf / expf. It must differ from the reference (glibc's) ONLY through them: the
    mismatch fraction is small, and switching both sides to the parity libm removes it entirely (the tests above)."""
    s = with_textures_and_sky(get_scene("atrium_small"))
    w, h, frames = 96, 64, 4
    cam = make_camera(s.data, w, h, aperture=0.1)
    rng = make_rng_table(w * h, 5)
    got, rng_got, _ = _gpu_frames(s, w, h, cam, rng, frames, max_depth=REFERENCE_MAX_DEPTH, microfacet_model=0)
    rng_ref, want = rng.copy(), np.zeros((h, w, 4), np.float32)
    o = oracle.Oracle(s)
    for _ in range(frames):
        o.ref_glsl_render(cam.as_array(), w, h, rng_ref, want, 0)
    differ = 1.0 - float(np.isclose(got[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2).mean())
    assert differ < 5e-3, differ
    assert rel_mse(got[..., :3], want[..., :3]) < 1e-3


# ------------------------------------------------------------------------------------------------ BASELINE.json sizes
@functools.lru_cache(maxsize=None)
def big_scene(which):
    """configs[1] (atrium, 262 144 triangles) / configs[2] (glossy interior, ~1.0 M triangles, textures on), as bench.py builds them."""
    if which == "atrium":
        return Scene(scenes.atrium(target_tris=262144, seed=2, aperture=0.0, glossy=False))
    return with_textures_and_sky(Scene(scenes.atrium(target_tris=1000000, seed=3, aperture=0.5, glossy=True)))


def _big_renderer(which, w, h, **kw):
    s = big_scene(which)
    r = Renderer().Initialize(w, h, s, bvh="build" if s.nodes is None else "given", **kw)    # the library's own SBVH builder
    return s, r


@needs_glsl
def test_headline_1080p_frame_equals_reference_main():
    """BASELINE.json configs[1] at full size: one 1920 x 1080 frame of the 262 144-triangle atrium, depth 5, on the library's own
    SBVH tree, against the reference's main() on all host cores (ref_glsl_render_frames_mp). Bit-identical image and RNG table
    under the parity libm; the default (fast-libm) build differs on < 1e-3 of the pixels."""
    w, h, depth = 1920, 1080, 5
    s, r = _big_renderer("atrium", w, h, max_depth=depth, microfacet_model=0, parity_libm=True, seed=1)
    rng = r.GetRNG()
    cam = make_camera(s.data, w, h)
    r.RenderFrame(cam, 1)
    got, rng_got, cnt = r.DownloadAccum(), r.GetRNG(), r.Counters()
    r.CleanUp()
    o = oracle.Oracle(s)
    rng_ref, want = rng.copy(), np.zeros((h, w, 4), np.float32)
    o.ref_glsl_render_mp(cam.as_array(), w, h, rng_ref, want, n_frames=1, max_depth=depth, model=4)
    eq = same_bits(got, want).all(axis=2)
    assert eq.all(), f"{(~eq).sum()} of {w * h} pixels differ"
    assert np.array_equal(rng_got, rng_ref)
    assert cnt["paths"] == w * h and (want[..., :3] > 0).mean() > 0.1
    # the fast-libm default against the stock reference (glibc): same frame, same table
    s, r = _big_renderer("atrium", w, h, max_depth=depth, microfacet_model=0, rng_state=rng)
    r.RenderFrame(cam, 1)
    got_d = r.DownloadAccum()
    r.CleanUp()
    rng_ref, want_d = rng.copy(), np.zeros((h, w, 4), np.float32)
    o.ref_glsl_render_mp(cam.as_array(), w, h, rng_ref, want_d, n_frames=1, max_depth=depth, model=0)
    # CUDA's sincosf differs from glibc's by 1 ulp on some arguments: many pixels differ in their last bits (same path, same
    # decisions), a few take a different decision somewhere along the path and end up with a different value
    bits = 1.0 - float(same_bits(got_d, want_d).all(axis=2).mean())
    frac = 1.0 - float(np.isclose(got_d[..., :3], want_d[..., :3], rtol=1e-4, atol=1e-6).all(axis=2).mean())
    print(f"default libm, 1080p, 1 frame: {bits:.2e} of the pixels differ in some bit, {frac:.2e} by more than 1e-4 relative")
    assert frac < 1e-3 and bits < 0.1


@needs_glsl
def test_bathroom_beckmann_thin_lens_frames_equal_reference_main():
    """BASELINE.json configs[2]: ~1.0 M triangles, Beckmann, lens radius 0.25, textures + cube-map sky on; two 480 x 272 frames."""
    w, h, depth, frames = 480, 272, 5, 2
    s, r = _big_renderer("bathroom", w, h, max_depth=depth, microfacet_model=1, parity_libm=True, seed=9)
    assert s.tris.shape[0] > 900000
    rng = r.GetRNG()
    cam = make_camera(s.data, w, h)
    assert abs(cam.plane.lens_radius - 0.25) < 1e-6
    r.RenderFrame(cam, frames)
    got, rng_got = r.DownloadAccum(), r.GetRNG()
    r.CleanUp()
    rng_ref, want = rng.copy(), np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).ref_glsl_render_mp(cam.as_array(), w, h, rng_ref, want, n_frames=frames, max_depth=depth, model=6)
    eq = same_bits(got, want).all(axis=2)
    assert eq.all(), f"{(~eq).sum()} of {w * h} pixels differ"
    assert np.array_equal(rng_got, rng_ref)


@needs_glsl
@pytest.mark.parametrize("which", ["atrium", "bathroom"])
def test_million_rays_bit_exact_on_own_trees(which):
    """1 M random + camera rays, closest hit and any hit, on the library's own SBVH tree of both scenes, against the reference's
    traversal shader source compiled as C++: ids, (t, u, v) bits, flags."""
    s, r = _big_renderer(which, 64, 64)
    g = oracle.RefGlsl(s.nodes, s.refs, s.tris)
    rays, anyr = mixed_rays(s, 1 << 20, seed=51), mixed_rays(s, 1 << 20, seed=52, any_hit=True)
    got, gflag = r.TraceClosest(rays), r.TraceAny(anyr)
    r.CleanUp()
    want, wflag = g.trace_closest(rays), g.trace_any(anyr)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(gflag, wflag)
    assert 0.2 < (want[:, 3].view(np.int32) >= 0).mean() and 0.02 < wflag.mean() < 0.98


@pytest.mark.parametrize("name,model,w,h,spp", [("atrium_small", 0, 96, 64, 1024), ("cornell", 1, 64, 64, 1024)])
def test_converged_images_match_the_oracle(name, model, w, h, spp):
    """north_star's converged-image gate at >= 1024 spp, GGX and Beckmann: relMSE(CUDA, oracle) <= 2 x the oracle's own
    seed-to-seed noise floor at equal spp (fast-libm default build, independent RNG tables)."""
    s = get_scene(name)
    cam = make_camera(s.data, w, h)
    r = Renderer().Initialize(w, h, s, max_depth=5, microfacet_model=model, seed=101)
    r.RenderFrame(cam, spp)
    got = r.DownloadAccum()[..., :3] / spp
    r.CleanUp()
    o = oracle.Oracle(s)
    imgs = []
    for seed in (1, 2):
        rng, acc = make_rng_table(w * h, seed), np.zeros((h, w, 4), np.float32)
        o.render(cam.as_array(), w, h, rng, acc, n_frames=spp, max_depth=5, model=model)
        imgs.append(acc[..., :3] / spp)
    floor = rel_mse(imgs[0], imgs[1])
    err = 0.5 * (rel_mse(got, imgs[0]) + rel_mse(got, imgs[1]))
    print(f"{name}: relMSE vs oracle {err:.3e}, oracle seed-to-seed floor {floor:.3e} at {spp} spp")
    assert err <= 2.0 * floor
