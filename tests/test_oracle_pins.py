"""CPU: pins the oracle (and the product's host-side twins) to the REFERENCE's own C++.

Two routes to the same truth:
  * tests/golden/ref_pins.npz -- outputs of the reference's C++ recorded by tests/golden/make_golden.py (always run);
  * oracle/_ref/libref_shim.so -- the reference's C++ itself, when it has been built here (skipped otherwise).
"""
import ctypes as C
import os

import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import Camera
from common import GOLDEN, SCENES, get_scene
from gpupathtracer_b200 import Scene

PINS = np.load(os.path.join(GOLDEN, "ref_pins.npz"))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def test_hybrid_taus_bit_exact_vs_reference_golden():
    # Renderer.cpp:764-782 == Random.glsl:47-70
    for i in range(PINS["taus_states"].shape[0]):
        draws, after = oracle.hybrid_taus(PINS["taus_states"][i], 64)
        assert np.array_equal(draws.view(np.uint32), PINS["taus_draws"][i].view(np.uint32))
        assert np.array_equal(after, PINS["taus_after"][i])
    assert float(PINS["taus_draws"].min()) >= 0.0 and float(PINS["taus_draws"].max()) <= 1.0


def test_hybrid_taus_can_return_one():
    # float(uint) rounds to nearest: integers >= 2^32 - 128 map to exactly 1.0f (SURVEY.md quirk list)
    from oracle import lib
    st = np.array([129, 129, 129, 129], np.uint32)
    v = np.float32(2.3283064365387e-10) * np.float32(np.uint32(0xFFFFFFFF))
    assert v == np.float32(1.0)


def test_slab_test_bit_exact_vs_reference_golden():
    # math/AABB.cpp:102-113 == BVH.glsl:366-391 on NaN-free inputs
    l = oracle.lib()
    n = PINS["slab_lo"].shape[0]
    dist = np.zeros(2, np.float32)
    bad = 0
    for i in range(n):
        hit = l.or_slab_test(_p(PINS["slab_lo"][i].copy()), _p(PINS["slab_hi"][i].copy()), _p(PINS["slab_iorg"][i].copy()),
                             _p(PINS["slab_idir"][i].copy()), float(PINS["slab_tcur"][i]), _p(dist))
        if hit != PINS["slab_hit"][i] or not np.array_equal(dist.view(np.uint32), PINS["slab_dist"][i].view(np.uint32)):
            bad += 1
    assert bad == 0
    assert 0.05 < PINS["slab_hit"].mean() < 0.95


def test_image_plane_bit_exact_vs_reference_golden():
    # math/Camera.cpp:4-22; both the oracle's and the PRODUCT's host implementation (pt_camera_image_plane)
    for i in range(PINS["cam_in"].shape[0]):
        a = PINS["cam_in"][i]
        want = PINS["cam_out"][i][:19]
        got_o = oracle.image_plane(float(a[0]), float(a[1]), float(a[2]), float(a[3]), a[4:7], a[7:10])
        cam = Camera(float(a[0]), float(a[1]), float(a[2]), float(a[3]))
        cam.SetPosition(a[4:7]); cam.SetRotation(a[7:10]); cam.GenerateImagePlane()
        got_p = cam.as_array()
        assert np.array_equal(got_o.view(np.uint32), want.view(np.uint32)), i
        assert np.array_equal(got_p.view(np.uint32), want.view(np.uint32)), i


def test_tonemap_matches_reference_golden():
    # Renderer.cpp:980-994 is the Uncharted-2 curve only; Present.frag:30-36 adds /numSamples, exposure, gamma
    col = PINS["tm_in"]
    got = oracle.tonemap(col, 1)
    want = np.power(PINS_tonemap_with_exposure(col), np.float32(1.0 / 2.2))
    np.testing.assert_allclose(got, want, rtol=2e-6, atol=1e-7)


def PINS_tonemap_with_exposure(col):
    # the recorded reference curve is f(c); Present.frag evaluates f(1.68 * c). Re-evaluate through the live reference
    # when available, else through the recorded points' own formula (checked against them first).
    A, B, C_, D, E, F, W = (np.float32(x) for x in (0.15, 0.50, 0.10, 0.20, 0.02, 0.30, 11.2))

    def curve(c):
        c = (c * np.float32(2.0)).astype(np.float32)
        v = ((c * (A * c + C_ * B) + D * E) / (c * (A * c + B) + D * F)) - E / F
        white = ((W * (A * W + C_ * B) + D * E) / (W * (A * W + B) + D * F)) - E / F
        return (v / white).astype(np.float32)
    assert np.array_equal(curve(PINS["tm_in"]).view(np.uint32), PINS["tm_out"].view(np.uint32)), "formula != recorded reference"
    return curve((np.float32(1.68) * col).astype(np.float32))


@pytest.mark.parametrize("name", ["cornell_small", "soup_small"])
def test_oracle_regression_on_golden_tree(name):
    """Oracle traversal + frames over the REFERENCE builder's committed tree reproduce the committed results bit for bit."""
    s = Scene(SCENES[name]())
    z = np.load(os.path.join(GOLDEN, f"tree_{name}.npz"))
    s.set_bvh(z["nodes"], z["refs"])
    o = oracle.Oracle(s)
    t = np.load(os.path.join(GOLDEN, f"trace_{name}.npz"))
    hits, _ = o.trace_closest(t["rays"])
    assert np.array_equal(hits.view(np.uint32), t["hits"].view(np.uint32))
    flags, _ = o.trace_any(t["rays_any"])
    assert np.array_equal(flags, t["flags"])
    f = np.load(os.path.join(GOLDEN, f"frame_{name}.npz"))
    rng = f["rng0"].copy(); acc = np.zeros_like(f["accum"])
    c = o.render(f["cam"], acc.shape[1], acc.shape[0], rng, acc, n_frames=2, max_depth=66, n_threads=2)
    assert np.array_equal(rng, f["rng1"])
    assert np.array_equal(acc.view(np.uint32), f["accum"].view(np.uint32))
    assert [c["closest_rays"], c["shadow_rays"]] == list(f["counters"])


@pytest.mark.parametrize("name", ["cornell_small", "soup_small"])
def test_oracle_traversal_equals_reference_shader_golden(name):
    """tests/golden/trace_glsl_*.npz hold what the reference's OWN shader source -- IfIfClosestHit / IfIfAnyHit with GetNode,
    IntersectNodeFast, IntersectLeaf, IntersectTriangleMT, extracted from common/BVH.glsl + common/Geometry.glsl and compiled
    as C++ by oracle/Makefile -- answered for the committed ray batches on the reference builder's committed tree. The hand
    restatement in oracle/pt_oracle.cpp must give the same bits: (t, u, v), the triangle index and every any-hit flag."""
    s = Scene(SCENES[name]())
    z = np.load(os.path.join(GOLDEN, f"tree_{name}.npz"))
    s.set_bvh(z["nodes"], z["refs"])
    t = np.load(os.path.join(GOLDEN, f"trace_{name}.npz"))
    g = np.load(os.path.join(GOLDEN, f"trace_glsl_{name}.npz"))
    o = oracle.Oracle(s)
    hits, _ = o.trace_closest(t["rays"])
    assert np.array_equal(hits.view(np.uint32), g["hits"].view(np.uint32))
    flags, _ = o.trace_any(t["rays_any"])
    assert np.array_equal(flags, g["flags"])
    assert (g["hits"][:, 3].view(np.int32) >= 0).mean() > 0.3 and 0.05 < g["flags"].mean() < 0.95       # the batches exercise both outcomes


def _same_bits_or_both_nan(a, b):
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def _check_shading_against(got, got_rng, want, want_rng, model):
    """Bit equality of the 20 probe outputs and the RNG words. One documented exception: where a NaN reaches a GLSL max(),
    glm's ternary (the C++ compile of the shader) propagates it, IEEE maxNum -- the stated float mode, what NVIDIA's FMNMX
    does for GLSL max -- returns the other operand. That happens in exactly one place, Beckmann's pdf of a direction whose
    half vector has n.m = 0 (log 0 inside MicrofacetDistributionBeckmann, Microfacet.glsl:35-39), and only where the BSDF is 0."""
    assert np.array_equal(got_rng, want_rng)
    same = _same_bits_or_both_nan(got, want)
    if model == 0:
        assert same.all()
        return
    cols = [c for c in range(20) if c != 3]
    assert same[:, cols].all()
    odd = ~same[:, 3]
    assert np.isnan(want[odd, 3]).all() and (got[odd, 3] <= np.float32(1e-10)).all() and (got[odd, 0:3] == 0).all()
    assert odd.mean() < 0.35


@pytest.mark.parametrize("model", [0, 1])
def test_oracle_shading_equals_reference_shader_golden(model):
    """tests/golden/shade_glsl.npz holds what the reference's OWN shader functions -- the six assignments of
    ConstructMaterialInstance, ConstructSurfaceInteraction, SetIncomingDirection / SetMicrofacetDirection, ComputeBSDF,
    ProbabilityDensityDirection, GenerateImportanceSample, CalcDiffusePmf, MISWeight, HybridTaus, and everything of
    Microfacet.glsl they call, compiled as C++ by oracle/Makefile -- return for 4096 random (material, normal, view, light
    direction, RNG state) cases, called in Iterative.comp:236-288's order; model 1 is the copy with Microfacet.glsl's three model
    macros pointed at its Beckmann functions. oracle/pt_oracle.cpp's restatement must give the same bits."""
    z = np.load(os.path.join(GOLDEN, "shade_glsl.npz"))
    got, rng = oracle.shade_probe(z["inputs"], z["rng"], model)
    key = "beckmann" if model else "ggx"
    _check_shading_against(got, rng, z[key], z[key + "_rng"], model)
    assert np.isfinite(z[key][:, 15:18]).mean() > 0.6 and (z[key][:, 8] > 0).mean() > 0.9           # the cases are not degenerate
    st = np.array([1234567, 7654321, 999999, 424242], np.uint32)
    out = np.zeros(4096, np.float32)
    oracle.lib().or_hybrid_taus(_p(st), 4096, _p(out))
    assert np.array_equal(out.view(np.uint32), z["taus"].view(np.uint32)) and np.array_equal(st, z["taus_state"])


def _check_light_against(got, got_rng, want, want_rng, model):
    """Bit equality of ray, tmax, lightThroughput, light material and lightPdf. The Beckmann exception of
    _check_shading_against shows here as lightThroughput = 0 * NaN = NaN in the glm compile against exactly 0 in the stated
    float mode (the BSDF is 0 there), nowhere else."""
    assert np.array_equal(got_rng, want_rng)
    same = _same_bits_or_both_nan(got, want)
    if model == 0:
        assert same.all()
        return
    assert same[:, [0, 1, 2, 3, 4, 5, 6, 10, 11]].all()
    odd = ~same[:, 7:10].all(axis=1)
    assert np.isnan(want[odd, 7:10]).all() and (got[odd, 7:10] == 0).all() and odd.mean() < 0.35


def test_oracle_raygen_and_light_sampling_equal_reference_shader_golden():
    """tests/golden/path_glsl_cornell_small.npz: what Iterative.comp's own InitRay + RayIndexToMorton (pixel order, jitter,
    thin-lens sample, ray, the RNG words left behind) and GenerateLightSample + RandomLightVertex (CDF search, point on the
    emitter, shadow ray, tmax, MIS-weighted light throughput, light material), compiled as C++ by oracle/Makefile, return on the
    committed cornell_small scene. The oracle's restatements give the same bits."""
    s = Scene(SCENES["cornell_small"]())
    t = np.load(os.path.join(GOLDEN, "tree_cornell_small.npz"))
    s.set_bvh(t["nodes"], t["refs"])
    z = np.load(os.path.join(GOLDEN, "path_glsl_cornell_small.npz"))
    o = oracle.Oracle(s)
    for model, key in ((0, "ggx"), (1, "beckmann")):
        got, rng = o.light_probe(z["inputs"], z["rng"], model)
        _check_light_against(got, rng, z[key], z[key + "_rng"], model)
    assert np.isfinite(z["ggx"][:, 7:10]).all() and (z["ggx"][:, 7:10] > 0).any(axis=1).mean() > 0.1
    w, h = (int(v) for v in z["size"])
    table = z["table0"].copy()
    rays = o.raygen(z["cam"], w, h, table, 0, w * h)
    assert np.array_equal(rays[:, :8].view(np.uint32), z["rays"][:, :8].view(np.uint32))
    assert np.array_equal(table, z["table1"])
    assert np.all(z["rays"][:, 8:11] == 1.0) and np.all(z["rays"][:, 11] == np.float32(0.5) / np.float32(s.total_light_area))


def test_oracle_present_equals_reference_shader_golden():
    """tests/golden/present_glsl.npz: what Present.frag (division by the sample count, exposure, Uncharted-2 curve, gamma), compiled
    as C++ by oracle/Makefile, writes for an HDR accumulation image; or_tonemap gives the same bits."""
    z = np.load(os.path.join(GOLDEN, "present_glsl.npz"))
    for n, want in zip(z["counts"], z["rgb"]):
        got = oracle.tonemap(z["accum"][..., :3], int(n)).reshape(want.shape)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert z["rgb"].max() > 1.0 and z["rgb"].min() < 0.01


def test_oracle_equirect_coordinates_equal_reference_shader_golden():
    """tests/golden/equirect_glsl.npz: SampleSphericalMap(normalize(sampleCoords)) of EquirectangularConverter.frag compiled as C++
    on random directions and the six axes; the oracle's statement inside or_equirect_to_cube gives the same bits. (Its one
    deliberate difference is atan(-0, x < 0), which GLSL leaves open: the oracle and the CUDA kernel read -0 as +0.)"""
    z = np.load(os.path.join(GOLDEN, "equirect_glsl.npz"))
    got = oracle.equirect_uv(z["coords"])
    assert np.array_equal(got.view(np.uint32), z["uv"].view(np.uint32))
    minus_zero = z["coords"][6:7].copy(); minus_zero[0, 2] = -0.0
    assert np.array_equal(oracle.equirect_uv(minus_zero), got[6:7])


needs_glsl = pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built (no /root/reference)")


@needs_glsl
@pytest.mark.parametrize("name", ["cornell", "soup", "bunny", "atrium_small"])
def test_oracle_traversal_equals_live_reference_shader(name):
    """The same comparison live and larger: 100 000 closest-hit + 100 000 any-hit rays per scene (random, camera-like and
    surface-to-surface rays, finite and infinite tmax) through the reference's shader source compiled as C++ and through the
    oracle, on the tree the reference's own builder makes for the scene."""
    from common import get_scene, mixed_rays
    s = get_scene(name)
    g = oracle.RefGlsl(s.nodes, s.refs, s.tris)
    o = oracle.Oracle(s)
    rays = mixed_rays(s, 100000, seed=5)
    want = g.trace_closest(rays)
    got, _ = o.trace_closest(rays)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    anyr = mixed_rays(s, 100000, seed=6, any_hit=True)
    wflag = g.trace_any(anyr)
    gflag, _ = o.trace_any(anyr)
    assert np.array_equal(gflag, wflag)


@needs_glsl
@pytest.mark.parametrize("model", [0, 1])
def test_oracle_shading_equals_live_reference_shader(model):
    """The same comparison live on 200 000 fresh cases per model."""
    a, rng = oracle.shade_probe_inputs(200000, 31 + model)
    want, wrng = oracle.ref_glsl_shade_probe(a, rng, model)
    got, grng = oracle.shade_probe(a, rng, model)
    _check_shading_against(got, grng, want, wrng, model)


@needs_glsl
@pytest.mark.parametrize("name", ["cornell", "atrium_small"])
def test_oracle_raygen_and_light_sampling_equal_live_reference_shader(name):
    """Live: 100 000 light-sampling cases per model and a full 64 x 40 frame of InitRay per scene."""
    from common import get_scene, make_camera
    from gpupathtracer_b200 import make_rng_table
    s = get_scene(name)
    o = oracle.Oracle(s)
    a, rng = oracle.light_probe_inputs(s, 100000, 7)
    for model in (0, 1):
        want, wrng = oracle.ref_glsl_light_probe(s, a, rng, model)
        got, grng = o.light_probe(a, rng, model)
        _check_light_against(got, grng, want, wrng, model)
    w, h = 64, 40
    cam = make_camera(s.data, w, h, aperture=0.3).as_array()
    t1 = make_rng_table(w * h, 5); t2 = t1.copy()
    got = o.raygen(cam, w, h, t1, 0, w * h)
    want = oracle.ref_glsl_raygen(cam, w, h, s.total_light_area, t2, 0, w * h)
    assert np.array_equal(got[:, :8].view(np.uint32), want[:, :8].view(np.uint32)) and np.array_equal(t1, t2)


def _check_frame_against(got, got_rng, want, want_rng, model):
    """Whole frames: accumulation image and RNG table bit for bit. Beckmann keeps the exception of _check_shading_against: a
    pixel whose path met max(NaN, x) is NaN in the glm compile and finite in the stated float mode; the RNG words agree even
    there (the draws do not depend on it)."""
    assert np.array_equal(got_rng, want_rng)
    same = _same_bits_or_both_nan(got, want).all(axis=2)
    if model == 0:
        assert same.all() and np.isfinite(want).all()
        return
    assert np.isnan(want[~same]).any(axis=1).all() and (~same).mean() < 0.1


def test_oracle_frame_equals_reference_main_golden():
    """tests/golden/frame_glsl_cornell_small.npz: one frame of the reference's own main() (Iterative.comp:214-307 with
    GetInterpolatedVertex and ConstructMaterialInstance, compiled as C++ by oracle/Makefile) on the committed cornell_small tree
    with textures on every unit and a cube-map sky: accumulation image and the RNG table it leaves. or_render gives the same
    bits, which pins the glue between the pieces pinned above: MIS bookkeeping across bounces, termination, the sky term,
    accumulation, RNG write-back. The third frame is the shader's SUN_LIGHT build."""
    from common import with_textures_and_sky
    s0 = Scene(SCENES["cornell_small"]())
    t = np.load(os.path.join(GOLDEN, "tree_cornell_small.npz"))
    s0.set_bvh(t["nodes"], t["refs"])
    s = with_textures_and_sky(s0)
    z = np.load(os.path.join(GOLDEN, "frame_glsl_cornell_small.npz"))
    w, h = (int(v) for v in z["size"])
    o = oracle.Oracle(s)
    for model, key in ((0, "ggx"), (1, "beckmann")):
        table = z["table0"].copy(); accum = np.zeros((h, w, 4), np.float32)
        o.render(z["cam"], w, h, table, accum, n_frames=1, max_depth=66, model=model)
        _check_frame_against(accum, table, z[key], z[key + "_rng"], model)
    assert (z["ggx"][..., :3] > 0).any(axis=2).mean() > 0.9
    table = z["table0"].copy(); accum = np.zeros((h, w, 4), np.float32)                  # Beckmann with max() = IEEE maxNum: exact
    o.render(z["cam"], w, h, table, accum, n_frames=1, max_depth=66, model=1)
    _check_frame_against(accum, table, z["beckmann_maxnum"], z["beckmann_maxnum_rng"], 0)
    table = z["table0"].copy(); accum = np.zeros((h, w, 4), np.float32)
    o.render(z["cam"], w, h, table, accum, n_frames=1, max_depth=66, model=0, sun=(z["sun_dir"], float(z["sun_radius"])))
    _check_frame_against(accum, table, z["ggx_sun"], z["ggx_sun_rng"], 0)
    assert not np.array_equal(z["ggx_sun"], z["ggx"])


@needs_glsl
@pytest.mark.parametrize("name", ["cornell", "soup_small", "atrium_small"])
def test_oracle_frame_equals_live_reference_main(name):
    """Live: the plain scene and its textured / cube-map variant, both microfacet models, two frames accumulated (the second
    starts from the first's image and RNG table)."""
    from common import get_scene, make_camera, with_textures_and_sky
    from gpupathtracer_b200 import make_rng_table
    base = get_scene(name)
    w, h = 48, 32
    for s in (base, with_textures_and_sky(base)):
        o = oracle.Oracle(s)
        cam = make_camera(s.data, w, h, aperture=0.2).as_array()
        for model in (0, 1):
            t1 = make_rng_table(w * h, 11); t2 = t1.copy()
            a1 = np.zeros((h, w, 4), np.float32); a2 = a1.copy()
            for _ in range(2):
                o.render(cam, w, h, t1, a1, n_frames=1, max_depth=66, model=model)
                o.ref_glsl_render(cam, w, h, t2, a2, model)
            _check_frame_against(a1, t1, a2, t2, model)
        # Beckmann with GLSL max() meaning IEEE maxNum, as the hardware the shader runs on executes it: no exception left
        t1 = make_rng_table(w * h, 11); t2 = t1.copy()
        a1 = np.zeros((h, w, 4), np.float32); a2 = a1.copy()
        o.render(cam, w, h, t1, a1, n_frames=1, max_depth=66, model=1)
        o.ref_glsl_render(cam, w, h, t2, a2, 2)
        _check_frame_against(a1, t1, a2, t2, 0)


@needs_glsl
def test_beckmann_exception_is_only_the_nan_rule_of_max():
    """The reference's Beckmann source built with scalar min / max = IEEE minNum / maxNum (oracle/ref_glsl_ns.h REFGLSL_MAXNUM;
    the stated float mode, what NVIDIA's FMNMX executes for GLSL max) instead of glm's NaN-propagating ternaries: the shading
    probe (200 000 cases) and the light-sampling probe agree with the oracle bit for bit, with no exception."""
    from common import get_scene
    a, rng = oracle.shade_probe_inputs(200000, 3)
    got, grng = oracle.shade_probe(a, rng, 1)
    want, wrng = oracle.ref_glsl_shade_probe(a, rng, 2)
    assert _same_bits_or_both_nan(got, want).all() and np.array_equal(grng, wrng)
    s = get_scene("atrium_small")
    la, lrng = oracle.light_probe_inputs(s, 100000, 7)
    got, grng = oracle.Oracle(s).light_probe(la, lrng, 1)
    want, wrng = oracle.ref_glsl_light_probe(s, la, lrng, 2)
    assert _same_bits_or_both_nan(got, want).all() and np.array_equal(grng, wrng)


@needs_glsl
@pytest.mark.parametrize("name", ["cornell", "soup_small", "atrium_small"])
def test_oracle_sun_light_frame_equals_live_reference_main(name):
    """The shader's SUN_LIGHT build (Iterative.comp:89-110,116-126: coin flip between the emitters and a sun disk one unit away,
    unoccluded sun rays credit the sky material's emission) against or_render's sun branch: whole frames, bit for bit."""
    from common import get_scene, make_camera
    from gpupathtracer_b200 import make_rng_table
    s = get_scene(name)
    o = oracle.Oracle(s)
    w, h = 48, 32
    cam = make_camera(s.data, w, h, aperture=0.2).as_array()
    d = np.array([2.0, 9.0, 12.0], np.float32)
    sun = (d / np.float32(np.sqrt((d * d).sum(dtype=np.float32))), 0.05)
    for model in (0, 1):
        t1 = make_rng_table(w * h, 11); t2 = t1.copy()
        a1 = np.zeros((h, w, 4), np.float32); a2 = a1.copy()
        o.render(cam, w, h, t1, a1, n_frames=1, max_depth=66, model=model, sun=sun)
        o.ref_glsl_render(cam, w, h, t2, a2, model, sun=sun)
        _check_frame_against(a1, t1, a2, t2, model)
        plain = np.zeros((h, w, 4), np.float32)
        o.render(cam, w, h, make_rng_table(w * h, 11), plain, n_frames=1, max_depth=66, model=model)
        assert not np.array_equal(plain, a1)                                   # the branch really is another program


@needs_glsl
@pytest.mark.parametrize("depth", [66, 5])
def test_reference_main_on_all_cores_is_the_same_frame(depth):
    """ref_glsl_render_frames_mp (what bench.py's reference arm times): several processes draining one shared ray counter, three
    accumulated frames, give the bits of the single invocation and of or_render -- also at BASELINE's depth-5 configuration,
    where the literal of the shader's `i > 64` is the only thing replaced."""
    from common import get_scene, make_camera
    from gpupathtracer_b200 import make_rng_table
    s = get_scene("atrium_small")
    o = oracle.Oracle(s)
    w, h = 64, 48
    cam = make_camera(s.data, w, h, aperture=0.2).as_array()
    t1 = make_rng_table(w * h, 3); t2 = t1.copy(); t3 = t1.copy()
    a1 = np.zeros((h, w, 4), np.float32); a2 = a1.copy(); a3 = a1.copy()
    o.render(cam, w, h, t1, a1, n_frames=3, max_depth=depth, model=0)
    o.ref_glsl_render_mp(cam, w, h, t2, a2, n_frames=3, max_depth=depth, model=0, n_procs=4)
    _check_frame_against(a1, t1, a2, t2, 0)
    if depth == 66:
        for _ in range(3):
            o.ref_glsl_render(cam, w, h, t3, a3, 0)
        _check_frame_against(a3, t3, a2, t2, 0)


@needs_glsl
def test_oracle_equirect_coordinates_equal_live_reference_shader():
    p = np.random.default_rng(5).normal(size=(500000, 3)).astype(np.float32)
    p[::7, 1] = 0.0
    p = p[~((p[:, 2] == 0) & np.signbit(p[:, 2]))]
    assert np.array_equal(oracle.equirect_uv(p).view(np.uint32), oracle.ref_glsl_equirect_uv(p).view(np.uint32))


@needs_glsl
def test_oracle_present_equals_live_reference_shader():
    g = np.random.default_rng(11)
    a = np.ones((64, 64, 4), np.float32)
    a[..., :3] = (g.uniform(0, 1, size=(64, 64, 3)) ** 6 * 5000).astype(np.float32)
    for n in (1, 3, 1000):
        want = oracle.ref_glsl_present(a, n)
        got = oracle.tonemap(a[..., :3], n).reshape(want.shape)
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


# ----------------------------------------------------------------------------------- live reference (this container)
needs_ref = pytest.mark.skipif(not oracle.have_ref(), reason="oracle/_ref/libref_shim.so not built (no /root/reference)")


@needs_ref
def test_live_reference_matches_golden_pins():
    r = oracle.ref()
    st = PINS["taus_states"][0].copy(); out = np.zeros(64, np.float32)
    r.ref_hybrid_taus(_p(st), 64, _p(out))
    assert np.array_equal(out.view(np.uint32), PINS["taus_draws"][0].view(np.uint32))


@needs_ref
@pytest.mark.parametrize("name", ["cornell_small", "soup_small"])
def test_live_reference_builder_matches_golden_tree(name):
    s = Scene(SCENES[name]())
    nodes, refs = oracle.ref_build_sbvh(s.tris_abs)
    z = np.load(os.path.join(GOLDEN, f"tree_{name}.npz"))
    assert np.array_equal(nodes.view(np.uint32), z["nodes"].view(np.uint32)) and np.array_equal(refs, z["refs"])


@needs_ref
def test_closest_hit_agrees_with_brute_force_on_reference_tree():
    """IfIfClosestHit over the reference tree vs testing every triangle: ids may differ only where the reference's
    own (non-watertight) slab test culls a valid hit or where two triangles tie in t."""
    s = get_scene("soup_small")
    o = oracle.Oracle(s)
    lo, hi = s.bounds()
    from gpupathtracer_b200 import scenes
    rays = scenes.random_rays(20000, lo, hi, seed=5)
    a, _ = o.trace_closest(rays)
    b = o.trace_closest_bruteforce(rays)
    mism = a[:, 3].view(np.int32) != b[:, 3].view(np.int32)
    assert mism.mean() < 1e-3
    # every mismatch is a tie or a reference-culling miss: brute force never finds a FARTHER hit
    assert np.all(b[mism, 0] <= a[mism, 0])
