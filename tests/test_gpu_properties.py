"""-m gpu: convergence against the oracle with DIFFERENT random numbers, and size-independent properties at the
BASELINE.json sizes (where the CPU oracle would take minutes, the properties carry the check)."""
import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import Renderer, make_rng_table, scenes, Scene
from common import get_scene, make_camera, rel_mse

pytestmark = pytest.mark.gpu


def test_converged_image_within_twice_the_oracles_own_noise():
    """BASELINE.json: 'converged images at high spp must agree within a stated relative-MSE tolerance'. Stated
    tolerance: relMSE(CUDA seed A, oracle seed B) <= 2 x relMSE(oracle seed C, oracle seed B) at equal spp
    -- i.e. the CUDA image is statistically indistinguishable from a third oracle run."""
    s = get_scene("cornell")
    w, h, spp = 48, 48, 192
    cam = make_camera(s.data, w, h)
    o = oracle.Oracle(s)

    def oracle_image(seed):
        img = np.zeros((h, w, 4), np.float32)
        o.render(cam.as_array(), w, h, make_rng_table(w * h, seed), img, n_frames=spp, max_depth=5)
        return img[..., :3] / spp

    b, c = oracle_image(101), oracle_image(202)
    r = Renderer().Initialize(w, h, s, max_depth=5, rng_state=make_rng_table(w * h, 303))
    r.RenderFrame(cam, spp)
    a = r.DownloadAccum()[..., :3] / spp
    r.CleanUp()
    floor = rel_mse(c, b)
    got = rel_mse(a, b)
    assert floor > 0 and got <= 2.0 * floor, f"relMSE {got:.3e} vs oracle-vs-oracle noise floor {floor:.3e}"
    # and the means agree: no energy gained or lost
    assert abs(np.nanmean(a) / np.nanmean(b) - 1.0) < 0.02


@pytest.fixture(scope="module")
def atrium_full():
    return Scene(scenes.atrium(target_tris=262144, seed=2))


def test_full_size_frame_properties(atrium_full):
    """1920x1080, 262 144 triangles, depth 5 (BASELINE.json configs[1]): every pixel receives exactly one sample per
    frame, alpha stays 1, ray counters obey the path structure, the image does not depend on the slice count, and two
    consecutive batches equal one batch of the same length."""
    s = atrium_full
    w, h = 1920, 1080
    cam = make_camera(s.data, w, h)
    outs = []
    for n_slices, split in ((1, False), (3, True)):
        r = Renderer().Initialize(w, h, s, max_depth=5, seed=77, bvh="build" if s.nodes is None else "given", n_slices=n_slices)
        if split:
            r.RenderFrame(cam, 1); r.RenderFrame(cam, 2)
        else:
            r.RenderFrame(cam, 3)
        outs.append((r.DownloadAccum(), r.Counters(), r.GetRNG()))
        r.CleanUp()
    (a, ca, ga), (b, cb, gb) = outs
    assert np.array_equal(a.view(np.uint32), b.view(np.uint32)) and np.array_equal(ga, gb)
    assert ca["paths"] == cb["paths"] == 3 * w * h
    assert np.all(a[..., 3] == 1.0)
    assert np.isfinite(a[..., :3]).mean() > 0.999999 and np.nanmin(a[..., :3]) >= 0.0
    # every path traces at least one and at most 5 closest-hit rays; a shadow ray follows every hit, never a miss
    assert 3 * w * h <= ca["closest_rays"] <= 5 * 3 * w * h
    assert 0 < ca["shadow_rays"] <= ca["closest_rays"]


def test_full_size_batch_closest_vs_any_consistency(atrium_full):
    """4 M incoherent rays: an any-hit query with tmax just past the closest hit must be occluded, one with tmax just
    short of it must be free -- the two kernels share nothing but the tree and the triangle test."""
    s = atrium_full
    lo, hi = s.bounds()
    n = 1 << 22
    rays = scenes.random_rays(n, lo, hi, seed=9)
    r = Renderer().Initialize(64, 64, s, max_depth=5, seed=1, bvh="build" if s.nodes is None else "given")
    hits = r.TraceClosest(rays)
    t = hits[:, 0]
    hit = hits[:, 3].view(np.int32) >= 0
    assert 0.9 < hit.mean() <= 1.0                                  # closed hall
    past = rays.copy(); past[:, 7] = np.where(hit, t * 1.001, 1e20)
    short = rays.copy(); short[:, 7] = np.where(hit, t * 0.999, 1e20)
    f_past, f_short = r.TraceAny(past), r.TraceAny(short)
    r.CleanUp()
    # hits a hair away from the origin are excluded: there the 0.1 % margin on tmax (1e-7 .. 1e-6) is below the absolute
    # rounding error of the reference's slab arithmetic (-o/d + p/d cancels at scene scale), which may cull the leaf
    far = hit & (t > 0.05)
    assert far.mean() > 0.9
    assert (f_past[far] == 1).mean() > 1 - 1e-5
    assert (f_short[far] == 0).mean() > 1 - 1e-5
    assert (f_past[~hit] == 0).all() and (f_short[~hit] == 0).all()
    # idempotence: tracing the same batch again gives the same bits
    r2 = Renderer().Initialize(64, 64, s, max_depth=5, seed=1)
    assert np.array_equal(r2.TraceClosest(rays).view(np.uint32), hits.view(np.uint32))
    r2.CleanUp()


@pytest.mark.parametrize("builder", ["device", "ploc", "sah"])
@pytest.mark.parametrize("name", ["cornell", "soup", "bunny"])
def test_device_built_bvh(name, builder):
    """pt_build_bvh_device (Morton LBVH built by CUDA kernels), pt_build_bvh_device_ploc (the same pipeline with the
    hierarchy from parallel locally-ordered clustering) and pt_build_bvh_device_sah (top-down binned SAH): the tree comes back in the reference's serialised form,
    references every triangle exactly once, its boxes enclose their subtrees, the CUDA walk of it is bit-exact against
    the ORACLE walking the same tree, and it finds the same surfaces as the reference builder's tree."""
    s_ref = get_scene(name)
    s = Scene(s_ref.data)
    r = Renderer().Initialize(64, 64, s, bvh=builder)
    st = r.bvh_stats
    nodes, refs = s.nodes, s.refs
    n = s.tris.shape[0]
    d0 = nodes[:, 3].copy().view(np.int32)
    assert refs.shape[0] == n and np.array_equal(np.sort(np.abs(refs)), np.arange(n))
    if builder == "sah":
        # leaves of several triangles where the SAH says so: a leaf per end-of-leaf marker (triangle 0 cannot carry one), every
        # leaf node points at the first entry of its run, the runs tile the reference list
        leaves = int((d0[1:] <= 0).sum())
        assert nodes.shape[0] == 2 * leaves - 1 and st.n_leaves == leaves and leaves <= n
        starts = np.sort(-d0[1:][d0[1:] <= 0])
        ends = np.nonzero((refs < 0) | ((refs == 0) & np.isin(np.arange(n) + 1, np.append(starts, n))))[0]
        assert starts[0] == 0 and np.array_equal(starts[1:], ends[:-1] + 1) and ends[-1] == n - 1
        if name != "soup":
            assert leaves < n                                    # quads and coplanar fans stay together
    else:
        assert nodes.shape[0] == 2 * n - 1 and st.n_leaves == n
        assert (refs <= 0).all()                                 # one triangle per leaf: every entry ends its leaf
    # boxes: every node's box contains both children's boxes
    inner = np.nonzero(d0 > 0)[0]
    for c in (0, 1):
        ch = d0[inner] + c
        assert (nodes[ch, 0:3] >= nodes[inner, 0:3]).all() and (nodes[ch, 4:7] <= nodes[inner, 4:7]).all()
    # larger-area child first (BVH.cpp:2240-2247)
    ext = nodes[:, 4:7] - nodes[:, 0:3]
    area = ext[:, 0] * (ext[:, 1] + ext[:, 2]) + ext[:, 1] * ext[:, 2]
    assert (area[d0[inner]] >= area[d0[inner] + 1]).all()
    from common import mixed_rays
    rays = mixed_rays(s, 100000)
    got = r.TraceClosest(rays)
    r.CleanUp()
    want, _ = oracle.Oracle(s).trace_closest(rays)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    other, _ = oracle.Oracle(s_ref).trace_closest(rays)
    same = got[:, 3].view(np.int32) == other[:, 3].view(np.int32)
    assert same.mean() > 1 - 1e-4                                # a different tree: only box-culling edge cases and ties differ
    assert np.array_equal(got[same, :3].view(np.uint32), other[same, :3].view(np.uint32))


def test_trace_primary_is_the_first_bounce_in_pixel_order():
    """pt_trace_primary_async = InitRay for every pixel + closest hit, hit records by pixel: equal, bit for bit, to tracing the
    rays pt_raygen_probe reports for the same RNG table (the probe advances the table, the primary trace does not), and
    leaves the table where it was (calling it twice gives the same hits)."""
    s = get_scene("cornell")
    w, h = 96, 64
    cam = make_camera(s.data, w, h, aperture=0.3)                                # lens radius > 0: the lens sample is part of the ray
    r = Renderer().Initialize(w, h, s, max_depth=5, rng_state=make_rng_table(w * h, 11))
    r.lib.pt_set_camera(r.ctx, cam.plane)
    first = r.TracePrimary()
    again = r.TracePrimary()
    assert np.array_equal(first.view(np.uint32), again.view(np.uint32))
    probe = r.RaygenProbe(0, w * h)
    px, py = probe[:, 0].copy().view(np.int32), probe[:, 1].copy().view(np.int32)
    rays = np.zeros((w * h, 8), np.float32)
    rays[:, 0:3] = probe[:, 2:5]; rays[:, 4:7] = probe[:, 5:8]; rays[:, 7] = 1e20
    want = r.TraceClosest(rays)
    r.CleanUp()
    assert np.array_equal(first[py * w + px].view(np.uint32), want.view(np.uint32))
    assert (first[:, 3].view(np.int32) >= 0).mean() > 0.5                        # it is looking at the box


def test_host_batches_larger_than_one_pipeline_chunk():
    """pt_trace_closest / pt_trace_any stream host batches through the device in 2 M-ray chunks (H2D, kernel, D2H
    overlapped, double-buffered staging): 2.5 chunks with a ragged tail must give the oracle's bits."""
    from common import mixed_rays
    s = get_scene("cornell")
    n = (1 << 21) * 2 + 777_777
    rays = mixed_rays(s, n, seed=31)
    anyr = mixed_rays(s, n, seed=32, any_hit=True)
    o = oracle.Oracle(s)
    want, _ = o.trace_closest(rays)
    wflag, _ = o.trace_any(anyr)
    r = Renderer().Initialize(64, 64, s)
    got = r.TraceClosest(rays)
    gflag = r.TraceAny(anyr)
    c = r.Counters()
    r.CleanUp()
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(gflag, wflag)
    assert c["closest_rays"] == n and c["shadow_rays"] == n


def test_ray_ordering_does_not_change_any_result(monkeypatch):
    """Big batches may be traced in binned order (origin cell + direction octant, launch_ray_order): forced on (it is
    automatic only for trees larger than L2) the hits and flags must be the bits of the natural order and of the oracle,
    device pointers and chunked host batches alike, including origins outside the root box and a ragged size."""
    from common import mixed_rays
    s = get_scene("soup")
    n = 300_001
    rays = mixed_rays(s, n, seed=41)
    rays[:1000, 0:3] *= 50.0                                     # far outside the scene
    anyr = mixed_rays(s, n, seed=42, any_hit=True)
    want, _ = oracle.Oracle(s).trace_closest(rays)
    wflag, _ = oracle.Oracle(s).trace_any(anyr)
    out = {}
    for mode in ("-1", "1"):
        monkeypatch.setenv("PT_RAY_ORDER", mode)
        r = Renderer().Initialize(64, 64, s)
        out[mode] = (r.TraceClosest(rays), r.TraceAny(anyr), r.Counters()["kernel_launches"])
        r.CleanUp()
    for mode in ("-1", "1"):
        assert np.array_equal(out[mode][0].view(np.uint32), want.view(np.uint32))
        assert np.array_equal(out[mode][1], wflag)
    assert out["1"][2] == out["-1"][2] + 8                       # 4 binning kernels per ordered launch


def test_device_seeded_rng_table_meets_hybrid_taus_requirements():
    """pt_seed_rng (k_seed_rng) stands in for Renderer.cpp:545-553 (uniform_int_distribution(129, UINT32_MAX) per word): every
    word >= 129 (the Tausworthe lanes need > 128), states distinct per pixel, the same table for the same seed and another
    one for another seed, words spread over the whole range."""
    from common import get_scene
    s = get_scene("cornell_small")
    tabs = []
    for seed in (5, 5, 6):
        r = Renderer().Initialize(96, 64, s, max_depth=5, seed=seed)
        tabs.append(r.GetRNG())
        r.CleanUp()
    t = tabs[0]
    assert t.shape == (96 * 64, 4) and t.dtype == np.uint32 and (t >= 129).all()
    assert np.array_equal(tabs[0], tabs[1]) and not np.array_equal(tabs[0], tabs[2])
    assert len({tuple(row) for row in t}) == t.shape[0]
    assert abs(float(t.mean()) / 2**32 - 0.5) < 0.01 and (t >> 31).mean() > 0.45


def test_inconsistent_tables_are_refused_and_leave_the_context_usable():
    """The device indexes materials by matId >> 1 and triangles by the light entries without bounds checks, so uploads that
    disagree are refused on the host; a failed pt_set_materials leaves the previous set in place."""
    import ctypes as C
    from gpupathtracer_b200.api import PtError, _ptr, pt_texture
    s = get_scene("cornell_small")
    w, h = 32, 32
    cam = make_camera(s.data, w, h)
    r = Renderer().Initialize(w, h, s, max_depth=5, seed=3)
    r.RenderFrame(cam, 1)
    before = r.DownloadAccum()
    lib, ctx = r.lib, r.ctx
    # fewer materials than the triangles name
    assert lib.pt_set_materials(ctx, _ptr(s.materials), 1, None, 0) == -1 and b"matId" in lib.pt_last_error()
    # a texture without texels
    bad = (pt_texture * 1)(); bad[0].width, bad[0].height = 4, 4
    assert lib.pt_set_materials(ctx, _ptr(s.materials), s.materials.shape[0], C.cast(bad, C.c_void_p), 1) == -1
    # a light entry beyond the triangle array
    lights = s.lights.copy(); lights.view(np.uint32)[0, 1] = s.tris.shape[0]
    assert lib.pt_set_lights(ctx, _ptr(lights), lights.shape[0], C.c_float(s.total_light_area)) == -1
    # triangles that name a material beyond the table
    tris = s.tris.copy(); tris.view(np.uint32)[5, 18] = 2 * s.materials.shape[0]
    assert lib.pt_set_triangles(ctx, _ptr(tris), tris.shape[0]) == -1
    # ray batches stop short of the int cursor's range
    assert lib.pt_trace_closest_device(ctx, C.c_void_p(16), (1 << 30) + 1, C.c_void_p(16)) == -1
    # nothing above changed the context: the same frame renders again
    r2 = Renderer().Initialize(w, h, s, max_depth=5, seed=3)
    r2.RenderFrame(cam, 2)
    r.RenderFrame(cam, 1)
    assert np.array_equal(np.nan_to_num(r.DownloadAccum(), nan=-1.0), np.nan_to_num(r2.DownloadAccum(), nan=-1.0))
    assert not np.array_equal(before, r.DownloadAccum())
    r.CleanUp(); r2.CleanUp()


def test_chained_frames_and_per_slice_present_change_nothing():
    """The interactive loop (RenderFrame(1) + PresentAsync per frame) chains frames slice by slice without joining them and lets
    every slice resolve its own pixels; a camera change between frames must reach exactly the frames rendered after it. Same
    accumulation bits, RNG table and presented bytes as the joined path (PT_LAZY_JOIN=0 semantics: one batch call / blocking
    Present), for 1, 3 and 4 slices including a ragged last slice, and across a camera-ring wrap (> 64 frames)."""
    s = get_scene("cornell")
    w, h, frames = 72, 40, 70
    cams = [make_camera(s.data, w, h, aperture=a) for a in (0.0, 0.4)]
    rng = make_rng_table(w * h, 33)
    # reference run: every frame joined (blocking Present after each)
    r0 = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng, n_slices=1)
    want8 = []
    for i in range(frames):
        r0.RenderFrame(cams[(i // 5) & 1], 1)
        want8.append(r0.Present(as_u8=True))
    want, want_rng = r0.DownloadAccum(), r0.GetRNG()
    r0.CleanUp()
    for n_slices in (1, 3, 4):
        r = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng, n_slices=n_slices)
        bufs = [np.zeros((h, w, 3), np.uint8) for _ in range(frames)]
        for i in range(frames):
            r.RenderFrame(cams[(i // 5) & 1], 1)
            r.PresentAsync(bufs[i].ctypes.data)
            if i % 2 == 1:
                r.Synchronize()                    # at most two presents in flight (double-buffered target)
        r.Synchronize()
        assert np.array_equal(np.nan_to_num(r.DownloadAccum(), nan=-1.0), np.nan_to_num(want, nan=-1.0)), n_slices
        assert np.array_equal(r.GetRNG(), want_rng)
        for i in range(frames):
            assert np.array_equal(bufs[i], want8[i]), (n_slices, i)
        assert r.Counters()["paths"] == frames * w * h
        r.CleanUp()


def test_device_visit_counters_equal_the_oracles():
    """pt_set_visit_counting: pair tests and triangle tests per ray, counted by the instrumented traversal kernels on the rays they
    trace, equal the instrumented oracle's on the same batch -- closest-hit rays exactly (same walk, ray by ray). Any-hit rays are
    walked near child first by the kernel and in the reference's fixed order by the oracle (same flags, pt_trace.cuh): an occluded
    ray stops at a different leaf, so those counts agree only roughly (15 %)."""
    from common import mixed_rays
    s = get_scene("atrium_small")
    rays, anyr = mixed_rays(s, 100000, seed=5), mixed_rays(s, 100000, seed=6, any_hit=True)
    o = oracle.Oracle(s)
    want_hits, cc = o.trace_closest(rays)
    _, ca = o.trace_any(anyr)
    r = Renderer().Initialize(64, 64, s)
    plain = r.TraceClosest(rays)
    r.SetVisitCounting(True)
    got = r.TraceClosest(rays)
    r.TraceAny(anyr)
    r.SetVisitCounting(False)
    c = r.Counters()
    r.TraceClosest(rays)                                   # not counted any more
    c2 = r.Counters()
    r.CleanUp()
    assert np.array_equal(got.view(np.uint32), plain.view(np.uint32)) and np.array_equal(got.view(np.uint32), want_hits.view(np.uint32))
    assert c["counted_closest_rays"] == rays.shape[0] and c["counted_shadow_rays"] == anyr.shape[0]
    assert c["closest_pair_tests"] == cc["node_pairs"] and c["closest_tri_tests"] == cc["tri_tests"]
    assert abs(c["shadow_pair_tests"] - ca["node_pairs"]) <= 0.15 * ca["node_pairs"]
    assert abs(c["shadow_tri_tests"] - ca["tri_tests"]) <= 0.15 * ca["tri_tests"]
    assert c2["closest_pair_tests"] == c["closest_pair_tests"] and c2["counted_closest_rays"] == c["counted_closest_rays"]


@pytest.mark.parametrize("builder", ["split", "split_lbvh"])
@pytest.mark.parametrize("name", ["cornell", "atrium_small", "bunny"])
def test_device_built_bvh_with_spatial_splits(name, builder, monkeypatch):
    """pt_build_bvh_device_split: long triangles are cut into clipped references before the sort (the device builders' spatial
    splits). The tree keeps the reference's serialised form with one leaf per reference, every triangle is referenced at least
    once, every reference box lies inside its triangle's box, parents contain children, the CUDA walk is bit-exact against the
    oracle walking the same tree, and it finds the same surfaces as the reference builder's tree (closest AND any hit)."""
    from common import mixed_rays
    monkeypatch.setenv("PT_SPLIT_GRID", "24")                    # coarse scenes: make sure plenty of triangles are cut
    s_ref = get_scene(name)
    s = Scene(s_ref.data)
    r = Renderer().Initialize(64, 64, s, bvh=builder)
    st = r.bvh_stats
    nodes, refs = s.nodes, s.refs
    n, m = s.tris.shape[0], refs.shape[0]
    assert m > n and nodes.shape[0] == 2 * m - 1 and st.n_leaves == m and st.n_refs == m
    assert (refs <= 0).all() and np.array_equal(np.unique(np.abs(refs)), np.arange(n))      # one reference per leaf; nobody lost
    d0 = nodes[:, 3].copy().view(np.int32)
    inner = np.nonzero(d0 > 0)[0]
    for c in (0, 1):
        ch = d0[inner] + c
        assert (nodes[ch, 0:3] >= nodes[inner, 0:3]).all() and (nodes[ch, 4:7] <= nodes[inner, 4:7]).all()
    leaves = np.nonzero(d0 <= 0)[0]
    leaves = leaves[leaves > 0]
    tri = np.abs(refs[-d0[leaves]])
    t = s.tris[tri]                                              # the device's own triangle box: p0, p0 + e1, p0 + e2 in binary32
    p = np.stack([t[:, 0:3], t[:, 0:3] + t[:, 3:6], t[:, 0:3] + t[:, 6:9]], axis=1)
    assert (nodes[leaves, 0:3] >= p.min(axis=1)).all() and (nodes[leaves, 4:7] <= p.max(axis=1)).all()
    ext = nodes[leaves, 4:7] - nodes[leaves, 0:3]
    assert (ext >= 0).all()
    rays, anyr = mixed_rays(s, 200000), mixed_rays(s, 200000, seed=13, any_hit=True)
    got, gflag = r.TraceClosest(rays), r.TraceAny(anyr)
    r.CleanUp()
    o = oracle.Oracle(s)
    want, _ = o.trace_closest(rays)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    assert np.array_equal(gflag, o.trace_any(anyr)[0])
    o_ref = oracle.Oracle(s_ref)
    other, _ = o_ref.trace_closest(rays)
    same = got[:, 3].view(np.int32) == other[:, 3].view(np.int32)
    assert same.mean() > 1 - 1e-4
    assert np.array_equal(got[same, :3].view(np.uint32), other[same, :3].view(np.uint32))
    assert (gflag == o_ref.trace_any(anyr)[0]).mean() > 1 - 1e-4


def _degenerate_soups():
    import dataclasses
    from gpupathtracer_b200 import scenes
    base = scenes.random_soup(600, seed=21)
    out = {}
    for n in (2, 3, 16, 17, 18, 33):                              # around the builder's small-subtree threshold (16)
        ids = base.mat_ids[:n].copy(); ids[0] = 6                 # material 3 of the soup is its emitter: a scene needs one
        out[f"n{n}"] = dataclasses.replace(base, verts=base.verts[:n].copy(), mat_ids=ids)
    same = base.verts.copy(); same[:500] = same[0]                # 500 copies of one triangle: every centroid in one bin on all axes
    out["copies"] = dataclasses.replace(base, verts=same)
    flat = base.verts.copy(); flat[:, :, 1] = 0.25                # everything in the plane y = 0.25: one axis has zero extent
    out["coplanar"] = dataclasses.replace(base, verts=flat)
    return out


@pytest.mark.parametrize("case", ["n2", "n3", "n16", "n17", "n18", "n33", "copies", "coplanar"])
def test_device_sah_builder_on_degenerate_input(case):
    """pt_build_bvh_device_sah where the binned search has nothing to choose from: tiny scenes around the small-subtree
    threshold, hundreds of identical triangles (no plane separates them: halved by position), a scene with no extent on one
    axis. The tree must still reference every triangle once, parents must contain children, and the CUDA walk of it must equal
    the oracle's walk of the same tree bit for bit."""
    from common import mixed_rays
    s = Scene(_degenerate_soups()[case])
    r = Renderer().Initialize(64, 64, s, bvh="sah")
    nodes, refs = s.nodes, s.refs
    n = s.tris.shape[0]
    d0 = nodes[:, 3].copy().view(np.int32)
    leaves = int((d0[1:] <= 0).sum())
    assert nodes.shape[0] == 2 * leaves - 1 and refs.shape[0] == n and 2 <= leaves <= n and r.bvh_stats.n_leaves == leaves
    assert np.array_equal(np.sort(np.abs(refs)), np.arange(n))
    if case == "copies":
        assert leaves < n - 400                                  # identical triangles share leaves (up to 16 each)
    inner = np.nonzero(d0 > 0)[0]
    for c in (0, 1):
        ch = d0[inner] + c
        assert (nodes[ch, 0:3] >= nodes[inner, 0:3]).all() and (nodes[ch, 4:7] <= nodes[inner, 4:7]).all()
    assert r.bvh_stats.max_depth < 96
    rays = mixed_rays(s, 20000, seed=5)
    got = r.TraceClosest(rays)
    r.CleanUp()
    want, _ = oracle.Oracle(s).trace_closest(rays)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("builder", ["build", "sah"])
def test_reinsertion_optimised_tree_on_the_device(builder):
    """pt_optimize_bvh on a tree built by the host builder / by the device SAH builder (read back, optimised on the host, installed
    again): the CUDA walk of the new tree equals the oracle's walk of it bit for bit, it finds the same surfaces as before with no
    more pair visits (device counters), and whole frames under the parity libm are still bit-identical to the oracle on that tree."""
    from common import mixed_rays
    s = Scene(get_scene("atrium_small").data)
    w, h = 64, 48
    rng = make_rng_table(w * h, 5)
    r = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng, bvh=builder, parity_libm=True)
    rays = mixed_rays(s, 150000, seed=17)
    r.SetVisitCounting(True)
    before = r.TraceClosest(rays)
    v0 = r.Counters()
    sah0 = float(r.bvh_stats.sah_cost)
    st = r.OptimizeBVH(4, 0.25)
    assert st.sah_cost < sah0 and st.max_depth < 96 and st.n_refs == s.refs.shape[0]
    got = r.TraceClosest(rays)
    v1 = r.Counters()
    r.SetVisitCounting(False)
    want, _ = oracle.Oracle(s).trace_closest(rays)               # the scene object follows the installed tree
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    same = got[:, 3].view(np.int32) == before[:, 3].view(np.int32)
    assert same.mean() > 1 - 1e-4 and np.array_equal(got[same].view(np.uint32), before[same].view(np.uint32))
    pairs0 = v0["closest_pair_tests"]; pairs1 = v1["closest_pair_tests"] - v0["closest_pair_tests"]
    assert pairs1 <= pairs0
    cam = make_camera(s.data, w, h, aperture=0.1)
    r.RenderFrame(cam, 2)
    img, rng_got = r.DownloadAccum(), r.GetRNG()
    r.CleanUp()
    rng_o, ref = rng.copy(), np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam.as_array(), w, h, rng_o, ref, n_frames=2, max_depth=5, parity_libm=True)
    eq = (img.view(np.uint32) == ref.view(np.uint32)) | (np.isnan(img) & np.isnan(ref))
    assert eq.all() and np.array_equal(rng_got, rng_o)
