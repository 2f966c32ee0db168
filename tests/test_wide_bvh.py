"""PT_BVH_WIDE / PT_BVH_WIDE4: the host-side collapse (CPU, decoded and walked here in numpy) and the CUDA traversal
(-m gpu) against the oracle. The wide layouts keep the triangles and the triangle test; only box culling differs."""
import numpy as np
import pytest

import oracle
from gpupathtracer_b200.api import collapse_bvh_host, PT_BVH_WIDE, PT_BVH_WIDE4
from common import get_scene, mixed_rays, make_camera, rel_mse


def decode(node_bytes, width):
    """-> dict of arrays following WideNodeHost / WideNode4Host (csrc/pt_host.h)."""
    nb = node_bytes
    n = nb.shape[0]
    out = {"p": nb[:, 0:12].copy().view(np.float32).reshape(n, 3), "e": nb[:, 12:15].astype(np.int32),
           "imask": nb[:, 15].astype(np.uint32), "child_base": nb[:, 16:20].copy().view(np.uint32).reshape(n),
           "tri_base": nb[:, 20:24].copy().view(np.uint32).reshape(n), "meta": nb[:, 24:24 + width].astype(np.uint32)}
    o = 24 + width
    out["qlo"] = nb[:, o:o + 3 * width].reshape(n, 3, width).astype(np.float64)
    out["qhi"] = nb[:, o + 3 * width:o + 6 * width].reshape(n, 3, width).astype(np.float64)
    return out


def walk(t, width, o, d, tmax):
    """Conservative float64 walk of the wide tree: the set of leaf-stream entries a ray can reach within [0, tmax]."""
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / d.astype(np.float64)
    reach, stack = [], [0]
    while stack:
        i = stack.pop()
        scale = np.ldexp(1.0, t["e"][i] - 127)
        lo = t["p"][i].astype(np.float64)[:, None] + scale[:, None] * t["qlo"][i]
        hi = t["p"][i].astype(np.float64)[:, None] + scale[:, None] * t["qhi"][i]
        with np.errstate(invalid="ignore"):
            t0 = (lo - o[:, None]) * inv[:, None]; t1 = (hi - o[:, None]) * inv[:, None]
        tn = np.fmax.reduce(np.fmin(t0, t1), axis=0); tf = np.fmin.reduce(np.fmax(t0, t1), axis=0)
        hit = (np.fmax(tn, 0.0) <= np.fmin(tf, tmax * (1 + 1e-5)) * (1 + 1e-9) + 1e-12)   # tmax = the float32 t of the oracle hit: allow its rounding
        for k in range(width):
            m = int(t["meta"][i, k])
            if m == 0:
                continue
            if (m & 31) >= 24:                       # inner: slot = (m & 31) - 24; children stored in slot order
                slot = (m & 31) - 24
                rel = bin(int(t["imask"][i]) & ((1 << slot) - 1)).count("1")
                if hit[k]:
                    stack.append(int(t["child_base"][i]) + rel)
            elif hit[k]:
                cnt = bin(m >> 5).count("1")
                reach.extend(range(int(t["tri_base"][i]) + (m & 31), int(t["tri_base"][i]) + (m & 31) + cnt))
    return reach


@pytest.mark.parametrize("name", ["cornell", "soup_small", "atrium_small"])
@pytest.mark.parametrize("width", [8, 4])
def test_collapse_is_complete_and_conservative(name, width):
    s = get_scene(name)
    nb, ids, depth = collapse_bvh_host(s.nodes, s.refs, s.tris, width)
    assert nb.shape[1] == (96 if width == 8 else 64) and depth < 30
    t = decode(nb, width)
    # every triangle the reference tree references is in the wide leaf stream
    want = np.unique(np.abs(s.refs))
    assert np.array_equal(np.unique(ids), want)
    # children blocks are disjoint and cover all nodes but the root exactly once
    n_inner = np.array([bin(int(m)).count("1") for m in t["imask"]])
    assert n_inner.sum() == nb.shape[0] - 1
    starts = t["child_base"][n_inner > 0].astype(np.int64)
    order = np.argsort(starts)
    assert np.array_equal(starts[order][1:], (starts[order] + n_inner[n_inner > 0][order])[:-1]) and starts[order][0] == 1
    # leaf slots: 1..3 triangles, offsets inside the node's 24-entry window
    meta = t["meta"]
    leaf = (meta != 0) & ((meta & 31) < 24)
    cnt = np.vectorize(lambda m: bin(int(m) >> 5).count("1"))(meta)
    assert (cnt[leaf] >= 1).all() and (cnt[leaf] <= 3).all() and ((meta[leaf] & 31) + cnt[leaf] <= 24).all()
    # a conservative walk reaches the triangle the oracle reports, for every ray
    rays = mixed_rays(s, 240, seed=5)
    hits, _ = oracle.Oracle(s).trace_closest(rays)
    for r, h in zip(rays, hits):
        tri = int(h[3:4].view(np.int32)[0])
        if tri < 0:
            continue
        reach = walk(t, width, r[0:3].astype(np.float64), r[4:7], float(h[0]))
        assert tri in set(int(ids[k]) for k in reach), "oracle's hit triangle is not reachable in the wide tree"


def _agree(s, layout, n=200000):
    from gpupathtracer_b200 import Renderer
    rays = mixed_rays(s, n)
    anyr = mixed_rays(s, n, any_hit=True)
    o = oracle.Oracle(s)
    want, _ = o.trace_closest(rays)
    wflag, _ = o.trace_any(anyr)
    r = Renderer().Initialize(64, 64, s, bvh_layout=layout)
    got = r.TraceClosest(rays); gflag = r.TraceAny(anyr)
    r.CleanUp()
    return rays, want, got, wflag, gflag


@pytest.mark.gpu
@pytest.mark.parametrize("layout", [PT_BVH_WIDE, PT_BVH_WIDE4])
@pytest.mark.parametrize("name", ["cornell", "soup", "bunny", "atrium_small"])
def test_wide_traversal_against_oracle(name, layout):
    """Stated bar for the wide layouts: triangle ids and any-hit flags agree with the reference walk on all but
    <= 2e-5 of the rays (measured: 0 to 5e-7); where the ids agree, (t, u, v) are BIT-identical (same triangle test);
    where they differ, the wide walk found a hit at least as close (it reaches a superset of the triangles)."""
    s = get_scene(name)
    rays, want, got, wflag, gflag = _agree(s, layout)
    wid, gid = want[:, 3].view(np.int32), got[:, 3].view(np.int32)
    same = wid == gid
    assert (~same).mean() <= 2e-5, f"{(~same).sum()} id mismatches"
    assert np.array_equal(got[same, :3].view(np.uint32), want[same, :3].view(np.uint32))
    assert (got[~same, 0] <= want[~same, 0]).all()
    assert (wflag != gflag).mean() <= 2e-5
    assert (gflag[wflag != gflag] == 1).all()          # only ever MORE occlusion found, never less


@pytest.mark.gpu
@pytest.mark.parametrize("layout", [PT_BVH_WIDE, PT_BVH_WIDE4])
def test_wide_frames_match_reference_layout(layout):
    from gpupathtracer_b200 import Renderer, make_rng_table
    s = get_scene("atrium_small")
    w, h = 64, 64
    cam = make_camera(s.data, w, h)
    rng = make_rng_table(w * h, 13)
    imgs = []
    for lay in (0, layout):
        r = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng, bvh_layout=lay)
        r.RenderFrame(cam, 4)
        imgs.append(r.DownloadAccum())
        r.CleanUp()
    fin = np.isfinite(imgs[0]).all(axis=2) & np.isfinite(imgs[1]).all(axis=2)
    same = (imgs[0][fin] == imgs[1][fin]).all(axis=1).mean()
    assert same > 0.995, f"only {same:.5f} of pixels are bit-identical between layouts"
    assert rel_mse(imgs[1][..., :3], imgs[0][..., :3]) < 1e-3
