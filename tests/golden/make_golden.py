"""Generates the committed golden fixtures under tests/golden/ (run in the BUILD container, where /root/reference
exists and oracle/_ref/libref_shim.so can be built):

    python tests/golden/make_golden.py

ref_pins.npz   outputs of the REFERENCE's own C++ (HybridTaus, AABB::Intersect, Camera::GenerateImagePlane,
               ComputeTonemapUncharted2, AABB::SurfaceArea) on seeded inputs -> pins the oracle without /root/reference
tree_<scene>.npz   the reference builder's (BuildBinnedSpatial) nodes + reference list for the small test scenes
trace_<scene>.npz  oracle closest/any results on seeded ray batches over that tree (regression + GPU parity input)
frame_<scene>.npz  oracle accumulation buffer + RNG table after 2 frames (regression + GPU parity input)
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
sys.path.insert(0, os.path.dirname(HERE))
import oracle  # noqa: E402
from common import SCENES, make_camera, mixed_rays  # noqa: E402
from gpupathtracer_b200 import Scene, make_rng_table  # noqa: E402


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def ref_pins():
    r = oracle.ref()
    assert r is not None, "build oracle/_ref first (make -C oracle ref)"
    rng = np.random.default_rng(2024)
    out = {}
    # HybridTaus: 8 states x 64 draws
    states = rng.integers(129, 2 ** 32, size=(8, 4), dtype=np.uint64).astype(np.uint32)
    draws = np.zeros((8, 64), np.float32); after = states.copy()
    for i in range(8):
        r.ref_hybrid_taus(_p(after[i:i + 1]), 64, _p(draws[i:i + 1]))
    out.update(taus_states=states, taus_draws=draws, taus_after=after)
    # slab test: 4096 random boxes/rays (NaN-free), incl. flat boxes and negative directions
    n = 4096
    lo = rng.uniform(-5, 5, size=(n, 3)).astype(np.float32)
    ext = rng.uniform(0, 3, size=(n, 3)).astype(np.float32); ext[::7, rng.integers(0, 3)] = 0
    hi = (lo + ext).astype(np.float32)
    org = rng.uniform(-8, 8, size=(n, 3)).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    aim = rng.uniform(size=n) < 0.7                      # most rays aim at a point of (or just outside) their box
    tgt = lo + ext * rng.uniform(-0.1, 1.1, size=(n, 3)).astype(np.float32)
    d[aim] = (tgt - org)[aim]
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    idir = (np.float32(1.0) / d).astype(np.float32)
    iorg = (-org * idir).astype(np.float32)
    tcur = np.where(rng.uniform(size=n) < 0.5, 1e20, rng.uniform(0, 10, size=n)).astype(np.float32)
    hit = np.zeros(n, np.int32); dist = np.zeros((n, 2), np.float32)
    for i in range(n):
        hit[i] = r.ref_aabb_intersect(_p(lo[i:i + 1]), _p(hi[i:i + 1]), _p(iorg[i:i + 1]), _p(idir[i:i + 1]), float(tcur[i]), _p(dist[i:i + 1]))
    area = np.array([r.ref_aabb_surface_area(_p(lo[i:i + 1]), _p(hi[i:i + 1])) for i in range(n)], np.float32)
    out.update(slab_lo=lo, slab_hi=hi, slab_iorg=iorg, slab_idir=idir, slab_tcur=tcur, slab_hit=hit, slab_dist=dist, slab_area=area)
    # camera: 64 poses
    m = 64
    cam_in = np.zeros((m, 10), np.float32); cam_out = np.zeros((m, 22), np.float32)
    for i in range(m):
        aspect, fov, focus, ap = rng.uniform(1, 2), rng.uniform(0.3, 1.5), rng.uniform(0.5, 100), rng.uniform(0, 1) * (i % 2)
        pos = rng.uniform(-10, 10, 3).astype(np.float32); rot = np.array([rng.uniform(-3, 3), rng.uniform(-1.4, 1.4), 0], np.float32)
        cam_in[i] = [aspect, fov, focus, ap, *pos, *rot]
        a = cam_in[i]
        r.ref_camera_image_plane(float(a[0]), float(a[1]), float(a[2]), float(a[3]), _p(a[4:7].copy()), _p(a[7:10].copy()), _p(cam_out[i:i + 1]))
    out.update(cam_in=cam_in, cam_out=cam_out)
    # tonemap
    col = rng.uniform(0, 8, size=(256, 3)).astype(np.float32); tm = np.zeros_like(col)
    for i in range(256):
        r.ref_tonemap_uncharted2(_p(col[i:i + 1]), _p(tm[i:i + 1]))
    out.update(tm_in=col, tm_out=tm)
    np.savez_compressed(os.path.join(HERE, "ref_pins.npz"), **out)


def scene_fixtures():
    for name in ("cornell_small", "soup_small"):
        s = Scene(SCENES[name]())
        nodes, refs = oracle.ref_build_sbvh(s.tris_abs)
        np.savez_compressed(os.path.join(HERE, f"tree_{name}.npz"), nodes=nodes, refs=refs)
        s.set_bvh(nodes, refs)
        o = oracle.Oracle(s)
        rays = mixed_rays(s, 4096, seed=21)
        hits, _ = o.trace_closest(rays)
        rays_any = mixed_rays(s, 4096, seed=23, any_hit=True)
        flags, _ = o.trace_any(rays_any)
        np.savez_compressed(os.path.join(HERE, f"trace_{name}.npz"), rays=rays, hits=hits, rays_any=rays_any, flags=flags)
        # the same batches through the reference's own traversal SHADER SOURCE compiled as C++ (oracle/ref_glsl.cpp):
        # these are the reference's answers, recorded so that the pin holds where /root/reference is absent
        g = oracle.RefGlsl(nodes, refs, s.tris)
        np.savez_compressed(os.path.join(HERE, f"trace_glsl_{name}.npz"), hits=g.trace_closest(rays, 1), flags=g.trace_any(rays_any, 1))
        w = h = 32
        cam = make_camera(s.data, w, h).as_array()
        rng = make_rng_table(w * h, 77); rng0 = rng.copy()
        acc = np.zeros((h, w, 4), np.float32)
        c = o.render(cam, w, h, rng, acc, n_frames=2, max_depth=66, n_threads=1)
        np.savez_compressed(os.path.join(HERE, f"frame_{name}.npz"), cam=cam, rng0=rng0, rng1=rng, accum=acc,
                            counters=np.array([c["closest_rays"], c["shadow_rays"]], np.int64))


def shading_fixture():
    """The shading probe through the reference's own shader functions compiled as C++ (oracle/ref_glsl.cpp), both microfacet
    models, 4096 random cases: the reference's answers for tests/test_oracle_pins.py where /root/reference is absent."""
    a, rng = oracle.shade_probe_inputs(4096, 2024)
    g0, r0 = oracle.ref_glsl_shade_probe(a, rng, 0)
    g1, r1 = oracle.ref_glsl_shade_probe(a, rng, 1)
    taus, st = oracle.ref_glsl_hybrid_taus(np.array([1234567, 7654321, 999999, 424242], np.uint32), 4096)
    np.savez_compressed(os.path.join(HERE, "shade_glsl.npz"), inputs=a, rng=rng, ggx=g0, ggx_rng=r0, beckmann=g1, beckmann_rng=r1,
                        taus=taus, taus_state=st)


def path_fixture():
    """Ray generation (InitRay + RayIndexToMorton) and light sampling (GenerateLightSample + RandomLightVertex) of Iterative.comp
    compiled as C++ (oracle/ref_glsl.cpp) on the committed cornell_small tree / scene: the reference's answers."""
    s = Scene(SCENES["cornell_small"]())
    z = np.load(os.path.join(HERE, "tree_cornell_small.npz"))
    s.set_bvh(z["nodes"], z["refs"])
    a, rng = oracle.light_probe_inputs(s, 4096, 99)
    l0, r0 = oracle.ref_glsl_light_probe(s, a, rng, 0)
    l1, r1 = oracle.ref_glsl_light_probe(s, a, rng, 1)
    w, h = 64, 40
    cam = make_camera(s.data, w, h, aperture=0.3).as_array()
    table = make_rng_table(w * h, 5); table0 = table.copy()
    rays = oracle.ref_glsl_raygen(cam, w, h, s.total_light_area, table, 0, w * h)
    np.savez_compressed(os.path.join(HERE, "path_glsl_cornell_small.npz"), inputs=a, rng=rng, ggx=l0, ggx_rng=r0, beckmann=l1,
                        beckmann_rng=r1, cam=cam, size=np.array([w, h]), table0=table0, table1=table, rays=rays)


def frame_fixture():
    """One frame of the reference's main() (Iterative.comp:214-307 compiled as C++) on the committed cornell_small tree, textured
    and under a cube-map sky (tests/common.py with_textures_and_sky), both microfacet models, and the SUN_LIGHT build."""
    from common import with_textures_and_sky
    s0 = Scene(SCENES["cornell_small"]())
    z = np.load(os.path.join(HERE, "tree_cornell_small.npz"))
    s0.set_bvh(z["nodes"], z["refs"])
    s = with_textures_and_sky(s0)
    o = oracle.Oracle(s)
    w, h = 40, 24
    cam = make_camera(s.data, w, h, aperture=0.2).as_array()
    table0 = make_rng_table(w * h, 17)
    out = {}
    for model, key in ((0, "ggx"), (1, "beckmann"), (2, "beckmann_maxnum")):
        table = table0.copy(); accum = np.zeros((h, w, 4), np.float32)
        o.ref_glsl_render(cam, w, h, table, accum, model)
        out[key], out[key + "_rng"] = accum, table
    d = np.array([2.0, 9.0, 12.0], np.float32)
    out["sun_dir"], out["sun_radius"] = d / np.float32(np.sqrt((d * d).sum(dtype=np.float32))), np.float32(0.05)
    table = table0.copy(); accum = np.zeros((h, w, 4), np.float32)
    o.ref_glsl_render(cam, w, h, table, accum, 0, sun=(out["sun_dir"], 0.05))
    out["ggx_sun"], out["ggx_sun_rng"] = accum, table
    np.savez_compressed(os.path.join(HERE, "frame_glsl_cornell_small.npz"), cam=cam, size=np.array([w, h]), table0=table0, **out)


def present_inputs():
    g = np.random.default_rng(3)
    a = np.ones((16, 32, 4), np.float32)
    a[..., :3] = (g.uniform(0, 1, size=(16, 32, 3)) ** 4 * 400).astype(np.float32)
    a[0, 0, :3] = 0
    a[0, 1, :3] = (1e6, 1e-8, 3.0)
    return a


def equirect_inputs(n, seed):
    p = np.random.default_rng(seed).normal(size=(n, 3)).astype(np.float32)
    p[:6] = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
    p[6] = [-1, 0.3, 0.0]
    return p


def equirect_fixture():
    p = equirect_inputs(4096, 1)
    np.savez_compressed(os.path.join(HERE, "equirect_glsl.npz"), coords=p, uv=oracle.ref_glsl_equirect_uv(p))


def present_fixture():
    """Present.frag compiled as C++ on an HDR accumulation image at three sample counts."""
    a = present_inputs()
    np.savez_compressed(os.path.join(HERE, "present_glsl.npz"), accum=a, counts=np.array([1, 7, 64]),
                        rgb=np.stack([oracle.ref_glsl_present(a, n) for n in (1, 7, 64)]))


if __name__ == "__main__":
    shading_fixture()
    present_fixture()
    equirect_fixture()
    frame_fixture()
    path_fixture()
    ref_pins()
    scene_fixtures()
    print(sorted(f for f in os.listdir(HERE) if f.endswith(".npz")))
