"""CPU: the C-ABI library loads and exports everything include/b200pt.h declares; host-side logic (scene assembly,
light table, builder, error behaviour without a GPU). No device compute is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import (LIB_PATH, SYMBOLS, Scene, build_bvh_host, load_library, scenes, assemble_triangles,
                                build_light_table, presubtract_edges)
from gpupathtracer_b200.api import pt_config
from common import SCENES, get_scene

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "b200pt.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = load_library()
    names = header_symbols()
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/b200pt.h but not exported by {LIB_PATH}"
    bound = {n for n, _, _ in SYMBOLS}
    assert set(names) == bound, f"python binding and header disagree: {set(names) ^ bound}"
    assert lib.pt_abi_version() == 3


def test_no_cpu_fallback_create_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = load_library()
    cfg = pt_config(width=64, height=64, max_depth=5, microfacet_model=0, device=0, bvh_layout=0)
    ctx = C.c_void_p()
    rc = lib.pt_create(C.byref(cfg), C.byref(ctx))
    assert rc == -2 and not ctx.value                      # PT_ERR_NO_DEVICE
    assert b"no CPU path" in lib.pt_last_error()


def test_create_rejects_bad_config():
    lib = load_library()
    ctx = C.c_void_p()
    for kw in (dict(width=65, height=64), dict(width=64, height=0), dict(width=64, height=64, max_depth=0),
               dict(width=64, height=64, max_depth=5, microfacet_model=7)):
        base = dict(width=64, height=64, max_depth=5, microfacet_model=0, device=0, bvh_layout=0)
        base.update(kw)
        assert lib.pt_create(C.byref(pt_config(**base)), C.byref(ctx)) == -1     # PT_ERR_INVALID
        assert lib.pt_last_error()


def test_assemble_triangles_follows_scene_cpp():
    # Scene.cpp:263-292: geometric normal, flipped to agree with the mean vertex normal; material of vertex 0
    v = np.zeros((2, 3, 8), np.float32)
    v[:, 0, 0:3] = (0, 0, 0); v[:, 1, 0:3] = (1, 0, 0); v[:, 2, 0:3] = (0, 1, 0)
    v[0, :, 3:6] = (0, 0, 1)          # agrees with +z
    v[1, :, 3:6] = (0, 0, -1)         # disagrees -> flipped
    v[:, :, 6:8] = [[0, 0], [1, 0], [0, 1]]
    t = assemble_triangles(v, np.array([2, 4], np.uint32))
    assert np.allclose(t[0, 15:18], (0, 0, 1)) and np.allclose(t[1, 15:18], (0, 0, -1))
    assert t[:, 18].view(np.uint32).tolist() == [2, 4]
    assert np.array_equal(t[0, 0:9], [0, 0, 0, 1, 0, 0, 0, 1, 0])
    p = presubtract_edges(np.array([[1, 2, 3, 2, 2, 3, 1, 4, 3] + [0] * 11], np.float32))
    assert np.array_equal(p[0, 0:9], [1, 2, 3, 1, 0, 0, 0, 2, 0])          # Scene.cpp:334-337


def test_light_table_is_sorted_cumulative_and_matches_numpy():
    s = Scene(scenes.random_soup(n=3000, seed=11))
    lights, total = build_light_table(s.tris_abs, s.materials)
    idx = lights[:, 1].view(np.uint32)
    mat = s.tris_abs[idx, 18].view(np.uint32)
    assert np.all(s.materials[mat // 2, 7].view(np.int32) == 1)            # emissive only (Scene.cpp:303)
    cum = lights[:, 0]
    assert np.all(np.diff(cum) >= 0) and cum[-1] == np.float32(total)
    areas = np.diff(np.concatenate([[0], cum.astype(np.float64)]))
    assert np.all(np.diff(areas) >= -1e-5 * areas.max())                   # ascending area (Scene.cpp:322-325)
    p = s.tris_abs[idx, 0:9].reshape(-1, 3, 3).astype(np.float64)
    ref_area = 0.5 * np.linalg.norm(np.cross(p[:, 1] - p[:, 0], p[:, 2] - p[:, 0]), axis=1)
    np.testing.assert_allclose(areas, ref_area, rtol=2e-3, atol=1e-6)       # Heron in fp32 vs cross product in fp64


@pytest.mark.parametrize("name", ["cornell", "soup", "atrium_small"])
def test_own_builder_tree_is_valid_and_equivalent(name):
    """The library's SBVH builder (pt_bvh_build_host) emits the reference's serialised form; traversed by the ORACLE it
    must return the same closest hits as the reference builder's tree, at no worse SAH cost."""
    s = Scene(SCENES[name]())
    nodes, refs, st = build_bvh_host(s.tris, n_threads=4)
    n = nodes.shape[0]
    d0 = nodes[:, 3].view(np.int32)
    assert n % 2 == 1 and d0[0] == 1 and st["max_depth"] < 60
    inner = d0 > 0
    assert np.all(d0[inner] % 2 == 1) and np.all(d0[inner] + 1 < n)
    # every node except the root is referenced exactly once
    kids = np.concatenate([d0[inner], d0[inner] + 1])
    assert np.array_equal(np.sort(kids), np.arange(1, n))
    # children boxes lie inside their parent's box
    for k in (0, 1):
        c = d0[inner] + k
        assert np.all(nodes[c, 0:3] >= nodes[inner, 0:3] - 1e-6) and np.all(nodes[c, 4:7] <= nodes[inner, 4:7] + 1e-6)
    # every triangle is referenced, every leaf run is terminated
    assert set(np.abs(refs).tolist()) == set(range(s.tris.shape[0]))
    leaf_off = -d0[~inner]
    assert np.all(leaf_off < refs.shape[0])
    # determinism across thread counts
    n2, r2, _ = build_bvh_host(s.tris, n_threads=1)
    assert np.array_equal(nodes.view(np.uint32), n2.view(np.uint32)) and np.array_equal(refs, r2)
    rays = scenes.random_rays(30000, *s.bounds(), seed=9)
    s.set_bvh(nodes, refs)
    mine, cm = oracle.Oracle(s).trace_closest(rays)
    if oracle.have_ref():
        rn, rr = oracle.ref_build_sbvh(s.tris_abs)
        s.set_bvh(rn, rr)
        ref, cr = oracle.Oracle(s).trace_closest(rays)
        same = mine[:, 3].view(np.int32) == ref[:, 3].view(np.int32)
        assert same.mean() > 0.9995
        assert np.allclose(mine[same, 0], ref[same, 0], rtol=1e-6)
        # quality: no more traversal work than the reference's tree (+5 %)
        assert cm["node_pairs"] <= 1.05 * cr["node_pairs"]
    bf = oracle.Oracle(s).trace_closest_bruteforce(rays[:3000])
    assert (mine[:3000, 3].view(np.int32) == bf[:, 3].view(np.int32)).mean() > 0.999


def test_degenerate_inputs_for_builder():
    lib = load_library()
    # coincident triangles still give an inner root
    t = np.zeros((4, 20), np.float32); t[:, 3] = 1; t[:, 7] = 1
    nodes, refs, st = build_bvh_host(t)
    assert nodes.shape[0] >= 3 and nodes[0, 3].view(np.int32) == 1
    with pytest.raises(Exception):
        build_bvh_host(np.zeros((1, 20), np.float32))       # root would be a leaf (BVH.glsl:627 assumes an inner root)


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the arm the driver times beside the CUDA one) runs without a GPU and prints one JSON line
    with the contract's keys; it times the reference's own code when oracle/_ref was built, the restatement otherwise."""
    import json, subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--workload", "bunny",
                          "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "Mrays/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": "Mrays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["steps"] == 1 and d["warmup"] == 0 and d["n_gpus"] == 1 and d["vs_baseline"] is None


def test_no_reference_text_is_left_in_the_tree():
    """oracle/Makefile extracts line ranges of the reference at build time; they are make intermediates and must be gone once
    the checker libraries are linked: oracle/_ref holds built libraries only."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    d = os.path.join(root, "oracle", "_ref")
    if not os.path.isdir(d):
        pytest.skip("oracle/_ref not built")
    left = [f for f in os.listdir(d) if not f.endswith(".so")]
    assert left == [], left


@pytest.mark.parametrize("name,tree", [("atrium_small", "own"), ("soup", "own"), ("cornell", "reference")])
def test_reinsertion_optimisation_keeps_the_tree_valid_and_the_hits(name, tree):
    """pt_bvh_optimize_host (the pass the reference carries as dead code, ReinsertionOptimize): subtrees move, leaves and reference
    list do not. The result is still the reference's serialised form (every node referenced once, children adjacent at an odd index,
    larger child first, boxes nested), it has the same leaves, a SAH cost that is no higher, and the ORACLE walking it returns the same
    hits bit for bit with fewer node visits."""
    import ctypes as C
    from gpupathtracer_b200 import load_library
    from gpupathtracer_b200.api import _ptr
    lib = load_library()
    s = Scene(SCENES[name]())
    if tree == "reference":
        if not oracle.have_ref():
            pytest.skip("oracle/_ref not built")
        nodes, refs = oracle.ref_build_sbvh(s.tris_abs)
    else:
        nodes, refs, _ = build_bvh_host(s.tris, n_threads=4)
    rays = scenes.random_rays(40000, *s.bounds(), seed=21)
    s.set_bvh(nodes.copy(), refs.copy())
    want, c0 = oracle.Oracle(s).trace_closest(rays)
    opt = np.ascontiguousarray(nodes.copy())
    before, after = C.c_float(), C.c_float()
    moved = lib.pt_bvh_optimize_host(_ptr(opt), opt.shape[0], _ptr(refs), refs.shape[0], 4, 0.25, C.byref(before), C.byref(after))
    assert moved > 0 and after.value <= before.value
    n = opt.shape[0]
    d0, d0_in = opt[:, 3].copy().view(np.int32), nodes[:, 3].copy().view(np.int32)
    inner = d0 > 0
    assert n == nodes.shape[0] and d0[0] == 1 and np.all(d0[inner] % 2 == 1)
    assert np.array_equal(np.sort(np.concatenate([d0[inner], d0[inner] + 1])), np.arange(1, n))
    assert np.array_equal(np.sort(d0[~inner]), np.sort(d0_in[d0_in <= 0]))                     # the same leaves
    leaf_boxes = lambda a, d: a[d <= 0][np.argsort(d[d <= 0], kind="stable")]
    assert np.array_equal(leaf_boxes(opt, d0)[:, [0, 1, 2, 4, 5, 6]], leaf_boxes(nodes, d0_in)[:, [0, 1, 2, 4, 5, 6]])
    for k in (0, 1):
        c = d0[inner] + k
        assert np.all(opt[c, 0:3] >= opt[inner, 0:3]) and np.all(opt[c, 4:7] <= opt[inner, 4:7])
    ext = opt[:, 4:7] - opt[:, 0:3]
    area = ext[:, 0] * (ext[:, 1] + ext[:, 2]) + ext[:, 1] * ext[:, 2]
    assert np.all(area[d0[inner]] >= area[d0[inner] + 1])                                       # BVH.cpp:2240-2247
    s.set_bvh(opt, refs.copy())
    got, c1 = oracle.Oracle(s).trace_closest(rays)
    ids_same = got[:, 3].view(np.int32) == want[:, 3].view(np.int32)
    assert ids_same.mean() > 1 - 1e-4                        # a different tree: only ties / box-grazing rays may differ
    assert np.array_equal(got[ids_same].view(np.uint32), want[ids_same].view(np.uint32))
    assert c1["node_pairs"] <= c0["node_pairs"]
    # doing nothing is allowed too: a tree below 7 nodes is returned as it is
    tiny = np.ascontiguousarray(nodes[:3].copy())
    assert lib.pt_bvh_optimize_host(_ptr(tiny), 3, _ptr(refs), refs.shape[0], 1, 0.5, None, None) == 0
    assert lib.pt_bvh_optimize_host(_ptr(opt), n, _ptr(refs), refs.shape[0], 0, 0.5, None, None) < 0      # bad argument
