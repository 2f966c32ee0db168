"""The parity libm (gpupathtracer_b200/csrc/pt_parity_libm.h): one definition of GLSL's sin / cos / log / exp / pow for the CUDA
kernels (pt_config.parity_flags & PT_PARITY_LIBM), the CPU oracle (or_set_parity_libm) and the reference's shader source compiled
as C++ (model codes 4 / 6). CPU half: it is a faithful libm, and the oracle and the reference's main() agree bit for bit with it
switched on. The GPU half (device == host bits, frames == reference frames) is tests/test_gpu_exact.py."""
import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import make_rng_table
from common import get_scene, make_camera, with_textures_and_sky


def _ulps(a, b):
    ia = a.view(np.int32).astype(np.int64); ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7fffffff), ia); ib = np.where(ib < 0, -(ib & 0x7fffffff), ib)
    return np.abs(ia - ib)


def libm_cases(n=400000, seed=0):
    """Arguments the shading path produces (phi in [0, 2 pi], log of (0, 1], Beckmann exponents) plus wide ranges and specials."""
    g = np.random.default_rng(seed)
    phi = (2.0 * np.float32(np.pi) * g.uniform(0, 1, n).astype(np.float32)).astype(np.float32)
    wide = g.uniform(-1000, 1000, n).astype(np.float32)
    unit = g.uniform(0, 1, n).astype(np.float32)
    pos = np.exp(g.uniform(-80, 80, n)).astype(np.float32)
    ex = g.uniform(-110, 95, n).astype(np.float32)
    special = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 1e-45, 3.4e38, -3.4e38, 1e-38, 88.73, -103.98, 0.5, 2.0], np.float32)
    return {"sin": np.concatenate([phi, wide, special]), "cos": np.concatenate([phi, wide, special]),
            "log": np.concatenate([unit, pos, special]), "exp": np.concatenate([ex, -unit * 20, special]),
            "pow": (np.concatenate([unit * 4, pos[: n // 4], special]), np.concatenate([np.full(n, 1 / 2.2, np.float32), g.uniform(-3, 3, n // 4).astype(np.float32), special[::-1]]))}


@pytest.mark.parametrize("fn", ["sin", "cos", "log", "exp", "pow"])
def test_parity_libm_is_a_faithful_libm(fn):
    """<= 1 ulp of binary32 from the binary64 libm's result rounded once; equal to it on all but a vanishing fraction."""
    c = libm_cases()[fn]
    x, y = c if fn == "pow" else (c, None)
    got = oracle.parity_libm_probe(fn, x, y)
    with np.errstate(all="ignore"):
        xd = x.astype(np.float64)
        want = {"sin": np.sin, "cos": np.cos, "log": np.log, "exp": np.exp}[fn](xd) if fn != "pow" else np.power(xd, y.astype(np.float64))
        want = want.astype(np.float32)
    dom = np.isfinite(x) & (np.abs(x) < 2e5) if fn in ("sin", "cos") else np.ones(x.shape, bool)
    if fn == "pow":
        dom = (x >= 0) & np.isfinite(x) & np.isfinite(y) & ~((x == 0) & (y <= 0))      # GLSL leaves the rest undefined
    both_nan = np.isnan(got) & np.isnan(want)
    u = _ulps(got, want)
    ok = both_nan | (u <= 1)
    assert ok[dom].all(), f"{fn}: {x[dom][~ok[dom]][:5]} -> {got[dom][~ok[dom]][:5]} vs {want[dom][~ok[dom]][:5]}"
    assert ((u > 0) & ~both_nan)[dom].mean() < 1e-5


def test_parity_libm_specials():
    p = oracle.parity_libm_probe
    z = np.array([0.0], np.float32)
    assert p("log", z)[0] == -np.inf and np.isnan(p("log", -np.ones(1, np.float32))[0]) and p("log", np.array([np.inf], np.float32))[0] == np.inf
    assert p("exp", np.array([-200.0], np.float32))[0] == 0.0 and p("exp", np.array([100.0], np.float32))[0] == np.inf
    assert p("exp", z)[0] == 1.0 and p("sin", z)[0] == 0.0 and p("cos", z)[0] == 1.0
    assert p("pow", z, np.array([0.4545], np.float32))[0] == 0.0 and p("pow", np.array([2.0], np.float32), np.array([10.0], np.float32))[0] == 1024.0
    assert np.isnan(p("sin", np.array([np.inf], np.float32))[0]) and np.isnan(p("pow", -np.ones(1, np.float32), np.ones(1, np.float32))[0])


@pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built")
@pytest.mark.parametrize("name,model,code", [("cornell", 0, 4), ("atrium_small", 1, 6), ("soup_small", 1, 6)])
def test_oracle_equals_reference_main_under_parity_libm(name, model, code):
    """or_render with the parity libm == the reference's main() compiled with the parity libm (textures, cube-map sky, thin lens),
    image and RNG table; and the switch matters: a few pixels differ from the glibc build."""
    s = with_textures_and_sky(get_scene(name))
    w, h, frames = 64, 48, 2
    cam = make_camera(s.data, w, h, aperture=0.2).as_array()
    o = oracle.Oracle(s)
    rng0 = make_rng_table(w * h, 3)
    rng_pl, acc_pl = rng0.copy(), np.zeros((h, w, 4), np.float32)
    o.render(cam, w, h, rng_pl, acc_pl, n_frames=frames, model=model, parity_libm=True)
    rng_g, acc_g = rng0.copy(), np.zeros((h, w, 4), np.float32)
    for _ in range(frames):
        o.ref_glsl_render(cam, w, h, rng_g, acc_g, code)
    assert np.array_equal(acc_pl.view(np.uint32), acc_g.view(np.uint32)) and np.array_equal(rng_pl, rng_g)
    rng_d, acc_d = rng0.copy(), np.zeros((h, w, 4), np.float32)
    o.render(cam, w, h, rng_d, acc_d, n_frames=frames, model=model)
    differ = (acc_d.view(np.uint32) != acc_pl.view(np.uint32)).any(axis=2).mean()
    assert differ < 0.05                                   # 1-2 ulp differences of glibc's float functions reach few pixels
    assert oracle.lib().or_get_parity_libm() == 0          # the switch does not leak out of the call


@pytest.mark.skipif(not oracle.have_glsl(), reason="oracle/_ref/libref_glsl.so not built")
def test_present_under_parity_libm():
    g = np.random.default_rng(2)
    a = g.uniform(0, 40, size=(16, 24, 4)).astype(np.float32)
    a[0, 0, :3] = 0.0
    want = oracle.ref_glsl_present(a, 7, parity_libm=True)
    oracle.set_parity_libm(True)
    try:
        got = oracle.tonemap(a[..., :3], 7).reshape(16, 24, 3)
    finally:
        oracle.set_parity_libm(False)
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    np.testing.assert_allclose(oracle.ref_glsl_present(a, 7), want, rtol=3e-7, atol=0)
