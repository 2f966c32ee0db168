"""Writes tests/golden/env/sky.hdr (a small synthetic Radiance sky: horizon gradient + sun) and sky_stbi.npy = what
stbi_loadf of the stb_image vendored in the reference (extern/soil2-cmake/src/SOIL2/stb_image.h) returns for it.
Needs oracle/_ref (i.e. /root/reference); run from the repo root:  python tools/make_env_golden.py"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle                                  # noqa: E402
from test_environment import write_hdr         # noqa: E402


def float_to_rgbe(img):
    m = img.max(axis=2)
    mant, exp = np.frexp(m)
    scale = np.where(m > 1e-32, mant * 256.0 / np.maximum(m, 1e-32), 0.0)
    out = np.zeros(img.shape[:2] + (4,), np.uint8)
    out[..., :3] = np.clip(img * scale[..., None], 0, 255).astype(np.uint8)
    out[..., 3] = np.where(m > 1e-32, exp + 128, 0).astype(np.uint8)
    return out


def main():
    w, h = 64, 32
    u = (np.arange(w) + 0.5) / w
    v = (np.arange(h) + 0.5) / h
    U, V = np.meshgrid(u, v)
    sky = np.zeros((h, w, 3))
    up = np.clip(V, 0, 1)
    sky[..., 0] = 0.4 + 0.5 * up
    sky[..., 1] = 0.6 + 0.3 * up
    sky[..., 2] = 1.2 - 0.6 * up
    sun = np.exp(-((U - 0.3) ** 2 + (V - 0.75) ** 2) / 0.002)
    sky += 400.0 * sun[..., None] * np.array([1.0, 0.9, 0.7])
    sky[V < 0.35] = np.array([0.08, 0.07, 0.05])          # ground: long runs for the RLE
    rgbe = float_to_rgbe(sky)
    out = os.path.join(ROOT, "tests", "golden", "env")
    os.makedirs(out, exist_ok=True)
    write_hdr(os.path.join(out, "sky.hdr"), rgbe, rle=True)
    assert oracle.have_stbi(), "build oracle/_ref first (make -C oracle ref)"
    ref = oracle.ref_stbi_loadf(os.path.join(out, "sky.hdr"))
    np.save(os.path.join(out, "sky_stbi.npy"), ref)
    print("wrote", out, ref.shape, float(ref.max()))


if __name__ == "__main__":
    main()
