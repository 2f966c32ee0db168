"""Writes tests/golden/png/tiles.png (4-bit palette + tRNS, Adam7) and tiles_stbi.npy = stbi_load(.., 4) of the stb_image
vendored in the reference's SOIL2 for it. Needs oracle/_ref; run from the repo root:  python tools/make_png_golden.py"""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle                              # noqa: E402
from test_images import write_png          # noqa: E402

rng = np.random.default_rng(12)
w, h = 48, 32
idx = ((np.arange(h)[:, None] // 8 + np.arange(w)[None, :] // 8) % 6).astype(np.int64)[..., None]
idx[rng.random((h, w)) < 0.05] = 7
pal = rng.integers(0, 256, size=(8, 3), dtype=np.uint8)
out = os.path.join(ROOT, "tests", "golden", "png")
os.makedirs(out, exist_ok=True)
p = os.path.join(out, "tiles.png")
write_png(p, idx, 3, 4, palette=np.concatenate([pal, np.zeros((8, 3), np.uint8)]), trns=bytes([255, 255, 200, 128]), interlace=True)
assert oracle.have_stbi(), "build oracle/_ref first (make -C oracle ref)"
np.save(os.path.join(out, "tiles_stbi.npy"), oracle.ref_stbi_load(p))
print("wrote", p)
