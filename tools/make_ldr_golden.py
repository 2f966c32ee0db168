"""Writes tests/golden/ldr/*.{tga,bmp} (a subset of tests/ldr_cases.py at 19 x 11 pixels) and, beside each, what stbi_load(.., 4)
of the stb_image vendored in the reference's SOIL2 returns for it. Needs oracle/_ref; run from the repo root."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle                                   # noqa: E402
import ldr_cases                                # noqa: E402
from test_images import _test_image             # noqa: E402

out = os.path.join(ROOT, "tests", "golden", "ldr")
os.makedirs(out, exist_ok=True)
assert oracle.have_stbi(), "build oracle/_ref first (make -C oracle ref)"
rgb = _test_image(19, 11, 9)
keep_pillow = ("rgb_rle.tga", "grey.tga", "palette_rle.tga", "grey_alpha.tga", "palette16.bmp", "rgba.bmp")
for name, (img, opts) in ldr_cases.pillow_cases(rgb).items():
    if name in keep_pillow:
        img.save(os.path.join(out, name), **opts)
for name, data in ldr_cases.handmade_cases(rgb).items():
    if name in ("rgb565_bitfields.bmp", "bgra_zero_alpha.bmp", "os2_palette.bmp", "rgb15_rle.tga", "cmap24_idx16_start2.tga"):
        open(os.path.join(out, name), "wb").write(data)
for f in sorted(os.listdir(out)):
    if f.endswith((".tga", ".bmp")):
        np.save(os.path.join(out, f + "_stbi.npy"), oracle.ref_stbi_load(os.path.join(out, f)))
        print("wrote", f, os.path.getsize(os.path.join(out, f)))
