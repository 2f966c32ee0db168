"""Writes tests/golden/jpg/{420,grey,422_restart,411,progressive_420,progressive_444_restart}.jpg (Pillow / OpenCV encoders) and, beside each, what stbi_load(.., 4)
of the stb_image vendored in the reference's SOIL2 returns for it. Needs oracle/_ref; run from the repo root."""
import os
import sys
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle                                   # noqa: E402
from test_images import _test_image             # noqa: E402
from PIL import Image                           # noqa: E402
import cv2                                      # noqa: E402

out = os.path.join(ROOT, "tests", "golden", "jpg")
os.makedirs(out, exist_ok=True)
assert oracle.have_stbi(), "build oracle/_ref first (make -C oracle ref)"
Image.fromarray(_test_image(40, 24, 1)).save(os.path.join(out, "420.jpg"), "JPEG", quality=75, subsampling=2)
Image.fromarray(_test_image(21, 13, 2)).convert("L").save(os.path.join(out, "grey.jpg"), "JPEG", quality=85)
Image.fromarray(_test_image(33, 20, 3)).save(os.path.join(out, "422_restart.jpg"), "JPEG", quality=60, subsampling=1, restart_marker_rows=1)
Image.fromarray(_test_image(45, 28, 5)).save(os.path.join(out, "progressive_420.jpg"), "JPEG", quality=70, subsampling=2, progressive=True)
Image.fromarray(_test_image(30, 22, 6)).save(os.path.join(out, "progressive_444_restart.jpg"), "JPEG", quality=90, subsampling=0, progressive=True,
                                             restart_marker_blocks=5)
cv2.imwrite(os.path.join(out, "411.jpg"), _test_image(37, 19, 4)[..., ::-1],
            [cv2.IMWRITE_JPEG_QUALITY, 88, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, cv2.IMWRITE_JPEG_SAMPLING_FACTOR_411])
for f in sorted(os.listdir(out)):
    if f.endswith(".jpg"):
        np.save(os.path.join(out, f[:-4] + "_stbi.npy"), oracle.ref_stbi_load(os.path.join(out, f)))
        print("wrote", f, os.path.getsize(os.path.join(out, f)))
