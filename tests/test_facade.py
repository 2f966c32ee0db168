"""The C++ facade (include/b200pt_renderer.hpp) and the reference's main() without the window (tools/headless_main.cpp,
Program.cpp:55-152): compiles against the C ABI with plain g++, fails loudly without a GPU, renders scene.txt -> PNG on one."""
import os
import subprocess
import numpy as np
import pytest

from gpupathtracer_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "gpupathtracer_b200")


def _build(tmp_path):
    api.load_library()                               # builds libb200pt.so if it is missing
    exe = str(tmp_path / "headless")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    cmd = [cxx, "-std=c++17", "-Wall", "-Werror", os.path.join(ROOT, "tools", "headless_main.cpp"), "-I", os.path.join(ROOT, "include"),
           "-L", LIBDIR, "-lb200pt", f"-Wl,-rpath,{LIBDIR}", "-o", exe]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return exe


def test_headless_main_builds_and_refuses_to_run_without_a_gpu(tmp_path):
    exe = _build(tmp_path)
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by test_headless_main_renders_scene_txt")
    r = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "obj", "scene.txt"), "1", str(tmp_path / "o.png")],
                       capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 1 and not (tmp_path / "o.png").exists()
    assert "device" in r.stderr.lower() or "cuda" in r.stderr.lower()       # PT_ERR_NO_DEVICE: there is no CPU path


@pytest.mark.gpu
def test_headless_main_renders_scene_txt(tmp_path):
    exe = _build(tmp_path)
    out = str(tmp_path / "room.png")
    r = subprocess.run([exe, os.path.join(ROOT, "tests", "golden", "obj", "scene.txt"), "8", out], capture_output=True, text=True, cwd=ROOT)
    assert r.returncode == 0, r.stderr
    assert "8 samples per pixel" in r.stdout
    img = api.load_image(out)                        # our own PNG reader (held against stb_image in test_images.py)
    assert img.shape == (720, 1280, 4)               # Program.cpp:22-23 / Renderer.cpp:698
    rgb = img[..., :3]
    assert np.isfinite(rgb).all() and 0.02 < float(rgb.mean()) < 0.98 and float(rgb.std()) > 0.01
