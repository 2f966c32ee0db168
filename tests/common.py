"""Shared helpers for the parity tests: scenes with the REFERENCE builder's tree, cameras, comparisons."""
import functools
import os
import numpy as np

import oracle
from gpupathtracer_b200 import Camera, Scene, scenes

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

SCENES = {
    "cornell": lambda: scenes.cornell(n=4),
    "cornell_small": lambda: scenes.cornell(n=1),
    "soup": lambda: scenes.random_soup(n=20000, seed=7),
    "soup_small": lambda: scenes.random_soup(n=600, seed=3),
    "bunny": lambda: scenes.bunny_like(seed=1),
    "atrium_small": lambda: scenes.atrium(target_tris=20000, seed=2),
}


@functools.lru_cache(maxsize=None)
def get_scene(name, tree="reference"):
    """Scene + BVH. tree='reference': the reference's own BuildBinnedSpatial via oracle/_ref (falls back to the
    committed golden tree when oracle/_ref is absent and a fixture exists)."""
    s = Scene(SCENES[name]())
    if tree == "reference":
        if oracle.have_ref():
            nodes, refs = oracle.ref_build_sbvh(s.tris_abs)
        else:
            f = os.path.join(GOLDEN, f"tree_{name}.npz")
            if not os.path.exists(f):
                import pytest
                pytest.skip(f"oracle/_ref not built and no golden tree for {name}")
            z = np.load(f)
            nodes, refs = z["nodes"], z["refs"]
        s.set_bvh(nodes, refs)
    return s


def make_camera(data, width, height, aperture=None):
    c = data.camera
    cam = Camera(width / height, c["fov"], c["focus"], c["aperture"] if aperture is None else aperture)
    cam.SetPosition(c["pos"])
    cam.SetRotation(c["rot"])
    cam.GenerateImagePlane()
    return cam


def camera_rays(scene, width, height, n, seed=0):
    """Primary-like rays through random pixels (coherent origin, pinhole)."""
    cam = make_camera(scene.data, width, height, aperture=0.0).as_array()
    rng = np.random.default_rng(seed)
    s = rng.uniform(0, 1, size=(n, 1)).astype(np.float32)
    t = rng.uniform(0, 1, size=(n, 1)).astype(np.float32)
    d = cam[0:3] + s * cam[3:6] + t * cam[6:9]
    d = d / np.linalg.norm(d, axis=1, keepdims=True)
    rays = np.zeros((n, 8), dtype=np.float32)
    rays[:, 0:3] = cam[9:12]
    rays[:, 4:7] = d.astype(np.float32)
    rays[:, 7] = 1e20
    return rays


def mixed_rays(scene, n, seed=11, any_hit=False):
    lo, hi = scene.bounds()
    pad = 0.05 * (hi - lo)
    a = scenes.random_rays(n // 2, lo - pad, hi + pad, seed=seed, any_hit=any_hit)
    b = camera_rays(scene, 256, 256, n - n // 2, seed=seed + 1)
    if any_hit:
        diag = float(np.linalg.norm(hi - lo))
        b[:, 7] = np.random.default_rng(seed + 2).uniform(0, diag, size=b.shape[0]).astype(np.float32)
    return np.concatenate([a, b], axis=0)


def rel_mse(a, b, eps=1e-2):
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    ok = np.isfinite(a) & np.isfinite(b)
    return float(np.mean(((a - b) ** 2 / (b ** 2 + eps))[ok]))


def with_textures_and_sky(scene, seed=5):
    """A copy of `scene` whose materials (all but the sky material 0) look their albedo and roughness/metallic up in small
    random textures (odd sizes, a 1 x 1 one, some left constant) and whose environment is a random 8 x 8 cube map: every
    texture unit main() can touch is exercised."""
    import copy
    s = copy.copy(scene)
    g = np.random.default_rng(seed)
    s.textures = [g.uniform(0.05, 0.95, size=(hh, ww, 4)).astype(np.float32) for hh, ww in ((8, 8), (5, 7), (16, 4), (1, 1))]
    m = np.array(scene.materials, dtype=np.float32, copy=True)
    mi = m.view(np.int32)
    for k in range(1, m.shape[0]):
        mi[k, 3] = k % 5 - 1
        mi[k, 11] = (k * 3) % 5 - 1
    s.materials = m
    s.env_cubemap = g.uniform(0.0, 3.0, size=(6, 8, 8, 4)).astype(np.float32)
    s.env_cubemap[..., 3] = 1.0
    return s
