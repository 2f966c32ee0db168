"""OBJ/MTL + scene.txt loading (csrc/pt_obj.cpp) against the reference's parser.

The reference parses with tinyobjloader and then applies LoadOBJ's rules (Scene.cpp:120-231). `expected_from_tinyobj`
restates those rules on tinyobj's raw output: live when oracle/_ref/libref_tinyobj.so exists (built from
/root/reference/extern/tinyobjloader in place), else from tests/golden/obj_room_tinyobj.npz recorded by make_golden.py."""
import glob
import os
import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import api

HERE = os.path.dirname(os.path.abspath(__file__))
ROOM = os.path.join(HERE, "golden", "obj", "room.obj")
TINYOBJ_MODELS = "/root/reference/extern/tinyobjloader/models"


def expected_from_tinyobj(t, folder):
    """LoadOBJ, Scene.cpp:153-231 + CreateMatInstance :86-118 + sky :236-241, on tinyobj's raw result."""
    n_faces = t["face_mat"].shape[0]
    verts = np.zeros((n_faces, 3, 8), np.float32)
    ci = t["corner_idx"].reshape(n_faces, 3, 3)
    verts[:, :, 0:3] = t["v"][ci[:, :, 0]]
    has_n = ci[:, :, 1] >= 0
    verts[:, :, 3:6] = np.where(has_n[..., None], t["vn"][np.maximum(ci[:, :, 1], 0)] if len(t["vn"]) else 0, (0.0, 1.0, 0.0))
    has_t = ci[:, :, 2] >= 0
    verts[:, :, 6:8] = np.where(has_t[..., None], t["vt"][np.maximum(ci[:, :, 2], 0)] if len(t["vt"]) else 0, 0.0)
    mat_ids = np.zeros(n_faces, np.uint32)
    sf = t["shape_first_face"]
    for s in range(len(sf) - 1):
        first = int(t["face_mat"][sf[s]])                      # material_ids[0] of the shape, Scene.cpp:182
        mat_ids[sf[s]:sf[s + 1]] = 0 if first < 0 else 2 * (first + 1)
    m = t["materials"]
    mats = np.zeros((m.shape[0] + 1, 12), np.float32)
    mats[0, 4:7] = (750.0, 650.0, 475.0)
    mats[0, 9] = 1.0
    iv = mats.view(np.int32)
    iv[:, 3] = -1; iv[:, 11] = -1
    iv[0, 7] = 1
    for k in range(m.shape[0]):
        kd, ks, ke, ns = m[k, 0:3], m[k, 3:6], m[k, 6:9], m[k, 9]
        mats[k + 1, 0:3] = kd
        mats[k + 1, 4:7] = ke
        iv[k + 1, 7] = 1 if np.float32(ke[0] + ke[1]) + ke[2] > np.float32(1e-5) else 0
        mats[k + 1, 9] = np.sqrt(np.float32(2.0) / (ns + np.float32(2.0)))
        mats[k + 1, 10] = 1.0 if ks.max() > np.float32(0.3) else 0.0
    tex = [""] + [(folder + n if n else "") for n in t["texnames"]]
    return verts, mat_ids, mats, tex


def _tinyobj(path):
    if oracle.have_tinyobj() and os.path.exists(path):
        return oracle.ref_tinyobj_load(path)
    if os.path.abspath(path) == os.path.abspath(ROOM):
        z = np.load(os.path.join(HERE, "golden", "obj_room_tinyobj.npz"), allow_pickle=False)
        t = {k: z[k] for k in z.files}
        t["texnames"] = [s for s in str(t["texnames"]).split("\n")]
        return t
    pytest.skip("oracle/_ref/libref_tinyobj.so not built")


def _compare(path):
    got = api.load_obj(path)
    want = expected_from_tinyobj(_tinyobj(path), os.path.dirname(path) + "/")
    assert got["verts"].shape == want[0].shape, "triangle count differs from tinyobj's triangulation"
    assert np.array_equal(got["verts"].view(np.uint32), want[0].view(np.uint32)), "vertex records differ"
    assert np.array_equal(got["mat_ids"], want[1])
    assert np.array_equal(got["materials"].view(np.uint32), want[2].view(np.uint32)), "material records differ"
    assert got["texture_paths"] == want[3]
    return got


def test_room_fixture_matches_tinyobj_and_reference_rules():
    got = _compare(ROOM)
    # 5 quads + lamp quad + pentagon (3) + skew quad (2) + 1 triangle
    assert got["verts"].shape[0] == 2 * 5 + 2 + 3 + 2 + 1
    assert got["n_shapes"] == 6
    assert "nosuchmaterial" in got["warnings"]
    # a shape keeps its FIRST face's material: the "walls" shape is all red although two faces say checker
    assert (got["mat_ids"][4:10] == 2 * 2).all()
    # unknown material -> sky (0); repeated `newmtl white` keeps the first record for lookups but still adds a material
    assert got["materials"].shape[0] == 1 + 6 and (got["mat_ids"][-3:-1] == 0).all()
    assert got["texture_paths"][5].endswith("obj/checker.ppm")


@pytest.mark.skipif(not (oracle.have_tinyobj() and os.path.isdir(TINYOBJ_MODELS)), reason="needs /root/reference + oracle/_ref")
def test_tinyobjloader_model_zoo_in_place():
    """Every triangle/quad model of the reference's vendored tinyobjloader test models, parsed where it lies."""
    n_ok = 0
    for path in sorted(glob.glob(os.path.join(TINYOBJ_MODELS, "*.obj"))):
        try:
            t = oracle.ref_tinyobj_load(path)
        except RuntimeError:
            continue
        if t["face_mat"].shape[0] == 0 or t["corner_idx"][:, 0].min() < 0:
            continue                                   # line/point-only files, or faces tinyobj itself flags invalid
        if t["corner_idx"][:, 1].max() >= len(t["vn"]) or t["corner_idx"][:, 2].max() >= len(t["vt"]):
            continue                                   # normal / texcoord index past the arrays: LoadOBJ reads out of bounds
        with open(path, errors="ignore") as f:
            big_polys = any(len(l.split()) > 5 for l in f if l.startswith("f "))
        if big_polys:
            continue                                   # > 4 corners: ear clipping order is implementation-defined
        _compare(path)
        n_ok += 1
    assert n_ok >= 15, f"only {n_ok} models compared"


def test_scene_txt_and_environment():
    path, env, pos, rot = api.read_scene_txt(os.path.join(HERE, "golden", "obj", "scene.txt"))
    assert path == "tests/golden/obj/room.obj" and env == "GENERATE COLOR WHITE"
    assert pos == (0.0, 1.0, np.float32(0.9)) and rot == (0.0, 0.0, 0.0)
    assert api.parse_environment("GENERATE COLOR WHITE") == (1.0, 1.0, 1.0)
    assert api.parse_environment("GENERATE COLOR BLACK") == (0.0, 0.0, 0.0)
    assert api.parse_environment("GENERATE COLOR PURPLE") == (1.0, 0.0, 0.0)          # "RED for error", Renderer.cpp:251
    with pytest.raises(api.PtError) as e:
        api.parse_environment("res/sky/ibl/NarrowPath_3k.hdr")
    assert e.value.code == -5
    with pytest.raises(api.PtError) as e:
        api.load_obj("/nonexistent/file.obj")
    assert e.value.code == -4


def test_image_readers_roundtrip(tmp_path):
    """map_Kd decoders: the same 3x2 image as PPM, BMP (bottom-up) and TGA (raw + RLE) must give the same texels."""
    import ctypes as C
    import struct
    img = np.array([[[255, 0, 0], [0, 255, 0], [0, 0, 255]], [[10, 20, 30], [40, 50, 60], [70, 80, 90]]], np.uint8)
    h, w = img.shape[:2]
    (tmp_path / "a.ppm").write_bytes(b"P6 %d %d 255\n" % (w, h) + img.tobytes())
    row = lambda y: img[y, :, ::-1].tobytes() + b"\0" * ((4 - (3 * w) % 4) % 4)
    bmp = b"".join(row(y) for y in range(h - 1, -1, -1))
    (tmp_path / "a.bmp").write_bytes(b"BM" + struct.pack("<IHHI", 54 + len(bmp), 0, 0, 54) +
                                     struct.pack("<IiiHHIIiiII", 40, w, h, 1, 24, 0, len(bmp), 2835, 2835, 0, 0) + bmp)
    hdr = lambda typ, desc: struct.pack("<BBBHHBHHHHBB", 0, 0, typ, 0, 0, 0, 0, 0, w, h, 24, desc)
    (tmp_path / "a.tga").write_bytes(hdr(2, 0x20) + img[:, :, ::-1].tobytes())
    rle = b"".join(bytes([0x80]) + img[y, x, ::-1].tobytes() for y in range(h - 1, -1, -1) for x in range(w))
    (tmp_path / "b.tga").write_bytes(hdr(10, 0x00) + rle)
    for name in ("a.ppm", "a.bmp", "a.tga", "b.tga"):
        obj = tmp_path / (name + ".obj")
        (tmp_path / (name + ".mtl")).write_text(f"newmtl m\nKd 1 1 1\nKe 1 1 1\nmap_Kd {name}\n")
        obj.write_text(f"mtllib {name}.mtl\nv 0 0 0\nv 1 0 0\nv 0 1 0\nusemtl m\nf 1 2 3\n")
        got = api.load_obj(str(obj))
        assert got["texture_paths"][1] == str(tmp_path / name)


def test_room_scene_data_feeds_the_oracle():
    """scene_data_from_obj -> Scene (assembly, light table) -> own builder -> oracle: the textured wall is sampled."""
    from gpupathtracer_b200 import Scene, build_bvh_host
    data = api.scene_data_from_obj(ROOM)
    assert len(data.textures) == 1 and data.textures[0].shape == (4, 4, 4)
    assert np.allclose(data.textures[0][0, 0, :3], np.array([230, 230, 230]) / 255.0)
    s = Scene(data)
    # the lamp quad (0.36) + the quad whose material is unknown: it falls back to the sky material, which IS emissive
    # (Scene.cpp:236-241), so the reference samples it as a light too
    assert s.lights.shape[0] == 4 and s.total_light_area > 0.36
    nodes, refs, _ = build_bvh_host(s.tris)
    s.set_bvh(nodes, refs)
    # (an exactly axis-parallel ray misses everything in the reference's slab test: 0 * inf; tilt it)
    d = np.array([0.01, 1.0, 0.02]); d /= np.linalg.norm(d)
    rays = np.zeros((1, 8), np.float32); rays[0, 0:3] = (0.1, 1, 0.05); rays[0, 4:7] = d; rays[0, 7] = 1e20
    hit, _ = oracle.Oracle(s).trace_closest(rays)
    assert abs(hit[0, 0] - 0.99 / d[1]) < 1e-5                                       # upwards: the lamp at y = 1.99


@pytest.mark.gpu
def test_load_scene_from_files_matches_array_upload():
    """pt_load_scene (parse + assemble + textures + lights + BVH, all inside the library) must render the same image,
    bit for bit, as uploading the same scene array by array through the Python mirror; and both match the oracle."""
    from gpupathtracer_b200 import Renderer, Scene, Camera, make_rng_table
    path, env, pos, rot = api.read_scene_txt(os.path.join(HERE, "golden", "obj", "scene.txt"))
    w = h = 64
    cam = Camera(1.0, np.radians(60.0), 1.5, 0.0)
    cam.SetPosition(pos); cam.SetRotation(rot); cam.GenerateImagePlane()
    rng = make_rng_table(w * h, 3)
    a = Renderer().InitializeFromFile(w, h, ROOM, env, max_depth=5)
    a.SetRNG(rng); a.RenderFrame(cam, 4)
    img_a = a.DownloadAccum()
    nodes, refs = a.GetBVH()
    a.CleanUp()
    s = Scene(api.scene_data_from_obj(ROOM, env_color=api.parse_environment(env)))
    s.set_bvh(nodes, refs)
    b = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng)
    b.RenderFrame(cam, 4)
    img_b = b.DownloadAccum()
    b.CleanUp()
    assert np.array_equal(img_a.view(np.uint32), img_b.view(np.uint32))
    want = np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam.as_array(), w, h, rng.copy(), want, n_frames=4, max_depth=5)
    close = np.isclose(img_a[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2)
    assert close.mean() > 0.995
    assert img_a[..., :3].max() > 0


@pytest.mark.gpu
@pytest.mark.parametrize("texture", ["jpg/420.jpg", "png/tiles.png"])
def test_room_with_jpeg_and_png_textures(tmp_path, texture):
    """The same room with its map_Kd pointing at a JPEG / PNG: pt_load_scene decodes it (pt_image.cpp) and the frames equal
    those of the array upload fed by the Python-side decode of the same file, and the oracle's."""
    import shutil
    from gpupathtracer_b200 import Renderer, Scene, Camera, make_rng_table
    src = os.path.join(HERE, "golden", "obj")
    for f in ("room.obj",):
        shutil.copy(os.path.join(src, f), tmp_path / f)
    tex = os.path.basename(texture)
    shutil.copy(os.path.join(HERE, "golden", texture), tmp_path / tex)
    mtl = open(os.path.join(src, "room.mtl")).read().replace("checker.ppm", tex)
    (tmp_path / "room.mtl").write_text(mtl)
    obj = str(tmp_path / "room.obj")
    data = api.scene_data_from_obj(obj)
    assert len(data.textures) == 1 and data.textures[0].shape[0] > 4       # the real image, not the 4x4 checker
    w = h = 64
    cam = Camera(1.0, np.radians(60.0), 1.5, 0.0)
    cam.SetPosition((0.0, 1.0, 0.9)); cam.SetRotation((0.0, 0.0, 0.0)); cam.GenerateImagePlane()
    rng = make_rng_table(w * h, 5)
    a = Renderer().InitializeFromFile(w, h, obj, "GENERATE COLOR WHITE", max_depth=5)
    a.SetRNG(rng); a.RenderFrame(cam, 4)
    img_a = a.DownloadAccum()
    nodes, refs = a.GetBVH()
    a.CleanUp()
    s = Scene(data)
    s.set_bvh(nodes, refs)
    b = Renderer().Initialize(w, h, s, max_depth=5, rng_state=rng)
    b.RenderFrame(cam, 4)
    img_b = b.DownloadAccum()
    b.CleanUp()
    assert np.array_equal(img_a.view(np.uint32), img_b.view(np.uint32))
    want = np.zeros((h, w, 4), np.float32)
    oracle.Oracle(s).render(cam.as_array(), w, h, rng.copy(), want, n_frames=4, max_depth=5)
    close = np.isclose(img_a[..., :3], want[..., :3], rtol=1e-4, atol=1e-6).all(axis=2)
    assert close.mean() > 0.995
