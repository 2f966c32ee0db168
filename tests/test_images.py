"""PNG reader (map_Kd textures, cube faces: SOIL_load_image(.., SOIL_LOAD_RGBA) = stbi_load(.., 4), Texture.cpp:146-156,:262)
and PNG writer (Renderer::SaveScreenshot, Renderer.cpp:697-705). CPU only.

The PNG files are produced here from numpy arrays by a small encoder (zlib from the standard library, every filter type,
Adam7), so the expected pixels are known without any decoder; when oracle/_ref is built the reference's own stb_image is a
second, byte-for-byte witness, and its output for the committed fixture travels as tests/golden/png/*.npy."""
import os
import struct
import zlib
import numpy as np
import pytest

import oracle
from gpupathtracer_b200 import api
from common import GOLDEN

PNG = os.path.join(GOLDEN, "png")


# ------------------------------------------------------------------------------------------------ test-side PNG encoder
def _chunk(kind, data):
    return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data) & 0xffffffff)


def _pack_rows(samples, depth):
    """samples (h, w, c) of unsigned ints at `depth` bits -> list of bytes rows."""
    h, w, c = samples.shape
    rows = []
    for y in range(h):
        flat = samples[y].reshape(-1)
        if depth == 16:
            rows.append(flat.astype(">u2").tobytes())
        elif depth == 8:
            rows.append(flat.astype(np.uint8).tobytes())
        else:
            bits = np.zeros(((flat.size * depth + 7) // 8) * 8, np.uint8)
            for k in range(depth):
                bits[np.arange(flat.size) * depth + k] = (flat >> (depth - 1 - k)) & 1
            rows.append(np.packbits(bits).tobytes())
    return rows


def _filter_rows(rows, bpp, kinds):
    out, prev = bytearray(), None
    for y, row in enumerate(rows):
        kind = kinds[y % len(kinds)]
        cur = np.frombuffer(row, np.uint8).astype(np.int32)
        up = np.frombuffer(prev, np.uint8).astype(np.int32) if prev is not None else np.zeros_like(cur)
        left = np.concatenate([np.zeros(bpp, np.int32), cur[:-bpp]]) if cur.size > bpp else np.zeros_like(cur)
        ul = np.concatenate([np.zeros(bpp, np.int32), up[:-bpp]]) if cur.size > bpp else np.zeros_like(cur)
        if kind == 0:
            pred = np.zeros_like(cur)
        elif kind == 1:
            pred = left
        elif kind == 2:
            pred = up
        elif kind == 3:
            pred = (left + up) // 2
        else:
            p = left + up - ul
            pa, pb, pc = np.abs(p - left), np.abs(p - up), np.abs(p - ul)
            pred = np.where((pa <= pb) & (pa <= pc), left, np.where(pb <= pc, up, ul))
        out.append(kind)
        out += ((cur - pred) & 255).astype(np.uint8).tobytes()
        prev = row
    return bytes(out)


def write_png(path, samples, color, depth, palette=None, trns=None, interlace=False, kinds=(0, 1, 2, 3, 4), level=6,
              idat_split=0, extra=()):
    h, w, c = samples.shape
    bpp = max(1, c * depth // 8)
    body = b""
    if interlace:
        xs, ys, dx, dy = (0, 4, 0, 2, 0, 1, 0), (0, 0, 4, 0, 2, 0, 1), (8, 8, 4, 4, 2, 2, 1), (8, 8, 8, 4, 4, 2, 2)
        for p in range(7):
            sub = samples[ys[p]::dy[p], xs[p]::dx[p]]
            if sub.shape[0] and sub.shape[1]:
                body += _filter_rows(_pack_rows(sub, depth), bpp, kinds)
    else:
        body = _filter_rows(_pack_rows(samples, depth), bpp, kinds)
    z = zlib.compress(body, level)
    out = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, color, 0, 0, 1 if interlace else 0))
    for kind, data in extra:
        out += _chunk(kind, data)
    if palette is not None:
        out += _chunk(b"PLTE", np.asarray(palette, np.uint8).tobytes())
    if trns is not None:
        out += _chunk(b"tRNS", trns)
    if idat_split:
        for i in range(0, len(z), idat_split):
            out += _chunk(b"IDAT", z[i:i + idat_split])
    else:
        out += _chunk(b"IDAT", z)
    out += _chunk(b"IEND", b"")
    with open(path, "wb") as f:
        f.write(out)


def expected_rgba8(samples, color, depth, palette=None, trns=None):
    h, w, c = samples.shape
    maxv = (1 << depth) - 1

    def to8(v):
        return (v >> 8) if depth == 16 else (v if depth == 8 else v * (255 // maxv))
    out = np.full((h, w, 4), 255, np.uint8)
    if color == 3:
        pal = np.asarray(palette, np.uint8).reshape(-1, 3)
        out[..., :3] = pal[samples[..., 0]]
        if trns is not None:
            alpha = np.full(256, 255, np.uint8); alpha[:len(trns)] = np.frombuffer(trns, np.uint8)
            out[..., 3] = alpha[samples[..., 0]]
    elif color in (0, 4):
        out[..., 0] = out[..., 1] = out[..., 2] = to8(samples[..., 0])
        if color == 4:
            out[..., 3] = to8(samples[..., 1])
        elif trns is not None:
            key = struct.unpack(">H", trns[:2])[0]
            out[..., 3] = np.where(samples[..., 0] == key, 0, 255)
    else:
        out[..., :3] = to8(samples[..., :3])
        if color == 6:
            out[..., 3] = to8(samples[..., 3])
        elif trns is not None:
            key = np.array(struct.unpack(">HHH", trns[:6]))
            out[..., 3] = np.where((samples[..., :3] == key).all(axis=2), 0, 255)
    return out


CASES = [  # (name, color, depth, channels, interlace, w, h)
    ("rgb8", 2, 8, 3, False, 37, 19), ("rgba8", 6, 8, 4, False, 16, 16), ("grey8", 0, 8, 1, False, 33, 9),
    ("ga8", 4, 8, 2, False, 12, 21), ("rgb16", 2, 16, 3, False, 11, 13), ("rgba16", 6, 16, 4, True, 23, 17),
    ("grey1", 0, 1, 1, False, 43, 7), ("grey2", 0, 2, 1, True, 21, 11), ("grey4", 0, 4, 1, False, 17, 5),
    ("grey16", 0, 16, 1, False, 9, 9), ("pal8", 3, 8, 1, False, 31, 14), ("pal4", 3, 4, 1, True, 29, 10),
    ("pal1", 3, 1, 1, False, 19, 6), ("rgb8_adam7", 2, 8, 3, True, 40, 27), ("rgb8_tiny", 2, 8, 3, True, 1, 1),
]


def _make_case(name, color, depth, ch, w, h):
    rng = np.random.default_rng(abs(hash(name)) % (2 ** 31) if False else sum(map(ord, name)))
    samples = rng.integers(0, 1 << depth, size=(h, w, ch), dtype=np.int64)
    samples[:, : w // 2] = samples[:, :1]              # runs, so that the deflate stream has matches
    palette = trns = None
    if color == 3:
        n = 1 << depth
        palette = rng.integers(0, 256, size=(n, 3), dtype=np.uint8)
        trns = bytes(rng.integers(0, 256, size=n // 2 + 1, dtype=np.uint8))
    elif color == 0 and name in ("grey8", "grey16", "grey4"):
        trns = struct.pack(">H", int(samples[0, 0, 0]))
    elif color == 2 and name in ("rgb8", "rgb16"):
        trns = struct.pack(">HHH", *[int(v) for v in samples[0, 0, :3]])
    return samples, palette, trns


@pytest.mark.parametrize("name,color,depth,ch,interlace,w,h", CASES)
def test_png_reader(tmp_path, name, color, depth, ch, interlace, w, h):
    samples, palette, trns = _make_case(name, color, depth, ch, w, h)
    p = str(tmp_path / f"{name}.png")
    write_png(p, samples, color, depth, palette=palette, trns=trns, interlace=interlace,
              idat_split=97 if name in ("rgb8", "pal8") else 0, extra=((b"gAMA", struct.pack(">I", 45455)),))
    want = expected_rgba8(samples, color, depth, palette, trns)
    got = api.load_image(p)
    assert got.shape == (h, w, 4)
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0))
    if oracle.have_stbi():
        assert np.array_equal(oracle.ref_stbi_load(p), want)


@pytest.mark.parametrize("level", [0, 1, 9])
def test_png_reader_deflate_block_kinds(tmp_path, level):
    """level 0 = stored blocks, 1 = mostly fixed/dynamic Huffman with short matches, 9 = dynamic with long matches."""
    rng = np.random.default_rng(level)
    samples = np.zeros((64, 300, 3), np.int64)
    samples[:, :, 0] = np.arange(300)[None, :] % 7 * 30
    samples[::3] = rng.integers(0, 256, size=samples[::3].shape)
    p = str(tmp_path / "t.png")
    write_png(p, samples, 2, 8, level=level, kinds=(4,))
    assert np.array_equal(api.load_image(p)[..., :3], samples.astype(np.float32) / np.float32(255.0))


def test_png_reader_refuses_damaged_files(tmp_path):
    samples = np.random.default_rng(1).integers(0, 256, size=(8, 8, 3), dtype=np.int64)
    p = str(tmp_path / "ok.png")
    write_png(p, samples, 2, 8)
    data = open(p, "rb").read()
    for name, blob in [("short.png", data[:len(data) // 2]), ("nosig.png", b"\x89PNX" + data[4:]),
                       ("noidat.png", data[:33] + _chunk(b"IEND", b""))]:
        q = tmp_path / name
        q.write_bytes(blob)
        with pytest.raises(api.PtError):
            api.load_image(str(q))


def test_png_writer_round_trip(tmp_path):
    rng = np.random.default_rng(2)
    for (w, h) in [(1, 1), (5, 3), (300, 200), (1280, 18)]:       # 300x200x3+200 > 65535: several stored blocks
        img = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
        p = str(tmp_path / f"w{w}.png")
        api.save_png(p, img)
        data = open(p, "rb").read()
        # independent decode: walk the chunks, check every CRC, inflate with zlib, expect filter byte 0 on every row
        assert data[:8] == b"\x89PNG\r\n\x1a\n"
        pos, idat, kinds = 8, b"", []
        while pos < len(data):
            n, kind = struct.unpack(">I4s", data[pos:pos + 8])
            body = data[pos + 8:pos + 8 + n]
            assert struct.unpack(">I", data[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(kind + body) & 0xffffffff
            kinds.append(kind)
            if kind == b"IHDR":
                assert struct.unpack(">IIBBBBB", body) == (w, h, 8, 2, 0, 0, 0)
            if kind == b"IDAT":
                idat += body
            pos += 12 + n
        assert kinds[0] == b"IHDR" and kinds[-1] == b"IEND"
        raw = np.frombuffer(zlib.decompress(idat), np.uint8).reshape(h, 3 * w + 1)
        assert np.all(raw[:, 0] == 0) and np.array_equal(raw[:, 1:].reshape(h, w, 3), img)
        assert np.array_equal(api.load_image(p)[..., :3], img.astype(np.float32) / np.float32(255.0))
        if oracle.have_stbi():
            assert np.array_equal(oracle.ref_stbi_load(p)[..., :3], img)


def test_png_committed_fixture():
    """tests/golden/png/tiles.png (palette, Adam7) with the reference's stb_image result recorded beside it."""
    got = api.load_image(os.path.join(PNG, "tiles.png"))
    want = np.load(os.path.join(PNG, "tiles_stbi.npy"))
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0))


# ------------------------------------------------------------------------------------------------ JPEG
JPG = os.path.join(GOLDEN, "jpg")


def _test_image(w, h, seed):
    """Smooth colour gradients + hard edges + noise: exercises DC prediction, long zero runs and busy blocks."""
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.zeros((h, w, 3), np.float64)
    img[..., 0] = 128 + 100 * np.sin(x / 7.0) * np.cos(y / 5.0)
    img[..., 1] = (x * 255.0 / max(w - 1, 1))
    img[..., 2] = (y * 255.0 / max(h - 1, 1))
    img[(x // 6 + y // 4) % 5 == 0] = (250, 10, 30)
    img += rng.normal(0, 12, size=img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


def _pil():
    return pytest.importorskip("PIL.Image")


JPEG_CASES = [  # (name, size, mode, PIL save options)
    ("q75_420", (64, 48), "RGB", dict(quality=75, subsampling=2)),
    ("q95_444", (33, 17), "RGB", dict(quality=95, subsampling=0)),
    ("q30_422", (57, 23), "RGB", dict(quality=30, subsampling=1)),
    ("q85_420_odd", (17, 33), "RGB", dict(quality=85, subsampling=2)),
    ("tiny_420", (1, 1), "RGB", dict(quality=90, subsampling=2)),
    ("narrow_420", (2, 40), "RGB", dict(quality=90, subsampling=2)),
    ("optimised_tables", (80, 60), "RGB", dict(quality=60, subsampling=2, optimize=True)),
    ("grey", (45, 31), "L", dict(quality=80)),
    ("cmyk", (24, 24), "CMYK", dict(quality=90)),
    ("restart_rows", (96, 64), "RGB", dict(quality=75, subsampling=2, restart_marker_rows=1)),
    ("restart_blocks", (96, 64), "RGB", dict(quality=75, subsampling=0, restart_marker_blocks=3)),
    ("q100_444", (40, 40), "RGB", dict(quality=100, subsampling=0)),
    # progressive (SOF2): libjpeg's default script = interleaved DC, per-component AC bands, successive approximation
    ("prog_420", (64, 48), "RGB", dict(quality=75, subsampling=2, progressive=True)),
    ("prog_444_q95", (33, 17), "RGB", dict(quality=95, subsampling=0, progressive=True)),
    ("prog_422_q30", (57, 23), "RGB", dict(quality=30, subsampling=1, progressive=True)),
    ("prog_420_odd", (17, 33), "RGB", dict(quality=85, subsampling=2, progressive=True)),
    ("prog_tiny", (1, 1), "RGB", dict(quality=90, subsampling=2, progressive=True)),
    ("prog_grey", (45, 31), "L", dict(quality=80, progressive=True)),
    ("prog_cmyk", (24, 24), "CMYK", dict(quality=90, progressive=True)),
    ("prog_restart", (96, 64), "RGB", dict(quality=75, subsampling=2, progressive=True, restart_marker_rows=1)),
    ("prog_restart_blocks", (96, 64), "RGB", dict(quality=60, subsampling=0, progressive=True, restart_marker_blocks=3)),
    ("prog_q100", (40, 40), "RGB", dict(quality=100, subsampling=0, progressive=True, optimize=True)),
    ("prog_big", (200, 150), "RGB", dict(quality=50, subsampling=2, progressive=True)),
]


@pytest.mark.parametrize("name,size,mode,opts", JPEG_CASES)
def test_jpeg_reader_matches_stb(tmp_path, name, size, mode, opts):
    Image = _pil()
    if not oracle.have_stbi():
        pytest.skip("oracle/_ref not built: covered by test_jpeg_committed_fixtures")
    w, h = size
    rgb = _test_image(w, h, seed=len(name))
    im = Image.fromarray(rgb).convert(mode)
    p = str(tmp_path / f"{name}.jpg")
    try:
        im.save(p, "JPEG", **opts)
    except TypeError:
        pytest.skip("this Pillow does not know " + ", ".join(opts))
    want = oracle.ref_stbi_load(p)
    assert want is not None
    got = api.load_image(p)
    assert got.shape == (h, w, 4)
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0))
    if mode == "RGB" and opts.get("quality", 0) >= 95:
        assert np.abs(want[..., :3].astype(int) - rgb.astype(int)).mean() < 6      # and it is the picture we encoded


@pytest.mark.parametrize("factor", [411, 440, 422, 420, 444])
def test_jpeg_reader_other_sampling_factors(tmp_path, factor):
    cv2 = pytest.importorskip("cv2")
    if not oracle.have_stbi() or not hasattr(cv2, "IMWRITE_JPEG_SAMPLING_FACTOR"):
        pytest.skip("needs oracle/_ref and an OpenCV that can choose the sampling factor")
    rgb = _test_image(70, 50, seed=factor)
    p = str(tmp_path / f"s{factor}.jpg")
    flag = getattr(cv2, f"IMWRITE_JPEG_SAMPLING_FACTOR_{factor}")
    assert cv2.imwrite(p, rgb[..., ::-1], [cv2.IMWRITE_JPEG_QUALITY, 88, cv2.IMWRITE_JPEG_SAMPLING_FACTOR, flag])
    want = oracle.ref_stbi_load(p)
    got = api.load_image(p)
    assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0))


def test_jpeg_reader_refuses_damaged(tmp_path):
    Image = _pil()
    rgb = _test_image(32, 32, seed=1)
    q = str(tmp_path / "base.jpg")
    Image.fromarray(rgb).save(q, "JPEG", quality=80)
    data = open(q, "rb").read()
    cut = tmp_path / "cut.jpg"
    cut.write_bytes(data[:len(data) // 3])
    with pytest.raises(api.PtError):
        api.load_image(str(cut))
    # a progressive file cut inside its scans: stb (and this reader) decode what is there; no crash, same picture
    p = str(tmp_path / "prog.jpg")
    Image.fromarray(rgb).save(p, "JPEG", quality=80, progressive=True)
    data = open(p, "rb").read()
    cut2 = tmp_path / "cut_prog.jpg"
    cut2.write_bytes(data[:len(data) * 2 // 3] + b"\xff\xd9")
    if oracle.have_stbi():
        want = oracle.ref_stbi_load(str(cut2))
        if want is not None:
            assert np.array_equal(api.load_image(str(cut2)), want.astype(np.float32) / np.float32(255.0))


def test_jpeg_committed_fixtures():
    """tests/golden/jpg/*.jpg with the reference's stb_image result recorded beside each (tools/make_jpg_golden.py)."""
    names = sorted(f[:-4] for f in os.listdir(JPG) if f.endswith(".jpg"))
    assert len(names) >= 3
    for n in names:
        got = api.load_image(os.path.join(JPG, n + ".jpg"))
        want = np.load(os.path.join(JPG, n + "_stbi.npy"))
        assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0)), n


def test_tga_and_bmp_readers_match_stb(tmp_path):
    """.tga (types 1/2/3/9/10/11: grey, grey + alpha, 15/16/24/32-bit colour, colour maps, RLE, both origins) and .bmp (4/8-bit
    palettes, 16/24/32 bits, BI_BITFIELDS, V4 headers, OS/2 headers, top-down, the all-zero-alpha rule) written by Pillow or
    packed by hand in tests/ldr_cases.py decode to the bytes stbi_load(.., 4) of the reference's stb_image gives -- Crytek's
    Sponza ships its textures as .tga."""
    import ldr_cases
    if not oracle.have_stbi():
        pytest.skip("oracle/_ref not built: covered by test_tga_bmp_committed_fixtures")
    rgb = _test_image(37, 23, seed=3)
    files = {}
    try:
        for name, (img, opts) in ldr_cases.pillow_cases(rgb).items():
            p = str(tmp_path / name); img.save(p, **opts); files[name] = p
    except ImportError:
        pass
    for name, data in ldr_cases.handmade_cases(rgb).items():
        p = tmp_path / name; p.write_bytes(data); files[name] = str(p)
    # binary .ppm / .pgm: stb copies the bytes (no rescaling by the maximum value), grey goes to all three channels
    files["plain.ppm"] = str(tmp_path / "plain.ppm"); (tmp_path / "plain.ppm").write_bytes(b"P6\n# comment\n37 23\n255\n" + rgb.tobytes())
    files["max100.ppm"] = str(tmp_path / "max100.ppm"); (tmp_path / "max100.ppm").write_bytes(b"P6 37 23 100\n" + (rgb // 3).tobytes())
    files["grey.pgm"] = str(tmp_path / "grey.pgm"); (tmp_path / "grey.pgm").write_bytes(b"P5\n37\t23\n255\n" + rgb[..., 1].tobytes())
    assert len(files) >= 14
    for name, p in files.items():
        want = oracle.ref_stbi_load(p)
        assert want is not None, name
        got = api.load_image(p)
        assert got.shape == (23, 37, 4), name
        assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0)), name
    # what stb refuses is refused: 1-bit and RLE-compressed .bmp, a colour-mapped .tga of a true-colour type
    from PIL import Image
    bad = {"mono.bmp": None, "typed_wrong.tga": None}
    Image.fromarray(rgb).convert("1").save(str(tmp_path / "mono.bmp"))
    (tmp_path / "typed_wrong.tga").write_bytes(bytes([0, 1, 2, 0, 0, 4, 0, 24, 0, 0, 0, 0, 2, 0, 2, 0, 24, 0]) + bytes(24))
    for name in bad:
        assert oracle.ref_stbi_load(str(tmp_path / name)) is None, name
        with pytest.raises(api.PtError):
            api.load_image(str(tmp_path / name))


def test_tga_bmp_committed_fixtures():
    """tests/golden/ldr/*.{tga,bmp} with stb_image's result recorded beside each (tools/make_ldr_golden.py)."""
    d = os.path.join(GOLDEN, "ldr")
    names = sorted(f for f in os.listdir(d) if f.endswith((".tga", ".bmp")))
    assert len(names) >= 6
    for n in names:
        got = api.load_image(os.path.join(d, n))
        want = np.load(os.path.join(d, n + "_stbi.npy"))
        assert np.array_equal(got, want.astype(np.float32) / np.float32(255.0)), n


def test_ldr_files_through_loadf_get_stbs_gamma(tmp_path):
    """The reference's ".jpg" environment is read with stbi_loadf (Renderer.cpp:262-266): 8-bit decode, then
    pow(byte / 255, 2.2) per colour channel. pt_image_load_hdr gives the same floats, bit for bit."""
    Image = _pil()
    if not oracle.have_stbi():
        pytest.skip("oracle/_ref not built")
    rgb = _test_image(48, 24, seed=9)
    p = str(tmp_path / "env.jpg")
    Image.fromarray(rgb).save(p, "JPEG", quality=92, subsampling=0)
    got = api.load_hdr(p)
    want = oracle.ref_stbi_loadf(p)
    assert got.shape == want.shape and np.array_equal(got, want)
    assert got[..., :3].max() <= 1.0 and np.all(got[..., 3] == 1.0)


def test_damaged_headers_are_refused_not_trusted(tmp_path):
    """Image headers drive allocations and table sizes; a damaged one must come back as an error code, never as a write outside
    a buffer or an exception through the C ABI (BMP palette size from a pixel offset smaller than the headers;
    PNG / TGA / PPM dimensions; deflate output larger than the image)."""
    import struct
    import zlib
    from ldr_cases import _bmp
    rgb = _test_image(8, 8, 3)
    idx = (rgb[..., 0] >> 4).astype(np.uint8)
    pal = b"".join(bytes([i, i // 2, 255 - i, 0]) for i in range(256))
    good = _bmp(8, 8, 8, [idx[y].tobytes() for y in range(8)], palette=pal)
    files = {}
    for name, off in (("bmp_offset_0.bmp", 0), ("bmp_offset_small.bmp", 30), ("bmp_offset_huge.bmp", 0x80000010), ("bmp_offset_neg.bmp", 0xfffffff0)):
        files[name] = good[:10] + struct.pack("<I", off) + good[14:]
    os2 = _bmp(8, 8, 8, [idx[y].tobytes() for y in range(8)], hsz=12, palette=pal[:3 * 68])
    files["os2_offset_0.bmp"] = os2[:10] + struct.pack("<I", 0) + os2[14:]
    files["tga_huge.tga"] = struct.pack("<BBBHHBHHHHBB", 0, 0, 2, 0, 0, 0, 0, 0, 65535, 65535, 32, 0) + bytes(64)
    files["ppm_huge.ppm"] = b"P6\n99999999999999999999 2\n255\n" + bytes(64)
    files["ppm_big.ppm"] = b"P6\n60000 60000\n255\n" + bytes(64)

    def chunk(kind, data):
        return struct.pack(">I", len(data)) + kind + data + struct.pack(">I", zlib.crc32(kind + data) & 0xffffffff)
    sig = bytes([137, 80, 78, 71, 13, 10, 26, 10])
    files["png_huge.png"] = sig + chunk(b"IHDR", struct.pack(">IIBBBBB", 1 << 23, 1 << 23, 8, 6, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(bytes(64))) + chunk(b"IEND", b"")
    # a 4 x 4 image whose IDAT inflates to 8 MB
    files["png_bomb.png"] = sig + chunk(b"IHDR", struct.pack(">IIBBBBB", 4, 4, 8, 2, 0, 0, 0)) + chunk(b"IDAT", zlib.compress(bytes(8 << 20), 9)) + chunk(b"IEND", b"")
    for name, data in files.items():
        p = tmp_path / name
        p.write_bytes(data)
        with pytest.raises(api.PtError):
            api.load_image(str(p))
    # the undamaged file still decodes
    (tmp_path / "good.bmp").write_bytes(good)
    assert api.load_image(str(tmp_path / "good.bmp")).shape == (8, 8, 4)
