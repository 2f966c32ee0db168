"""Generators of .tga / .bmp test files (Pillow + hand-packed headers) shared by tests/test_images.py."""
import struct
import numpy as np


def pillow_cases(rgb):
    from PIL import Image
    im = Image.fromarray(rgb)
    a = (rgb[..., 0] // 2 + 64).astype(np.uint8)
    rgba = Image.fromarray(np.dstack([rgb, a]))
    return {
        "rgb.tga": (im, {}), "rgb_rle.tga": (im, dict(compression="tga_rle")),
        "rgba.tga": (rgba, {}), "rgba_rle.tga": (rgba, dict(compression="tga_rle")),
        "grey.tga": (im.convert("L"), {}), "grey_rle.tga": (im.convert("L"), dict(compression="tga_rle")),
        "palette.tga": (im.convert("P"), {}), "palette_rle.tga": (im.convert("P"), dict(compression="tga_rle")),
        "grey_alpha.tga": (rgba.convert("LA"), {}), "rgb_bottom_up.tga": (im, dict(orientation=-1)),
        "rgb.bmp": (im, {}), "rgba.bmp": (rgba, {}), "grey.bmp": (im.convert("L"), {}), "palette.bmp": (im.convert("P"), {}),
        "palette16.bmp": (im.quantize(16), dict(bits=4)),
    }


def _bmp(w, h, bpp, rows, hsz=40, compress=0, masks=b"", palette=b"", top_down=False, v4_masks=None):
    """rows: list of h byte strings (top row first), already packed; padded to 4 bytes here."""
    body = b"".join(r + b"\0" * ((-len(r)) & 3) for r in (rows if top_down else rows[::-1]))
    if hsz == 12:
        info = struct.pack("<IHHHH", 12, w, h, 1, bpp)
    else:
        info = struct.pack("<IiiHHIIiiII", hsz, w, -h if top_down else h, 1, bpp, compress, len(body), 2835, 2835, 0, 0)
        if hsz >= 108:
            info += struct.pack("<IIII", *v4_masks) + b"\0" * (hsz - 56)
    off = 14 + len(info) + len(masks) + len(palette)
    return b"BM" + struct.pack("<IHHI", off + len(body), 0, 0, off) + info + masks + palette + body


def handmade_cases(rgb):
    h, w = rgb.shape[:2]
    r, g, b = (rgb[..., k].astype(np.uint32) for k in range(3))
    out = {}
    p555 = ((r >> 3) << 10 | (g >> 3) << 5 | (b >> 3)).astype("<u2")
    p565 = ((r >> 3) << 11 | (g >> 2) << 5 | (b >> 3)).astype("<u2")
    out["rgb555.bmp"] = _bmp(w, h, 16, [p555[y].tobytes() for y in range(h)])
    out["rgb565_bitfields.bmp"] = _bmp(w, h, 16, [p565[y].tobytes() for y in range(h)], compress=3, masks=struct.pack("<III", 0xF800, 0x07E0, 0x001F))
    bgra = np.dstack([rgb[..., ::-1], np.zeros((h, w), np.uint8)])
    out["bgra_zero_alpha.bmp"] = _bmp(w, h, 32, [bgra[y].tobytes() for y in range(h)])
    out["bgr_top_down.bmp"] = _bmp(w, h, 24, [rgb[y, :, ::-1].tobytes() for y in range(h)], top_down=True)
    argb = (np.uint32(0x80) << 24 | r << 16 | g << 8 | b).astype("<u4")
    out["v4_argb.bmp"] = _bmp(w, h, 32, [argb[y].tobytes() for y in range(h)], hsz=108, compress=3, v4_masks=(0xFF0000, 0xFF00, 0xFF, 0xFF000000))
    rgb10 = ((r * 4) << 20 | (g * 4) << 10 | (b * 4)).astype("<u4")
    out["v4_rgb10.bmp"] = _bmp(w, h, 32, [rgb10[y].tobytes() for y in range(h)], hsz=108, compress=3, v4_masks=(0x3FF00000, 0xFFC00, 0x3FF, 0))
    idx = ((r >> 6) << 4 | (g >> 6) << 2 | (b >> 6)).astype(np.uint8)
    # stb sizes an OS/2 palette as (offset - 14 - 24) / 3 -- 24, not the 12-byte header -- so 4 spare entries keep 0..63 inside it
    pal3 = b"".join(bytes([(i & 3) * 85, ((i >> 2) & 3) * 85, ((i >> 4) & 3) * 85]) for i in range(68))
    out["os2_palette.bmp"] = _bmp(w, h, 8, [idx[y].tobytes() for y in range(h)], hsz=12, palette=pal3)
    # TGA by hand: 16-bit RGB (raw + RLE packets of one pixel), 16-bit colour map with 8-bit indices, a first-entry index, id field
    hdr = lambda idlen, cmap, typ, ps, pl, pb, bpp, desc: struct.pack("<BBBHHBHHHHBB", idlen, cmap, typ, ps, pl, pb, 0, 0, w, h, bpp, desc)
    out["rgb16.tga"] = hdr(0, 0, 2, 0, 0, 0, 16, 0x20) + p555.tobytes()
    out["rgb15_rle.tga"] = hdr(5, 0, 10, 0, 0, 0, 15, 0x00) + b"hello" + b"".join(bytes([0x80 | 0]) + v.tobytes() for v in p555.ravel())
    pal16 = np.array([((i & 3) * 10) << 10 | (((i >> 2) & 3) * 10) << 5 | ((i >> 4) & 3) * 10 for i in range(64)], "<u2")
    out["cmap16.tga"] = hdr(0, 1, 1, 0, 64, 16, 8, 0x20) + pal16.tobytes() + idx.tobytes()
    pal24 = b"".join(bytes([((i >> 4) & 3) * 85, ((i >> 2) & 3) * 85, (i & 3) * 85]) for i in range(64))
    out["cmap24_idx16_start2.tga"] = hdr(0, 1, 9, 2, 64, 24, 16, 0x00) + b"\xAA\xBB" + pal24 + \
        b"".join(bytes([0x00]) + struct.pack("<H", int(v) + (200 if k % 7 == 0 else 0)) for k, v in enumerate(idx.ravel()))
    return out
