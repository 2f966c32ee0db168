// pt_image.cpp -- Radiance .hdr (RGBE) reader for the environment path; PNG and JPEG readers (map_Kd, cube faces, .jpg environment),
// PNG writer (SaveScreenshot).
//
// The reference reads its equirectangular environment with stbi_loadf(path, &w, &h, &c, 4) (Renderer.cpp:262-266; stb_image
// comes with SOIL2, extern/soil2-cmake/src/SOIL2/stb_image.h). This restates what that call yields for a Radiance file:
//   * signature "#?RADIANCE\n" or "#?RGBE\n", header lines up to the first empty one, one of them must be
//     "FORMAT=32-bit_rle_rgbe"; resolution line "-Y <h> +X <w>" only (other orientations are refused)
//   * scanlines 8 <= w < 32768 in the new run-length form (2, 2, w>>8, w&255, then the four byte planes as runs/dumps);
//     a scanline that does not start that way switches the reader to flat 4-byte pixels -- and, like stb, restarts the
//     flat read at pixel 1 of row 0 (only right when it happens on the first row; kept for identical output)
//   * pixel = byte * 2^(e - 136), alpha 1, e == 0 -> black; rows in file order (no flip); reads past the end yield 0
// oracle/ref_stbi.c compiles stb_image in place as the checker (tests/test_environment.py).
#include "pt_host.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace pt {
namespace {

struct ByteReader {
    const std::vector<unsigned char>& d; size_t at = 0;
    int get() { return at < d.size() ? d[at++] : 0; }
    bool eof() const { return at >= d.size(); }
    // one header line without the newline; over-long lines are cut at 1022 characters and skipped to their end
    std::string token() {
        std::string t;
        char c = (char)get();
        while (!eof() && c != '\n') {
            t.push_back(c);
            if (t.size() == 1023) { while (!eof() && get() != '\n') {} break; }
            c = (char)get();
        }
        return t;
    }
};

inline void rgbe_to_float(const unsigned char* in, float* out) {
    if (in[3] != 0) {
        float f = (float)std::ldexp(1.0f, (int)in[3] - (128 + 8));
        out[0] = in[0] * f; out[1] = in[1] * f; out[2] = in[2] * f;
    } else out[0] = out[1] = out[2] = 0.0f;
    out[3] = 1.0f;
}

}  // namespace

bool load_hdr_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba, std::string& err) {
    std::vector<unsigned char> buf;
    {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) { err = "cannot open " + path; return false; }
        unsigned char tmp[65536]; size_t n;
        while ((n = std::fread(tmp, 1, sizeof(tmp), f)) > 0) buf.insert(buf.end(), tmp, tmp + n);
        std::fclose(f);
    }
    const char* sig0 = "#?RADIANCE\n"; const char* sig1 = "#?RGBE\n";
    if (!((buf.size() >= 11 && !std::memcmp(buf.data(), sig0, 11)) || (buf.size() >= 7 && !std::memcmp(buf.data(), sig1, 7)))) {
        err = path + ": not a Radiance HDR file"; return false;
    }
    ByteReader r{buf};
    r.token();
    bool valid = false;
    for (;;) {
        std::string t = r.token();
        if (t.empty()) break;
        if (t == "FORMAT=32-bit_rle_rgbe") valid = true;
    }
    if (!valid) { err = path + ": unsupported HDR format"; return false; }
    std::string res = r.token();
    if (res.compare(0, 3, "-Y ") != 0) { err = path + ": unsupported HDR data layout"; return false; }
    char* end = nullptr;
    const char* p = res.c_str() + 3;
    long hh = std::strtol(p, &end, 10);
    while (*end == ' ') end++;
    if (std::strncmp(end, "+X ", 3) != 0) { err = path + ": unsupported HDR data layout"; return false; }
    long ww = std::strtol(end + 3, nullptr, 10);
    if (ww <= 0 || hh <= 0 || ww > (1 << 20) || hh > (1 << 20) || (double)ww * (double)hh > 1.0e9) { err = path + ": bad HDR size"; return false; }
    w = (int)ww; h = (int)hh;
    rgba.assign((size_t)w * h * 4, 0.0f);

    auto flat_from = [&](int j, int i) {
        for (; j < h; j++, i = 0)
            for (; i < w; i++) {
                unsigned char px[4];
                for (int k = 0; k < 4; k++) px[k] = (unsigned char)r.get();
                rgbe_to_float(px, rgba.data() + ((size_t)j * w + i) * 4);
            }
    };
    if (w < 8 || w >= 32768) { flat_from(0, 0); return true; }
    std::vector<unsigned char> line((size_t)w * 4);
    for (int j = 0; j < h; j++) {
        int c1 = r.get(), c2 = r.get(), len = r.get();
        if (c1 != 2 || c2 != 2 || (len & 0x80)) {
            unsigned char px[4] = {(unsigned char)c1, (unsigned char)c2, (unsigned char)len, (unsigned char)r.get()};
            rgbe_to_float(px, rgba.data());
            flat_from(0, 1);
            return true;
        }
        len = (len << 8) | r.get();
        if (len != w) { err = path + ": corrupt HDR (scanline length)"; return false; }
        for (int k = 0; k < 4; k++) {
            int i = 0, left;
            while ((left = w - i) > 0) {
                int count = r.get();
                if (count > 128) {
                    int value = r.get();
                    count -= 128;
                    if (count > left) { err = path + ": corrupt HDR (run)"; return false; }
                    for (int z = 0; z < count; z++) line[(size_t)(i++) * 4 + k] = (unsigned char)value;
                } else {
                    if (count > left) { err = path + ": corrupt HDR (dump)"; return false; }
                    if (count == 0 && r.eof()) { err = path + ": truncated HDR"; return false; }
                    for (int z = 0; z < count; z++) line[(size_t)(i++) * 4 + k] = (unsigned char)r.get();
                }
            }
        }
        for (int i = 0; i < w; i++) rgbe_to_float(line.data() + (size_t)i * 4, rgba.data() + ((size_t)j * w + i) * 4);
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ PNG
// map_Kd textures and cube faces reach the reference as SOIL_load_image(path, .., SOIL_LOAD_RGBA) = stbi_load(.., 4)
// (Texture.cpp:146-156,:262). For a PNG that means: all five colour types, bit depths 1-16 (16-bit samples keep their
// high byte), palette + tRNS alpha, a tRNS colour key for grey / RGB (compared before the 16 -> 8 reduction), Adam7
// interlace, gAMA and every other ancillary chunk ignored, grey replicated to RGB, missing alpha = 255. CRCs are not
// verified (stb does not either). SaveScreenshot writes a PNG (SOIL_save_screenshot, Renderer.cpp:697-705): the
// writer below emits 8-bit RGB, filter 0, stored deflate blocks -- valid for every reader, no compressor needed.
namespace {

struct BitReader {
    const unsigned char* p; size_t n, pos = 0; uint32_t buf = 0; int cnt = 0; bool bad = false;
    int bits(int need) {
        while (cnt < need) {
            if (pos >= n) { bad = true; return 0; }
            buf |= (uint32_t)p[pos++] << cnt; cnt += 8;
        }
        int v = (int)(buf & ((1u << need) - 1u));
        buf >>= need; cnt -= need;
        return v;
    }
};
// canonical Huffman code described by the number of codes per length and the symbols in code order
struct Huffman {
    uint16_t count[16]; uint16_t symbol[288];
    void build(const unsigned char* len, int n) {
        std::memset(count, 0, sizeof(count));
        for (int i = 0; i < n; i++) count[len[i]]++;
        uint16_t offs[16]; offs[1] = 0;
        for (int l = 1; l < 15; l++) offs[l + 1] = (uint16_t)(offs[l] + count[l]);
        for (int i = 0; i < n; i++) if (len[i]) symbol[offs[len[i]]++] = (uint16_t)i;
        count[0] = 0;
    }
    int decode(BitReader& br) const {
        int code = 0, first = 0, index = 0;
        for (int l = 1; l <= 15; l++) {
            code |= br.bits(1);
            if (br.bad) return -1;
            int c = count[l];
            if (code - c < first) return symbol[index + (code - first)];
            index += c; first += c; first <<= 1; code <<= 1;
        }
        return -1;
    }
};

// max_out: the decoder stops with an error once more than this has been produced (a PNG knows its raw size: no zip bombs)
bool inflate_zlib(const unsigned char* data, size_t n, std::vector<unsigned char>& out, size_t max_out = (size_t)-1) {
    if (n < 2 || (data[0] & 15) != 8 || ((data[0] << 8) | data[1]) % 31 != 0 || (data[1] & 32)) return false;
    BitReader br{data + 2, n - 2};
    static const uint16_t lbase[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    static const uint16_t lext[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
    static const uint16_t dbase[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
    static const uint16_t dext[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};
    for (;;) {
        int last = br.bits(1), type = br.bits(2);
        if (br.bad) return false;
        if (type == 0) {
            br.buf = 0; br.cnt = 0;                                   // to the byte boundary
            if (br.pos + 4 > br.n) return false;
            unsigned len = br.p[br.pos] | (br.p[br.pos + 1] << 8), nlen = br.p[br.pos + 2] | (br.p[br.pos + 3] << 8);
            br.pos += 4;
            if ((len ^ 0xffffu) != nlen || br.pos + len > br.n) return false;
            if (out.size() + len > max_out) return false;
            out.insert(out.end(), br.p + br.pos, br.p + br.pos + len);
            br.pos += len;
        } else if (type == 1 || type == 2) {
            Huffman lit, dist;
            unsigned char lens[320];
            if (type == 1) {
                int i = 0;
                for (; i < 144; i++) lens[i] = 8;
                for (; i < 256; i++) lens[i] = 9;
                for (; i < 280; i++) lens[i] = 7;
                for (; i < 288; i++) lens[i] = 8;
                lit.build(lens, 288);
                for (i = 0; i < 30; i++) lens[i] = 5;
                dist.build(lens, 30);
            } else {
                int nlen = br.bits(5) + 257, ndist = br.bits(5) + 1, ncode = br.bits(4) + 4;
                if (br.bad || nlen > 286 || ndist > 30) return false;
                static const unsigned char order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
                unsigned char cl[19]; std::memset(cl, 0, sizeof(cl));
                for (int i = 0; i < ncode; i++) cl[order[i]] = (unsigned char)br.bits(3);
                Huffman lencode; lencode.build(cl, 19);
                int idx = 0;
                while (idx < nlen + ndist) {
                    int sym = lencode.decode(br);
                    if (sym < 0) return false;
                    if (sym < 16) lens[idx++] = (unsigned char)sym;
                    else {
                        int prev = 0, rep;
                        if (sym == 16) { if (idx == 0) return false; prev = lens[idx - 1]; rep = 3 + br.bits(2); }
                        else if (sym == 17) rep = 3 + br.bits(3);
                        else rep = 11 + br.bits(7);
                        if (br.bad || idx + rep > nlen + ndist) return false;
                        while (rep--) lens[idx++] = (unsigned char)prev;
                    }
                }
                if (lens[256] == 0) return false;
                lit.build(lens, nlen);
                dist.build(lens + nlen, ndist);
            }
            for (;;) {
                int sym = lit.decode(br);
                if (sym < 0) return false;
                if (sym < 256) { if (out.size() >= max_out) return false; out.push_back((unsigned char)sym); }
                else if (sym == 256) break;
                else {
                    sym -= 257;
                    if (sym >= 29) return false;
                    int len = lbase[sym] + br.bits(lext[sym]);
                    int ds = dist.decode(br);
                    if (ds < 0 || ds >= 30) return false;
                    size_t d = dbase[ds] + (size_t)br.bits(dext[ds]);
                    if (br.bad || d > out.size() || out.size() + (size_t)len > max_out) return false;
                    size_t from = out.size() - d;
                    for (int k = 0; k < len; k++) out.push_back(out[from + k]);
                }
            }
        } else return false;
        if (last) break;
    }
    return true;
}

inline uint32_t be32(const unsigned char* p) { return ((uint32_t)p[0] << 24) | ((uint32_t)p[1] << 16) | ((uint32_t)p[2] << 8) | p[3]; }

// undo the scanline filters of one (sub-)image: `raw` = h rows of (1 + stride) bytes; bpp = bytes per complete pixel
bool unfilter(unsigned char* raw, size_t stride, int h, int bpp) {
    for (int y = 0; y < h; y++) {
        unsigned char* cur = raw + (size_t)y * (stride + 1);
        const unsigned char* up = y ? cur - (stride + 1) + 1 : nullptr;
        int f = cur[0];
        unsigned char* x = cur + 1;
        if (f > 4) return false;
        for (size_t i = 0; i < stride; i++) {
            int a = i >= (size_t)bpp ? x[i - bpp] : 0, b = up ? up[i] : 0, c = (up && i >= (size_t)bpp) ? up[i - bpp] : 0;
            int pred = 0;
            if (f == 1) pred = a;
            else if (f == 2) pred = b;
            else if (f == 3) pred = (a + b) >> 1;
            else if (f == 4) { int pp = a + b - c, pa = std::abs(pp - a), pb = std::abs(pp - b), pc = std::abs(pp - c); pred = (pa <= pb && pa <= pc) ? a : (pb <= pc ? b : c); }
            x[i] = (unsigned char)(x[i] + pred);
        }
    }
    return true;
}

}  // namespace

bool decode_png_rgba8(const std::vector<unsigned char>& b, int& w, int& h, std::vector<unsigned char>& rgba) {
    static const unsigned char sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
    if (b.size() < 8 + 25 || std::memcmp(b.data(), sig, 8) != 0) return false;
    size_t pos = 8;
    int depth = 0, color = 0, interlace = 0;
    std::vector<unsigned char> idat, plte;
    unsigned char pal_alpha[256]; std::memset(pal_alpha, 255, sizeof(pal_alpha));
    bool has_key = false; unsigned key[3] = {0, 0, 0};
    bool first = true, done = false;
    while (!done && pos + 8 <= b.size()) {
        uint32_t len = be32(&b[pos]); const unsigned char* type = &b[pos + 4]; const unsigned char* d = &b[pos + 8];
        if (pos + 12 + (size_t)len > b.size()) return false;
        auto is = [&](const char* t) { return std::memcmp(type, t, 4) == 0; };
        if (first != is("IHDR")) return false;
        if (is("IHDR")) {
            if (len != 13) return false;
            w = (int)be32(d); h = (int)be32(d + 4); depth = d[8]; color = d[9]; interlace = d[12];
            if (w <= 0 || h <= 0 || w > (1 << 24) || h > (1 << 24) || (double)w * h > 1.0e9 || d[10] || d[11] || interlace > 1) return false;
            if (depth != 1 && depth != 2 && depth != 4 && depth != 8 && depth != 16) return false;
            if (color != 0 && color != 2 && color != 3 && color != 4 && color != 6) return false;
            if (color == 3 && depth == 16) return false;
            if ((color == 2 || color == 4 || color == 6) && depth < 8) return false;
            first = false;
        } else if (is("PLTE")) { if (len > 768 || len % 3) return false; plte.assign(d, d + len); }
        else if (is("tRNS")) {
            if (color == 3) { if (len > 256) return false; std::memcpy(pal_alpha, d, len); }
            else if (color == 0 && len >= 2) { has_key = true; key[0] = key[1] = key[2] = ((unsigned)d[0] << 8) | d[1]; if (depth <= 8) key[0] &= 255; }
            else if (color == 2 && len >= 6) { has_key = true; for (int k = 0; k < 3; k++) { key[k] = ((unsigned)d[2 * k] << 8) | d[2 * k + 1]; if (depth <= 8) key[k] &= 255; } }
        } else if (is("IDAT")) idat.insert(idat.end(), d, d + len);
        else if (is("IEND")) done = true;
        pos += 12 + (size_t)len;
    }
    if (first || idat.empty() || (color == 3 && plte.empty())) return false;
    const int channels = color == 0 ? 1 : color == 2 ? 3 : color == 3 ? 1 : color == 4 ? 2 : 4;
    const int bits_pp = channels * depth, bpp = std::max(1, bits_pp / 8);
    std::vector<unsigned char> raw;
    // filter byte + packed row for every row; the seven Adam7 passes together add at most 7 more bytes per row and 8 short rows
    const size_t raw_max = ((size_t)w * bits_pp / 8 + 9) * ((size_t)h + 8);
    raw.reserve(std::min(raw_max, idat.size() * 64 + 4096));
    if (!inflate_zlib(idat.data(), idat.size(), raw, raw_max)) return false;
    rgba.assign((size_t)w * h * 4, 255);
    const unsigned maxv = depth == 16 ? 65535u : (1u << depth) - 1u;
    auto emit = [&](int x, int y, const unsigned* s) {           // s: channel samples at the file's bit depth
        unsigned char* o = &rgba[((size_t)y * w + x) * 4];
        auto to8 = [&](unsigned v) -> unsigned char { return depth == 16 ? (unsigned char)(v >> 8) : depth == 8 ? (unsigned char)v : (unsigned char)(v * (255u / maxv)); };
        if (color == 3) {
            size_t i = s[0];
            if (3 * i + 2 < plte.size()) { o[0] = plte[3 * i]; o[1] = plte[3 * i + 1]; o[2] = plte[3 * i + 2]; o[3] = pal_alpha[i]; }
            else { o[0] = o[1] = o[2] = 0; o[3] = 255; }
        } else if (color == 0 || color == 4) {
            o[0] = o[1] = o[2] = to8(s[0]);
            o[3] = color == 4 ? to8(s[1]) : (has_key && s[0] == key[0] ? 0 : 255);
        } else {
            o[0] = to8(s[0]); o[1] = to8(s[1]); o[2] = to8(s[2]);
            o[3] = color == 6 ? to8(s[3]) : (has_key && s[0] == key[0] && s[1] == key[1] && s[2] == key[2] ? 0 : 255);
        }
    };
    static const int xs[7] = {0, 4, 0, 2, 0, 1, 0}, ys[7] = {0, 0, 4, 0, 2, 0, 1}, dx[7] = {8, 8, 4, 4, 2, 2, 1}, dy[7] = {8, 8, 8, 4, 4, 2, 2};
    size_t at = 0;
    for (int pass = 0; pass < (interlace ? 7 : 1); pass++) {
        int pw = interlace ? (w - xs[pass] + dx[pass] - 1) / dx[pass] : w, ph = interlace ? (h - ys[pass] + dy[pass] - 1) / dy[pass] : h;
        if (pw <= 0 || ph <= 0) continue;
        size_t stride = ((size_t)pw * bits_pp + 7) / 8;
        if (at + (stride + 1) * ph > raw.size()) return false;
        if (!unfilter(&raw[at], stride, ph, bpp)) return false;
        for (int y = 0; y < ph; y++) {
            const unsigned char* row = &raw[at + (size_t)y * (stride + 1) + 1];
            for (int x = 0; x < pw; x++) {
                unsigned smp[4] = {0, 0, 0, 0};
                for (int c = 0; c < channels; c++) {
                    size_t k = (size_t)x * channels + c;
                    if (depth == 16) smp[c] = ((unsigned)row[2 * k] << 8) | row[2 * k + 1];
                    else if (depth == 8) smp[c] = row[k];
                    else { size_t bit = k * depth; smp[c] = (row[bit >> 3] >> (8 - depth - (bit & 7))) & maxv; }
                }
                emit(interlace ? xs[pass] + x * dx[pass] : x, interlace ? ys[pass] + y * dy[pass] : y, smp);
            }
        }
        at += (stride + 1) * ph;
    }
    return true;
}

// ------------------------------------------------------------------------------------------------ JPEG
// Baseline / extended-sequential / progressive Huffman JPEG (SOF0, SOF1, SOF2; 8 bit; 1, 3 or 4 components; any sampling factors; restart
// intervals) decoded to RGBA8 exactly as stbi_load(.., 4) of the reference's stb_image v2.15 does, which pins four things
// the JPEG standard leaves open:
//   * the inverse DCT: the integer "islow" scheme with 12-bit constants computed as (int)(c * 4096 + 0.5) -- truncation,
//     so negative constants round towards zero --, column pass kept at 2 extra bits (+512 >> 10, or dc << 2 when the
//     column's AC terms are all zero), row pass + 65536 + (128 << 17) >> 17, clamped to 0..255
//   * chroma upsampling: 2x horizontally / vertically / both with the (3 * near + far) triangle filter and stb's rounding
//     (+2 >> 2 once, +8 >> 4 for the 2x2 case computed from the vertically filtered row), edge samples replicated; any
//     other factor by sample replication; the far row of the last image row is the last row that carries image data
//   * YCbCr -> RGB in 20-bit fixed point with constants ((int)(c * 4096.0f + 0.5f)) << 8 and the green Cb term masked to
//     its upper 16 bits; components named 'R','G','B', or an Adobe segment with transform 0 in a file without JFIF
//     header, are taken as RGB; four components as CMYK / YCCK per the Adobe transform, multiplied out with
//     (x * k + 128 + ((x * k + 128) >> 8)) >> 8
//   * the decode covers whole MCUs; the picture is the top-left width x height of the padded planes
// Progressive files (SOF2) go through the same back end: spectral selection and successive approximation scans (DC
// first / refinement, AC first / refinement with EOB runs) accumulate un-dequantised coefficients, transformed at EOI.
namespace {

struct JpegHuff {
    unsigned char size[257]; unsigned short code[256]; unsigned char value[256];
    int maxcode[18], delta[17]; int n = 0;
    bool build(const int* counts) {                               // ITU T.81 Annex C
        int k = 0;
        for (int i = 0; i < 16; i++) for (int j = 0; j < counts[i]; j++) { if (k >= 256) return false; size[k++] = (unsigned char)(i + 1); }
        size[k] = 0; n = k;
        int c = 0; k = 0;
        for (int l = 1; l <= 16; l++) {
            delta[l] = k - c;
            if (size[k] == l) {
                while (size[k] == l) code[k++] = (unsigned short)c++;
                if (c - 1 >= (1 << l)) return false;
            }
            maxcode[l] = c << (16 - l);
            c <<= 1;
        }
        maxcode[17] = 0x7fffffff;
        return true;
    }
};

struct JpegBits {
    const unsigned char* p; size_t n, pos = 0;
    unsigned buf = 0; int cnt = 0; int marker = 0xff; bool nomore = false;   // marker 0xff: none pending
    int byte() { return pos < n ? p[pos++] : 0; }
    bool eof() const { return pos >= n; }
    void fill() {
        do {
            int b = nomore ? 0 : byte();
            if (b == 0xff) {
                int c = byte();
                while (c == 0xff) c = byte();
                if (c != 0) { marker = c; nomore = true; return; }
            }
            buf |= (unsigned)b << (24 - cnt);
            cnt += 8;
        } while (cnt <= 24);
    }
    int decode(const JpegHuff& h) {                               // Annex F.2.2.3, 16 bits looked at at once
        if (cnt < 16) fill();
        unsigned top = buf >> 16;
        int l = 1;
        while (l <= 16 && (int)top >= h.maxcode[l]) l++;
        if (l > 16 || l > cnt) return -1;
        int idx = (int)((buf >> (32 - l)) & ((1u << l) - 1u)) + h.delta[l];
        if (idx < 0 || idx >= h.n) return -1;
        buf <<= l; cnt -= l;
        return h.value[idx];
    }
    int receive_extend(int s) {                                   // F.2.2.1 RECEIVE + EXTEND
        if (cnt < s) fill();
        int v = (int)(buf >> (32 - s));
        buf <<= s; cnt -= s;
        return v < (1 << (s - 1)) ? v - (1 << s) + 1 : v;
    }
    int get_bits(int n) {                                         // n unsigned bits (progressive EOB runs)
        if (cnt < n) fill();
        int v = (int)(buf >> (32 - n));
        buf <<= n; cnt -= n;
        return v;
    }
    bool get_bit() {
        if (cnt < 1) fill();
        bool v = (buf & 0x80000000u) != 0;
        buf <<= 1; cnt--;
        return v;
    }
    void reset() { buf = 0; cnt = 0; nomore = false; marker = 0xff; }
};

const unsigned char kDezigzag[64 + 15] = {
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21, 28,
    35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62, 63,
    63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63, 63};

inline int jfix(float c) { return (int)(c * 4096 + 0.5); }
inline unsigned char jclamp(int x) { return (unsigned char)(x < 0 ? 0 : x > 255 ? 255 : x); }

// one 8-point pass of the "islow" inverse DCT; even part in x[0..3], odd part in t[0..3]
inline void idct_1d(int s0, int s1, int s2, int s3, int s4, int s5, int s6, int s7, int x[4], int t[4]) {
    int p2 = s2, p3 = s6;
    int p1 = (p2 + p3) * jfix(0.5411961f);
    int t2 = p1 + p3 * jfix(-1.847759065f);
    int t3 = p1 + p2 * jfix(0.765366865f);
    p2 = s0; p3 = s4;
    int t0 = (p2 + p3) * 4096, t1 = (p2 - p3) * 4096;
    x[0] = t0 + t3; x[3] = t0 - t3; x[1] = t1 + t2; x[2] = t1 - t2;
    t0 = s7; t1 = s5; t2 = s3; t3 = s1;
    p3 = t0 + t2; int p4 = t1 + t3; p1 = t0 + t3; p2 = t1 + t2;
    int p5 = (p3 + p4) * jfix(1.175875602f);
    t0 = t0 * jfix(0.298631336f); t1 = t1 * jfix(2.053119869f); t2 = t2 * jfix(3.072711026f); t3 = t3 * jfix(1.501321110f);
    p1 = p5 + p1 * jfix(-0.899976223f); p2 = p5 + p2 * jfix(-2.562915447f);
    p3 = p3 * jfix(-1.961570560f); p4 = p4 * jfix(-0.390180644f);
    t[3] = t3 + p1 + p4; t[2] = t2 + p2 + p3; t[1] = t1 + p2 + p4; t[0] = t0 + p1 + p3;
}
void idct_block(unsigned char* out, int stride, const short d[64]) {
    int v[64];
    for (int i = 0; i < 8; i++) {
        if (d[i + 8] == 0 && d[i + 16] == 0 && d[i + 24] == 0 && d[i + 32] == 0 && d[i + 40] == 0 && d[i + 48] == 0 && d[i + 56] == 0) {
            int dc = d[i] * 4;
            for (int r = 0; r < 8; r++) v[i + 8 * r] = dc;
        } else {
            int x[4], t[4];
            idct_1d(d[i], d[i + 8], d[i + 16], d[i + 24], d[i + 32], d[i + 40], d[i + 48], d[i + 56], x, t);
            for (int k = 0; k < 4; k++) x[k] += 512;
            v[i] = (x[0] + t[3]) >> 10; v[i + 56] = (x[0] - t[3]) >> 10;
            v[i + 8] = (x[1] + t[2]) >> 10; v[i + 48] = (x[1] - t[2]) >> 10;
            v[i + 16] = (x[2] + t[1]) >> 10; v[i + 40] = (x[2] - t[1]) >> 10;
            v[i + 24] = (x[3] + t[0]) >> 10; v[i + 32] = (x[3] - t[0]) >> 10;
        }
    }
    for (int r = 0; r < 8; r++) {
        const int* w = v + 8 * r; unsigned char* o = out + (size_t)r * stride;
        int x[4], t[4];
        idct_1d(w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], x, t);
        for (int k = 0; k < 4; k++) x[k] += 65536 + (128 << 17);
        o[0] = jclamp((x[0] + t[3]) >> 17); o[7] = jclamp((x[0] - t[3]) >> 17);
        o[1] = jclamp((x[1] + t[2]) >> 17); o[6] = jclamp((x[1] - t[2]) >> 17);
        o[2] = jclamp((x[2] + t[1]) >> 17); o[5] = jclamp((x[2] - t[1]) >> 17);
        o[3] = jclamp((x[3] + t[0]) >> 17); o[4] = jclamp((x[3] - t[0]) >> 17);
    }
}

struct JpegComp {
    int id = 0, h = 1, v = 1, tq = 0, hd = 0, ha = 0, dc_pred = 0;
    int x = 0, y = 0, w2 = 0, h2 = 0;                             // effective size, padded plane size
    std::vector<unsigned char> data;
    std::vector<short> coeff;                                     // progressive files: 64 coefficients per block of the padded plane
};

inline unsigned char blinn8(unsigned x, unsigned y) { unsigned t = x * y + 128; return (unsigned char)((t + (t >> 8)) >> 8); }

}  // namespace

bool decode_jpeg_rgba8(const std::vector<unsigned char>& file, int& width, int& height, std::vector<unsigned char>& rgba, std::string& err) {
    JpegBits br{file.data(), file.size()};
    auto get16 = [&]() { int a = br.byte(); return (a << 8) | br.byte(); };
    auto next_marker = [&]() -> int {                             // 0xff: no marker at this position
        if (br.marker != 0xff) { int m = br.marker; br.marker = 0xff; return m; }
        int x = br.byte();
        if (x != 0xff) return 0xff;
        while (x == 0xff) x = br.byte();
        return x;
    };
    if (next_marker() != 0xd8) { err = "not a JPEG file"; return false; }
    JpegHuff hdc[4], hac[4]; bool have_dc[4] = {}, have_ac[4] = {};
    unsigned short dequant[4][64] = {};
    JpegComp comp[4]; int ncomp = 0, order[4] = {}, scan_n = 0;
    int restart = 0, adobe = -1, rgb_ids = 0; bool jfif = false, have_frame = false;
    int hmax = 1, vmax = 1, mcux = 0, mcuy = 0;
    auto table_marker = [&](int m) -> bool {
        if (m == 0xdd) { if (get16() != 4) return false; restart = get16(); return true; }
        if (m == 0xdb) {
            int L = get16() - 2;
            while (L > 0) {
                int q = br.byte(), p = q >> 4, t = q & 15;
                if ((p != 0 && p != 1) || t > 3) return false;
                for (int i = 0; i < 64; i++) dequant[t][kDezigzag[i]] = (unsigned short)(p ? get16() : br.byte());
                L -= p ? 129 : 65;
            }
            return L == 0;
        }
        if (m == 0xc4) {
            int L = get16() - 2;
            while (L > 0) {
                int q = br.byte(), tc = q >> 4, th = q & 15, counts[16], n = 0;
                if (tc > 1 || th > 3) return false;
                for (int i = 0; i < 16; i++) { counts[i] = br.byte(); n += counts[i]; }
                if (n > 256) return false;
                JpegHuff& h = tc ? hac[th] : hdc[th];
                if (!h.build(counts)) return false;
                for (int i = 0; i < n; i++) h.value[i] = (unsigned char)br.byte();
                (tc ? have_ac : have_dc)[th] = true;
                L -= 17 + n;
            }
            return L == 0;
        }
        if ((m >= 0xe0 && m <= 0xef) || m == 0xfe) {
            int L = get16();
            if (L < 2) return false;
            L -= 2;
            if (m == 0xe0 && L >= 5) {
                const char tag[5] = {'J', 'F', 'I', 'F', 0}; bool ok = true;
                for (int i = 0; i < 5; i++) if (br.byte() != (unsigned char)tag[i]) ok = false;
                L -= 5;
                if (ok) jfif = true;
            } else if (m == 0xee && L >= 12) {
                const char tag[6] = {'A', 'd', 'o', 'b', 'e', 0}; bool ok = true;
                for (int i = 0; i < 6; i++) if (br.byte() != (unsigned char)tag[i]) ok = false;
                L -= 6;
                if (ok) { br.byte(); get16(); get16(); adobe = br.byte(); L -= 6; }
            }
            br.pos = std::min(br.n, br.pos + (size_t)L);
            return true;
        }
        return false;
    };
    // ---- tables and frame header
    int m = next_marker();
    while (!(m == 0xc0 || m == 0xc1 || m == 0xc2)) {
        if (m == 0xff || !table_marker(m)) { err = "corrupt JPEG (marker before the frame header)"; return false; }
        m = next_marker();
        while (m == 0xff) { if (br.eof()) { err = "corrupt JPEG (no frame header)"; return false; } m = next_marker(); }
    }
    const bool progressive = m == 0xc2;
    {
        int Lf = get16(), prec = br.byte();
        height = get16(); width = get16(); ncomp = br.byte();
        if (Lf < 11 || prec != 8 || height <= 0 || width <= 0 || (ncomp != 1 && ncomp != 3 && ncomp != 4) || Lf != 8 + 3 * ncomp) { err = "unsupported JPEG frame header"; return false; }
        for (int i = 0; i < ncomp; i++) {
            comp[i].id = br.byte();
            if (ncomp == 3 && comp[i].id == "RGB"[i]) rgb_ids++;
            int q = br.byte();
            comp[i].h = q >> 4; comp[i].v = q & 15; comp[i].tq = br.byte();
            if (comp[i].h < 1 || comp[i].h > 4 || comp[i].v < 1 || comp[i].v > 4 || comp[i].tq > 3) { err = "corrupt JPEG (component)"; return false; }
            hmax = std::max(hmax, comp[i].h); vmax = std::max(vmax, comp[i].v);
        }
        if ((double)width * height > 1.0e9) { err = "JPEG too large"; return false; }
        mcux = (width + hmax * 8 - 1) / (hmax * 8); mcuy = (height + vmax * 8 - 1) / (vmax * 8);
        for (int i = 0; i < ncomp; i++) {
            comp[i].x = (width * comp[i].h + hmax - 1) / hmax; comp[i].y = (height * comp[i].v + vmax - 1) / vmax;
            comp[i].w2 = mcux * comp[i].h * 8; comp[i].h2 = mcuy * comp[i].v * 8;
            comp[i].data.assign((size_t)comp[i].w2 * comp[i].h2, 0);
            if (progressive) comp[i].coeff.assign((size_t)comp[i].w2 * comp[i].h2, 0);
        }
        have_frame = true;
    }
    // ---- scans
    auto decode_block = [&](JpegComp& c, short data[64]) -> bool {
        if (!have_dc[c.hd] || !have_ac[c.ha]) return false;
        const unsigned short* dq = dequant[c.tq];
        int t = br.decode(hdc[c.hd]);
        if (t < 0 || t > 15) return false;
        std::memset(data, 0, 64 * sizeof(short));
        int diff = t ? br.receive_extend(t) : 0;
        c.dc_pred += diff;
        data[0] = (short)(c.dc_pred * dq[0]);
        int k = 1;
        do {
            int rs = br.decode(hac[c.ha]);
            if (rs < 0) return false;
            int s = rs & 15, r = rs >> 4;
            if (s == 0) { if (rs != 0xf0) break; k += 16; }
            else { k += r; int zig = kDezigzag[k++]; data[zig] = (short)(br.receive_extend(s) * dq[zig]); }
        } while (k < 64);
        return true;
    };
    // Progressive scans (SOF2): coefficients accumulate over the scans in comp[].coeff, un-dequantised, and are
    // dequantised + transformed once at the end -- the arithmetic (shifts in 16-bit shorts, refinement of the sign-carrying
    // magnitude, EOB runs) as stb_image v2.15 does it.
    int spec_start = 0, spec_end = 63, succ_high = 0, succ_low = 0, eob_run = 0;
    auto prog_dc = [&](JpegComp& c, short* data) -> bool {
        if (spec_end != 0) return false;                          // "can't merge dc and ac"
        if (br.cnt < 16) br.fill();
        if (succ_high == 0) {
            if (!have_dc[c.hd]) return false;
            std::memset(data, 0, 64 * sizeof(short));
            int t = br.decode(hdc[c.hd]);
            if (t < 0 || t > 15) return false;
            int diff = t ? br.receive_extend(t) : 0;
            c.dc_pred += diff;
            data[0] = (short)(c.dc_pred * (1 << succ_low));
        } else if (br.get_bit()) data[0] = (short)(data[0] + (short)(1 << succ_low));
        return true;
    };
    auto prog_ac = [&](JpegComp& c, short* data) -> bool {
        if (spec_start == 0 || !have_ac[c.ha]) return false;
        const JpegHuff& h = hac[c.ha];
        if (succ_high == 0) {
            if (eob_run) { eob_run--; return true; }
            int k = spec_start;
            do {
                int rs = br.decode(h);
                if (rs < 0) return false;
                int s2 = rs & 15, r = rs >> 4;
                if (s2 == 0) {
                    if (r < 15) { eob_run = 1 << r; if (r) eob_run += br.get_bits(r); eob_run--; break; }
                    k += 16;
                } else {
                    k += r;
                    int zig = kDezigzag[k++];
                    data[zig] = (short)(br.receive_extend(s2) * (1 << succ_low));
                }
            } while (k <= spec_end);
        } else {
            const short bit = (short)(1 << succ_low);
            auto refine = [&](short* q) {                         // one correction bit for an already non-zero coefficient
                if (br.get_bit() && (*q & bit) == 0) *q = (short)(*q > 0 ? *q + bit : *q - bit);
            };
            if (eob_run) {
                eob_run--;
                for (int k = spec_start; k <= spec_end; k++) { short* q = &data[kDezigzag[k]]; if (*q != 0) refine(q); }
            } else {
                int k = spec_start;
                do {
                    int rs = br.decode(h);
                    if (rs < 0) return false;
                    int s2 = rs & 15, r = rs >> 4;
                    if (s2 == 0) {
                        if (r < 15) { eob_run = (1 << r) - 1; if (r) eob_run += br.get_bits(r); r = 64; }   // r = 64: to the end of the band
                    } else {
                        if (s2 != 1) return false;
                        s2 = br.get_bit() ? bit : -bit;
                    }
                    while (k <= spec_end) {
                        short* q = &data[kDezigzag[k++]];
                        if (*q != 0) refine(q);
                        else { if (r == 0) { *q = (short)s2; break; } r--; }
                    }
                } while (k <= spec_end);
            }
        }
        return true;
    };
    m = next_marker();
    while (m != 0xd9) {
        if (m == 0xda) {
            int Ls = get16(); scan_n = br.byte();
            if (scan_n < 1 || scan_n > 4 || scan_n > ncomp || Ls != 6 + 2 * scan_n) { err = "corrupt JPEG (scan header)"; return false; }
            for (int i = 0; i < scan_n; i++) {
                int id = br.byte(), q = br.byte(), which = 0;
                while (which < ncomp && comp[which].id != id) which++;
                if (which == ncomp || (q >> 4) > 3 || (q & 15) > 3) { err = "corrupt JPEG (scan component)"; return false; }
                comp[which].hd = q >> 4; comp[which].ha = q & 15; order[i] = which;
            }
            spec_start = br.byte(); spec_end = br.byte(); int a = br.byte();
            succ_high = a >> 4; succ_low = a & 15;
            if (progressive) {
                if (spec_start > 63 || spec_end > 63 || spec_start > spec_end || succ_high > 13 || succ_low > 13) { err = "corrupt JPEG (progressive scan header)"; return false; }
            } else if (spec_start != 0 || a != 0) { err = "corrupt JPEG (spectral selection in a sequential scan)"; return false; }
            eob_run = 0;
            br.reset();
            for (int i = 0; i < 4; i++) comp[i].dc_pred = 0;
            int todo = restart ? restart : 0x7fffffff;
            short data[64];
            bool stop = false;
            auto after_mcu = [&]() {                              // restart interval bookkeeping; true: leave the scan
                if (--todo > 0) return false;
                if (br.cnt < 24) br.fill();
                if (!(br.marker >= 0xd0 && br.marker <= 0xd7)) return true;
                br.reset();
                for (int i = 0; i < 4; i++) comp[i].dc_pred = 0;
                eob_run = 0;
                todo = restart ? restart : 0x7fffffff;
                return false;
            };
            if (progressive && scan_n == 1) {
                JpegComp& c = comp[order[0]];
                int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3, cw = c.w2 >> 3;
                for (int j = 0; j < bh && !stop; j++)
                    for (int i = 0; i < bw && !stop; i++) {
                        short* blk = c.coeff.data() + 64 * ((size_t)i + (size_t)j * cw);
                        if (!(spec_start == 0 ? prog_dc(c, blk) : prog_ac(c, blk))) { err = "corrupt JPEG (progressive scan data)"; return false; }
                        stop = after_mcu();
                    }
            } else if (progressive) {                             // interleaved progressive scans carry DC bits only
                for (int j = 0; j < mcuy && !stop; j++)
                    for (int i = 0; i < mcux && !stop; i++) {
                        for (int k = 0; k < scan_n; k++) {
                            JpegComp& c = comp[order[k]];
                            const int cw = c.w2 >> 3;
                            for (int y = 0; y < c.v; y++)
                                for (int x = 0; x < c.h; x++) {
                                    short* blk = c.coeff.data() + 64 * ((size_t)(i * c.h + x) + (size_t)(j * c.v + y) * cw);
                                    if (!prog_dc(c, blk)) { err = "corrupt JPEG (progressive scan data)"; return false; }
                                }
                        }
                        stop = after_mcu();
                    }
            } else if (scan_n == 1) {
                JpegComp& c = comp[order[0]];
                int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3;
                for (int j = 0; j < bh && !stop; j++)
                    for (int i = 0; i < bw && !stop; i++) {
                        if (!decode_block(c, data)) { err = "corrupt JPEG (entropy-coded data)"; return false; }
                        idct_block(c.data.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, data);
                        stop = after_mcu();
                    }
            } else {
                for (int j = 0; j < mcuy && !stop; j++)
                    for (int i = 0; i < mcux && !stop; i++) {
                        for (int k = 0; k < scan_n; k++) {
                            JpegComp& c = comp[order[k]];
                            for (int y = 0; y < c.v; y++)
                                for (int x = 0; x < c.h; x++) {
                                    if (!decode_block(c, data)) { err = "corrupt JPEG (entropy-coded data)"; return false; }
                                    idct_block(c.data.data() + (size_t)c.w2 * ((j * c.v + y) * 8) + (i * c.h + x) * 8, c.w2, data);
                                }
                        }
                        stop = after_mcu();
                    }
            }
            if (br.marker == 0xff) {                              // padding after the scan: look for the next 0xff xx
                while (!br.eof()) { if (br.byte() == 255) { br.marker = br.byte(); break; } }
            }
        } else if (m == 0xdc) { get16(); get16(); }
        else if (m == 0xff || !table_marker(m)) { err = "corrupt JPEG (marker)"; return false; }
        m = next_marker();
    }
    if (!have_frame) return false;
    if (progressive) {                                            // stbi__jpeg_finish: dequantise + inverse DCT of the blocks that carry image data
        for (int n = 0; n < ncomp; n++) {
            JpegComp& c = comp[n];
            const unsigned short* dq = dequant[c.tq];
            int bw = (c.x + 7) >> 3, bh = (c.y + 7) >> 3, cw = c.w2 >> 3;
            for (int j = 0; j < bh; j++)
                for (int i = 0; i < bw; i++) {
                    short* blk = c.coeff.data() + 64 * ((size_t)i + (size_t)j * cw);
                    for (int k = 0; k < 64; k++) blk[k] = (short)(blk[k] * dq[k]);
                    idct_block(c.data.data() + (size_t)c.w2 * j * 8 + i * 8, c.w2, blk);
                }
        }
    }
    // ---- upsample + colour conversion, row by row
    const bool is_rgb = ncomp == 3 && (rgb_ids == 3 || (adobe == 0 && !jfif));
    rgba.assign((size_t)width * height * 4, 255);
    struct Up { int hs, vs, ystep, ypos, w_lores; const unsigned char *line0, *line1; std::vector<unsigned char> buf; };
    Up up[4];
    for (int k = 0; k < ncomp; k++) {
        up[k].hs = hmax / comp[k].h; up[k].vs = vmax / comp[k].v;
        up[k].ystep = up[k].vs >> 1; up[k].ypos = 0;
        up[k].w_lores = (width + up[k].hs - 1) / up[k].hs;
        up[k].line0 = up[k].line1 = comp[k].data.data();
        up[k].buf.assign((size_t)up[k].w_lores * std::max(up[k].hs, 1) + 8, 0);
    }
    const int c_r = ((int)(1.40200f * 4096.0f + 0.5f)) << 8, c_g1 = ((int)(0.71414f * 4096.0f + 0.5f)) << 8;
    const int c_g2 = ((int)(0.34414f * 4096.0f + 0.5f)) << 8, c_b = ((int)(1.77200f * 4096.0f + 0.5f)) << 8;
    for (int j = 0; j < height; j++) {
        const unsigned char* row[4] = {nullptr, nullptr, nullptr, nullptr};
        for (int k = 0; k < ncomp; k++) {
            Up& u = up[k];
            const bool bottom = u.ystep >= (u.vs >> 1);
            const unsigned char* near = bottom ? u.line1 : u.line0; const unsigned char* far = bottom ? u.line0 : u.line1;
            unsigned char* o = u.buf.data();
            const int w = u.w_lores;
            if (u.hs == 1 && u.vs == 1) row[k] = near;
            else if (u.hs == 1 && u.vs == 2) { for (int i = 0; i < w; i++) o[i] = (unsigned char)((3 * near[i] + far[i] + 2) >> 2); row[k] = o; }
            else if (u.hs == 2 && u.vs == 1) {
                if (w == 1) o[0] = o[1] = near[0];
                else {
                    o[0] = near[0]; o[1] = (unsigned char)((near[0] * 3 + near[1] + 2) >> 2);
                    int i = 1;
                    for (; i < w - 1; i++) { int n3 = 3 * near[i] + 2; o[2 * i] = (unsigned char)((n3 + near[i - 1]) >> 2); o[2 * i + 1] = (unsigned char)((n3 + near[i + 1]) >> 2); }
                    o[2 * i] = (unsigned char)((near[w - 2] * 3 + near[w - 1] + 2) >> 2); o[2 * i + 1] = near[w - 1];
                }
                row[k] = o;
            } else if (u.hs == 2 && u.vs == 2) {
                if (w == 1) o[0] = o[1] = (unsigned char)((3 * near[0] + far[0] + 2) >> 2);
                else {
                    int t1 = 3 * near[0] + far[0];
                    o[0] = (unsigned char)((t1 + 2) >> 2);
                    for (int i = 1; i < w; i++) {
                        int t0 = t1; t1 = 3 * near[i] + far[i];
                        o[2 * i - 1] = (unsigned char)((3 * t0 + t1 + 8) >> 4); o[2 * i] = (unsigned char)((3 * t1 + t0 + 8) >> 4);
                    }
                    o[2 * w - 1] = (unsigned char)((t1 + 2) >> 2);
                }
                row[k] = o;
            } else { for (int i = 0; i < w; i++) for (int q = 0; q < u.hs; q++) o[i * u.hs + q] = near[i]; row[k] = o; }
            if (++u.ystep >= u.vs) {
                u.ystep = 0; u.line0 = u.line1;
                if (++u.ypos < comp[k].y) u.line1 += comp[k].w2;
            }
        }
        unsigned char* out = &rgba[(size_t)j * width * 4];
        auto ycc = [&](int i, unsigned char* o) {
            int yf = (row[0][i] << 20) + (1 << 19), cr = row[2][i] - 128, cb = row[1][i] - 128;
            int r = yf + cr * c_r, g = yf + cr * -c_g1 + (int)((unsigned)(cb * -c_g2) & 0xffff0000u), b = yf + cb * c_b;
            o[0] = jclamp(r >> 20); o[1] = jclamp(g >> 20); o[2] = jclamp(b >> 20);
        };
        for (int i = 0; i < width; i++) {
            unsigned char* o = out + 4 * i;
            if (ncomp == 1) o[0] = o[1] = o[2] = row[0][i];
            else if (ncomp == 3) { if (is_rgb) { o[0] = row[0][i]; o[1] = row[1][i]; o[2] = row[2][i]; } else ycc(i, o); }
            else if (adobe == 0) { unsigned k = row[3][i]; o[0] = blinn8(row[0][i], k); o[1] = blinn8(row[1][i], k); o[2] = blinn8(row[2][i], k); }
            else if (adobe == 2) { unsigned k = row[3][i]; ycc(i, o); o[0] = blinn8(255u - o[0], k); o[1] = blinn8(255u - o[1], k); o[2] = blinn8(255u - o[2], k); }
            else ycc(i, o);
            o[3] = 255;
        }
    }
    return true;
}

bool write_png_rgb8(const char* path, int w, int h, const unsigned char* rgb_top_down) {
    static uint32_t crc_table[256]; static bool have_table = false;
    if (!have_table) {
        for (uint32_t i = 0; i < 256; i++) { uint32_t c = i; for (int k = 0; k < 8; k++) c = (c & 1) ? 0xedb88320u ^ (c >> 1) : c >> 1; crc_table[i] = c; }
        have_table = true;
    }
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    auto chunk = [&](const char* type, const std::vector<unsigned char>& d) {
        unsigned char hd[8] = {(unsigned char)(d.size() >> 24), (unsigned char)(d.size() >> 16), (unsigned char)(d.size() >> 8), (unsigned char)d.size(),
                               (unsigned char)type[0], (unsigned char)type[1], (unsigned char)type[2], (unsigned char)type[3]};
        uint32_t c = 0xffffffffu;
        for (int i = 4; i < 8; i++) c = crc_table[(c ^ hd[i]) & 255] ^ (c >> 8);
        for (unsigned char v : d) c = crc_table[(c ^ v) & 255] ^ (c >> 8);
        c ^= 0xffffffffu;
        unsigned char tail[4] = {(unsigned char)(c >> 24), (unsigned char)(c >> 16), (unsigned char)(c >> 8), (unsigned char)c};
        std::fwrite(hd, 1, 8, f); if (!d.empty()) std::fwrite(d.data(), 1, d.size(), f); std::fwrite(tail, 1, 4, f);
    };
    static const unsigned char sig[8] = {137, 80, 78, 71, 13, 10, 26, 10};
    std::fwrite(sig, 1, 8, f);
    std::vector<unsigned char> ihdr = {(unsigned char)(w >> 24), (unsigned char)(w >> 16), (unsigned char)(w >> 8), (unsigned char)w,
                                       (unsigned char)(h >> 24), (unsigned char)(h >> 16), (unsigned char)(h >> 8), (unsigned char)h, 8, 2, 0, 0, 0};
    chunk("IHDR", ihdr);
    // scanlines with filter byte 0, wrapped in stored deflate blocks
    const size_t stride = 3 * (size_t)w + 1, total = stride * h;
    std::vector<unsigned char> raw(total);
    for (int y = 0; y < h; y++) { raw[y * stride] = 0; std::memcpy(&raw[y * stride + 1], rgb_top_down + 3 * (size_t)y * w, 3 * (size_t)w); }
    std::vector<unsigned char> z; z.reserve(total + total / 65535 * 5 + 16);
    z.push_back(0x78); z.push_back(0x01);
    uint32_t s1 = 1, s2 = 0;
    for (size_t off = 0; off < total || off == 0; ) {
        size_t n = std::min<size_t>(65535, total - off);
        z.push_back(off + n >= total ? 1 : 0);
        z.push_back((unsigned char)(n & 255)); z.push_back((unsigned char)(n >> 8));
        z.push_back((unsigned char)(~n & 255)); z.push_back((unsigned char)((~n >> 8) & 255));
        for (size_t i = 0; i < n; i++) { unsigned char v = raw[off + i]; z.push_back(v); s1 = (s1 + v) % 65521u; s2 = (s2 + s1) % 65521u; }
        off += n;
        if (n == 0) break;
    }
    z.push_back((unsigned char)(s2 >> 8)); z.push_back((unsigned char)s2); z.push_back((unsigned char)(s1 >> 8)); z.push_back((unsigned char)s1);
    chunk("IDAT", z);
    chunk("IEND", {});
    return std::fclose(f) == 0;
}

// stbi_loadf(path, .., 4) for a file that is NOT Radiance (the ".jpg" environment branch, Renderer.cpp:262-266): the 8-bit
// decode goes through stbi__ldr_to_hdr -- colour = (float)(pow(byte / 255.0f, 2.2f) * 1.0f) evaluated in double (SOIL2.c is
// C: pow is the double one, its arguments the float quotient and the float constant 2.2f), alpha = byte / 255.0f.
bool load_ldr_as_hdr_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba, std::string& err) {
    std::vector<unsigned char> file;
    {
        FILE* f = std::fopen(path.c_str(), "rb");
        if (!f) { err = "cannot open " + path; return false; }
        unsigned char tmp[65536]; size_t n;
        while ((n = std::fread(tmp, 1, sizeof(tmp), f)) > 0) file.insert(file.end(), tmp, tmp + n);
        std::fclose(f);
    }
    std::vector<unsigned char> px;
    bool ok = false;
    if (file.size() > 3 && file[0] == 0xff && file[1] == 0xd8) ok = decode_jpeg_rgba8(file, w, h, px, err);
    else if (file.size() > 8 && file[0] == 137 && file[1] == 'P') { ok = decode_png_rgba8(file, w, h, px); if (!ok) err = "cannot decode PNG"; }
    else err = "neither JPEG nor PNG";
    if (!ok) { err = path + ": " + err; return false; }
    static float lut[256]; static bool have = false;
    if (!have) { for (int i = 0; i < 256; i++) lut[i] = (float)(std::pow((double)(i / 255.0f), (double)2.2f) * (double)1.0f); have = true; }
    rgba.resize(px.size());
    for (size_t i = 0; i < px.size(); i += 4) {
        rgba[i] = lut[px[i]]; rgba[i + 1] = lut[px[i + 1]]; rgba[i + 2] = lut[px[i + 2]];
        rgba[i + 3] = px[i + 3] / 255.0f;
    }
    return true;
}

}  // namespace pt

static int image_load_hdr(const char* path, int32_t* width, int32_t* height, float** rgba);
extern "C" int pt_image_load_hdr(const char* path, int32_t* width, int32_t* height, float** rgba) {
    try { return image_load_hdr(path, width, height, rgba); }
    catch (const std::exception& e) { pt::set_last_error((std::string("pt_image_load_hdr: ") + e.what()).c_str()); return PT_ERR_LIMIT; }
}
static int image_load_hdr(const char* path, int32_t* width, int32_t* height, float** rgba) {
    if (!path || !width || !height || !rgba) return PT_ERR_INVALID;
    int w = 0, h = 0; std::vector<float> px; std::string err;
    // stbi_loadf: Radiance files as they are, anything else through the 8-bit decoders and the gamma-2.2 expansion
    bool radiance = false;
    if (FILE* f = std::fopen(path, "rb")) { char sig[3] = {0, 0, 0}; size_t got = std::fread(sig, 1, 2, f); std::fclose(f); radiance = got == 2 && sig[0] == '#' && sig[1] == '?'; }
    const bool ok = radiance ? pt::load_hdr_rgba32f(path, w, h, px, err) : pt::load_ldr_as_hdr_rgba32f(path, w, h, px, err);
    if (!ok) { pt::set_last_error(err.c_str()); return PT_ERR_IO; }
    *rgba = (float*)std::malloc(px.size() * sizeof(float));
    if (!*rgba) return PT_ERR_LIMIT;
    std::memcpy(*rgba, px.data(), px.size() * sizeof(float));
    *width = w; *height = h;
    return PT_OK;
}

extern "C" int pt_image_save_png(const char* path, int32_t width, int32_t height, const uint8_t* rgb_top_down) {
    if (!path || !rgb_top_down || width < 1 || height < 1) return PT_ERR_INVALID;
    if (!pt::write_png_rgb8(path, width, height, rgb_top_down)) { pt::set_last_error((std::string("cannot write ") + path).c_str()); return PT_ERR_IO; }
    return PT_OK;
}
