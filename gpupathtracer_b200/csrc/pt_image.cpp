// pt_image.cpp -- Radiance .hdr (RGBE) reader for the environment path; PNG reader (map_Kd, cube faces) and writer (SaveScreenshot).
//
// The reference reads its equirectangular environment with stbi_loadf(path, &w, &h, &c, 4) (Renderer.cpp:262-266; stb_image
// comes with SOIL2, extern/soil2-cmake/src/SOIL2/stb_image.h). This restates what that call yields for a Radiance file:
//   * signature "#?RADIANCE\n" or "#?RGBE\n", header lines up to the first empty one, one of them must be
//     "FORMAT=32-bit_rle_rgbe"; resolution line "-Y <h> +X <w>" only (other orientations are refused)
//   * scanlines 8 <= w < 32768 in the new run-length form (2, 2, w>>8, w&255, then the four byte planes as runs/dumps);
//     a scanline that does not start that way switches the reader to flat 4-byte pixels -- and, like stb, restarts the
//     flat read at pixel 1 of row 0 (only right when it happens on the first row; kept for identical output)
//   * pixel = byte * 2^(e - 136), alpha 1, e == 0 -> black; rows in file order (no flip); reads past the end yield 0
// oracle/ref_stbi.cpp compiles stb_image in place as the checker (tests/test_environment.py).
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

bool inflate_zlib(const unsigned char* data, size_t n, std::vector<unsigned char>& out) {
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
                if (sym < 256) out.push_back((unsigned char)sym);
                else if (sym == 256) break;
                else {
                    sym -= 257;
                    if (sym >= 29) return false;
                    int len = lbase[sym] + br.bits(lext[sym]);
                    int ds = dist.decode(br);
                    if (ds < 0 || ds >= 30) return false;
                    size_t d = dbase[ds] + (size_t)br.bits(dext[ds]);
                    if (br.bad || d > out.size()) return false;
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
            if (w <= 0 || h <= 0 || w > (1 << 24) || h > (1 << 24) || d[10] || d[11] || interlace > 1) return false;
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
    raw.reserve(((size_t)w * bits_pp / 8 + 2) * h);
    if (!inflate_zlib(idat.data(), idat.size(), raw)) return false;
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

}  // namespace pt

extern "C" int pt_image_load_hdr(const char* path, int32_t* width, int32_t* height, float** rgba) {
    if (!path || !width || !height || !rgba) return PT_ERR_INVALID;
    int w = 0, h = 0; std::vector<float> px; std::string err;
    if (!pt::load_hdr_rgba32f(path, w, h, px, err)) { pt::set_last_error(err.c_str()); return PT_ERR_IO; }
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
