// pt_image.cpp -- Radiance .hdr (RGBE) reader for the environment path.
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
