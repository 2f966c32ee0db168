// oracle/ref_stbi.c -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// The reference decodes images with the stb_image that ships inside SOIL2 (extern/soil2-cmake/src/SOIL2/stb_image.h):
// stbi_loadf for the equirectangular environment (Renderer.cpp:265), SOIL_load_image -> stbi_load for map_Kd and cube
// faces (Texture.cpp:146-156, :262). This harness compiles that UNMODIFIED header where it lies under /root/reference
// (oracle/Makefile, target `ref`) so the tests can hold libb200pt's own decoders (csrc/pt_image.cpp, csrc/pt_obj.cpp)
// against it byte for byte. Compiled as C, like SOIL2.c, the one translation unit of the reference that carries the
// implementation (so that e.g. pow() in stbi__ldr_to_hdr is the double-precision one).
// (oracle/Makefile feeds the compiler stb_image.h, implementation enabled, in front of this file)
#include <stdlib.h>
#include <string.h>

// RGBA32F, rows in file order; returns 0 on success. *out is malloc()ed (ref_stbi_free).
int ref_stbi_loadf(const char* path, int* w, int* h, float** out) {
    int c = 0;
    float* d = stbi_loadf(path, w, h, &c, 4);
    if (!d) return -1;
    size_t n = (size_t)*w * *h * 4 * sizeof(float);
    *out = (float*)malloc(n);
    memcpy(*out, d, n);
    stbi_image_free(d);
    return 0;
}
// RGBA8, rows in file order
int ref_stbi_load(const char* path, int* w, int* h, unsigned char** out) {
    int c = 0;
    unsigned char* d = stbi_load(path, w, h, &c, 4);
    if (!d) return -1;
    size_t n = (size_t)*w * *h * 4;
    *out = (unsigned char*)malloc(n);
    memcpy(*out, d, n);
    stbi_image_free(d);
    return 0;
}
void ref_stbi_free(void* p) { free(p); }
