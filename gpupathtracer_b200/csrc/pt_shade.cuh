// pt_shade.cuh -- RNG, material, microfacet BSDF, light sampling and ray generation device code.
// Replaces (reference paths relative to src/shaders/): common/Random.glsl:47-70 (HybridTaus),
// common/Material.glsl:39-130, common/Microfacet.glsl:13-55,96-193, common/BSDF.glsl:8-21,
// common/MIS.glsl:6-8, Iterative.comp:34-82 (RandomLightVertex), :113-147 (GenerateLightSample),
// :163-210 (RayIndexToMorton, InitRay).
#pragma once
#include "pt_math.cuh"
#include "pt_layout.h"
#include "pt_parity_libm.h"

namespace pt {

// The template parameter MODEL of the shading functions carries two switches: bit 0 = microfacet model (0 Trowbridge-Reitz,
// 1 Beckmann), bit 1 = PT_PARITY_LIBM (transcendentals through pt_parity_libm.h instead of CUDA's libm).
#define PT_MODEL_BECKMANN(M) (((M) & 1) != 0)
#define PT_MODEL_PLIBM(M) (((M) & 2) != 0)
template <bool PL> __device__ __forceinline__ void sincos_sel(float x, float* s, float* c) {
    if (PL) ptpl::sincosf_(x, s, c); else sincosf(x, s, c);
}
template <bool PL> __device__ __forceinline__ float log_sel(float x) { return PL ? ptpl::logf_(x) : logf(x); }
template <bool PL> __device__ __forceinline__ float exp_sel(float x) { return PL ? ptpl::expf_(x) : expf(x); }
template <bool PL> __device__ __forceinline__ float pow_sel(float x, float y) { return PL ? ptpl::powf_(x, y) : powf(x, y); }

// ---------------------------------------------------------------- HybridTaus (Random.glsl:47-70)
struct Rng {
    uint32_t x, y, z, w;
    __device__ __forceinline__ static uint32_t taus(uint32_t& z, int s1, int s2, int s3, uint32_t m) {
        uint32_t b = (((z << s1) ^ z) >> s2);
        z = (((z & m) << s3) ^ b);
        return z;
    }
    __device__ __forceinline__ float next() {
        uint32_t i = taus(x, 13, 19, 12, 4294967294U) ^ taus(y, 2, 25, 4, 4294967288U) ^
                     taus(z, 3, 11, 17, 4294967280U);
        w = 1664525U * w + 1013904223U;
        i ^= w;
        return 2.3283064365387e-10f * __uint2float_rn(i);   // may return exactly 1.0f, like the shader
    }
};

// ---------------------------------------------------------------- material (Material.glsl)
struct Material {
    V3 albedo, emission, reflectance;
    float metallic, roughness, roughness2;
};
struct Interaction {
    V3 normal, outgoing, incoming, microfacet;
    float ndm, ndo, ndi, ndm2, idm;
    V3 tbn_t, tbn_b;    // tbn_n == normal
};

// GL_LINEAR + GL_REPEAT lookup of an RGBA32F image (Texture.cpp:123-131), float weights
__device__ inline float4 sample_texture(const TextureRec& t, float u, float v) {
    float x = u * (float)t.width - 0.5f, y = v * (float)t.height - 0.5f;
    float fx = floorf(x), fy = floorf(y);
    float ax = x - fx, ay = y - fy;
    int ix = (int)fx, iy = (int)fy;
    int x0 = ix % t.width; if (x0 < 0) x0 += t.width;
    int x1 = (ix + 1) % t.width; if (x1 < 0) x1 += t.width;
    int y0 = iy % t.height; if (y0 < 0) y0 += t.height;
    int y1 = (iy + 1) % t.height; if (y1 < 0) y1 += t.height;
    float4 a = __ldg(t.texels + (size_t)y0 * t.width + x0), b = __ldg(t.texels + (size_t)y0 * t.width + x1);
    float4 c = __ldg(t.texels + (size_t)y1 * t.width + x0), d = __ldg(t.texels + (size_t)y1 * t.width + x1);
    float4 r;
    r.x = mixf(mixf(a.x, b.x, ax), mixf(c.x, d.x, ax), ay);
    r.y = mixf(mixf(a.y, b.y, ax), mixf(c.y, d.y, ax), ay);
    r.z = mixf(mixf(a.z, b.z, ax), mixf(c.z, d.z, ax), ay);
    r.w = mixf(mixf(a.w, b.w, ax), mixf(c.w, d.w, ax), ay);
    return r;
}

// texture(samplerCube(sky), d).xyz, Iterative.comp:251. Face and (s, t) by the OpenGL cube-map table (major axis; ties,
// which GL leaves to the implementation, go to x, then y), then GL_LINEAR with CLAMP_TO_EDGE inside that face
// (Renderer.cpp:318-322; seamless filtering is never enabled), float weights. The reference's CPU twin of the lookup is
// TextureCubemap::Sample (Texture.cpp:278-340, nearest texel, y negated).
__device__ inline V3 sample_environment(const SceneDev& s, V3 d) {
    if (s.env_size == 0) return v3(s.env[0], s.env[1], s.env[2]);
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int face; float sc, tc, ma;
    if (ax >= ay && ax >= az) { face = d.x > 0.0f ? 0 : 1; ma = ax; sc = d.x > 0.0f ? -d.z : d.z; tc = -d.y; }
    else if (ay >= az)        { face = d.y > 0.0f ? 2 : 3; ma = ay; sc = d.x; tc = d.y > 0.0f ? d.z : -d.z; }
    else                      { face = d.z > 0.0f ? 4 : 5; ma = az; sc = d.z > 0.0f ? d.x : -d.x; tc = -d.y; }
    const int n = s.env_size;
    float u = 0.5f * (sc / ma + 1.0f), v = 0.5f * (tc / ma + 1.0f);
    float x = u * (float)n - 0.5f, y = v * (float)n - 0.5f;
    float fx = floorf(x), fy = floorf(y);
    float wx = x - fx, wy = y - fy;
    int x0 = min(max((int)fx, 0), n - 1), x1 = min(max((int)fx + 1, 0), n - 1);
    int y0 = min(max((int)fy, 0), n - 1), y1 = min(max((int)fy + 1, 0), n - 1);
    const float4* f = s.env_faces + (size_t)face * n * n;
    float4 a = __ldg(f + (size_t)y0 * n + x0), b = __ldg(f + (size_t)y0 * n + x1);
    float4 c = __ldg(f + (size_t)y1 * n + x0), e = __ldg(f + (size_t)y1 * n + x1);
    return v3(mixf(mixf(a.x, b.x, wx), mixf(c.x, e.x, wx), wy),
              mixf(mixf(a.y, b.y, wx), mixf(c.y, e.y, wx), wy),
              mixf(mixf(a.z, b.z, wx), mixf(c.z, e.z, wx), wy));
}

// ConstructMaterialInstance, Material.glsl:39-54
__device__ inline Material make_material(const SceneDev& s, uint32_t matId, float tu, float tv) {
    const MaterialRec* m = s.materials + (matId >> 1);
    float4 d0 = __ldg(&m->albedo_tex), e = __ldg(&m->emission_flag), d1 = __ldg(&m->props_tex);
    int at = __float_as_int(d0.w), pt_ = __float_as_int(d1.w);
    if (at >= 0 && at < s.n_textures) d0 = sample_texture(s.textures[at], tu, tv);
    if (pt_ >= 0 && pt_ < s.n_textures) d1 = sample_texture(s.textures[pt_], tu, tv);
    Material r;
    r.emission = xyz(e);
    r.albedo = xyz(d0);
    r.roughness = gmax(d1.y * d1.y, 1e-4f);
    r.roughness2 = r.roughness * r.roughness;
    r.metallic = d1.z;
    r.reflectance = mix(v3(0.04f), r.albedo, r.metallic);
    return r;
}

// ConstructSurfaceInteraction(n, o) + ConstructTBN, Material.glsl:71-76,99-109
__device__ inline Interaction make_interaction(V3 n, V3 o) {
    Interaction it;
    it.normal = n; it.outgoing = o;
    it.ndo = nndot(n, o);
    V3 normcrs = (fabsf(n.y) > 0.99f) ? v3(1.0f, 0.0f, 0.0f) : v3(0.0f, 1.0f, 0.0f);
    it.tbn_t = normalize(cross(normcrs, n));
    it.tbn_b = cross(it.tbn_t, n);
    it.incoming = v3(0.0f); it.microfacet = v3(0.0f);
    it.ndm = it.ndi = it.ndm2 = it.idm = 0.0f;
    return it;
}
__device__ __forceinline__ V3 tbn_mul(const Interaction& it, V3 v) {
    return v3(it.tbn_t.x * v.x + it.tbn_b.x * v.y + it.normal.x * v.z,
              it.tbn_t.y * v.x + it.tbn_b.y * v.y + it.normal.y * v.z,
              it.tbn_t.z * v.x + it.tbn_b.z * v.y + it.normal.z * v.z);
}
__device__ __forceinline__ void finish_dirs(Interaction& it) {
    it.ndi = nndot(it.normal, it.incoming);
    it.ndm = nndot(it.normal, it.microfacet);
    it.ndm2 = it.ndm * it.ndm;
    it.idm = nndot(it.incoming, it.microfacet);
}
__device__ __forceinline__ void set_incoming(Interaction& it, V3 i) {      // Material.glsl:112-120
    it.incoming = i;
    it.microfacet = normalize(it.outgoing + it.incoming);
    finish_dirs(it);
}
__device__ __forceinline__ void set_microfacet(Interaction& it, V3 m) {    // Material.glsl:122-130
    it.microfacet = m;
    it.incoming = reflect(-it.outgoing, it.microfacet);
    finish_dirs(it);
}

// ---------------------------------------------------------------- microfacet (Microfacet.glsl)
__device__ __forceinline__ V3 fresnel(V3 f0, float ndo) {                  // :96-99
    float x = 1.0f - ndo;
    return f0 + (v3(1.0f) - f0) * (x * x * x * x * x);
}
template <int MODEL>
__device__ __forceinline__ float D_micro(const Material& m, const Interaction& it) {
    if (PT_MODEL_BECKMANN(MODEL)) {                                        // Beckmann :35-39
        float sub = 2.0f * log_sel<PT_MODEL_PLIBM(MODEL)>(sqrtf(PT_PI) * m.roughness * it.ndm);
        float add = (it.ndm2 - 1.0f) / (it.ndm2 * m.roughness2);
        return exp_sel<PT_MODEL_PLIBM(MODEL)>(add - sub);
    } else {                                                               // Trowbridge-Reitz :13-16
        float divsor = (m.roughness2 - 1.0f) * it.ndm2 + 1.0f;
        return m.roughness2 / gmax(PT_PI * divsor * divsor, 1e-20f);
    }
}
template <int MODEL>
__device__ __forceinline__ float pdf_micro(const Material& m, const Interaction& it) {   // :29-31 / :53-55
    return gmax(D_micro<MODEL>(m, it) * it.ndm / (4.0f * it.idm), 1e-10f);
}
__device__ __forceinline__ float vis_ggx(const Material& m, float ndo) {   // :117-119
    return 1.0f / gmax(ndo + sqrtf(m.roughness2 * (1.0f - m.roughness2) * ndo * ndo), 1e-5f);
}
__device__ __forceinline__ float vis_smith(const Material& m, const Interaction& it) {   // :121-123
    return vis_ggx(m, it.ndi) * vis_ggx(m, it.ndo) / 4.0f;
}
__device__ __forceinline__ V3 diffuse_energy(const Material& m, float ndi, float ndo) {  // :140-142
    return (1.0f - m.metallic) * (v3(1.0f) - fresnel(m.reflectance, ndi)) * (v3(1.0f) - fresnel(m.reflectance, ndo));
}
__device__ __forceinline__ float diffuse_pmf(const Material& m, const Interaction& it) { // :156-161
    V3 e = diffuse_energy(m, 0.5f, it.ndo);
    float lum = dot(e, v3(1.0f / 3.0f));
    return clampf(lum, 0.0f, 1.0f);
}
template <int MODEL>
__device__ inline V3 bsdf(const Material& m, const Interaction& it) {      // BSDF.glsl:8-21
    if (dot(it.normal, it.incoming) < 0.0f || dot(it.normal, it.outgoing) < 0.0f) return v3(0.0f);
    V3 specular = fresnel(m.reflectance, it.idm) * D_micro<MODEL>(m, it) * vis_smith(m, it);
    V3 diffuse = m.albedo / PT_PI * diffuse_energy(m, it.ndi, it.ndo);
    return specular + diffuse;
}
template <int MODEL>
__device__ inline float pdf_direction(const Material& m, const Interaction& it) {   // :163-170
    float dp = diffuse_pmf(m, it);
    float sp = 1.0f - dp;
    float pdfDiffuse = dp * (it.ndi / PT_PI);
    float pdfSpecular = sp * pdf_micro<MODEL>(m, it);
    return pdfDiffuse + pdfSpecular;
}
__device__ __forceinline__ float mis_weight(float top, float bottom) { return 1.0f / (1.0f + bottom / top); }

// GenerateImportanceSample, Microfacet.glsl:172-193
template <int MODEL>
__device__ inline V3 importance_sample(const Material& m, Interaction& it, Rng& rng, float& pdfSample, float& pdfMIS) {
    float dp = diffuse_pmf(m, it);
    float sp = 1.0f - dp;
    bool diffuse = rng.next() < dp;
    float rx = rng.next(), ry = rng.next();
    float phi = 2.0f * PT_PI * ry;
    float sphi, cphi;
    sincos_sel<PT_MODEL_PLIBM(MODEL)>(phi, &sphi, &cphi);
    if (diffuse) {
        float radius = sqrtf(rx);
        float z = sqrtf(1.0f - rx);
        set_incoming(it, tbn_mul(it, v3(radius * sphi, radius * cphi, z)));
        pdfSample = dp * (it.ndi / PT_PI);
        pdfMIS = sp * pdf_micro<MODEL>(m, it);
    } else {
        float z2, radius;
        if (PT_MODEL_BECKMANN(MODEL)) {
            float g = -m.roughness2 * log_sel<PT_MODEL_PLIBM(MODEL)>(1.0f - rx);
            z2 = 1.0f / (1.0f + g);
            radius = sqrtf(1.0f - z2);
        } else {
            z2 = gmax((1.0f - rx) / (rx * (m.roughness2 - 1.0f) + 1.0f), 0.0f);
            radius = sqrtf(gmax(1.0f - z2, 0.0f));
        }
        float z = sqrtf(z2);
        set_microfacet(it, tbn_mul(it, v3(radius * sphi, radius * cphi, z)));
        float pdfDiffuse = dp * (it.ndi / PT_PI);
        pdfSample = sp * pdf_micro<MODEL>(m, it);
        pdfMIS = pdfDiffuse;
    }
    return it.incoming;
}

// ---------------------------------------------------------------- light sampling
// RandomLightVertex, Iterative.comp:34-82. `data` is the LAST FETCHED entry, as in the shader.
__device__ inline void random_light_vertex(const SceneDev& s, Rng& rng, V3& pos, V3& nrm, uint32_t& mat) {
    float selectedArea = rng.next() * s.total_light_area;
    float2 data = __ldg(s.lights);
    int a = 0, b = s.n_lights - 1;
    while (b > a) {
        int m = (a + b) / 2;
        data = __ldg(s.lights + m);
        if (data.x > selectedArea) b = m; else a = m + 1;
    }
    int idx = __float_as_int(data.y);
    const float4* t = s.tris + 5 * (size_t)idx;
    float4 t0 = __ldg(t), t1 = __ldg(t + 1), t2 = __ldg(t + 2), t3 = __ldg(t + 3), t4 = __ldg(t + 4);
    V3 p0 = v3(t0.x, t0.y, t0.z);
    V3 p1 = v3(t0.w, t1.x, t1.y) + p0;
    V3 p2 = v3(t1.z, t1.w, t2.x) + p0;
    float rx = rng.next(), ry = rng.next();
    float sr = sqrtf(rx);
    float u = 1.0f - sr;
    float v = ry * sr;
    float w = 1.0f - u - v;
    pos = p0 * u + p1 * v + p2 * w;
    nrm = v3(t3.w, t4.x, t4.y);
    mat = __float_as_uint(t4.z);
}

// RandomSunVertex, Iterative.comp:84-110: the sun is a disk of radius sunRadius one unit away along sunDir, sampled
// uniformly by area in the frame ConstructTBN(-sunDir); its material is the sky's (matId 0), whose emission
// (25 * (30, 26, 19), Scene.cpp:239) is what the NEE sample picks up.
template <bool PL>
__device__ inline void random_sun_vertex(const SceneDev& s, Rng& rng, V3 vertex, V3& pos, V3& nrm, uint32_t& mat) {
    V3 sd = v3(s.sun_dir[0], s.sun_dir[1], s.sun_dir[2]);
    V3 n = -sd;
    V3 normcrs = (fabsf(n.y) > 0.99f) ? v3(1.0f, 0.0f, 0.0f) : v3(0.0f, 1.0f, 0.0f);     // ConstructTBN, Material.glsl:71-76
    V3 t = normalize(cross(normcrs, n));
    V3 b = cross(t, n);
    float rx = rng.next(), ry = rng.next();
    float radius = s.sun_radius * sqrtf(rx);
    float phi = 2.0f * PT_PI * ry;
    float sn, cs; sincos_sel<PL>(phi, &sn, &cs);
    V3 d = v3(radius * cs, radius * sn, 0.0f);
    V3 off = v3(t.x * d.x + b.x * d.y + n.x * d.z, t.y * d.x + b.y * d.y + n.y * d.z, t.z * d.x + b.z * d.y + n.z * d.z);
    pos = vertex + sd + off;
    nrm = n;
    mat = 0u;
}

// ---------------------------------------------------------------- ray generation
__device__ __forceinline__ int compact1by1(int x) {          // Iterative.comp:154-161
    x &= 0x55555555;
    x = (x ^ (x >> 1)) & 0x33333333;
    x = (x ^ (x >> 2)) & 0x0f0f0f0f;
    x = (x ^ (x >> 4)) & 0x00ff00ff;
    x = (x ^ (x >> 8)) & 0x0000ffff;
    return x;
}
__device__ __forceinline__ void ray_index_to_morton(int idx, int width, int& px, int& py) {   // :163-172
    int i = idx % 64;
    int j = idx / 64;
    int x = j % (width / 8);
    int y = j / (width / 8);
    px = compact1by1(i) + 8 * x;
    py = compact1by1(i >> 1) + 8 * y;
}
__device__ __forceinline__ V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// InitRay, Iterative.comp:174-210
template <bool PL>
__device__ inline void init_ray(const CameraRec& cam, int width, int height, int px, int py, Rng& rng, V3& o, V3& d) {
    float jx = rng.next(), jy = rng.next();
    float s = ((float)px + jx) / (float)width;
    float t = ((float)py + jy) / (float)height;
    float rx = rng.next(), ry = rng.next();
    float phi = 2.0f * PT_PI * ry;
    float sphi, cphi;
    sincos_sel<PL>(phi, &sphi, &cphi);
    float rdx = cam.lens_radius * rx * cphi;     // `lens_radius * r * vec2(cos, sin)`: r, not sqrt(r.x)
    float rdy = cam.lens_radius * ry * sphi;
    V3 offset = ld3(cam.u) * rdx + ld3(cam.v) * rdy;
    o = ld3(cam.position) + offset;
    d = normalize(ld3(cam.lower_left) + s * ld3(cam.horizontal) + t * ld3(cam.vertical) - offset);
}

}  // namespace pt
