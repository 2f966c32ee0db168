// oracle/pt_oracle.cpp -- TEST INFRASTRUCTURE ONLY. See pt_oracle.h for scope, parity status
// and the stated float mode. Every function cites the reference file:line it restates
// (paths relative to /root/reference/src).
#include "pt_oracle.h"
#include "pt_parity_libm.h"     // gpupathtracer_b200/csrc: the ONE definition of sin / cos / log / exp / pow shared by every compile in parity mode
#include <cmath>
#include <cstring>
#include <cstdlib>
#include <thread>
#include <vector>
#include <atomic>
#include <algorithm>

namespace {

// GLSL's sin / cos / log / exp / pow: glibc's float functions by default; pt_parity_libm.h after or_set_parity_libm(1)
// (set once before a run; the worker threads only read it).
int g_parity_libm = 0;
inline float o_sin(float x) { return g_parity_libm ? ptpl::sinf_(x) : sinf(x); }
inline float o_cos(float x) { return g_parity_libm ? ptpl::cosf_(x) : cosf(x); }
inline float o_log(float x) { return g_parity_libm ? ptpl::logf_(x) : logf(x); }
inline float o_exp(float x) { return g_parity_libm ? ptpl::expf_(x) : expf(x); }
inline float o_pow(float x, float y) { return g_parity_libm ? ptpl::powf_(x, y) : powf(x, y); }

// ---------------------------------------------------------------- GLSL-like vector math
struct V3 { float x, y, z; };
inline V3 v3(float x, float y, float z) { return V3{x, y, z}; }
inline V3 v3(float s) { return V3{s, s, s}; }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator*(V3 a, V3 b) { return v3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline V3 operator/(V3 a, V3 b) { return v3(a.x / b.x, a.y / b.y, a.z / b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return v3(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, float s) { return v3(a.x / s, a.y / s, a.z / s); }
inline V3 operator-(V3 a) { return v3(-a.x, -a.y, -a.z); }
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// glm::cross operand order (detail/func_geometric.inl): x.y*y.z - y.y*x.z, ...
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline float length(V3 a) { return sqrtf(dot(a, a)); }
inline float distance(V3 a, V3 b) { return length(a - b); }
inline V3 normalize(V3 a) { return a * (1.0f / sqrtf(dot(a, a))); }
inline float gmin(float a, float b) { return fminf(a, b); }
inline float gmax(float a, float b) { return fmaxf(a, b); }
inline float clampf(float x, float lo, float hi) { return gmin(gmax(x, lo), hi); }
inline float mixf(float x, float y, float a) { return x * (1.0f - a) + y * a; }
inline V3 mix(V3 x, V3 y, float a) { return v3(mixf(x.x, y.x, a), mixf(x.y, y.y, a), mixf(x.z, y.z, a)); }
inline V3 reflect(V3 I, V3 N) { return I - N * dot(N, I) * 2.0f; }
inline float nndot(V3 a, V3 b) { return gmax(dot(a, b), 0.0f); }   // Util.glsl:6
inline float avdot(V3 a, V3 b) { return fabsf(dot(a, b)); }        // Util.glsl:7
inline int fbs(float f) { int i; std::memcpy(&i, &f, 4); return i; }
inline float sbf(int i) { float f; std::memcpy(&f, &i, 4); return f; }
const float kPi = 3.141592653589793238462643383279f;             // Constants.glsl:4

struct Ray { V3 origin, direction; };

// ---------------------------------------------------------------- RNG  (Random.glsl:47-70)
struct Rng {
    uint32_t x, y, z, w;
    static inline uint32_t taus(uint32_t& z, int s1, int s2, int s3, uint32_t m) {
        uint32_t b = (((z << s1) ^ z) >> s2);
        z = (((z & m) << s3) ^ b);
        return z;
    }
    static inline uint32_t lcg(uint32_t& z, uint32_t a, uint32_t c) { z = a * z + c; return z; }
    inline float next() {
        uint32_t i = taus(x, 13, 19, 12, 4294967294U) ^ taus(y, 2, 25, 4, 4294967288U) ^
                     taus(z, 3, 11, 17, 4294967280U) ^ lcg(w, 1664525U, 1013904223U);
        return 2.3283064365387e-10f * (float)i;   // float(uint) rounds to nearest: can return exactly 1.0f
    }
};

// ---------------------------------------------------------------- traversal
struct Hit { float t, u, v; int tri; };
struct Cnt { uint64_t pairs = 0, leaves = 0, tris = 0, nulls = 0; std::vector<uint8_t>* ops = nullptr; };   // ops: optional per-step trace (0 = pair test, 1 = triangle test)

// IntersectNodeFast, BVH.glsl:379-391. iray = {o' = -o/d, 1/d}.
inline float slab(const float* node, const Ray& iray, float tcur, bool& hit) {
    V3 nmin = v3(node[0], node[1], node[2]), nmax = v3(node[4], node[5], node[6]);
    V3 a = nmin * iray.direction + iray.origin;
    V3 b = nmax * iray.direction + iray.origin;
    V3 tmin = v3(gmin(a.x, b.x), gmin(a.y, b.y), gmin(a.z, b.z));
    V3 tmax = v3(gmax(a.x, b.x), gmax(a.y, b.y), gmax(a.z, b.z));
    float t_entry = gmax(tmin.x, gmax(tmin.y, tmin.z));
    float t_exit = gmin(tmax.x, gmin(tmax.y, gmin(tmax.z, tcur)));
    hit = (t_entry <= t_exit && t_exit > 0.0f);
    return t_entry;
}

// IntersectTriangleMT, Geometry.glsl:198-225 (tri = 20 floats, 5 x vec4 layout).
inline bool tri_mt(const float* tri, const Ray& ray, Hit& best, int index) {
    V3 p0 = v3(tri[0], tri[1], tri[2]);
    V3 v01 = v3(tri[3], tri[4], tri[5]);
    V3 v02 = v3(tri[6], tri[7], tri[8]);
    V3 p = cross(ray.direction, v02);
    float det = dot(v01, p);
    float idet = 1.0f / det;
    V3 tv = ray.origin - p0;
    float u = dot(tv, p) * idet;
    V3 q = cross(tv, v01);
    float v = dot(ray.direction, q) * idet;
    float t = dot(v02, q) * idet;
    if (u > 0.0f && u < 1.0f && v > 0.0f && u + v < 1.0f && t < best.t && t > 0.0f) {
        best.t = t; best.u = u; best.v = v; best.tri = index;
        return true;
    }
    return false;
}

// IntersectLeaf, BVH.glsl:420-437. The whole leaf is always tested. A reference equal to 0
// cannot carry the end marker (-0 == 0, BVH.cpp:2117), so such a leaf runs on into the next
// one exactly like the shader; the walk is clamped at the end of the list where the shader
// would read out of bounds.
inline void leaf(const OrScene& s, int leafv, const Ray& ray, Hit& h, bool& result, Cnt& c) {
    int i = -leafv;
    bool iterating = true;
    c.leaves++;
    if (c.ops) c.ops->push_back(2);   // leaf entered
    while (iterating) {
        if (i >= s.n_refs) break;
        int index = s.refs[i++];
        if (index < 0) { index = -index; iterating = false; }
        c.tris++;
        if (c.ops) c.ops->push_back(1);
        bool hit = tri_mt(s.tris + 20 * (size_t)index, ray, h, index);
        result = result || hit;
    }
}

const int kStack = 96;   // reference: 27 (BVH.glsl:489); deeper trees are UB there, defined here (libb200pt accepts depth < 96)

// IfIfClosestHit, BVH.glsl:634-767 (SPECULATIVE_TRAVERSAL off).
bool closest_hit(const OrScene& s, const Ray& ray, Hit& h, Cnt& c) {
    Ray iray;
    iray.direction = v3(1.0f) / ray.direction;
    iray.origin = -ray.origin * iray.direction;
    int current = 1;             // RootFirstChild(), BVH.glsl:627
    int stack[kStack]; int next = 0;
    bool result = false;
    while (true) {
        if (!(current <= 0)) {
            const float* c0 = s.nodes + 8 * (size_t)current;
            const float* c1 = c0 + 8;
            c.pairs++;
            if (c.ops) c.ops->push_back(0);
            int sub0 = fbs(c0[3]); bool hit0; float d0 = slab(c0, iray, h.t, hit0);
            int sub1 = fbs(c1[3]); bool hit1; float d1 = slab(c1, iray, h.t, hit1);
            if (hit0 && hit1) {
                if (d0 > d1) { current = sub1; if (next < kStack) stack[next++] = sub0; }
                else { current = sub0; if (next < kStack) stack[next++] = sub1; }
            } else if (hit0 != hit1) {
                current = hit0 ? sub0 : sub1;
            } else {
                if (next == 0) break;
                current = stack[--next];
            }
        }
        if (current <= 0) {
            leaf(s, current, ray, h, result, c);
            if (next == 0) break;
            current = stack[--next];
        }
    }
    return result;
}

// IfIfAnyHit, BVH.glsl:769-854: no near/far ordering; returns after the first LEAF that hit.
bool any_hit(const OrScene& s, const Ray& ray, Hit& h, Cnt& c) {
    Ray iray;
    iray.direction = v3(1.0f) / ray.direction;
    iray.origin = -ray.origin * iray.direction;
    int current = 1;
    int stack[kStack]; int next = 0;
    bool result = false;
    while (true) {
        if (!(current <= 0)) {
            const float* c0 = s.nodes + 8 * (size_t)current;
            const float* c1 = c0 + 8;
            c.pairs++;
            if (c.ops) c.ops->push_back(0);
            int sub0 = fbs(c0[3]); bool hit0; slab(c0, iray, h.t, hit0);
            int sub1 = fbs(c1[3]); bool hit1; slab(c1, iray, h.t, hit1);
            if (hit0 && hit1) {
                current = sub0; if (next < kStack) stack[next++] = sub1;
            } else if (hit0 != hit1) {
                current = hit0 ? sub0 : sub1;
            } else {
                if (next == 0) break;
                current = stack[--next];
            }
        }
        if (current <= 0) {
            leaf(s, current, ray, h, result, c);
            if (result) return true;
            if (next == 0) break;
            current = stack[--next];
        }
    }
    return result;
}

template <class F>
void parallel_for(int64_t n, int n_threads, F f) {
    int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (n < 1024) nt = 1;
    std::atomic<int64_t> next(0);
    const int64_t chunk = 4096;
    auto work = [&](int tid) {
        while (true) {
            int64_t b = next.fetch_add(chunk);
            if (b >= n) break;
            f(tid, b, std::min(n, b + chunk));
        }
    };
    if (nt == 1) { work(0); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < nt; i++) th.emplace_back(work, i);
    for (auto& t : th) t.join();
}

void add_counters(OrCounters* c, const Cnt& k, uint64_t closest, uint64_t shadow, uint64_t paths,
                  const Cnt* shadowShare = nullptr) {
    if (!c) return;
    // callers serialise this
    c->node_pairs += k.pairs; c->leaf_visits += k.leaves; c->tri_tests += k.tris;
    if (shadowShare) { c->shadow_node_pairs += shadowShare->pairs; c->shadow_tri_tests += shadowShare->tris; c->null_shadow_rays += shadowShare->nulls; }
    c->closest_rays += closest; c->shadow_rays += shadow; c->paths += paths;
}

// ---------------------------------------------------------------- materials / shading
struct Material {            // Material.glsl:24-37
    V3 albedo, emission, reflectance;
    float metallic, roughness, roughness2;
};
struct Interaction {         // Material.glsl:57-70
    V3 normal, outgoing, incoming, microfacet;
    float ndm, ndo, ndi, ndm2, idm;
    V3 tbn_t, tbn_b, tbn_n;
};

// GL_LINEAR + GL_REPEAT lookup of an RGBA32F image (Texture.cpp:123-131); float weights.
void sample_texture(const OrTexture& t, float u, float v, float out[4]) {
    float x = u * (float)t.width - 0.5f, y = v * (float)t.height - 0.5f;
    float fx = floorf(x), fy = floorf(y);
    float ax = x - fx, ay = y - fy;
    auto wrap = [](int i, int n) { int m = i % n; return m < 0 ? m + n : m; };
    int x0 = wrap((int)fx, t.width), x1 = wrap((int)fx + 1, t.width);
    int y0 = wrap((int)fy, t.height), y1 = wrap((int)fy + 1, t.height);
    for (int c = 0; c < 4; c++) {
        float a = t.rgba[4 * ((size_t)y0 * t.width + x0) + c], b = t.rgba[4 * ((size_t)y0 * t.width + x1) + c];
        float d = t.rgba[4 * ((size_t)y1 * t.width + x0) + c], e = t.rgba[4 * ((size_t)y1 * t.width + x1) + c];
        out[c] = mixf(mixf(a, b, ax), mixf(d, e, ax), ay);
    }
}

// GL_LINEAR + CLAMP_TO_EDGE lookup in a w x h RGBA32F image, float weights
void sample_clamped(const float* img, int w, int h, float u, float v, float out[4]) {
    float x = u * (float)w - 0.5f, y = v * (float)h - 0.5f;
    float fx = floorf(x), fy = floorf(y);
    float ax = x - fx, ay = y - fy;
    auto clampi = [](int i, int n) { return i < 0 ? 0 : (i > n - 1 ? n - 1 : i); };
    int x0 = clampi((int)fx, w), x1 = clampi((int)fx + 1, w);
    int y0 = clampi((int)fy, h), y1 = clampi((int)fy + 1, h);
    for (int c = 0; c < 4; c++) {
        float a = img[4 * ((size_t)y0 * w + x0) + c], b = img[4 * ((size_t)y0 * w + x1) + c];
        float d = img[4 * ((size_t)y1 * w + x0) + c], e = img[4 * ((size_t)y1 * w + x1) + c];
        out[c] = mixf(mixf(a, b, ax), mixf(d, e, ax), ay);
    }
}

// texture(samplerCube(fbu(materialInstance[0].xy)), ray.direction).xyz, Iterative.comp:251. OpenGL 4.3 core, 8.13
// "Cube Map Texture Selection", table 8.19:  major axis  target  sc   tc   ma
//                                              +rx        +X     -rz  -ry  rx      -rx  -X  +rz  -ry  rx
//                                              +ry        +Y     +rx  +rz  ry      -ry  -Y  +rx  -rz  ry
//                                              +rz        +Z     +rx  -ry  rz      -rz  -Z  -rx  -ry  rz
// s = (sc/|ma| + 1)/2, t = (tc/|ma| + 1)/2; then an ordinary 2-D lookup in that face with the cubemap's sampler
// state (LINEAR, CLAMP_TO_EDGE: Renderer.cpp:318-322; GL_TEXTURE_CUBE_MAP_SEAMLESS is never enabled). Ties between
// axes are implementation-defined in GL; here x wins over y wins over z.
V3 sample_environment(const OrScene& s, V3 d) {
    if (s.env_size == 0) return v3(s.env_color[0], s.env_color[1], s.env_color[2]);
    float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z);
    int face; float sc, tc, ma;
    if (ax >= ay && ax >= az) { face = d.x > 0.0f ? 0 : 1; ma = ax; sc = d.x > 0.0f ? -d.z : d.z; tc = -d.y; }
    else if (ay >= az)        { face = d.y > 0.0f ? 2 : 3; ma = ay; sc = d.x; tc = d.y > 0.0f ? d.z : -d.z; }
    else                      { face = d.z > 0.0f ? 4 : 5; ma = az; sc = d.z > 0.0f ? d.x : -d.x; tc = -d.y; }
    float u = 0.5f * (sc / ma + 1.0f), v = 0.5f * (tc / ma + 1.0f);
    float out[4];
    sample_clamped(s.env_faces + 4 * (size_t)face * s.env_size * s.env_size, s.env_size, s.env_size, u, v, out);
    return v3(out[0], out[1], out[2]);
}

// ConstructMaterialInstance, Material.glsl:39-54
Material make_material(const OrScene& s, uint32_t matId, float tu, float tv) {
    const float* m = s.materials + 12 * (size_t)(matId >> 1);
    Material r;
    r.emission = v3(m[4], m[5], m[6]);
    float d0[4] = {m[0], m[1], m[2], 1.0f}, d1[4] = {m[8], m[9], m[10], 1.0f};
    int at = fbs(m[3]), pt = fbs(m[11]);
    if (at >= 0 && at < s.n_textures) sample_texture(s.textures[at], tu, tv, d0);
    if (pt >= 0 && pt < s.n_textures) sample_texture(s.textures[pt], tu, tv, d1);
    r.albedo = v3(d0[0], d0[1], d0[2]);
    r.roughness = gmax(d1[1] * d1[1], 1e-4f);
    r.roughness2 = r.roughness * r.roughness;
    r.metallic = d1[2];
    r.reflectance = mix(v3(0.04f), r.albedo, r.metallic);
    return r;
}

// ConstructTBN, Material.glsl:71-76
void make_tbn(V3 n, Interaction& it) {
    V3 normcrs = (fabsf(n.y) > 0.99f) ? v3(1, 0, 0) : v3(0, 1, 0);
    it.tbn_t = normalize(cross(normcrs, n));
    it.tbn_b = cross(it.tbn_t, n);
    it.tbn_n = n;
}
inline V3 tbn_mul(const Interaction& it, V3 v) {   // mat3 * vec3, column-major
    return v3(it.tbn_t.x * v.x + it.tbn_b.x * v.y + it.tbn_n.x * v.z,
              it.tbn_t.y * v.x + it.tbn_b.y * v.y + it.tbn_n.y * v.z,
              it.tbn_t.z * v.x + it.tbn_b.z * v.y + it.tbn_n.z * v.z);
}
// ConstructSurfaceInteraction(n, o), Material.glsl:99-109
Interaction make_interaction(V3 n, V3 o) {
    Interaction it{};
    it.normal = n; it.outgoing = o;
    it.ndo = nndot(n, o);
    make_tbn(n, it);
    return it;
}
inline void finish_dirs(Interaction& it) {
    it.ndi = nndot(it.normal, it.incoming);
    it.ndm = nndot(it.normal, it.microfacet);
    it.ndm2 = it.ndm * it.ndm;
    it.idm = nndot(it.incoming, it.microfacet);
}
// SetIncomingDirection, Material.glsl:112-120
void set_incoming(Interaction& it, V3 i) {
    it.incoming = i;
    it.microfacet = normalize(it.outgoing + it.incoming);
    finish_dirs(it);
}
// SetMicrofacetDirection, Material.glsl:122-130
void set_microfacet(Interaction& it, V3 m) {
    it.microfacet = m;
    it.incoming = reflect(-it.outgoing, it.microfacet);
    finish_dirs(it);
}

// FresnelSchlick, Microfacet.glsl:96-99
inline V3 fresnel(V3 f0, float ndo) {
    float x = 1.0f - ndo;
    return f0 + (v3(1.0f) - f0) * (x * x * x * x * x);
}
// Microfacet.glsl:13-16 / :35-39
inline float D_tr(const Material& m, const Interaction& it) {
    float divsor = (m.roughness2 - 1.0f) * it.ndm2 + 1.0f;
    return m.roughness2 / gmax(kPi * divsor * divsor, 1e-20f);
}
inline float D_beckmann(const Material& m, const Interaction& it) {
    float sub = 2.0f * o_log(sqrtf(kPi) * m.roughness * it.ndm);
    float add = (it.ndm2 - 1.0f) / (it.ndm2 * m.roughness2);
    return o_exp(add - sub);
}
inline float D_micro(int model, const Material& m, const Interaction& it) {
    return model == 1 ? D_beckmann(m, it) : D_tr(m, it);
}
// ProbabilityDensity{TrowbridgeReitz,Beckmann}, Microfacet.glsl:29-31 / :53-55
inline float pdf_micro(int model, const Material& m, const Interaction& it) {
    return gmax(D_micro(model, m, it) * it.ndm / (4.0f * it.idm), 1e-10f);
}
// VisibilityGGX / VisibilitySmith, Microfacet.glsl:117-123 (the r2*(1-r2) product is the reference's)
inline float vis_ggx(const Material& m, float ndo) {
    return 1.0f / gmax(ndo + sqrtf(m.roughness2 * (1.0f - m.roughness2) * ndo * ndo), 1e-5f);
}
inline float vis_smith(const Material& m, const Interaction& it) {
    return vis_ggx(m, it.ndi) * vis_ggx(m, it.ndo) / 4.0f;
}
// DiffuseEnergyConservation, Microfacet.glsl:140-142
inline V3 diffuse_energy(const Material& m, float ndi, float ndo) {
    return (1.0f - m.metallic) * (v3(1.0f) - fresnel(m.reflectance, ndi)) * (v3(1.0f) - fresnel(m.reflectance, ndo));
}
// CalcDiffusePmf, Microfacet.glsl:156-161 (returns diffEnergy, with ndi forced to 0.5)
inline float diffuse_pmf(const Material& m, const Interaction& it) {
    V3 e = diffuse_energy(m, 0.5f, it.ndo);
    float lum = dot(e, v3(1.0f / 3.0f));     // AverageLuminance, Util.glsl:9-11
    return clampf(lum, 0.0f, 1.0f);
}
// ComputeBSDF, BSDF.glsl:8-21
V3 bsdf(int model, const Material& m, const Interaction& it) {
    if (dot(it.normal, it.incoming) < 0.0f || dot(it.normal, it.outgoing) < 0.0f) return v3(0.0f);
    V3 specular = fresnel(m.reflectance, it.idm) * D_micro(model, m, it) * vis_smith(m, it);
    V3 diffuse = m.albedo / kPi * diffuse_energy(m, it.ndi, it.ndo);
    return specular + diffuse;
}
// ProbabilityDensityDirection, Microfacet.glsl:163-170
float pdf_direction(int model, const Material& m, const Interaction& it) {
    float dp = diffuse_pmf(m, it);
    float sp = 1.0f - dp;
    float pdfDiffuse = dp * (it.ndi / kPi);
    float pdfSpecular = sp * pdf_micro(model, m, it);
    return pdfDiffuse + pdfSpecular;
}
inline float mis_weight(float top, float bottom) { return 1.0f / (1.0f + bottom / top); }  // MIS.glsl:6-8

// GenerateImportanceSample, Microfacet.glsl:172-193 (+ samplers :19-27, :42-51, :148-154)
V3 importance_sample(int model, const Material& m, Interaction& it, Rng& rng, float& pdfSample, float& pdfMIS) {
    float dp = diffuse_pmf(m, it);
    float sp = 1.0f - dp;
    if (rng.next() < dp) {
        float rx = rng.next(), ry = rng.next();
        float radius = sqrtf(rx);
        float phi = 2.0f * kPi * ry;
        float z = sqrtf(1.0f - rx);
        set_incoming(it, tbn_mul(it, v3(radius * o_sin(phi), radius * o_cos(phi), z)));
        pdfSample = dp * (it.ndi / kPi);
        pdfMIS = sp * pdf_micro(model, m, it);
    } else {
        float rx = rng.next(), ry = rng.next();
        float z2;
        if (model == 1) {
            float g = -m.roughness2 * o_log(1.0f - rx);
            z2 = 1.0f / (1.0f + g);
        } else {
            z2 = gmax((1.0f - rx) / (rx * (m.roughness2 - 1.0f) + 1.0f), 0.0f);
        }
        float z = sqrtf(z2);
        float phi = 2.0f * kPi * ry;
        float radius = (model == 1) ? sqrtf(1.0f - z2) : sqrtf(gmax(1.0f - z2, 0.0f));
        set_microfacet(it, tbn_mul(it, v3(radius * o_sin(phi), radius * o_cos(phi), z)));
        float pdfDiffuse = dp * (it.ndi / kPi);
        pdfSample = sp * pdf_micro(model, m, it);
        pdfMIS = pdfDiffuse;
    }
    return it.incoming;
}

// ---------------------------------------------------------------- light sampling
struct LightVertex { V3 position, normal; uint32_t mat; };

// RandomLightVertex, Iterative.comp:34-82 (BINARY_SEARCH). NOTE the shader keeps `data` from the
// LAST FETCHED table entry m, not from the converged index a; that is reproduced. With a single
// emitter the shader's `data` is never written (undefined); entry 0 is used here.
LightVertex random_light_vertex(const OrScene& s, Rng& rng) {
    float selectedArea = rng.next() * s.total_light_area;
    const float* data = s.lights;
    int a = 0, b = s.n_lights - 1;
    while (b > a) {
        int m = (a + b) / 2;
        data = s.lights + 2 * (size_t)m;
        if (data[0] > selectedArea) b = m; else a = m + 1;
    }
    int idx = fbs(data[1]);
    const float* t = s.tris + 20 * (size_t)idx;
    V3 p0 = v3(t[0], t[1], t[2]);
    V3 p1 = v3(t[3], t[4], t[5]) + p0;
    V3 p2 = v3(t[6], t[7], t[8]) + p0;
    float rx = rng.next(), ry = rng.next();
    float sr = sqrtf(rx);
    float u = 1.0f - sr;
    float v = ry * sr;
    float w = 1.0f - u - v;
    LightVertex lv;
    lv.position = p0 * u + p1 * v + p2 * w;
    lv.normal = v3(t[15], t[16], t[17]);
    std::memcpy(&lv.mat, t + 18, 4);
    return lv;
}

// RandomSunVertex, Iterative.comp:84-110
LightVertex random_sun_vertex(const OrConfig& cfg, V3 vertexPosition, Rng& rng) {
    V3 sunDir = v3(cfg.sun_dir[0], cfg.sun_dir[1], cfg.sun_dir[2]);
    Interaction tbn{};
    make_tbn(-sunDir, tbn);                       // mat3 sunTBN = ConstructTBN(-sunDir)
    float rx = rng.next(), ry = rng.next();       // vec2 r = rand2()
    float radius = cfg.sun_radius * sqrtf(rx);
    float phi = 2.0f * kPi * ry;
    V3 diskPosition = v3(radius * o_cos(phi), radius * o_sin(phi), 0.0f);
    LightVertex lv;
    lv.position = vertexPosition + sunDir + tbn_mul(tbn, diskPosition);
    lv.normal = -sunDir;
    lv.mat = 0;
    return lv;
}

// GenerateLightSample, Iterative.comp:113-147. `it` is passed by value as in the shader.
Ray light_sample(const OrScene& s, const OrConfig& cfg, V3 vpos, V3 vnormal, const Material& m, Interaction it, Rng& rng,
                 float& tmax, V3 throughput, V3& lightThroughput, uint32_t& lightMat) {
    const int model = cfg.microfacet_model;
    LightVertex lv;
    float lightPdf;
    if (cfg.sun_light) {                          // #ifdef SUN_LIGHT, :116-126
        float lightPick = rng.next();
        if (lightPick > 0.5f) {
            lv = random_light_vertex(s, rng);
            lightPdf = 0.5f / s.total_light_area;
            tmax = distance(lv.position, vpos) - 0.005f;
        } else {
            lv = random_sun_vertex(cfg, vpos, rng);
            float sunArea = kPi * cfg.sun_radius * cfg.sun_radius;     // :92
            lightPdf = 0.5f / sunArea;
            tmax = 1e20f;
        }
    } else {
        lv = random_light_vertex(s, rng);
        lightPdf = 1.0f / s.total_light_area;
        tmax = distance(lv.position, vpos) - 0.005f;
    }
    float radius = distance(lv.position, vpos);
    V3 lightDir = normalize(lv.position - vpos);
    set_incoming(it, lightDir);
    float bouncePdf = pdf_direction(model, m, it) * avdot(lv.normal, -lightDir) / (radius * radius);
    lightThroughput = throughput * bsdf(model, m, it) * it.ndi * avdot(lv.normal, -lightDir) *
                      mis_weight(lightPdf, bouncePdf) / (lightPdf * radius * radius);
    Ray r;
    r.origin = vpos + 0.001f * vnormal;
    r.direction = lightDir;
    lightMat = lv.mat;
    return r;
}

// ---------------------------------------------------------------- ray generation
inline int compact1by1(int x) {   // Iterative.comp:154-161
    x &= 0x55555555;
    x = (x ^ (x >> 1)) & 0x33333333;
    x = (x ^ (x >> 2)) & 0x0f0f0f0f;
    x = (x ^ (x >> 4)) & 0x00ff00ff;
    x = (x ^ (x >> 8)) & 0x0000ffff;
    return x;
}
inline void ray_index_to_morton(int idx, int width, int& px, int& py) {   // Iterative.comp:163-172
    int i = idx % 64;
    int mx = compact1by1(i), my = compact1by1(i >> 1);
    int j = idx / 64;
    int x = j % (width / 8);
    int y = j / (width / 8);
    px = mx + 8 * x; py = my + 8 * y;
}

struct Path {
    int px, py;
    Ray ray;
    V3 throughput, contribution, lastPosition, viewDir;
    float bxdfPdf0, bxdfPdf1, neePdf;
};

inline V3 ld3(const float* p) { return v3(p[0], p[1], p[2]); }

// InitRay, Iterative.comp:174-210 (the atomic ray counter is the caller's loop variable)
void init_ray(const OrCamera& cam, const OrConfig& cfg, float totalLightArea, int rayIndex, Rng& rng, Path& p) {
    ray_index_to_morton(rayIndex, cfg.width, p.px, p.py);
    float jx = rng.next(), jy = rng.next();
    float s = ((float)p.px + jx) / (float)cfg.width;
    float t = ((float)p.py + jy) / (float)cfg.height;
    float rx = rng.next(), ry = rng.next();
    float phi = 2.0f * kPi * ry;
    float rdx = cam.lens_radius * rx * o_cos(phi);     // `lens_radius * r * vec2(cos, sin)`: r, not sqrt(r.x)
    float rdy = cam.lens_radius * ry * o_sin(phi);
    V3 offset = ld3(cam.u) * rdx + ld3(cam.v) * rdy;
    p.ray.origin = ld3(cam.position) + offset;
    p.ray.direction = normalize(ld3(cam.lower_left) + s * ld3(cam.horizontal) + t * ld3(cam.vertical) - offset);
    p.throughput = v3(1.0f);
    p.contribution = v3(0.0f);
    p.bxdfPdf0 = 1.0f; p.bxdfPdf1 = 1.0f;
    p.neePdf = 0.5f / totalLightArea;
    p.viewDir = -p.ray.direction;
    p.lastPosition = p.ray.origin;
}

// One path = the body of main()'s while(true) until the path is stored. Iterative.comp:225-303
void trace_path(const OrScene& s, const OrCamera& cam, const OrConfig& cfg, int rayIndex, uint32_t* rngTable,
                float* accum, Cnt& cnt, Cnt& scnt, uint64_t& nClosest, uint64_t& nShadow) {
    Path p;
    ray_index_to_morton(rayIndex, cfg.width, p.px, p.py);
    size_t pix = (size_t)p.py * cfg.width + p.px;
    Rng rng;                                     // initRNG, Random.glsl:232-237
    std::memcpy(&rng, rngTable + 4 * pix, 16);
    init_ray(cam, cfg, s.total_light_area, rayIndex, rng, p);
    const int model = cfg.microfacet_model;
    const int lastBounce = cfg.max_depth - 2;    // reference: terminate when i > 64  <=> max_depth 66
    int i = 0;
    while (true) {
        Hit hit; hit.t = 1e20f; hit.u = 0; hit.v = 0; hit.tri = -1;
        nClosest++;
        bool miss = !closest_hit(s, p.ray, hit, cnt);
        if (miss) {
            // Iterative.comp:251-255,260-265: sky radiance is NOT MIS-weighted (throughput as is)
            p.contribution = p.contribution + p.throughput * sample_environment(s, p.ray.direction);
            break;
        }
        // GetInterpolatedVertex, Geometry.glsl:315-337 + Iterative.comp:236-240
        const float* tri = s.tris + 20 * (size_t)hit.tri;
        float w = 1.0f - hit.u - hit.v;
        V3 vpos = p.ray.origin + p.ray.direction * hit.t;
        V3 vnormal = v3(tri[15], tri[16], tri[17]);
        float tu = tri[11] * hit.u + tri[13] * hit.v + tri[9] * w;
        float tv = tri[12] * hit.u + tri[14] * hit.v + tri[10] * w;
        uint32_t matId; std::memcpy(&matId, tri + 18, 4);
        vpos = vpos + 0.003f * vnormal;

        Material mat = make_material(s, matId, tu, tv);
        Interaction it = make_interaction(vnormal, p.viewDir);

        V3 idt = p.throughput;
        if (i != 0) {                            // Iterative.comp:256-261
            float dist = distance(p.lastPosition, vpos);
            float oldMis = mis_weight(p.bxdfPdf0, p.bxdfPdf1);
            p.bxdfPdf0 *= 0.5f * avdot(vnormal, p.viewDir) / (dist * dist);
            p.bxdfPdf1 *= 0.5f * avdot(vnormal, p.viewDir) / (dist * dist);
            idt = (p.throughput / oldMis) * p.bxdfPdf0 / (p.bxdfPdf0 + p.bxdfPdf1 + p.neePdf);
        }
        p.contribution = p.contribution + idt * mat.emission;

        // NEE, Iterative.comp:272-280
        float tmax; V3 lightThroughput; uint32_t lightMat;
        Ray lightRay = light_sample(s, cfg, vpos, vnormal, mat, it, rng, tmax, p.throughput, lightThroughput, lightMat);
        Hit lh; lh.t = tmax; lh.u = lh.v = 0; lh.tri = -1;
        nShadow++;
        {
            // bookkeeping only (the reference always traces the ray): samples whose radiance is exactly zero whatever the
            // visibility (light below the horizon: n.i = 0) -- the CUDA path does not trace those
            const float* lm0 = s.materials + 12 * (size_t)(lightMat >> 1);
            V3 L = lightThroughput * v3(lm0[4], lm0[5], lm0[6]);
            if (L.x == 0.0f && L.y == 0.0f && L.z == 0.0f) scnt.nulls++;
        }
        if (!any_hit(s, lightRay, lh, scnt)) {
            const float* lm = s.materials + 12 * (size_t)(lightMat >> 1);
            p.contribution = p.contribution + lightThroughput * v3(lm[4], lm[5], lm[6]);
        }

        // BSDF sample + throughput, Iterative.comp:282-288
        p.ray.direction = importance_sample(model, mat, it, rng, p.bxdfPdf0, p.bxdfPdf1);
        p.throughput = p.throughput * (bsdf(model, mat, it) * it.ndi * mis_weight(p.bxdfPdf0, p.bxdfPdf1) / p.bxdfPdf0);
        p.viewDir = -p.ray.direction;
        p.lastPosition = vpos;
        p.ray.origin = vpos;

        // Russian roulette, Iterative.comp:290-301
        float continuation = gmin(gmax(p.throughput.x, gmax(p.throughput.y, p.throughput.z)), 1.0f);
        p.throughput = p.throughput / continuation;
        if (rng.next() > continuation || i > lastBounce) break;
        i++;
    }
    // imageStore(accum, pixel, vec4(load.xyz + contribution, 1)), Iterative.comp:263,295; freeRNG
    float* a = accum + 4 * pix;
    a[0] = a[0] + p.contribution.x; a[1] = a[1] + p.contribution.y; a[2] = a[2] + p.contribution.z; a[3] = 1.0f;
    std::memcpy(rngTable + 4 * pix, &rng, 16);
}

}  // namespace

// ================================================================= C entry points
extern "C" {

// Probe of the shading arithmetic on explicit inputs, in the order Iterative.comp:236-288 uses it (material from the two
// texture reads, interaction, NEE evaluation towards `light_dir`, BSDF sample, throughput factor). in: 16 floats per case
// {data0.rgb, data1.g, data1.b, normal.xyz, outgoing.xyz, light_dir.xyz, -, -} + 4 RNG words; out: 20 floats
// {bsdf_light.rgb, pdf_light, ndi_light, sample_dir.xyz, pdfSample, pdfMIS, bsdf_sample.rgb, ndi_sample, mis, factor.rgb, pmf, -}
// and the RNG words after the sample. model: 0 GGX (the shader's macros as shipped), 1 Beckmann.
void or_shade_probe(const float* in, uint32_t* rng_io, int64_t n, int model, float* out) {
    for (int64_t k = 0; k < n; k++) {
        const float* a = in + 16 * k; float* o = out + 20 * k; uint32_t* rs = rng_io + 4 * k;
        Material m{};
        m.albedo = v3(a[0], a[1], a[2]);
        m.roughness = gmax(a[3] * a[3], 1e-4f);
        m.roughness2 = m.roughness * m.roughness;
        m.metallic = a[4];
        m.reflectance = mix(v3(0.04f), m.albedo, m.metallic);
        V3 nrm = v3(a[5], a[6], a[7]), og = v3(a[8], a[9], a[10]), ld = v3(a[11], a[12], a[13]);
        Interaction it = make_interaction(nrm, og);
        Interaction li = it;
        set_incoming(li, ld);
        V3 fl = bsdf(model, m, li);
        o[0] = fl.x; o[1] = fl.y; o[2] = fl.z; o[3] = pdf_direction(model, m, li); o[4] = li.ndi;
        Rng rng; rng.x = rs[0]; rng.y = rs[1]; rng.z = rs[2]; rng.w = rs[3];
        float p0, p1;
        V3 d = importance_sample(model, m, it, rng, p0, p1);
        V3 fs = bsdf(model, m, it);
        float w = mis_weight(p0, p1);
        V3 factor = fs * it.ndi * w / p0;                                // Iterative.comp:284
        o[5] = d.x; o[6] = d.y; o[7] = d.z; o[8] = p0; o[9] = p1; o[10] = fs.x; o[11] = fs.y; o[12] = fs.z; o[13] = it.ndi; o[14] = w;
        o[15] = factor.x; o[16] = factor.y; o[17] = factor.z; o[18] = diffuse_pmf(m, it); o[19] = 0.0f;
        rs[0] = rng.x; rs[1] = rng.y; rs[2] = rng.z; rs[3] = rng.w;
    }
}

// GenerateLightSample (Iterative.comp:113-147) on explicit inputs, as main() calls it (:276). in: n x 20 floats {vertex
// position, vertex normal, outgoing, data0.rgb, data1.g, data1.b, throughput.rgb, ...}; out: n x 12 floats {ray origin, ray
// direction, tmax, lightThroughput.rgb, light material id (bits), lightPdf}; rng_io advanced. sun_light is taken from cfg.
void or_light_probe(const OrScene* s, const OrConfig* cfg, const float* in, uint32_t* rng_io, int64_t n, float* out) {
    for (int64_t k = 0; k < n; k++) {
        const float* a = in + 20 * k; float* o = out + 12 * k; uint32_t* rs = rng_io + 4 * k;
        V3 vpos = v3(a[0], a[1], a[2]), vn = v3(a[3], a[4], a[5]), og = v3(a[6], a[7], a[8]);
        Material m{};
        m.albedo = v3(a[9], a[10], a[11]);
        m.roughness = gmax(a[12] * a[12], 1e-4f);
        m.roughness2 = m.roughness * m.roughness;
        m.metallic = a[13];
        m.reflectance = mix(v3(0.04f), m.albedo, m.metallic);
        Interaction it = make_interaction(vn, og);
        Rng rng; rng.x = rs[0]; rng.y = rs[1]; rng.z = rs[2]; rng.w = rs[3];
        float tmax; V3 lt; uint32_t lmat;
        Ray r = light_sample(*s, *cfg, vpos, vn, m, it, rng, tmax, v3(a[14], a[15], a[16]), lt, lmat);
        o[0] = r.origin.x; o[1] = r.origin.y; o[2] = r.origin.z; o[3] = r.direction.x; o[4] = r.direction.y; o[5] = r.direction.z;
        o[6] = tmax; o[7] = lt.x; o[8] = lt.y; o[9] = lt.z; std::memcpy(&o[10], &lmat, 4);
        o[11] = cfg->sun_light ? 0.0f : 1.0f / s->total_light_area;
        rs[0] = rng.x; rs[1] = rng.y; rs[2] = rng.z; rs[3] = rng.w;
    }
}

// GL_LINEAR + GL_REPEAT lookup of scene texture `index` (what `texture(sampler2D, uv)` is to the shader); used by the compiled
// shader source in oracle/ref_glsl.cpp as its texture unit, so that GL's fixed-function filtering is stated once
void or_sample_texture(const OrScene* s, int index, float u, float v, float* rgba) {
    if (index < 0 || index >= s->n_textures) { rgba[0] = rgba[1] = rgba[2] = 0.0f; rgba[3] = 1.0f; return; }
    sample_texture(s->textures[index], u, v, rgba);
}

void or_set_parity_libm(int on) { g_parity_libm = on ? 1 : 0; }
int or_get_parity_libm(void) { return g_parity_libm; }
// pt_parity_libm.h as this compiler evaluates it: out[i] = fn(x[i] [, y[i]]), fn 0 sin, 1 cos, 2 log, 3 exp, 4 pow
void or_parity_libm_probe(int fn, const float* x, const float* y, int64_t n, float* out) {
    for (int64_t i = 0; i < n; i++) {
        float a = x[i];
        out[i] = fn == 0 ? ptpl::sinf_(a) : fn == 1 ? ptpl::cosf_(a) : fn == 2 ? ptpl::logf_(a) : fn == 3 ? ptpl::expf_(a) : ptpl::powf_(a, y[i]);
    }
}
void or_hybrid_taus(uint32_t state[4], int n, float* out) {
    Rng r{state[0], state[1], state[2], state[3]};
    for (int i = 0; i < n; i++) out[i] = r.next();
    state[0] = r.x; state[1] = r.y; state[2] = r.z; state[3] = r.w;
}

// Camera::Camera + SetPosition/SetRotation + GenerateImagePlane, Camera.cpp:4-22,33-36,45-47
void or_image_plane(float aspect, float fov, float focus, float aperture, const float pos[3], const float rot[3],
                    OrCamera* cam) {
    float yaw = rot[0], pitch = rot[1];
    float lens_radius = aperture / 2;
    V3 d = v3(cosf(pitch) * sinf(yaw), sinf(pitch), cosf(pitch) * -cosf(yaw));
    d = normalize(-d);
    float image_height = 2 * tanf(fov / 2);
    float image_width = aspect * image_height;
    V3 u = normalize(cross(v3(0, 1, 0), d));
    V3 v = cross(d, u);
    V3 horizontal = image_width * u * focus;
    V3 vertical = image_height * v * focus;
    V3 lower_left = -horizontal / 2.0f - vertical / 2.0f - d * focus;
    auto st = [](float* o, V3 a) { o[0] = a.x; o[1] = a.y; o[2] = a.z; };
    st(cam->lower_left, lower_left); st(cam->horizontal, horizontal); st(cam->vertical, vertical);
    st(cam->position, v3(pos[0], pos[1], pos[2])); st(cam->u, u); st(cam->v, v);
    cam->lens_radius = lens_radius;
}

int or_slab_test(const float bmin[3], const float bmax[3], const float iorg[3], const float idir[3], float tcur,
                 float dist2[2]) {
    float node[8] = {bmin[0], bmin[1], bmin[2], 0, bmax[0], bmax[1], bmax[2], 0};
    Ray ir; ir.origin = ld3(iorg); ir.direction = ld3(idir);
    bool hit; float te = slab(node, ir, tcur, hit);
    // t_exit recomputed for the probe only
    V3 a = ld3(bmin) * ir.direction + ir.origin, b = ld3(bmax) * ir.direction + ir.origin;
    dist2[0] = te;
    dist2[1] = gmin(gmax(a.x, b.x), gmin(gmax(a.y, b.y), gmin(gmax(a.z, b.z), tcur)));
    return hit ? 1 : 0;
}

// Present.frag:13-37 with exposure = kExposure 1.68 (Renderer.cpp:18)
void or_tonemap(const float rgb_sum[3], int num_samples, float out_rgb[3]) {
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f, W = 11.2f;
    float white = ((W * (A * W + C * B) + D * E) / (W * (A * W + B) + D * F)) - E / F;
    for (int c = 0; c < 3; c++) {
        float color = rgb_sum[c] / (float)num_samples;
        color = 1.68f * color;
        color *= 2.0f;
        color = ((color * (A * color + C * B) + D * E) / (color * (A * color + B) + D * F)) - E / F;
        color /= white;
        out_rgb[c] = o_pow(color, 1.0f / 2.2f);
    }
}

void or_sample_environment(const OrScene* s, const float* dirs, int64_t n, float* rgb) {
    for (int64_t i = 0; i < n; i++) {
        V3 c = sample_environment(*s, ld3(dirs + 3 * i));
        rgb[3 * i] = c.x; rgb[3 * i + 1] = c.y; rgb[3 * i + 2] = c.z;
    }
}

// LoadEnvironmnet, .hdr branch (Renderer.cpp:262-312). Per face i the reference rasterises a unit cube with
// captureProjection = perspective(90 deg, 1, 0.1, 10) and captureViews[i] = lookAt(0, centre_i, up_i) (:286-294);
// EquirectangularConverter.vert hands the fragment sampleCoords = -position, the fragment shader
// (EquirectangularConverter.frag:9-19) reads hdrTex at (atan(v.z, v.x), asin(v.y)) * (0.1591, 0.3183) + 0.5 of the
// normalised vector. The fragment at window (px + .5, py + .5) sees the cube point f + nx*s + ny*u with the lookAt basis
// s = normalize(cross(f, up)), u = cross(s, f) (glm::lookAtRH) and nx = (px + .5)/size*2 - 1; it lands in texel
// (px, py) of face i. hdrTex: GL_LINEAR, CLAMP_TO_EDGE (:269-272), row 0 of the stbi image at t = 0. The cubemap's
// storage is RGB32F (:282): alpha reads back as 1.
void or_equirect_to_cube(const float* rgba, int width, int height, int size, float* faces) {
    const float centre[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    const float upv[6][3] = {{0, -1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}, {0, -1, 0}, {0, -1, 0}};
    parallel_for((int64_t)6 * size, 0, [&](int, int64_t b, int64_t e) {
        for (int64_t row = b; row < e; row++) {
            int face = (int)(row / size), py = (int)(row % size);
            V3 f = ld3(centre[face]), up = ld3(upv[face]);
            V3 sv = cross(f, up), uv = cross(sv, f);
            for (int px = 0; px < size; px++) {
                float nx = ((float)px + 0.5f) / (float)size * 2.0f - 1.0f, ny = ((float)py + 0.5f) / (float)size * 2.0f - 1.0f;
                V3 p = f + sv * nx + uv * ny;
                V3 d = normalize(-p);
                if (d.z == 0.0f) d.z = 0.0f;      // -0 -> +0, see k_equirect_to_cube (GLSL leaves atan(-0, x < 0) open)
                float tu = atan2f(d.z, d.x) * 0.1591f + 0.5f, tv = asinf(d.y) * 0.3183f + 0.5f;
                float out[4];
                sample_clamped(rgba, width, height, tu, tv, out);
                float* o = faces + 4 * (((size_t)face * size + py) * size + px);
                o[0] = out[0]; o[1] = out[1]; o[2] = out[2]; o[3] = 1.0f;
            }
        }
    });
}

// the fragment shader's part of the above on explicit sampleCoords (= -position): n x 3 -> n x 2 texture coordinates
void or_equirect_uv(const float* sample_coords, int64_t n, float* uv) {
    for (int64_t i = 0; i < n; i++) {
        V3 d = normalize(ld3(sample_coords + 3 * i));
        if (d.z == 0.0f) d.z = 0.0f;
        uv[2 * i] = atan2f(d.z, d.x) * 0.1591f + 0.5f; uv[2 * i + 1] = asinf(d.y) * 0.3183f + 0.5f;
    }
}

static OrCounters g_dummy;

void or_trace_closest(const OrScene* s, const float* rays, int64_t n, float* hits, OrCounters* c, int n_threads) {
    std::atomic_flag lock = ATOMIC_FLAG_INIT;
    parallel_for(n, n_threads, [&](int, int64_t b, int64_t e) {
        Cnt k;
        for (int64_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            Ray ray; ray.origin = ld3(r); ray.direction = ld3(r + 4);
            Hit h; h.t = r[7]; h.u = 0; h.v = 0; h.tri = -1;
            closest_hit(*s, ray, h, k);
            float* o = hits + 4 * i;
            o[0] = h.t; o[1] = h.u; o[2] = h.v; o[3] = sbf(h.tri);
        }
        while (lock.test_and_set()) {}
        add_counters(c, k, (uint64_t)(e - b), 0, 0);
        lock.clear();
    });
}

void or_trace_any(const OrScene* s, const float* rays, int64_t n, uint8_t* flags, OrCounters* c, int n_threads) {
    std::atomic_flag lock = ATOMIC_FLAG_INIT;
    parallel_for(n, n_threads, [&](int, int64_t b, int64_t e) {
        Cnt k;
        for (int64_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            Ray ray; ray.origin = ld3(r); ray.direction = ld3(r + 4);
            Hit h; h.t = r[7]; h.u = 0; h.v = 0; h.tri = -1;
            flags[i] = any_hit(*s, ray, h, k) ? 1 : 0;
        }
        while (lock.test_and_set()) {}
        add_counters(c, k, 0, (uint64_t)(e - b), 0, &k);
        lock.clear();
    });
}

// Per-ray operation trace for scheduling studies (tools/simulate_warp.py): ops = concatenated step codes
// (0 = sibling-pair test, 1 = triangle test, 2 = leaf entered), offsets[n+1] = start of each ray's steps. Returns total steps
// (call with ops == NULL to size the buffer).
int64_t or_trace_ops(const OrScene* s, const float* rays, int64_t n, int any, uint8_t* ops, int64_t* offsets) {
    std::vector<uint8_t> v;
    Cnt k; k.ops = &v;
    int64_t total = 0;
    for (int64_t i = 0; i < n; i++) {
        const float* r = rays + 8 * i;
        Ray ray; ray.origin = ld3(r); ray.direction = ld3(r + 4);
        Hit h; h.t = r[7]; h.u = 0; h.v = 0; h.tri = -1;
        v.clear();
        if (any) any_hit(*s, ray, h, k); else closest_hit(*s, ray, h, k);
        if (offsets) offsets[i] = total;
        if (ops) std::memcpy(ops + total, v.data(), v.size());
        total += (int64_t)v.size();
    }
    if (offsets) offsets[n] = total;
    return total;
}

void or_trace_closest_bruteforce(const OrScene* s, const float* rays, int64_t n, float* hits, int n_threads) {
    parallel_for(n, n_threads, [&](int, int64_t b, int64_t e) {
        for (int64_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            Ray ray; ray.origin = ld3(r); ray.direction = ld3(r + 4);
            Hit h; h.t = r[7]; h.u = 0; h.v = 0; h.tri = -1;
            for (int t = 0; t < s->n_tris; t++) tri_mt(s->tris + 20 * (size_t)t, ray, h, t);
            float* o = hits + 4 * i;
            o[0] = h.t; o[1] = h.u; o[2] = h.v; o[3] = sbf(h.tri);
        }
    });
}

void or_render(const OrScene* s, const OrCamera* cam, const OrConfig* cfg, uint32_t* rng, float* accum, int n_frames,
               OrCounters* c) {
    const int64_t npx = (int64_t)cfg->width * cfg->height;
    std::atomic_flag lock = ATOMIC_FLAG_INIT;
    for (int f = 0; f < n_frames; f++) {
        // one dispatch: globalRayCount = width*height paths, one per pixel (Iterative.comp:149,175-178)
        parallel_for(npx, cfg->n_threads, [&](int, int64_t b, int64_t e) {
            Cnt k, ks; uint64_t nc = 0, ns = 0;
            for (int64_t i = b; i < e; i++) trace_path(*s, *cam, *cfg, (int)i, rng, accum, k, ks, nc, ns);
            k.pairs += ks.pairs; k.leaves += ks.leaves; k.tris += ks.tris;
            while (lock.test_and_set()) {}
            add_counters(c, k, nc, ns, (uint64_t)(e - b), &ks);
            lock.clear();
        });
    }
}

void or_raygen(const OrCamera* cam, const OrConfig* cfg, const float total_light_area, uint32_t* rng, int64_t first,
               int64_t n, float* out) {
    for (int64_t i = 0; i < n; i++) {
        Path p;
        int px, py; ray_index_to_morton((int)(first + i), cfg->width, px, py);
        size_t pix = (size_t)py * cfg->width + px;
        Rng r; std::memcpy(&r, rng + 4 * pix, 16);
        init_ray(*cam, *cfg, total_light_area, (int)(first + i), r, p);
        std::memcpy(rng + 4 * pix, &r, 16);
        float* o = out + 12 * i;
        o[0] = sbf(p.px); o[1] = sbf(p.py);
        o[2] = p.ray.origin.x; o[3] = p.ray.origin.y; o[4] = p.ray.origin.z;
        o[5] = p.ray.direction.x; o[6] = p.ray.direction.y; o[7] = p.ray.direction.z;
        o[8] = o[9] = o[10] = o[11] = 0.0f;
    }
}

}  // extern "C"
