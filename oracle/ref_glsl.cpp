// oracle/ref_glsl.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Harness around the reference's OWN shader source for the traversal path, compiled as C++ where it lies under
// /root/reference: oracle/Makefile extracts, at build time and into the git-ignored oracle/_ref/, these line ranges
//   src/shaders/common/Util.glsl      4-5      fbu / fbs
//   src/shaders/common/Geometry.glsl  8-11     Ray;  145-161 PackedCompactTriangle + ReadPackedCompactTriangle;
//                                     193-226  HitInfo + IntersectTriangleMT;  313 #define IntersectTriangle
//   src/shaders/common/BVH.glsl       300-302  BVHNode;  356-364 GetNode;  379-391 IntersectNodeFast;  420-437 IntersectLeaf;
//                                     622-627  the ifif macros;  634-767 IfIfClosestHit;  769-854 IfIfAnyHit
//   and, for the shading arithmetic, Constants.glsl 4; Util.glsl 6-11; Random.glsl 42-45, 47-75, 77-79, 155; MIS.glsl 6-8;
//   Material.glsl 23-37, 46-51, 57-130; Microfacet.glsl 9-193; BSDF.glsl 8-21
//   for the path itself, src/shaders/Iterative.comp 20-22, 24-25, 32, 34-82 (RandomLightVertex), 89-110 (the sun disk),
//   113-147 (GenerateLightSample), 154-172 (RayIndexToMorton), 174-210 (InitRay), 214-307 (main) with Geometry.glsl 92-98,
//   130-143, 163-182, 315-337 (GetInterpolatedVertex) and Material.glsl 39-54 (ConstructMaterialInstance)
//   and the two shaders either side of it: Present.frag 8-9, 11, 13-27, 29-37; EquirectangularConverter.frag 9-16
// and applies syntactic rewrites only (parameter qualifiers, swizzle spelling, ^^, float literal suffixes, two implicit
// conversions; one literal, the path-length bound of `i > 64`, becomes a variable); see the Makefile. Nothing of the
// reference is copied into the repository: the extracts are make intermediates, deleted once this file is compiled. This file supplies only what the GL runtime supplies to a shader: the buffer
// textures (samplerBuffer + texelFetch over plain float arrays), glm as the vector library (the reference's own, extern/glm)
// and entry points that loop over a ray batch; for main() also the material SSBO, bindless handles, the accumulation image,
// the RNG table and the shared ray counter (ref_glsl_render_frames_mp: one process per core, what bench.py's reference arm
// times). GL's fixed-function texture filtering is NOT here: the texture units call back into the oracle library. It exists to PIN oracle/pt_oracle.cpp's hand restatement of the same
// functions: tests/test_oracle_pins.py requires both to agree bit for bit.
//
// Two things the shader leaves open and this harness fixes, both outside the arithmetic:
//   * BVH_STACK_SIZE is 27 in the shader (BVH.glsl:489) and deeper trees overrun its array; here 96, like the product;
//   * the shader never stores the triangle INDEX of a hit (it copies the 80-byte triangle into HitInfo.intersected). The
//     harness feeds it a vertex buffer whose padding word (float 19 of each triangle, "uint padding", Geometry.glsl:142) holds
//     the triangle's index, so the accepted triangle names itself. No traversal arithmetic reads that word.
#define GLM_FORCE_SWIZZLE
#include <glm/glm.hpp>
#include "pt_parity_libm.h"     // gpupathtracer_b200/csrc: used only by the *_pl namespaces below (model codes 4 and 6)
#include <atomic>
#include <cstdint>
#include <cmath>
#include <sched.h>
#include <signal.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <unistd.h>
#include <cstring>
#include <thread>
#include <vector>

namespace refglsl {
using namespace glm;

// what `uniform samplerBuffer` + texelFetch are to the shader: RGBA32F texels (nodes, triangles) or R32F (reference list)
struct samplerBuffer { const float* p = nullptr; int comps = 4; int count = 0; };
inline vec4 texelFetch(const samplerBuffer& s, int i) {
    if (s.comps == 4) return vec4(s.p[4 * (size_t)i], s.p[4 * (size_t)i + 1], s.p[4 * (size_t)i + 2], s.p[4 * (size_t)i + 3]);
    if (s.comps == 2) return vec4(s.p[2 * (size_t)i], s.p[2 * (size_t)i + 1], 0.0f, 1.0f);     // RG32F: the light table
    return vec4(s.p[i], 0.0f, 0.0f, 1.0f);                       // GL fills missing components with (0, 0, 1)
}
inline int textureSize(const samplerBuffer& s) { return s.count; }
static samplerBuffer vertexTex, nodesTex, referenceTex;           // Geometry.glsl:6, BVH.glsl:353-354

#define BVH_STACK_SIZE 96
#include "ref_glsl_traversal.inc"

// ---- the rest of the GL runtime main() sees: bindless texture handles, the material SSBO, the accumulation image.
// A handle is an index into g_units; a unit is a scene texture (filtered by the oracle library's statement of GL_LINEAR /
// GL_REPEAT through a function pointer, so that fixed-function behaviour is stated once) or a constant colour, which is what
// the reference's 1 x 1 RGBA32F colour textures return for any coordinate (Texture.cpp:146-161). The environment cube map goes
// through the oracle library's lookup the same way (constant colour or six faces).
struct TexUnit { int scene_texture; vec4 constant; };
static std::vector<TexUnit> g_units;
typedef void (*tex2d_fn)(const void* scene, int index, float u, float v, float* rgba);        // or_sample_texture
typedef void (*cube_fn)(const void* scene, const float* dirs, int64_t n, float* rgb);          // or_sample_environment
static tex2d_fn g_tex2d = nullptr;
static cube_fn g_cube = nullptr;
static const void* g_gl_scene = nullptr;
struct sampler2D { uvec2 h; explicit sampler2D(uvec2 x) : h(x) {} };
struct samplerCube { uvec2 h; explicit samplerCube(uvec2 x) : h(x) {} };
inline vec4 texture(const sampler2D& s, vec2 uv) {
    const TexUnit& t = g_units[s.h.x];
    if (t.scene_texture < 0) return t.constant;
    if (!std::isfinite(uv.x) || !std::isfinite(uv.y)) return t.constant;   // main() looks material 0 up with a missed ray's unset hit record
    float o[4]; g_tex2d(g_gl_scene, t.scene_texture, uv.x, uv.y, o);
    return vec4(o[0], o[1], o[2], o[3]);
}
inline vec4 texture(const samplerCube&, vec3 d) {
    float in[3] = {d.x, d.y, d.z}, o[3]; g_cube(g_gl_scene, in, 1, o);
    return vec4(o[0], o[1], o[2], 1.0f);
}
static std::vector<vec4> g_material_ssbo;                          // `materialInstance[]`, Material.glsl:10-12 / Scene.cpp:296-299
struct image2D { float* p = nullptr; int w = 0; };
inline vec4 imageLoad(const image2D& im, ivec2 px) { const float* q = im.p + 4 * ((size_t)px.y * im.w + px.x); return vec4(q[0], q[1], q[2], q[3]); }
inline void imageStore(const image2D& im, ivec2 px, vec4 v) { float* q = im.p + 4 * ((size_t)px.y * im.w + px.x); q[0] = v.x; q[1] = v.y; q[2] = v.z; q[3] = v.w; }
static uvec3 gl_LocalInvocationID(0u);

}  // namespace refglsl

// ---- shading arithmetic + ray generation + light sampling, once per microfacet model (oracle/ref_glsl_ns.h): the extracted
// source as shipped (GGX) and the copy with Microfacet.glsl's three model macros pointed at its Beckmann functions
namespace refglsl_ggx {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_ggx
namespace refglsl_beckmann {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading_beckmann.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_beckmann
// ... and both again as the SUN_LIGHT build (GenerateLightSample's coin flip between the emitters and the sun disk)
#define REFGLSL_SUN
namespace refglsl_ggx_sun {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_ggx_sun
namespace refglsl_beckmann_sun {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading_beckmann.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_beckmann_sun
#undef REFGLSL_SUN
// ... and the Beckmann copy once more with scalar min / max = IEEE minNum / maxNum (see ref_glsl_ns.h): model code 2
#define REFGLSL_MAXNUM
namespace refglsl_beckmann_maxnum {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading_beckmann.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_beckmann_maxnum
#undef REFGLSL_MAXNUM

// ... and GGX / Beckmann-maxNum once more with the scalar transcendentals of pt_parity_libm.h (see ref_glsl_ns.h): model codes 4
// and 6 (= 0 | 4, 2 | 4), the builds whose frames the CUDA path must reproduce BIT for bit under PT_PARITY_LIBM
#define REFGLSL_PLIBM
namespace refglsl_ggx_pl {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_ggx_pl
#define REFGLSL_MAXNUM
namespace refglsl_beckmann_maxnum_pl {
using namespace glm;
using namespace refglsl;
#define REFGLSL_SHADING "ref_glsl_shading_beckmann.inc"
#include "ref_glsl_ns.h"
#undef REFGLSL_SHADING
}  // namespace refglsl_beckmann_maxnum_pl
#undef REFGLSL_MAXNUM
#undef REFGLSL_PLIBM

// The present pass (Present.frag): division by the sample count, exposure, Uncharted-2 curve, gamma. The harness is the
// texture unit behind `directAccum`, the fragment coordinate and the output variable.
namespace refglsl_present {
using namespace glm;
struct AccumTex { const float* p = nullptr; int w = 0; };
static AccumTex directAccum;                                        // Present.frag:7
inline vec4 texelFetch(const AccumTex& t, ivec2 px, int) { const float* q = t.p + 4 * ((size_t)px.y * t.w + px.x); return vec4(q[0], q[1], q[2], q[3]); }
static vec4 gl_FragCoord;
static vec3 color;                                                  // Present.frag:5
#define uniform
#include "ref_glsl_present.inc"
#undef uniform
}  // namespace refglsl_present

namespace refglsl_present_pl {                                     // the same with pow() from pt_parity_libm.h, component by component
using namespace glm;
struct AccumTex { const float* p = nullptr; int w = 0; };
static AccumTex directAccum;
inline vec4 texelFetch(const AccumTex& t, ivec2 px, int) { const float* q = t.p + 4 * ((size_t)px.y * t.w + px.x); return vec4(q[0], q[1], q[2], q[3]); }
static vec4 gl_FragCoord;
static vec3 color;
inline vec3 pow(vec3 x, vec3 y) { return vec3(ptpl::powf_(x.x, y.x), ptpl::powf_(x.y, y.y), ptpl::powf_(x.z, y.z)); }
#define uniform
#include "ref_glsl_present.inc"
#undef uniform
}  // namespace refglsl_present_pl

namespace refglsl_equirect {
using namespace glm;
#include "ref_glsl_equirect.inc"
}  // namespace refglsl_equirect

namespace {
template <class F>
void parallel_for(int64_t n, int n_threads, F f) {
    int nt = n_threads > 0 ? n_threads : (int)std::thread::hardware_concurrency();
    if (nt < 1 || n < 4096) nt = 1;
    std::atomic<int64_t> next(0);
    auto work = [&]() { for (;;) { int64_t b = next.fetch_add(2048); if (b >= n) break; f(b, std::min(n, b + 2048)); } };
    if (nt == 1) { work(); return; }
    std::vector<std::thread> th;
    for (int i = 0; i < nt; i++) th.emplace_back(work);
    for (auto& t : th) t.join();
}
std::vector<float> g_stamped;                                      // triangles with their index in the padding word

// One compile of the shader path per model code: 0 GGX as shipped, 1 Beckmann (Microfacet.glsl's macros), 2 Beckmann with scalar
// min / max = IEEE minNum / maxNum, 4 / 6 = 0 / 2 with the transcendentals of pt_parity_libm.h.
struct ModelBuild {
    void (*shade_probe)(const float*, uint32_t*, int64_t, float*);
    void (*light_probe)(const float*, uint32_t*, int64_t, float*);
    void (*bind_lights)(const float*, int, float);
    void (*render_frame)(const float*, int, int, float, uint32_t*, float*);
    void (*render_frame_shared)(const float*, int, int, float, uint32_t*, float*, glm::uint*);
    void (*raygen_probe)(const float*, int, int, float, uint32_t*, int64_t, int64_t, float*);
    const glm::vec4** material_instance;
    int* max_path_index;
};
#define REFGLSL_BUILD(ns) ModelBuild{ns::shade_probe, ns::light_probe, ns::bind_lights, ns::render_frame, ns::render_frame_shared, \
                                     ns::raygen_probe, &ns::materialInstance, &ns::maxPathIndex}
ModelBuild g_builds[] = {REFGLSL_BUILD(refglsl_ggx), REFGLSL_BUILD(refglsl_beckmann), REFGLSL_BUILD(refglsl_beckmann_maxnum),
                         REFGLSL_BUILD(refglsl_ggx_pl), REFGLSL_BUILD(refglsl_beckmann_maxnum_pl),
                         REFGLSL_BUILD(refglsl_ggx_sun), REFGLSL_BUILD(refglsl_beckmann_sun)};
const int kBuilds = (int)(sizeof(g_builds) / sizeof(g_builds[0]));
ModelBuild& build_of(int model) {
    switch (model) { case 1: return g_builds[1]; case 2: return g_builds[2]; case 4: return g_builds[3]; case 6: return g_builds[4]; default: return g_builds[0]; }
}
}  // namespace

extern "C" {

// nodes: n_nodes x 8 floats (NodeSerialized, NewLayout), refs: int32 reference list, tris: n_tris x 20 floats (pre-subtracted
// edges, as uploaded by Scene.cpp:334-347). Not re-entrant (the samplers are globals, as uniforms are).
void ref_glsl_bind(const float* nodes, const int32_t* refs, const float* tris, int64_t n_tris) {
    g_stamped.assign(tris, tris + 20 * (size_t)n_tris);
    for (int64_t i = 0; i < n_tris; i++) { int32_t id = (int32_t)i; std::memcpy(&g_stamped[20 * (size_t)i + 19], &id, 4); }
    refglsl::nodesTex.p = nodes; refglsl::nodesTex.comps = 4;
    refglsl::referenceTex.p = reinterpret_cast<const float*>(refs); refglsl::referenceTex.comps = 1;
    refglsl::vertexTex.p = g_stamped.data(); refglsl::vertexTex.comps = 4;
}

// rays: n x 8 floats {o.xyz, -, d.xyz, tmax}; hits: n x 4 floats {t, u, v, triangle index as int bits (-1 = miss)}.
// The caller of IfIfClosestHit in the shader starts with intersection.di.x = the ray's maximum distance (Iterative.comp:232).
void ref_glsl_trace_closest(const float* rays, int64_t n, float* hits, int n_threads) {
    parallel_for(n, n_threads, [&](int64_t b, int64_t e) {
        for (int64_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            refglsl::Ray ray; ray.origin = glm::vec3(r[0], r[1], r[2]); ray.direction = glm::vec3(r[4], r[5], r[6]);
            refglsl::HitInfo h; h.di = glm::vec4(r[7], 0.0f, 0.0f, 0.0f);
            bool any = refglsl::IfIfClosestHit(ray, h);
            int32_t id = -1;
            if (any) std::memcpy(&id, &h.intersected.data[4].w, 4);
            hits[4 * i] = h.di.x; hits[4 * i + 1] = h.di.y; hits[4 * i + 2] = h.di.z; std::memcpy(&hits[4 * i + 3], &id, 4);
        }
    });
}

void ref_glsl_trace_any(const float* rays, int64_t n, uint8_t* flags, int n_threads) {
    parallel_for(n, n_threads, [&](int64_t b, int64_t e) {
        for (int64_t i = b; i < e; i++) {
            const float* r = rays + 8 * i;
            refglsl::Ray ray; ray.origin = glm::vec3(r[0], r[1], r[2]); ray.direction = glm::vec3(r[4], r[5], r[6]);
            refglsl::HitInfo h; h.di = glm::vec4(r[7], 0.0f, 0.0f, 0.0f);
            flags[i] = refglsl::IfIfAnyHit(ray, h) ? 1 : 0;
        }
    });
}

// The probes of oracle/ref_glsl_ns.h; model 0 = GGX (the macros as shipped), 1 = Beckmann. Single-threaded: `state` is a global.
void ref_glsl_shade_probe(const float* in, uint32_t* rng_io, int64_t n, float* out) { refglsl_ggx::shade_probe(in, rng_io, n, out); }
void ref_glsl_shade_probe_beckmann(const float* in, uint32_t* rng_io, int64_t n, float* out) { refglsl_beckmann::shade_probe(in, rng_io, n, out); }
void ref_glsl_shade_probe_beckmann_maxnum(const float* in, uint32_t* rng_io, int64_t n, float* out) { refglsl_beckmann_maxnum::shade_probe(in, rng_io, n, out); }
void ref_glsl_shade_probe_model(const float* in, uint32_t* rng_io, int64_t n, int model, float* out) { build_of(model).shade_probe(in, rng_io, n, out); }
void ref_glsl_bind_lights(const float* lights2, int n_lights, float total_light_area) {
    for (int k = 0; k < kBuilds; k++) g_builds[k].bind_lights(lights2, n_lights, total_light_area);
}
void ref_glsl_light_probe(const float* in, uint32_t* rng_io, int64_t n, int model, float* out) { build_of(model).light_probe(in, rng_io, n, out); }
// materials: n x 12 floats as OrScene has them {colour.rgb, albedo texture index (int bits, -1 = none), emission.rgb, -,
// props.rgb, props texture index}; laid out for the shader as Scene.cpp:296-299 does: [albedo handle, props handle][emission, flag]
void ref_glsl_bind_materials(const float* materials12, int n, const void* gl_scene, refglsl::tex2d_fn tex2d, refglsl::cube_fn cube) {
    using namespace refglsl;
    g_units.clear(); g_material_ssbo.clear();
    g_tex2d = tex2d; g_cube = cube; g_gl_scene = gl_scene;
    for (int k = 0; k < n; k++) {
        const float* m = materials12 + 12 * (size_t)k;
        int at, pt; std::memcpy(&at, m + 3, 4); std::memcpy(&pt, m + 11, 4);
        g_units.push_back(TexUnit{at, glm::vec4(m[0], m[1], m[2], 1.0f)});
        g_units.push_back(TexUnit{pt, glm::vec4(m[8], m[9], m[10], 1.0f)});
        glm::uvec4 handles(2u * k, 0u, 2u * k + 1u, 0u);
        g_material_ssbo.push_back(glm::uintBitsToFloat(handles));
        g_material_ssbo.push_back(glm::vec4(m[4], m[5], m[6], m[7]));
    }
    for (int k = 0; k < kBuilds; k++) *g_builds[k].material_instance = g_material_ssbo.data();
}

// One dispatch of the reference's main() (Iterative.comp:214-307) over a whole frame: every pixel's path, accumulated into
// `accum` (h x w x 4 floats, `imageStore`), RNG words written back by freeRNG. One invocation loops over all rays, which the
// persistent shader allows (any number of invocations share the ray counter). Scene: ref_glsl_bind + _bind_lights + _bind_materials.
void ref_glsl_render_frame(const float* cam19, int w, int h, float total_light_area, uint32_t* table, float* accum, int model) {
    build_of(model).render_frame(cam19, w, h, total_light_area, table, accum);
}
// n_frames dispatches of main() by n_procs processes sharing the ray counter, as the GPU's invocations share the SSBO word:
// image, RNG table, one counter and one arrival word per frame live in a MAP_SHARED mapping, the children are forked once,
// each drains frame f's counter and waits for the others before frame f + 1 (a pixel's next frame reads what any process wrote).
// max_depth as pt_config has it (66 = the shader as shipped). Returns 0, or -1 when a child did not exit cleanly.
int ref_glsl_render_frames_mp(const float* cam19, int w, int h, float total_light_area, uint32_t* table, float* accum, int model,
                              int max_depth, int n_frames, int n_procs) {
    const size_t px = (size_t)w * h, words = 2 * (size_t)n_frames;
    const size_t bytes = ((words * 4 + 63) / 64) * 64 + px * 16 * 2;
    void* m = mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (m == MAP_FAILED) return -1;
    uint32_t* ctr = (uint32_t*)m;
    uint32_t* arrive = ctr + n_frames;
    uint32_t* stable = (uint32_t*)((char*)m + ((words * 4 + 63) / 64) * 64);
    float* simage = (float*)(stable + px * 4);
    std::memset(ctr, 0, words * 4);
    std::memcpy(stable, table, px * 16); std::memcpy(simage, accum, px * 16);
    ModelBuild& mb = build_of(model);
    *mb.max_path_index = max_depth - 2;
    if (n_procs < 1) n_procs = 1;
    std::vector<pid_t> kids;
    for (int k = 0; k < n_procs; k++) {
        pid_t pid = fork();
        if (pid == 0) {
            for (int f = 0; f < n_frames; f++) {
                mb.render_frame_shared(cam19, w, h, total_light_area, stable, simage, ctr + f);
                __atomic_fetch_add(arrive + f, 1u, __ATOMIC_ACQ_REL);
                while (__atomic_load_n(arrive + f, __ATOMIC_ACQUIRE) < (uint32_t)n_procs) sched_yield();
            }
            _exit(0);
        }
        if (pid < 0) break;
        kids.push_back(pid);
    }
    // a child that could not be started or died leaves the others waiting at the frame barrier: stop them all and report it
    int rc = (int)kids.size() == n_procs ? 0 : -1;
    size_t left = kids.size();
    std::vector<bool> done(kids.size(), false);
    while (left) {
        if (rc != 0) for (size_t k = 0; k < kids.size(); k++) if (!done[k]) kill(kids[k], SIGKILL);
        for (size_t k = 0; k < kids.size(); k++) {
            if (done[k]) continue;
            int st = 0;
            pid_t r = waitpid(kids[k], &st, rc != 0 ? 0 : WNOHANG);
            if (r == 0) continue;
            done[k] = true; left--;
            if (r < 0 || !WIFEXITED(st) || WEXITSTATUS(st) != 0) rc = -1;
        }
        if (left && rc == 0) usleep(500);
    }
    *mb.max_path_index = 64;
    if (rc == 0) { std::memcpy(table, stable, px * 16); std::memcpy(accum, simage, px * 16); }
    munmap(m, bytes);
    return rc;
}

// The same with the shader compiled as the SUN_LIGHT build (sunDir, sunRadius: Iterative.comp:89-90)
void ref_glsl_render_frame_sun(const float* cam19, int w, int h, float total_light_area, uint32_t* table, float* accum, int model,
                               const float* sun_dir3, float sun_radius) {
    if (model == 1) { refglsl_beckmann_sun::set_sun(sun_dir3, sun_radius); refglsl_beckmann_sun::render_frame(cam19, w, h, total_light_area, table, accum); }
    else { refglsl_ggx_sun::set_sun(sun_dir3, sun_radius); refglsl_ggx_sun::render_frame(cam19, w, h, total_light_area, table, accum); }
}

// Present.frag over a whole accumulation image (h x w x 4 float sums) -> h x w x 3 floats, the colour the fragment shader writes
// before the framebuffer quantises it
void ref_glsl_present(const float* accum, int w, int h, int num_samples, float exposure_uniform, float* rgb) {
    using namespace refglsl_present;
    directAccum.p = accum; directAccum.w = w; numSamples = num_samples; exposure = exposure_uniform;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            gl_FragCoord = glm::vec4(x + 0.5f, y + 0.5f, 0.0f, 1.0f);
            present_main();
            float* o = rgb + 3 * ((size_t)y * w + x);
            o[0] = color.x; o[1] = color.y; o[2] = color.z;
        }
}

void ref_glsl_present_pl(const float* accum, int w, int h, int num_samples, float exposure_uniform, float* rgb) {
    using namespace refglsl_present_pl;
    directAccum.p = accum; directAccum.w = w; numSamples = num_samples; exposure = exposure_uniform;
    for (int y = 0; y < h; y++)
        for (int x = 0; x < w; x++) {
            gl_FragCoord = glm::vec4(x + 0.5f, y + 0.5f, 0.0f, 1.0f);
            present_main();
            float* o = rgb + 3 * ((size_t)y * w + x);
            o[0] = color.x; o[1] = color.y; o[2] = color.z;
        }
}

// EquirectangularConverter.frag's main() without the fetch: where in hdrTex the fragment with these sampleCoords looks
void ref_glsl_equirect_uv(const float* sample_coords, int64_t n, float* uv) {
    for (int64_t i = 0; i < n; i++) {
        glm::vec3 c(sample_coords[3 * i], sample_coords[3 * i + 1], sample_coords[3 * i + 2]);
        glm::vec2 r = refglsl_equirect::SampleSphericalMap(glm::normalize(c));
        uv[2 * i] = r.x; uv[2 * i + 1] = r.y;
    }
}

void ref_glsl_raygen_probe(const float* cam19, int w, int h, float total_light_area, uint32_t* table, int64_t first, int64_t n, float* out) {
    refglsl_ggx::raygen_probe(cam19, w, h, total_light_area, table, first, n, out);
}
void ref_glsl_raygen_probe_model(const float* cam19, int w, int h, float total_light_area, uint32_t* table, int64_t first, int64_t n, int model, float* out) {
    build_of(model).raygen_probe(cam19, w, h, total_light_area, table, first, n, out);
}

// n draws of the shader's HybridTaus from one state (Random.glsl:47-70)
void ref_glsl_hybrid_taus(uint32_t st[4], int n, float* out) {
    using namespace refglsl_ggx;
    state = uvec4(st[0], st[1], st[2], st[3]);
    for (int i = 0; i < n; i++) out[i] = HybridTaus();
    st[0] = state.x; st[1] = state.y; st[2] = state.z; st[3] = state.w;
}

}  // extern "C"
