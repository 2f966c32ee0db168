// oracle/ref_glsl_ns.h -- TEST INFRASTRUCTURE ONLY. Included by oracle/ref_glsl.cpp once per microfacet model, inside a
// namespace, with REFGLSL_SHADING naming the extracted shading source (GGX as shipped / the Beckmann macro variant).
// Everything between the #include lines is what the GL runtime supplies to the shader (uniforms, SSBOs, atomics) or a
// probe that calls the shader's functions in the shader's order; the arithmetic is all in the .inc files.
#undef M_PI
#ifdef REFGLSL_MAXNUM
// GLSL leaves max(NaN, x) undefined; glm spells max as a ternary (NaN propagates), NVIDIA's FMNMX, which is what the shader runs
// on, is IEEE maxNum (the other operand). This build gives scalar min / max the latter meaning and nothing else: the
// float mode oracle/pt_oracle.h states. (Every min / max of the shading path is scalar.)
inline float max(float a, float b) { return fmaxf(a, b); }
inline float min(float a, float b) { return fminf(a, b); }
#endif
#ifdef REFGLSL_PLIBM
// GLSL pins no algorithm for its transcendental functions; this build gives the scalar sin / cos / log / exp / pow the shading path
// calls (Iterative.comp:101,192; Microfacet.glsl:25,36-49,153) the one definition of pt_parity_libm.h instead of glm's (= glibc's),
// and nothing else: the same header the CUDA kernels use under PT_PARITY_LIBM, so whole frames can be compared for equality.
inline float sin(float x) { return ptpl::sinf_(x); }
inline float cos(float x) { return ptpl::cosf_(x); }
inline float log(float x) { return ptpl::logf_(x); }
inline float exp(float x) { return ptpl::expf_(x); }
inline float pow(float x, float y) { return ptpl::powf_(x, y); }
#endif
#include REFGLSL_SHADING

// ConstructMaterialInstance (Material.glsl:39-54) reads two bindless textures; the harness supplies the two texels and runs the
// shader's own six assignments after them (Material.glsl:46-51, extracted like the rest)
inline MaterialInstance material_from_texels(vec4 data0, vec4 data1) {
    MaterialInstance material;
    material.emission = vec3(0.0f); material.materialType = 1u; material.ior = 0.0f;
#include "ref_glsl_material_body.inc"
    return material;
}

// ---- Iterative.comp's ray generation and light sampling
#define uniform
static uint g_ray_counter_local;                                    // the ray counter SSBO, Iterative.comp:150-152: this word, or one
static uint* g_ray_counter = &g_ray_counter_local;                  // in memory shared by several processes (ref_glsl_render_frames_mp)
#define globalNextRay (*g_ray_counter)
inline uint atomicAdd(uint& x, uint v) { return __atomic_fetch_add(&x, v, __ATOMIC_RELAXED); }
static int maxPathIndex = 64;                                       // the literal of `i > 64` (Iterative.comp:295), see oracle/Makefile
#define globalRayCount int(width * height)                          // Iterative.comp:149
static uint32_t* rngTable;                                          // `states[]`, Random.glsl (one uvec4 per pixel)
inline void initRNG(uint ridx) {                                    // Random.glsl:232-237 without the LD-sequence state the path never reads
    stateIdx = ridx;
    state = uvec4(rngTable[4 * (size_t)ridx], rngTable[4 * (size_t)ridx + 1], rngTable[4 * (size_t)ridx + 2], rngTable[4 * (size_t)ridx + 3]);
}
#ifdef REFGLSL_SUN
#define SUN_LIGHT                                                   // Renderer.cpp compiles the shader with this when the sun is on
#endif
#include "ref_glsl_path.inc"
#undef SUN_LIGHT
#undef uniform

// ---- main() itself (Iterative.comp:214-307) with GetInterpolatedVertex and ConstructMaterialInstance
inline void freeRNG() {                                             // Random.glsl:239-242, again without the LD-sequence word
    rngTable[4 * (size_t)stateIdx] = state.x; rngTable[4 * (size_t)stateIdx + 1] = state.y;
    rngTable[4 * (size_t)stateIdx + 2] = state.z; rngTable[4 * (size_t)stateIdx + 3] = state.w;
}
static const vec4* materialInstance;                                // the `samplers` SSBO, Material.glsl:10-12
static image2D accum;                                               // Iterative.comp:29
static uvec2 threadIdx;                                             // Iterative.comp:4
#include "ref_glsl_main.inc"

// in / out layouts: oracle/pt_oracle.cpp or_shade_probe
static void shade_probe(const float* in, uint32_t* rng_io, int64_t n, float* out) {
    for (int64_t k = 0; k < n; k++) {
        const float* a = in + 16 * k; float* o = out + 20 * k; uint32_t* rs = rng_io + 4 * k;
        MaterialInstance material = material_from_texels(vec4(a[0], a[1], a[2], 1.0f), vec4(0.0f, a[3], a[4], 1.0f));
        vec3 nrm(a[5], a[6], a[7]), og(a[8], a[9], a[10]), ld(a[11], a[12], a[13]);
        SurfaceInteraction interaction = ConstructSurfaceInteraction(nrm, og);
        SurfaceInteraction li = interaction;
        SetIncomingDirection(li, ld);
        vec3 fl = ComputeBSDF(material, li);
        o[0] = fl.x; o[1] = fl.y; o[2] = fl.z; o[3] = ProbabilityDensityDirection(material, li); o[4] = li.ndi;
        state = uvec4(rs[0], rs[1], rs[2], rs[3]);
        float bxdfPdf0, bxdfPdf1;
        vec3 d = GenerateImportanceSample(material, interaction, bxdfPdf0, bxdfPdf1);
        vec3 fs = ComputeBSDF(material, interaction);
        float w = MISWeight(bxdfPdf0, bxdfPdf1);
        vec3 factor = ComputeBSDF(material, interaction) * interaction.ndi * MISWeight(bxdfPdf0, bxdfPdf1) / bxdfPdf0;   // Iterative.comp:284
        o[5] = d.x; o[6] = d.y; o[7] = d.z; o[8] = bxdfPdf0; o[9] = bxdfPdf1; o[10] = fs.x; o[11] = fs.y; o[12] = fs.z; o[13] = interaction.ndi; o[14] = w;
        o[15] = factor.x; o[16] = factor.y; o[17] = factor.z; o[18] = CalcDiffusePmf(material, interaction); o[19] = 0.0f;
        rs[0] = state.x; rs[1] = state.y; rs[2] = state.z; rs[3] = state.w;
    }
}

// GenerateLightSample (Iterative.comp:113-147) as main() calls it (:276). in: n x 20 floats {vertex position, vertex normal,
// outgoing, data0.rgb, data1.g, data1.b, throughput.rgb, -...}; out: n x 12 floats {ray origin, ray direction, lightHit.di.x,
// lightThroughput.rgb, lightMatID (bits), lightPdf}; rng_io advanced. Light table / triangles: the bound scene.
static void light_probe(const float* in, uint32_t* rng_io, int64_t n, float* out) {
    for (int64_t k = 0; k < n; k++) {
        const float* a = in + 20 * k; float* o = out + 12 * k; uint32_t* rs = rng_io + 4 * k;
        Vertex vertex; vertex.Position = vec3(a[0], a[1], a[2]); vertex.Normal = vec3(a[3], a[4], a[5]); vertex.TextureCoordinate = vec2(0.0f); vertex.MatID = 0u;
        vec3 og(a[6], a[7], a[8]);
        MaterialInstance material = material_from_texels(vec4(a[9], a[10], a[11], 1.0f), vec4(0.0f, a[12], a[13], 1.0f));
        SurfaceInteraction interaction = ConstructSurfaceInteraction(vertex.Normal, og);
        vec3 throughput(a[14], a[15], a[16]);
        state = uvec4(rs[0], rs[1], rs[2], rs[3]);
        HitInfo lightHit; vec3 lightThroughput; uint lightMatID; float lightPdf;
        Ray lightRay = GenerateLightSample(vertex, material, interaction, lightHit, lightPdf, throughput, lightThroughput, lightMatID);
        o[0] = lightRay.origin.x; o[1] = lightRay.origin.y; o[2] = lightRay.origin.z;
        o[3] = lightRay.direction.x; o[4] = lightRay.direction.y; o[5] = lightRay.direction.z; o[6] = lightHit.di.x;
        o[7] = lightThroughput.x; o[8] = lightThroughput.y; o[9] = lightThroughput.z; std::memcpy(&o[10], &lightMatID, 4); o[11] = lightPdf;
        rs[0] = state.x; rs[1] = state.y; rs[2] = state.z; rs[3] = state.w;
    }
}

// InitRay (Iterative.comp:174-210) for ray indices [first, first + n); out: n x 12 floats as or_raygen {pixel.x, pixel.y (int
// bits), origin, direction, throughput.x, bxdfPdf0, bxdfPdf1, neePdf}; the pixel's RNG words in `table` are advanced (freeRNG's part)
static void set_camera(const float* cam19, int w, int h, float total_light_area, uint32_t* table) {
    lower_left = vec3(cam19[0], cam19[1], cam19[2]); horizontal = vec3(cam19[3], cam19[4], cam19[5]); vertical = vec3(cam19[6], cam19[7], cam19[8]);
    position = vec3(cam19[9], cam19[10], cam19[11]); u = vec3(cam19[12], cam19[13], cam19[14]); v = vec3(cam19[15], cam19[16], cam19[17]);
    lens_radius = cam19[18]; width = w; height = h; totalLightArea = total_light_area; rngTable = table;
}
static void set_sun(const float* dir3, float radius) { sunDir = vec3(dir3[0], dir3[1], dir3[2]); sunRadius = radius; }
// one dispatch of main(): a single invocation drains the ray counter
static void render_frame(const float* cam19, int w, int h, float total_light_area, uint32_t* table, float* image) {
    set_camera(cam19, w, h, total_light_area, table);
    accum.p = image; accum.w = w;
    globalNextRay = 0u;
    shader_main();
}
// the same dispatch with the ray counter in shared memory: every process that calls this with the same counter is one more
// invocation of the persistent shader draining it (pixels are disjoint, so image and RNG table can be shared as they are)
static void render_frame_shared(const float* cam19, int w, int h, float total_light_area, uint32_t* table, float* image, uint* counter) {
    set_camera(cam19, w, h, total_light_area, table);
    accum.p = image; accum.w = w;
    g_ray_counter = counter;
    shader_main();
    g_ray_counter = &g_ray_counter_local;
}
static void raygen_probe(const float* cam19, int w, int h, float total_light_area, uint32_t* table, int64_t first, int64_t n, float* out) {
    set_camera(cam19, w, h, total_light_area, table);
    for (int64_t k = 0; k < n; k++) {
        globalNextRay = (uint)(first + k);
        ivec2 pixel; Ray ray; vec3 throughput, contribution, lastPosition; float bxdfPdf0, bxdfPdf1, neePdf;
        bool ok = InitRay(pixel, ray, throughput, contribution, bxdfPdf0, bxdfPdf1, neePdf, lastPosition);
        float* o = out + 12 * k;
        if (!ok) { for (int q = 0; q < 12; q++) o[q] = 0.0f; continue; }
        std::memcpy(&o[0], &pixel.x, 4); std::memcpy(&o[1], &pixel.y, 4);
        o[2] = ray.origin.x; o[3] = ray.origin.y; o[4] = ray.origin.z; o[5] = ray.direction.x; o[6] = ray.direction.y; o[7] = ray.direction.z;
        o[8] = throughput.x; o[9] = bxdfPdf0; o[10] = bxdfPdf1; o[11] = neePdf;
        size_t pix = (size_t)pixel.y * w + pixel.x;
        table[4 * pix] = state.x; table[4 * pix + 1] = state.y; table[4 * pix + 2] = state.z; table[4 * pix + 3] = state.w;
    }
}

static void bind_lights(const float* lights2, int n_lights, float total_light_area) {
    lightTex.p = lights2; lightTex.comps = 2; lightTex.count = n_lights; totalLightArea = total_light_area;
}
