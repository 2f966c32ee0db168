/* oracle/pt_oracle.h -- TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement of the reference's live GLSL path-tracing path (saadam1n/GPUPathTracer,
 * src/shaders/Iterative.comp + common/*.glsl) used as the parity checker for the CUDA
 * kernels. Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
 * reference legs may load this library; the product (libb200pt.so) never does.
 *
 * PARITY STATUS: the reference ships no tests, golden vectors or known-answer fixtures
 * (SURVEY.md section 4), and its GLSL cannot execute on this box (no GL stack), so the
 * pins below come from compiling the reference's own sources in place, not from its tests. What IS pinned,
 * against the reference's own C++ compiled in place (oracle/_ref/libref_shim.so, see
 * oracle/Makefile and tests/test_oracle_vs_ref.py):
 *   - HybridTaus            vs Renderer.cpp:764-782   (bit-exact)
 *   - slab test             vs math/AABB.cpp:102-113  (bit-exact on NaN-free inputs)
 *   - image plane           vs math/Camera.cpp:6-22   (bit-exact)
 *   - Uncharted-2 tonemap   vs Renderer.cpp:980-994   (bit-exact)
 *   - the BVH it traverses  IS the reference builder's output (BVH.cpp:2197-2293)
 *   - closest_hit / any_hit (IfIfClosestHit, IfIfAnyHit, IntersectNodeFast, IntersectLeaf, IntersectTriangleMT)
 *                           vs the reference's own SHADER SOURCE for them, extracted from common/BVH.glsl +
 *                           common/Geometry.glsl and compiled as C++ with glm (oracle/ref_glsl.cpp ->
 *                           oracle/_ref/libref_glsl.so): same triangle index, (t,u,v) bits and flags, live and on
 *                           the committed vectors tests/golden/trace_glsl_*.npz
 *   - the shading arithmetic (material assignments, SurfaceInteraction, all of Microfacet.glsl, ComputeBSDF, MISWeight,
 *                           GenerateImportanceSample, HybridTaus) vs the same kind of compile of common/{Material,Microfacet,
 *                           BSDF,MIS,Random}.glsl, through or_shade_probe / ref_glsl_shade_probe: same bits, GGX and
 *                           Beckmann, live and on tests/golden/shade_glsl.npz (one NaN-in-max() exception)
 *   - init_ray / light_sample (InitRay + RayIndexToMorton, GenerateLightSample + RandomLightVertex) vs Iterative.comp's own
 *                           source for them compiled the same way (or_raygen, or_light_probe vs ref_glsl_raygen_probe,
 *                           ref_glsl_light_probe): same bits, vectors in tests/golden/path_glsl_cornell_small.npz
 *   - or_render (whole frames: main() of Iterative.comp:214-307, GetInterpolatedVertex, ConstructMaterialInstance, and the
 *                           SUN_LIGHT build) vs that source compiled the same way with the harness as its GL runtime
 *                           (ref_glsl_render_frame[_sun]): same bits in the accumulation image and the RNG table, plain and
 *                           textured / cube-map scenes, vectors in tests/golden/frame_glsl_cornell_small.npz
 *   - or_tonemap            vs Present.frag compiled the same way (ref_glsl_present), tests/golden/present_glsl.npz
 *   - the coordinate arithmetic of or_equirect_to_cube vs EquirectangularConverter.frag's SampleSphericalMap (ref_glsl_equirect_uv),
 *                           tests/golden/equirect_glsl.npz
 * The whole shader path is therefore pinned to the reference's code. What remains "parity unpinned" is GL's fixed-function
 * part only: GL_LINEAR / GL_REPEAT filtering behind texture() and the cube-map face selection (or_sample_texture,
 * or_sample_environment, which the compiled shader source calls as its texture units).
 *
 * Float mode (stated, because GLSL does not pin one): IEEE-754 binary32, every operation
 * separately rounded (built with -ffp-contract=off -fno-fast-math), division and sqrt
 * correctly rounded, GLSL min()/max() = fminf()/fmaxf() (IEEE minNum/maxNum, which is what
 * NVIDIA's FMNMX executes; identical to glm's ternaries whenever no operand is NaN),
 * normalize(v) = v * (1/sqrt(dot(v,v))) and mix(x,y,a) = x*(1-a)+y*a as in glm.
 */
#ifndef PT_ORACLE_H
#define PT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Scene tables exactly as the reference uploads them. */
typedef struct OrTexture {
    int width, height;     /* RGBA32F texels, row-major, GL_REPEAT + GL_LINEAR (Texture.cpp:123-131) */
    const float* rgba;
} OrTexture;

typedef struct OrScene {
    const float* nodes;    /* n_nodes x 8 floats: {min.xyz, data0(int bits), max.xyz, data1} BVH.cpp:2261-2277 */
    int n_nodes;
    const int32_t* refs;   /* reference list, negative = last of leaf (BVH.cpp:2107-2121) */
    int n_refs;
    const float* tris;     /* n_tris x 20 floats, 5 x vec4 layout, e1/e2 pre-subtracted (Scene.cpp:334-337) */
    int n_tris;
    /* material k (matId = 2k, Scene.cpp:153-158): 12 floats
     *   [albedo.rgb, albedo_tex(int bits, <0: constant colour)]
     *   [emission.rgb, isEmissive(int bits)]
     *   [props.r, props.g = roughness source, props.b = metallic, props_tex(int bits, <0: constant)]
     * a constant colour stands for the reference's 1x1 RGBA32F texture (Texture2D::SetColor). */
    const float* materials;
    int n_materials;
    const OrTexture* textures;
    int n_textures;
    const float* lights;   /* n_lights x {float cumulativeArea; uint32 triIndex} (Scene.cpp:296-331) */
    int n_lights;
    float total_light_area;
    float env_color[3];    /* "GENERATE COLOR x" constant cubemap (Renderer.cpp:242-260), used when env_size == 0 */
    const float* env_faces; /* cubemap: 6 x env_size x env_size x RGBA32F, +X -X +Y -Y +Z -Z, rows as uploaded */
    int env_size;
} OrScene;

/* The 7 camera uniforms of Shader::LoadCamera (Shader.cpp:74-84). */
typedef struct OrCamera {
    float lower_left[3], horizontal[3], vertical[3], position[3], u[3], v[3];
    float lens_radius;
} OrCamera;

typedef struct OrConfig {
    int width, height;     /* multiples of 8 (Iterative.comp:163-172) */
    int max_depth;         /* max closest-hit rays per path; reference = 66 (`i > 64`, Iterative.comp:293) */
    int microfacet_model;  /* 0 = Trowbridge-Reitz (live, Microfacet.glsl:90-92), 1 = Beckmann (:35-55) */
    int n_threads;         /* host threads; <=0: hardware_concurrency */
    int sun_light;         /* 1: the `#ifdef SUN_LIGHT` branch of GenerateLightSample (Iterative.comp:116-126) */
    float sun_dir[3];      /* uniform sunDir (Iterative.comp:89; Renderer.cpp:23: normalize(2, 69, 12)) */
    float sun_radius;      /* uniform sunRadius (Iterative.comp:90; Renderer.cpp:24-25: tan(radians(5))) */
} OrConfig;

typedef struct OrCounters {
    uint64_t closest_rays, shadow_rays;
    uint64_t node_pairs, leaf_visits, tri_tests;   /* summed over both ray kinds */
    uint64_t paths;
    uint64_t shadow_node_pairs, shadow_tri_tests;  /* the any-hit share of node_pairs / tri_tests */
    uint64_t null_shadow_rays;                     /* shadow rays (counted in shadow_rays) whose radiance is exactly 0 */
} OrCounters;

/* ---- scalar pieces ---- */
void or_hybrid_taus(uint32_t state[4], int n, float* out);            /* Random.glsl:47-70 */
/* GLSL pins no algorithm for sin / cos / log / exp / pow. Default: glibc's float functions. After or_set_parity_libm(1): the
 * binary64 series of gpupathtracer_b200/csrc/pt_parity_libm.h, the same header the CUDA kernels (pt_config.parity_flags &
 * PT_PARITY_LIBM) and the reference's shader source compiled as C++ (model code | 4) use, so whole frames compare for equality. */
void or_set_parity_libm(int on);
int or_get_parity_libm(void);
void or_parity_libm_probe(int fn, const float* x, const float* y, int64_t n, float* out);   /* fn 0 sin 1 cos 2 log 3 exp 4 pow */
/* shading arithmetic on explicit inputs (material, interaction, NEE evaluation, BSDF sample) in Iterative.comp:236-288's order;
 * in: n x 16 floats, rng_io: n x 4 words (advanced), out: n x 20 floats -- layout in pt_oracle.cpp */
void or_equirect_uv(const float* sample_coords, int64_t n, float* uv);   /* EquirectangularConverter.frag:9-19 without the fetch */
void or_sample_texture(const OrScene* s, int index, float u, float v, float* rgba);   /* texture(sampler2D, uv): LINEAR + REPEAT */
void or_shade_probe(const float* in, uint32_t* rng_io, int64_t n, int model, float* out);
/* GenerateLightSample (Iterative.comp:113-147) on explicit inputs: n x 20 floats in, n x 12 out -- layout in pt_oracle.cpp */
void or_light_probe(const OrScene* s, const OrConfig* cfg, const float* in, uint32_t* rng_io, int64_t n, float* out);
void or_image_plane(float aspect, float fov, float focus, float aperture, const float pos[3],
                    const float rot[3], OrCamera* cam);                /* Camera.cpp:4-22 */
int or_slab_test(const float bmin[3], const float bmax[3], const float iorg[3], const float idir[3],
                 float tcur, float dist2[2]);                          /* BVH.glsl:379-391 */
void or_tonemap(const float rgb_sum[3], int num_samples, float out_rgb[3]); /* Present.frag:13-37 */
/* texture(samplerCube, d).xyz as Iterative.comp:251 samples the sky: GL cube-map face table, GL_LINEAR, CLAMP_TO_EDGE */
void or_sample_environment(const OrScene* s, const float* dirs, int64_t n, float* rgb);
/* LoadEnvironmnet's six raster passes over EquirectangularConverter.vert/.frag (Renderer.cpp:262-312), per face texel */
void or_equirect_to_cube(const float* rgba, int width, int height, int size, float* faces);

/* ---- fixed ray batches: rays[n][8] = {o.xyz, tmin(ignored), d.xyz, tmax} ---- */
/* hits[n][4] = {t, u, v, triId as int bits}; miss: t = tmax given, triId = -1. BVH.glsl:634-767 */
void or_trace_closest(const OrScene* s, const float* rays, int64_t n, float* hits, OrCounters* c, int n_threads);
/* flags[n] = 1 if occluded within (0, tmax). BVH.glsl:769-854 */
void or_trace_any(const OrScene* s, const float* rays, int64_t n, uint8_t* flags, OrCounters* c, int n_threads);

/* ---- whole frames: n_frames x (1 path per pixel), Iterative.comp:214-307 ----
 * rng[px][4] persistent per-pixel HybridTaus state (read, advanced, written back: Random.glsl:232-242)
 * accum[px][4] running SUM of radiance (alpha forced to 1), Iterative.comp:263,295 */
void or_render(const OrScene* s, const OrCamera* cam, const OrConfig* cfg, uint32_t* rng, float* accum,
               int n_frames, OrCounters* c);

/* ---- per-stage probes for wavefront-kernel parity (same code paths as or_render) ---- */
/* ray generation for ray index range [first, first+n): out[n][12] =
 * {pixel.x, pixel.y (int bits), o.xyz, d.xyz, 4 pad}; advances rng of the touched pixels. Iterative.comp:174-210 */
void or_raygen(const OrCamera* cam, const OrConfig* cfg, const float total_light_area, uint32_t* rng,
               int64_t first, int64_t n, float* out);

/* Brute force over all triangles with the same triangle test (no BVH): used by tests to classify
 * mismatches of conservative (wide/quantised) traversal as "reference box-cull miss" or "tie". */
void or_trace_closest_bruteforce(const OrScene* s, const float* rays, int64_t n, float* hits, int n_threads);

/* per-ray step trace (0 = pair test, 1 = triangle test) for warp-scheduling studies; returns total steps */
int64_t or_trace_ops(const OrScene* s, const float* rays, int64_t n, int any, uint8_t* ops, int64_t* offsets);

#ifdef __cplusplus
}
#endif
#endif
