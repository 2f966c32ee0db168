/* include/b200pt.h -- C ABI of libb200pt.so, the B200-native replacement for the OpenGL hot path of
 * saadam1n/GPUPathTracer (the src/shaders wavefront stages + the src/core BVH build / flatten / scene
 * upload). Plain pointers and sizes only; no torch, no C++ types. Every entry point cites the
 * reference interface it replaces (paths relative to the reference's src/).
 *
 * The reference has no FFI: the seam is the four C++ classes Renderer / Scene / BoundingVolumeHierarchy /
 * Camera driven by main() (Program.cpp:69-152). INTEGRATION.md shows the binding a maintainer adds.
 *
 * Conventions: every function returning int returns PT_OK (0) or a negative pt_status and never calls
 * exit() (the reference prints and exit()s, e.g. Scene.cpp:134,258; BVH.cpp:2146). pt_last_error() returns
 * the message of the calling thread's last failure. A context is bound to one CUDA device, is driven by
 * one host thread at a time, and owns all its device memory; host arrays are copied during the call.
 * There is no CPU fallback: pt_create fails with PT_ERR_NO_DEVICE when no sm_100 device is usable.
 */
#ifndef B200PT_H
#define B200PT_H
#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200PT_ABI_VERSION 3

typedef enum pt_status {
    PT_OK = 0,
    PT_ERR_INVALID = -1,      /* bad argument / call order */
    PT_ERR_NO_DEVICE = -2,    /* no usable CUDA device (the product has no CPU path) */
    PT_ERR_CUDA = -3,         /* a CUDA runtime call failed; message has the cudaError string */
    PT_ERR_IO = -4,           /* file could not be read / written */
    PT_ERR_UNSUPPORTED = -5,
    PT_ERR_LIMIT = -6         /* e.g. BVH deeper than the traversal stack (reference: exit(-1), BVH.cpp:2144-2147) */
} pt_status;

typedef struct pt_context pt_context;

/* microfacet model: the reference executes Trowbridge-Reitz (Microfacet.glsl:90-92) and carries Beckmann
 * as an alternate (Microfacet.glsl:35-55); both are selectable at run time here. */
enum { PT_MICROFACET_GGX = 0, PT_MICROFACET_BECKMANN = 1 };

/* traversal layout: PT_BVH_REFERENCE keeps the reference tree's topology and fp32 boxes (results are
 * bit-identical to IfIfClosestHit/IfIfAnyHit). PT_BVH_WIDE / PT_BVH_WIDE4 collapse that tree into 8- / 4-wide nodes
 * with 8-bit quantised child boxes (96 B / 64 B per node): same triangles, same triangle test, conservative boxes --
 * (t,u,v) are bit-identical whenever the same triangle is reported; the reported triangle differs for the ~1e-6 of
 * rays whose hit the reference's fp32 slab test culls or that tie in t. Faster once the tree outgrows L2. */
enum { PT_BVH_REFERENCE = 0, PT_BVH_WIDE = 1, PT_BVH_WIDE4 = 2 };

typedef struct pt_config {
    int32_t width, height;        /* multiples of 8 (Iterative.comp:163-172); Program.cpp:22-23 hard-codes 1280x720 */
    int32_t max_depth;            /* max closest-hit rays per path; 66 == the reference's `i > 64` cap (Iterative.comp:293) */
    int32_t microfacet_model;     /* PT_MICROFACET_* */
    int32_t device;               /* CUDA device ordinal */
    int32_t bvh_layout;           /* PT_BVH_* used by pt_render_frames / pt_trace_* */
    int32_t n_slices;             /* a frame is rendered as this many pixel-disjoint slices on concurrent streams so that
                                     one slice's kernels fill the drain phase of another's; 0 = default (3). Results do
                                     not depend on it (every pixel keeps its own RNG sequence). */
    int32_t sun_light;            /* 1 = compile the reference's `#ifdef SUN_LIGHT` branch in (Iterative.comp:116-126): NEE flips a
                                     coin between the emitter table and a sun disk. Off in the reference as shipped. */
    int32_t trace_null_shadow;    /* 0 (default): an NEE sample whose radiance is exactly 0 whatever the visibility (light below
                                     the surface's horizon) is settled without a shadow ray -- same image bits, ~1/3 fewer
                                     shadow rays on the atrium. 1: trace it anyway, as the reference does, so that ray
                                     counts are comparable one to one with the reference's. */
    int32_t parity_flags;         /* PT_PARITY_*: switches that pin what GLSL leaves open so that results can be compared for
                                     EQUALITY with the reference's shader source compiled on a CPU (tests, parity runs); 0 = fast defaults */
    int32_t reserved[2];
} pt_config;

/* pt_config.parity_flags. PT_PARITY_LIBM: sin / cos / log / exp / pow of the shading path (InitRay Iterative.comp:191-192, the
 * importance samplers Microfacet.glsl:25,49,153, Beckmann :36-44, Present.frag:36) are evaluated by csrc/pt_parity_libm.h -- binary64
 * series built from IEEE + - * / only, one rounding to binary32 -- instead of CUDA's sincosf / logf / expf / powf, which differ from
 * glibc's by 1-2 ulp on some arguments. With it, frames are BIT-identical to the reference's main() compiled as C++ with the same
 * header (accumulation image and RNG table); without it < 0.1 % of pixels differ after a few frames. ~15 % slower shade stage. */
enum { PT_PARITY_LIBM = 1 };

/* The 7 camera uniforms of Shader::LoadCamera (Shader.cpp:74-84). */
typedef struct pt_camera {
    float lower_left[3], horizontal[3], vertical[3], position[3], u[3], v[3];
    float lens_radius;
} pt_camera;

/* Material k is addressed by matId = 2*k (Scene.cpp:153-158); k = 0 is the sky (Scene.cpp:236-241).
 * Replaces MaterialInstance{albedoHandle, propertiesHandle, emission, isEmissive} (Scene.h:12-17): bindless
 * texture handles become indices into the texture array, and the 1x1 RGBA32F "colour textures" of
 * Texture2D::SetColor (Scene.cpp:97-103) become inline constants (tex index < 0). */
typedef struct pt_material {
    float albedo[3];   int32_t albedo_tex;
    float emission[3]; int32_t is_emissive;
    float props[3];    int32_t props_tex;      /* props = (unused, roughness source g, metallic b), Scene.cpp:103 */
} pt_material;

typedef struct pt_texture {        /* RGBA32F, row-major, sampled GL_LINEAR + GL_REPEAT (Texture.cpp:123-131) */
    int32_t width, height;
    const float* rgba;
} pt_texture;

typedef struct pt_light {          /* LightTriangleInfo after the prefix sum, Scene.cpp:296-330 */
    float cumulative_area;
    uint32_t tri_index;
} pt_light;

typedef struct pt_counters {
    uint64_t closest_rays, shadow_rays, paths, frames;
    uint64_t kernel_launches;      /* kernels of this library launched since pt_create */
    uint64_t null_shadow_rays;     /* NEE samples whose radiance is exactly 0 (light below the horizon): the reference traces
                                      their shadow ray and adds 0, this library settles them without a ray; not part of
                                      shadow_rays, which counts rays actually traced */
    /* visit counts of the rays traced while pt_set_visit_counting was on (0 otherwise): sibling-pair tests (IntersectNodeFast x 2,
       BVH.glsl:692-705) and triangle tests (IntersectTriangleMT) of closest-hit and of shadow rays -- SURVEY 8d's N_inner / N_tri */
    uint64_t closest_pair_tests, closest_tri_tests, shadow_pair_tests, shadow_tri_tests;
    uint64_t counted_closest_rays, counted_shadow_rays;   /* the rays those visits belong to */
} pt_counters;

typedef struct pt_bvh_stats {
    int32_t n_nodes, n_refs, n_leaves, max_depth;
    float sah_cost;                /* CalculateCost, BVH.cpp:2174-2195 */
    float build_ms;                /* the build itself (host builder: wall time; device builder: CUDA-event time of its kernels) */
    float total_ms;                /* wall time of the call: build + re-layout + upload (+ read-back for the device builder) */
} pt_bvh_stats;

int pt_abi_version(void);
const char* pt_last_error(void);

/* ---- lifetime: Renderer::Initialize / CleanUp (Renderer.cpp:448-636, :638-647) ---- */
int pt_create(const pt_config* cfg, pt_context** out);
void pt_destroy(pt_context* ctx);

/* ---- scene tables: the uploads of Scene::LoadScene (Scene.cpp:339-354) ---- */
/* tris: n x 20 floats in the CompactTriangle order (math/Triangle.h:13-29) with position1/2 holding the
 * EDGES e1 = p1-p0, e2 = p2-p0 (Scene.cpp:334-337), i.e. exactly the bytes of the reference's vertexBuf. */
int pt_set_triangles(pt_context* ctx, const float* tris, int64_t n);
int pt_set_materials(pt_context* ctx, const pt_material* mats, int32_t n_mats, const pt_texture* tex, int32_t n_tex);
int pt_set_lights(pt_context* ctx, const pt_light* lights, int32_t n, float total_light_area); /* lightBuf + totalLightArea */
int pt_set_environment_color(pt_context* ctx, const float rgb[3]);   /* LoadEnvironmnet "GENERATE COLOR", Renderer.cpp:242-260 */
/* Environment cubemap sampled on a miss (Iterative.comp:251: texture(samplerCube, ray.direction)): six size x size
 * RGBA32F faces in the order +X -X +Y -Y +Z -Z, rows as they would be handed to glTexImage2D; looked up with the
 * OpenGL cube-map face table, GL_LINEAR, CLAMP_TO_EDGE per face (Renderer.cpp:318-322). */
/* uniforms sunDir / sunRadius (Iterative.comp:89-90). Defaults are the reference's constants (Renderer.cpp:23-26):
 * normalize(2, 69, 12) and tan(radians(5)). Only read when pt_config.sun_light is set. */
int pt_set_sun(pt_context* ctx, const float direction[3], float radius);
int pt_set_environment_cubemap(pt_context* ctx, const float* faces_rgba, int32_t size);
/* The .hdr branch of LoadEnvironmnet (Renderer.cpp:262-312) as one kernel: the six 90-degree raster passes over
 * EquirectangularConverter.vert/.frag written out per face texel. rgba = width x height RGBA32F, row 0 = top of the
 * file (stbi_loadf order); cube_size is 1024 in the reference. */
int pt_set_environment_equirect(pt_context* ctx, const float* rgba, int32_t width, int32_t height, int32_t cube_size);
int32_t pt_environment_size(pt_context* ctx);                        /* 0: constant colour */
int pt_download_environment(pt_context* ctx, float* faces_rgba);     /* 6 x size x size x 4 floats */
/* All of LoadEnvironmnet: "GENERATE COLOR x" | path ending in .hdr (Radiance RGBE, equirect -> 1024^2 cubemap) | text
 * file naming six face images (TextureCubemap::LoadTexture, Texture.cpp:245-276; faces as .ppm/.bmp/.tga).
 * ".jpg" goes through the JPEG decoder and stb's gamma-2.2 expansion (stbi_loadf on an 8-bit file), progressive files
 * included. */
int pt_load_environment(pt_context* ctx, const char* args);

/* ---- BVH: BoundingVolumeHierarchy::BuildBinnedSpatial (BVH.cpp:2197-2293) ---- */
/* Inject a tree in the reference's serialised form: nodes = n_nodes x {float3 min; int32 data0; float3 max;
 * int32 data1} (NewLayout, BVH.cpp:2261-2277), refs = the leaf reference list (ConvertIntoLeaf, :2107-2121).
 * Re-laid out on upload for 128-bit loads; topology, boxes and leaf order are preserved. */
int pt_set_bvh(pt_context* ctx, const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs);
/* Build with this library's own SAH/SBVH builder from the triangles given to pt_set_triangles and upload it. */
int pt_build_bvh(pt_context* ctx, pt_bvh_stats* stats);
/* Fast build ON THE DEVICE from the triangles given to pt_set_triangles: 30-bit Morton codes, radix sort, Karras'
 * parallel hierarchy, bottom-up boxes, one triangle per leaf; output in the same serialised form (readable with
 * pt_get_bvh, traversed by the same kernels, collapsible to the wide layouts). ~100x faster to build than the SBVH,
 * costlier to traverse (no SAH search, no spatial splits): for dynamic scenes and build-time-bound uses. */
int pt_build_bvh_device(pt_context* ctx, pt_bvh_stats* stats);
/* The same pipeline with the hierarchy built by parallel locally-ordered clustering (PLOC, Meister & Bittner 2018)
 * instead of Morton-bit splits: clusters of the Morton-sorted leaves merge with the neighbour within `radius` positions
 * (1 .. 256; 16 is a good default) that gives the smallest union surface area, round after round. A few ms instead of
 * well under one, and a tree whose SAH cost and traversal speed sit between the LBVH's and the host SBVH's. The
 * reference has no counterpart (its one builder is BuildBinnedSpatial on the host, BVH.cpp:2197-2293). Clustering does
 * not bound the depth: when the tree comes out deeper than the traversal stack (96 entries) the Morton hierarchy is
 * used instead, PT_OK is returned and pt_last_error() carries a note (10 M-triangle nested shells came out 64-66 deep). */
int pt_build_bvh_device_ploc(pt_context* ctx, int32_t radius, pt_bvh_stats* stats);
/* The device builders with SPATIAL SPLITS (FindBestSpatialSplit, BVH.cpp:1793-1984, by other means): before the Morton sort every
 * triangle whose box is longer than 1.5 cells of a split_grid^3 grid over the scene box is cut into slabs of its own box (<= 4096
 * pieces per triangle), clipped to each piece, empty pieces dropped; the hierarchy (ploc_radius 0 = Morton bits, else PLOC) is built
 * over the resulting references, so the tree has one leaf per reference and a triangle can sit in several leaves, as in the
 * reference's SBVH. Same serialised form. split_grid 2 .. 1024 (64 is a good start; PT_ERR_LIMIT if it would more than double the scene). */
int pt_build_bvh_device_split(pt_context* ctx, int32_t ploc_radius, int32_t split_grid, pt_bvh_stats* stats);
/* TOP-DOWN BINNED SAH on the device: the search the reference's builder runs (FindBestObjectSplit, BVH.cpp:1619-1763: bins per
 * axis, the cheapest of the planes between them by surface area x count; sweep for small nodes) without its spatial splits,
 * breadth first: per level one binning pass over the primitives of every node above 16 primitives (16 bins per axis), one thread
 * per node choosing the plane, one partition pass; subtrees of <= 16 primitives are finished by one thread each with an exact
 * sweep over every centroid plane, where a node also becomes a LEAF of several triangles by the reference's rule (AdjustSAH + leaf
 * test, BVH.cpp:2123-2131): unless Ct + Ci (A_l n_l + A_r n_r) / A < Ci n. Measured to trace like the host SBVH tree (1.00 x on four
 * scenes) for a 1.4-5.5 ms build. Same serialised form (n_leaves <= n_refs), same fallback to the Morton hierarchy (PT_OK + a note in
 * pt_last_error()) should the tree come out deeper than the traversal stack. The builders' scratch stays in the context (<= 512 MB),
 * so a rebuild per frame costs its kernels only. */
int pt_build_bvh_device_sah(pt_context* ctx, pt_bvh_stats* stats);
/* Reinsertion optimisation of the tree the context holds, whoever built it (Bittner, Hapala, Havran 2013, batched: the nodes with the
 * largest surface look -- branch and bound -- for the place in the tree where their subtree adds the least surface area, and move there).
 * The reference carries such a pass, ReinsertionOptimize, and never calls it (BVH.cpp; SURVEY.md 2 lists it as dead code). Leaves and
 * reference list are untouched, the serialised form is kept; `iterations` passes (1 .. 64) over the `batch_fraction` (0 .. 1] of nodes
 * with the largest surface. Host work (seconds at 1 M triangles); the tree is read back from the device if it was built there.
 * PT_ERR_LIMIT (tree unchanged) if the result would be deeper than the traversal stack. */
int pt_optimize_bvh(pt_context* ctx, int32_t iterations, float batch_fraction, pt_bvh_stats* stats);
/* The same on arrays, in place, no context: returns the number of subtrees moved (>= 0) or an error code (< 0). */
int pt_bvh_optimize_host(float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs, int32_t iterations, float batch_fraction,
                         float* sah_before, float* sah_after);
/* The same builder without a context or a device (pure host work): arrays are malloc()ed, release with pt_free. */
int pt_bvh_build_host(const float* tris, int64_t n, int32_t n_threads, float** nodes, int32_t* n_nodes,
                      int32_t** refs, int32_t* n_refs, pt_bvh_stats* stats);
void pt_free(void* p);
/* Collapse a reference-form tree into the compressed wide layout used by PT_BVH_WIDE (width 8, 96-B nodes) or
 * PT_BVH_WIDE4 (width 4, 64-B nodes) without a device: node_bytes = n_wide x node_size bytes exactly as uploaded
 * (layout: gpupathtracer_b200/csrc/pt_host.h WideNodeHost / WideNode4Host), tri_ids = triangle index per entry of the
 * wide leaf stream. The reference has no counterpart (its device tree is the binary NewLayout array, BVH.cpp:2261-2277);
 * north_star item (1) asks for it. Arrays are malloc()ed, release with pt_free. */
int pt_bvh_collapse_host(const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs, const float* tris,
                         int64_t n_tris, int32_t width, uint8_t** node_bytes, int64_t* n_wide, int32_t* node_size,
                         int32_t** tri_ids, int64_t* n_tri_ids, int32_t* depth);
/* SAH cost of a serialised tree with the reference's constants (CalculateCost, BVH.cpp:2174-2195; traversal 1.23,
 * intersection 5.33), relative to the root's surface area. */
float pt_bvh_sah_cost(const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs);
/* Read back the tree in the reference's serialised form (what nodesVec/referenceVec hold, BVH.cpp:2291-2292);
 * call with NULL arrays to query sizes. */
int pt_get_bvh(pt_context* ctx, float* nodes, int32_t* n_nodes, int32_t* refs, int32_t* n_refs);

/* ---- host-side scene assembly (no device work): Scene.cpp:263-337 ---- */
/* verts: n_tris x 3 x {pos.xyz, normal.xyz, uv.xy} (Vertex, math/Vertex.h), mat_ids: per triangle (first
 * vertex's, Scene.cpp:289). out_tris_abs: n x 20 floats with ABSOLUTE positions (the form the BVH builder
 * takes, Scene.cpp:294); geometric normal flipped to agree with the mean vertex normal (Scene.cpp:277-287). */
int pt_assemble_triangles(const float* verts, const uint32_t* mat_ids, int64_t n_tris, float* out_tris_abs);
/* emitter list sorted by area + running sum (Scene.cpp:296-331); returns count, writes total area. */
int pt_build_light_table(const float* tris_abs, int64_t n_tris, const pt_material* mats, int32_t n_mats,
                         pt_light* out_lights, int32_t capacity, float* total_area);
/* e1/e2 pre-subtraction in place (Scene.cpp:334-337). */
void pt_presubtract_edges(float* tris, int64_t n_tris);

/* ---- scene files: Scene::LoadScene / LoadOBJ (Scene.cpp:120-259), scene.txt (Program.cpp:70-84) ----
 * Own .obj/.mtl reader that follows tinyobjloader's behaviour for what the reference consumes (shapes at g/o, per-face
 * usemtl, quads cut along the shorter diagonal, missing normal (0,1,0) / texcoord (0,0)) and the reference's material
 * rules (matId = 2*(1 + .mtl index), sky = 0, every triangle of a shape takes the shape's FIRST face's material,
 * roughness = sqrt(2/(Ns+2)), metallic = max(Ks) > 0.3, emissive = sum(Ke) > 1e-5). Host only. */
typedef struct pt_scene_file pt_scene_file;
int pt_scene_file_load_obj(const char* obj_path, pt_scene_file** out);
void pt_scene_file_free(pt_scene_file* sf);
int64_t pt_scene_file_num_triangles(const pt_scene_file* sf);
const float* pt_scene_file_vertices(const pt_scene_file* sf);          /* n_tris x 3 x {pos.xyz, normal.xyz, uv.xy} */
const uint32_t* pt_scene_file_material_ids(const pt_scene_file* sf);   /* per triangle */
int32_t pt_scene_file_num_materials(const pt_scene_file* sf);          /* sky included */
const pt_material* pt_scene_file_materials(const pt_scene_file* sf);   /* albedo_tex = -1: see texture_path */
const char* pt_scene_file_texture_path(const pt_scene_file* sf, int32_t material);   /* folder + map_Kd, "" if none */
int32_t pt_scene_file_num_shapes(const pt_scene_file* sf);
const char* pt_scene_file_warnings(const pt_scene_file* sf);
/* map_Kd image -> RGBA32F texels (byte / 255, rows in file order; Texture2D::LoadTexture, Texture.cpp:146-156, decodes
 * through SOIL2 -- here: .jpg (baseline / extended sequential / progressive, as stb_image reads them), .png (all colour types / depths / Adam7), .ppm, 24/32-bit .bmp, true-colour .tga). *rgba is malloc()ed, release with pt_free. */
int pt_image_load(const char* path, int32_t* width, int32_t* height, float** rgba);
/* 8-bit RGB PNG, rows top-down (filter 0, stored deflate blocks): the format of Renderer::SaveScreenshot (Renderer.cpp:698). */
int pt_image_save_png(const char* path, int32_t width, int32_t height, const uint8_t* rgb_top_down);
/* stbi_loadf(path, .., 4) (Renderer.cpp:265): a Radiance .hdr as byte * 2^(e-136), alpha 1; a JPEG / PNG through the 8-bit
 * decode and pow(byte / 255, 2.2) (alpha byte / 255). Rows in file order. *rgba is malloc()ed, release with pt_free. */
int pt_image_load_hdr(const char* path, int32_t* width, int32_t* height, float** rgba);
/* scene.txt: line 1 scene path, line 2 environment string, then camera position xyz and rotation (yaw pitch roll). */
typedef struct pt_scene_txt {
    char scene_path[1024];
    char environment[1024];
    float camera_position[3], camera_rotation[3];
} pt_scene_txt;
int pt_scene_txt_read(const char* path, pt_scene_txt* out);
/* The "GENERATE COLOR BLACK|WHITE" branch of LoadEnvironmnet (Renderer.cpp:236-260; anything else after COLOR: red, as
 * the reference). Like the reference (find_first_of, :238) any string whose first character is one of G,E,N,R,A,T is
 * taken as a command; a command other than GENERATE COLOR returns PT_ERR_UNSUPPORTED (the reference leaves the
 * cubemap without storage). */
int pt_parse_environment(const char* args, float rgb[3]);
/* Everything Renderer::Initialize does with a scene path: parse, assemble triangles, light table, edge pre-subtraction,
 * textures (map_Kd as .ppm/.bmp/.tga; other formats keep the Kd colour), uploads, BVH build. environment goes through
 * pt_load_environment; NULL = "GENERATE COLOR WHITE". */
int pt_load_scene(pt_context* ctx, const char* scene_path, const char* environment, pt_bvh_stats* stats_or_null);

/* ---- camera: Camera::GenerateImagePlane (math/Camera.cpp:6-22) + Shader::LoadCamera ---- */
void pt_camera_image_plane(float aspect, float fov, float focus, float aperture, const float pos[3],
                           const float rot[3], pt_camera* out);
int pt_set_camera(pt_context* ctx, const pt_camera* cam);

/* ---- RNG tables: Renderer.cpp:545-553 (uvec4 per pixel, every lane >= 129) ---- */
int pt_set_rng(pt_context* ctx, const uint32_t* state4);          /* width*height*4 words */
int pt_seed_rng(pt_context* ctx, uint64_t seed);                   /* generated on the device; seed^rank decorrelates GPUs */
int pt_get_rng(pt_context* ctx, uint32_t* state4);

/* ---- rendering: Renderer::RenderFrame / ResetSamples / GetNumSamples (Renderer.cpp:651-695) ---- */
int pt_render_frames(pt_context* ctx, int32_t n_frames);           /* each frame = +1 sample for every pixel */
int pt_reset_samples(pt_context* ctx);
uint32_t pt_num_samples(pt_context* ctx);
int pt_synchronize(pt_context* ctx);
/* accum = running SUM, RGBA32F, alpha 1 (Iterative.comp:263,295). */
int pt_download_accum(pt_context* ctx, float* rgba);
void* pt_accum_device_ptr(pt_context* ctx);                        /* for NCCL reduce across ranks (width*height*4 floats) */
int pt_set_accum_device_ptr(pt_context* ctx, void* dev_rgba);      /* render into caller-owned device memory */
/* Checkpoint / resume of a long render (no counterpart in the reference): the per-pixel RNG table, the
 * accumulation buffer and numSamples, with a checksum; written beside `path` and renamed into place. After a load into a
 * context of the same size and scene, further frames are bit-identical to those of an uninterrupted run. One file per
 * GPU in a multi-GPU render (each rank owns its RNG table and partial sum). */
int pt_checkpoint_save(pt_context* ctx, const char* path);
int pt_checkpoint_load(pt_context* ctx, const char* path);
/* Renderer::Present (Present.frag:13-37): /numSamples, Uncharted-2, gamma 2.2 -> RGB float and/or 8-bit. */
int pt_resolve(pt_context* ctx, uint32_t num_samples, float* rgb_f32_or_null, uint8_t* rgb_u8_or_null);
/* Present without the stall of a swap: the resolve kernel is queued behind the frames rendered so far and its 8-bit image
 * is copied to host_rgb_u8 (width*height*3 bytes, pinned memory for a truly asynchronous copy) on a separate copy stream,
 * so the next RenderFrame overlaps the transfer. The image is complete after pt_synchronize. Device-side the target is
 * double-buffered: at most two presents are in flight. */
int pt_present_async(pt_context* ctx, uint32_t num_samples, uint8_t* host_rgb_u8);
int pt_save_screenshot(pt_context* ctx, const char* path);         /* Renderer::SaveScreenshot (Renderer.cpp:697-705); .png (as the reference) / .ppm / .pfm by extension */

/* ---- fixed ray batches (parity + microbenchmark entry points) ----
 * rays: n x 8 floats {o.xyz, unused, d.xyz, tmax}; hits: n x {t, u, v, triId(int32)}; miss = {tmax, 0, 0, -1}.
 * IfIfClosestHit (common/BVH.glsl:634-767) / IfIfAnyHit (:769-854) over host or device buffers. */
int pt_trace_closest(pt_context* ctx, const float* rays, int64_t n, float* hits);
int pt_trace_any(pt_context* ctx, const float* rays, int64_t n, uint8_t* flags);
int pt_trace_closest_device(pt_context* ctx, const void* d_rays, int64_t n, void* d_hits);
int pt_trace_any_device(pt_context* ctx, const void* d_rays, int64_t n, void* d_flags);

/* NO COUNTERPART IN THE REFERENCE -- added for callers that cannot call cudaMalloc (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid it). Same contract and same results as
 * pt_trace_closest_device (rays always traced in binned order), but the call allocates nothing: the scratch of the binned trace
 * (work cursor, bin counters, per-ray bin ids, permutation) lives in a file-scope __device__ array of fixed size inside the
 * library (csrc/pt_noalloc.cu), so the call issues only memsets and kernel launches on the context's stream (asynchronous, like
 * pt_trace_closest_device). Limits and caveats: at most PT_NOALLOC_MAX_RAYS rays per call (PT_ERR_LIMIT beyond); binary BVH layout
 * only; the array costs ~9.4 MB on every GPU that loads the library, until exit; it is shared by all calls and contexts on a GPU
 * and keeps the last call's contents (each call clears what must start at zero), so calls that overlap on one GPU race -- serialise them. */
#define PT_NOALLOC_MAX_RAYS (1 << 20)
int pt_trace_closest_device_noalloc(pt_context* ctx, const void* d_rays, int64_t n, void* d_hits);
int64_t pt_noalloc_max_rays(void);

/* One primary ray per pixel exactly as RenderFrame's first bounce traces it -- InitRay (Iterative.comp:174-210: pixel jitter +
 * thin-lens sample from the pixel's RNG state, which is NOT advanced) and IfIfClosestHit -- with the hit records
 * {t, u, v, triangle id (int bits, -1 = miss)} delivered in PIXEL order (y * width + x) to host memory. Asynchronous like
 * pt_present_async: the 16-byte-per-pixel copy runs on the copy stream from one of two device buffers, so back-to-back
 * calls overlap tracing and transfer; the data is complete after pt_synchronize. Camera and RNG table must be set.
 * BASELINE.json configs[0] ("1 spp primary closest-hit") without shipping 32 bytes per ray to the device first. */
int pt_trace_primary_async(pt_context* ctx, float* host_hits);

/* ---- per-stage probes (parity of raygen on its own; rng table is advanced like the shader does) ---- */
/* out: n x 12 floats {pixel.x, pixel.y (int bits), o.xyz, d.xyz, 0,0,0,0} for ray indices [first, first+n). */
int pt_raygen_probe(pt_context* ctx, int64_t first, int64_t n, float* out);
/* csrc/pt_parity_libm.h evaluated on the device: out[i] = fn(x[i] [, y[i]]) with fn 0 sin, 1 cos, 2 log, 3 exp, 4 pow (host arrays).
 * The CPU checkers compile the same header; tests require equal bits, which is what makes PT_PARITY_LIBM frames comparable
 * for equality with the reference's shader source (GLSL's sin / cos / log / exp / pow: Iterative.comp:191-192, Microfacet.glsl:25-49). */
int pt_parity_libm_probe(pt_context* ctx, int32_t fn, const float* x, const float* y_or_null, int64_t n, float* out);

int pt_get_counters(pt_context* ctx, pt_counters* out);
/* Instrumented traversal: while on, pt_render_frames / pt_trace_* run a variant of the binary-layout traversal kernels that also
 * counts node-pair and triangle tests per ray kind (same rays, same results; a few percent slower, so timed runs leave it off). */
int pt_set_visit_counting(pt_context* ctx, int on);
/* CUDA-event milliseconds of the last pt_render_frames / pt_trace_*_device call (device time only). */
float pt_last_device_ms(pt_context* ctx);
/* Per-stage device time of the last pt_render_frames call, from CUDA events recorded on the library's stream
 * between kernels: out = {raygen, extend, shade, shadow, tally} milliseconds summed over the call's frames. */
int pt_set_stage_timing(pt_context* ctx, int on);
int pt_get_stage_ms(pt_context* ctx, float out[5]);
/* cudaStream_t the library launches on (for callers that record their own events). */
void* pt_stream(pt_context* ctx);

/* ---- several GPUs: samples split across ranks, one NCCL reduce per batch ----
 * RenderFrame (Renderer.cpp:651-662) adds one sample to every pixel; with N GPUs every rank adds its own samples of every
 * pixel to a private sum (scene + BVH replicated, RNG tables decorrelated by rank) and a batch of frames ends with ONE
 * ncclReduce(sum) of the RGBA32F sums to rank 0, which keeps the running total. Two batch accumulators alternate and the
 * reduce runs on its own stream, so the reduce of batch k overlaps the rendering of batch k + 1. fp32 addition order differs
 * from a 1-GPU run: N-GPU images are compared by relMSE, not bitwise. NCCL is bound at run time (dlopen of libnccl.so.2,
 * the copy already in the process if there is one; PT_NCCL_LIB overrides); contexts without a communicator never load it.
 *
 * One process per GPU (torchrun, MPI): rank 0 calls pt_comm_unique_id, the caller's launcher moves the 128 bytes to the
 * other ranks, every rank calls pt_comm_init on its own context. */
#define PT_COMM_ID_BYTES 128
int pt_comm_unique_id(uint8_t id[PT_COMM_ID_BYTES]);
int pt_comm_init(pt_context* ctx, const uint8_t id[PT_COMM_ID_BYTES], int32_t rank, int32_t world);
int32_t pt_comm_rank(pt_context* ctx);
int32_t pt_comm_world(pt_context* ctx);
int pt_comm_nccl_version(void);                                   /* e.g. 22809; 0 when NCCL cannot be loaded */
uint64_t pt_rank_seed(uint64_t seed, int32_t rank);               /* the seed rank `rank` passes to pt_seed_rng (host arithmetic) */
void pt_split_samples(uint32_t total_spp, int32_t world, int32_t rank, uint32_t* first, uint32_t* count);  /* rank's share of a fixed total */
int pt_comm_seed_rng(pt_context* ctx, uint64_t seed);             /* pt_seed_rng(ctx, pt_rank_seed(seed, rank)) */
/* The tree is built once: the root rank's BVH (any builder) reaches the others as device-to-device broadcasts of the traversal
 * layout and the serialised form; pt_set_triangles must have been called on every rank. Binary layout only. */
int pt_comm_broadcast_bvh(pt_context* ctx, int32_t root);
/* n_frames on this rank (and on every other rank: a collective call), then the batch's reduce to rank 0, asynchronous. */
int pt_render_batch(pt_context* ctx, int32_t n_frames);
int pt_comm_reset(pt_context* ctx);                               /* ResetSamples for the running total */
uint32_t pt_comm_num_samples(pt_context* ctx);                    /* samples per pixel in the total: world x frames so far */
int pt_comm_synchronize(pt_context* ctx);
int pt_comm_download_total(pt_context* ctx, float* rgba);         /* rank 0: the running total (sum); other ranks: no-op */
int pt_comm_download_batch(pt_context* ctx, float* rgba);         /* this rank's own sum of the last batch (what it sent) */
void* pt_comm_total_device_ptr(pt_context* ctx);                  /* rank 0, else NULL */
float pt_comm_last_batch_ms(pt_context* ctx);                     /* last pt_render_batch on this rank: first kernel .. end of its reduce */
float pt_comm_last_reduce_ms(pt_context* ctx);                    /* CUDA-event time of the last ncclReduce on its stream */
int pt_comm_present_async(pt_context* ctx, uint8_t* host_rgb_u8); /* rank 0: Present of total / samples, as pt_present_async */

/* One process, one host thread, n GPUs (ncclCommInitAll): the form b200pt::Renderer::Initialize(.., n_gpus) uses.
 * device_ids NULL = 0 .. n-1. n = 1 is a plain context without NCCL. Load or upload the scene into pt_group_context(g, 0),
 * then pt_group_share_scene copies tables, textures, environment and the BVH as built to the other GPUs. */
typedef struct pt_group pt_group;
int pt_create_multi(const pt_config* cfg, const int32_t* device_ids, int32_t n, pt_group** out);
void pt_group_destroy(pt_group* g);
int32_t pt_group_size(pt_group* g);
pt_context* pt_group_context(pt_group* g, int32_t i);
int pt_group_share_scene(pt_group* g);
int pt_group_set_camera(pt_group* g, const pt_camera* cam);
int pt_group_seed_rng(pt_group* g, uint64_t seed);
int pt_group_render_batch(pt_group* g, int32_t n_frames_per_gpu);
int pt_group_reset_samples(pt_group* g);
uint32_t pt_group_num_samples(pt_group* g);
int pt_group_synchronize(pt_group* g);
int pt_group_download_total(pt_group* g, float* rgba);
int pt_group_present_async(pt_group* g, uint8_t* host_rgb_u8);
int pt_group_save_screenshot(pt_group* g, const char* path);      /* Renderer::SaveScreenshot of the combined image */

#ifdef __cplusplus
}
#endif
#endif
