// pt_api.cu -- implementation of the C ABI declared in include/b200pt.h.
// Owns the context (device buffers, stream, events) and the host->device re-layout of the scene.
#include "pt_context.h"
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <cstdarg>
#include <string>
#include <vector>
#include <chrono>
#include <cstdlib>
#include <climits>
#include <exception>

using namespace pt;

static thread_local std::string g_err;
namespace pt { void set_last_error(const char* msg) { g_err = msg ? msg : ""; } }

int pt_fail(int code, const char* fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof(buf), fmt, ap); va_end(ap);
    g_err = buf;
    return code;
}

// every entry point that touches device state the slices may still be working on: select the device, then make `stream`
// wait for the slices (a device-side wait; nothing to do unless chained frames are pending)
#define PT_ENTER(c) do { CK(cudaSetDevice((c)->cfg.device)); int rcj__ = pt_join_slices(c); if (rcj__ != PT_OK) return rcj__; } while (0)

void pt_invalidate_graphs(pt_context* c) {
    if (c->graph.exec) cudaGraphExecDestroy(c->graph.exec);
    c->graph.exec = nullptr; c->graph.launches = 0;
}

SceneDev pt_scene_dev(pt_context* c) {
    SceneDev s{};
    s.pairs = c->d_pairs.p; s.leaf_tris = c->d_leaf.p; s.n_leaf_tris = c->n_leaf; s.top_pairs = c->top_pairs;
    s.wide_nodes = c->d_wide.p; s.wide_tris = c->d_wide_tris.p;
    s.tris = c->d_tris.p; s.materials = c->d_mats.p; s.textures = c->d_texrec.p; s.n_textures = c->n_textures;
    s.lights = c->d_lights.p; s.n_lights = c->n_lights; s.total_light_area = c->total_light_area;
    s.env[0] = c->env[0]; s.env[1] = c->env[1]; s.env[2] = c->env[2];
    s.env_faces = c->d_env.p; s.env_size = c->env_size;
    s.visits = c->count_visits ? c->d_visits.p : nullptr;
    s.sun_dir[0] = c->sun_dir[0]; s.sun_dir[1] = c->sun_dir[1]; s.sun_dir[2] = c->sun_dir[2]; s.sun_radius = c->sun_radius;
    return s;
}

extern "C" {

int pt_abi_version(void) { return B200PT_ABI_VERSION; }
const char* pt_last_error(void) { return g_err.c_str(); }

int pt_create(const pt_config* cfg, pt_context** out) {
    if (!cfg || !out) return fail(PT_ERR_INVALID, "pt_create: null argument");
    *out = nullptr;
    if (cfg->width <= 0 || cfg->height <= 0 || (cfg->width % 8) || (cfg->height % 8))
        return fail(PT_ERR_INVALID, "pt_create: width/height must be positive multiples of 8 (got %dx%d)", cfg->width, cfg->height);
    if (cfg->max_depth < 1 || cfg->max_depth > 66)
        return fail(PT_ERR_INVALID, "pt_create: max_depth must be in [1, 66] (got %d)", cfg->max_depth);
    if (cfg->microfacet_model != PT_MICROFACET_GGX && cfg->microfacet_model != PT_MICROFACET_BECKMANN)
        return fail(PT_ERR_INVALID, "pt_create: unknown microfacet model %d", cfg->microfacet_model);
    if (cfg->bvh_layout != PT_BVH_REFERENCE && cfg->bvh_layout != PT_BVH_WIDE && cfg->bvh_layout != PT_BVH_WIDE4)
        return fail(PT_ERR_INVALID, "pt_create: unknown bvh_layout %d", cfg->bvh_layout);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0)
        return fail(PT_ERR_NO_DEVICE, "pt_create: no CUDA device (%s); libb200pt has no CPU path",
                    e != cudaSuccess ? cudaGetErrorString(e) : "device count 0");
    if (cfg->device < 0 || cfg->device >= ndev) return fail(PT_ERR_INVALID, "pt_create: device %d of %d", cfg->device, ndev);
    cudaDeviceProp prop;
    CK(cudaGetDeviceProperties(&prop, cfg->device));
    if (prop.major != 10)
        return fail(PT_ERR_NO_DEVICE, "pt_create: device %d is sm_%d%d; this library carries sm_100a code only", cfg->device, prop.major, prop.minor);
    CK(cudaSetDevice(cfg->device));
    pt_context* c = new pt_context();
    c->cfg = *cfg;
    {   // Renderer.cpp:23-26: sunDir = normalize(vec3(2, 40 + 29, 12)), sunRadius = tan(radians(5)), all in binary32
        const float v[3] = {2.0f, 40.0f + 29.0f, 12.0f};
        const float inv = 1.0f / sqrtf(v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        for (int k = 0; k < 3; k++) c->sun_dir[k] = v[k] * inv;
        c->sun_radius = tanf(5.0f * 0.01745329251994329576923690768489f);
    }
    c->l2_bytes = (size_t)prop.l2CacheSize;
    if (const char* e = getenv("PT_RAY_ORDER")) c->ray_order = atoi(e);
    if (const char* e = getenv("PT_TRACE_NULL_SHADOW")) c->cfg.trace_null_shadow = atoi(e);
    CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CK(cudaEventCreate(&c->ev0));
    CK(cudaEventCreate(&c->ev1));
    query_launch_dims(cfg->device, cfg->bvh_layout == PT_BVH_WIDE ? 8 : cfg->bvh_layout == PT_BVH_WIDE4 ? 4 : 0, &c->dims);
    const size_t npx = (size_t)cfg->width * cfg->height;
    {
        int want = cfg->n_slices > 0 ? cfg->n_slices : 3;
        if (const char* e = getenv("PT_SLICES")) want = atoi(e);
        if (want < 1) want = 1;
        if (want > 16) want = 16;
        size_t sz = (npx + want - 1) / want;
        sz = (sz + 63) / 64 * 64;                               // whole 8x8 Morton tiles (Iterative.comp:163-172)
        if (cfg->width % 8 == 0) {
            // whole tile ROWS (8 pixel rows): a slice is then one contiguous band of the accumulator, which is what lets a slice
            // clear and reduce its own pixels without waiting for the others (pt_comm.cu). 1080p and 4K split this way anyway.
            const size_t band = (size_t)cfg->width * 8;
            sz = (sz + band - 1) / band * band;
            c->slices_banded = true;
        }
        c->slice_size = (int)sz;
        c->n_slices = (int)((npx + sz - 1) / sz);
        c->slice_streams.resize(c->n_slices); c->slice_done.resize(2 * c->n_slices);
        for (int s = 0; s < c->n_slices; s++) {
            SliceStreams& ss = c->slice_streams[s];
            CK(cudaStreamCreateWithFlags(&ss.a, cudaStreamNonBlocking));
            CK(cudaStreamCreateWithFlags(&ss.b, cudaStreamNonBlocking));
            CK(cudaEventCreateWithFlags(&ss.ev_shade, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&ss.ev_shadow, cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->slice_done[2 * s], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->slice_done[2 * s + 1], cudaEventDisableTiming));
        }
        CK(cudaEventCreateWithFlags(&c->ev_fork, cudaEventDisableTiming));
        if (const char* e = getenv("PT_NO_GRAPH")) c->use_graph = atoi(e) == 0;
        if (const char* e = getenv("PT_LAZY_JOIN")) c->lazy_join = atoi(e) != 0;
        c->slice_presented.resize(c->n_slices);
        for (int s = 0; s < c->n_slices; s++) CK(cudaEventCreateWithFlags(&c->slice_presented[s], cudaEventDisableTiming));
        CK(c->d_cam.alloc(PT_CAM_RING));
        CK(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
        CK(cudaEventCreateWithFlags(&c->ev_resolved, cudaEventDisableTiming));
        for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&c->ev_copied[k], cudaEventDisableTiming));
        CK(cudaEventCreateWithFlags(&c->ev_primary, cudaEventDisableTiming));
        for (int k = 0; k < 2; k++) CK(cudaEventCreateWithFlags(&c->ev_primary_copied[k], cudaEventDisableTiming));
    }
    CK(c->d_rng.alloc(npx));
    CK(c->d_accum_own.alloc(npx));
    c->d_accum = c->d_accum_own.p;
    CK(cudaMemsetAsync(c->d_accum, 0, npx * sizeof(float4), c->stream));
    for (int q = 0; q < 2; q++) {
        for (int k = 0; k < 4; k++) CK(c->d_q[q][k].alloc(npx));
        CK(c->d_qrng[q].alloc(npx));
    }
    for (int k = 0; k < 4; k++) CK(c->d_shadow[k].alloc(npx));
    CK(c->d_hits.alloc(npx));
    for (int k = 0; k < 2; k++) CK(c->d_present[k].alloc(npx * 3));    // present targets up front: a cudaMalloc inside the frame loop drains the device
    CK(c->d_counters.alloc((size_t)2 * c->n_slices * (cfg->max_depth + 2)));   // 2 blocks per slice: frame parity
    CK(c->d_totals.alloc(4));
    CK(cudaMemsetAsync(c->d_totals.p, 0, 4 * sizeof(unsigned long long), c->stream));
    CK(c->d_cursor.alloc(4));
    CK(cudaStreamSynchronize(c->stream));
    *out = c;
    return PT_OK;
}

void pt_destroy(pt_context* c) {
    if (!c) return;
    pt_comm_release(c);
    cudaSetDevice(c->cfg.device);
    cudaStreamSynchronize(c->stream);
    c->d_wide.release(); c->d_wide_tris.release(); c->d_ser_nodes.release(); c->d_ser_refs.release(); c->d_build_ws.release();
    c->d_tris.release(); c->d_pairs.release(); c->d_leaf.release(); c->d_mats.release(); c->d_texrec.release();
    for (float4* t : c->d_texels) cudaFree(t);
    c->d_lights.release(); c->d_env.release(); c->d_rng.release(); c->d_accum_own.release();
    for (int q = 0; q < 2; q++) { for (int k = 0; k < 4; k++) c->d_q[q][k].release(); c->d_qrng[q].release(); }
    for (int k = 0; k < 4; k++) c->d_shadow[k].release();
    c->d_hits.release(); c->d_counters.release(); c->d_totals.release(); c->d_cursor.release(); c->d_visits.release();
    c->d_stage_rays.release(); c->d_stage_hits.release(); c->d_stage_flags.release();
    c->d_ray_bins.release(); c->d_ray_bin.release(); c->d_ray_perm.release();
    c->d_resolve_f32.release(); c->d_resolve_u8.release();
    for (cudaEvent_t e : c->stage_events) cudaEventDestroy(e);
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1);
    for (SliceStreams& ss : c->slice_streams) {
        cudaStreamSynchronize(ss.a); cudaStreamSynchronize(ss.b);
        cudaStreamDestroy(ss.a); cudaStreamDestroy(ss.b);
        cudaEventDestroy(ss.ev_shade); cudaEventDestroy(ss.ev_shadow);
    }
    for (cudaEvent_t e : c->slice_done) cudaEventDestroy(e);
    for (cudaEvent_t e : c->slice_presented) cudaEventDestroy(e);
    if (c->graph.exec) cudaGraphExecDestroy(c->graph.exec);
    c->d_cam.release();
    if (c->copy_stream) { cudaStreamSynchronize(c->copy_stream); cudaStreamDestroy(c->copy_stream); }
    if (c->in_stream) { cudaStreamSynchronize(c->in_stream); cudaStreamDestroy(c->in_stream); }
    for (int k = 0; k < 2; k++) {
        if (c->ev_h_in[k]) cudaEventDestroy(c->ev_h_in[k]);
        if (c->ev_h_done[k]) cudaEventDestroy(c->ev_h_done[k]);
        if (c->ev_h_out[k]) cudaEventDestroy(c->ev_h_out[k]);
    }
    if (c->ev_resolved) cudaEventDestroy(c->ev_resolved);
    if (c->ev_primary) cudaEventDestroy(c->ev_primary);
    for (int k = 0; k < 2; k++) { if (c->ev_primary_copied[k]) cudaEventDestroy(c->ev_primary_copied[k]); c->d_primary[k].release(); }
    for (int k = 0; k < 2; k++) { if (c->ev_copied[k]) cudaEventDestroy(c->ev_copied[k]); c->d_present[k].release(); }
    if (c->ev_fork) cudaEventDestroy(c->ev_fork);
    cudaStreamDestroy(c->stream);
    delete c;
}

// host mirror of a device-built tree (nodesVec / referenceVec of the reference), fetched when first needed
int pt_ensure_host_bvh(pt_context* c) {
    if (!c->host_bvh_stale) return PT_OK;
    HostBvh b;
    b.nodes.resize(8 * (size_t)c->ser_nodes); b.refs.resize((size_t)c->ser_refs);
    CK(cudaMemcpyAsync(b.nodes.data(), c->d_ser_nodes.p, b.nodes.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(b.refs.data(), c->d_ser_refs.p, b.refs.size() * 4, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    b.max_depth = c->h_bvh.max_depth; b.n_leaves = c->h_bvh.n_leaves;
    b.sah_cost = bvh_sah_cost(b.nodes.data(), (int)c->ser_nodes, b.refs.data(), (int)c->ser_refs);
    c->h_bvh = std::move(b);
    c->host_bvh_stale = false;
    return PT_OK;
}

// PT_BVH_WIDE: collapse the reference-form tree into 8-wide quantised nodes + their own leaf triangle stream
static int upload_wide(pt_context* c) {
    std::string msg;
    auto t0 = std::chrono::steady_clock::now();
    if (!build_wide_bvh(c->h_bvh, c->h_tris.data(), (int64_t)(c->h_tris.size() / 20), c->dims.wide, c->h_wide, msg))
        return fail(PT_ERR_LIMIT, "wide BVH: %s", msg.c_str());
    c->wide_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    static_assert(sizeof(WideNodeHost) == sizeof(WideNode), "wide node layout");
    const size_t nb = c->h_wide.bytes.size(), nt = c->h_wide.tri_ids.size();
    std::vector<LeafTri> ls(nt ? nt : 1);
    for (size_t i = 0; i < nt; i++) {
        int32_t id = c->h_wide.tri_ids[i];
        const float* t = c->h_tris.data() + 20 * (size_t)id;
        float idf; memcpy(&idf, &id, 4);
        ls[i].r[0] = make_float4(t[0], t[1], t[2], t[3]);
        ls[i].r[1] = make_float4(t[4], t[5], t[6], t[7]);
        ls[i].r[2] = make_float4(t[8], idf, 0.0f, 0.0f);
        ls[i].r[3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    CK(c->d_wide.alloc(nb));
    CK(c->d_wide_tris.alloc(ls.size()));
    CK(cudaMemcpyAsync(c->d_wide.p, c->h_wide.bytes.data(), nb, cudaMemcpyHostToDevice, c->stream));
    CK(cudaMemcpyAsync(c->d_wide_tris.p, ls.data(), ls.size() * sizeof(LeafTri), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PT_OK;
}

// (re)build the leaf triangle stream from host triangles + reference list
static int upload_leaf_stream(pt_context* c) {
    { int rc = pt_ensure_host_bvh(c); if (rc != PT_OK) return rc; }
    if (c->h_tris.empty() || c->h_bvh.refs.empty()) return PT_OK;
    const size_t nr = c->h_bvh.refs.size();
    const int64_t nt = (int64_t)(c->h_tris.size() / 20);
    std::vector<LeafTri> ls(nr);
    for (size_t i = 0; i < nr; i++) {
        int32_t r = c->h_bvh.refs[i];
        bool last = r < 0;
        int32_t idx = last ? -r : r;
        if (idx >= nt) return fail(PT_ERR_INVALID, "BVH reference %d out of range (%lld triangles)", idx, (long long)nt);
        const float* t = c->h_tris.data() + 20 * (size_t)idx;
        int32_t id = last ? ~idx : idx;
        float idf; memcpy(&idf, &id, 4);
        ls[i].r[0] = make_float4(t[0], t[1], t[2], t[3]);
        ls[i].r[1] = make_float4(t[4], t[5], t[6], t[7]);
        ls[i].r[2] = make_float4(t[8], idf, 0.0f, 0.0f);
        ls[i].r[3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    CK(c->d_leaf.alloc(nr));
    CK(cudaMemcpyAsync(c->d_leaf.p, ls.data(), nr * sizeof(LeafTri), cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->n_leaf = (int)nr;
    if (c->cfg.bvh_layout != PT_BVH_REFERENCE) return upload_wide(c);
    return PT_OK;
}

// The tables index each other on the device without bounds checks (make_material: matId >> 1; random_light_vertex: the light
// entry's triangle), so inconsistent uploads are refused here, whichever table arrives last.
static int check_mat_ids(const float* tris, int64_t n, size_t n_mats, const char* who) {
    if (!n_mats) return PT_OK;
    for (int64_t i = 0; i < n; i++) {
        uint32_t id; memcpy(&id, tris + 20 * i + 18, 4);
        if ((size_t)(id >> 1) >= n_mats) return fail(PT_ERR_INVALID, "%s: triangle %lld has matId %u, but there are %zu materials (matId = 2 * k, Scene.cpp:153-158)", who, (long long)i, id, n_mats);
    }
    return PT_OK;
}
static int check_light_tris(const pt_light* lights, int32_t n, int64_t n_tris, const char* who) {
    if (n_tris <= 0) return PT_OK;
    for (int32_t i = 0; i < n; i++)
        if ((int64_t)lights[i].tri_index >= n_tris) return fail(PT_ERR_INVALID, "%s: light %d names triangle %u of %lld", who, i, lights[i].tri_index, (long long)n_tris);
    return PT_OK;
}

int pt_set_triangles(pt_context* c, const float* tris, int64_t n) {
    if (!c || !tris || n <= 0) return fail(PT_ERR_INVALID, "pt_set_triangles: bad argument");
    if (n >= (1ll << 30)) return fail(PT_ERR_LIMIT, "pt_set_triangles: too many triangles");
    int rc = check_mat_ids(tris, n, c->h_mats.size(), "pt_set_triangles");
    if (rc == PT_OK) rc = check_light_tris(c->h_lights.data(), (int32_t)c->h_lights.size(), n, "pt_set_triangles");
    if (rc != PT_OK) return rc;
    PT_ENTER(c);
    pt_invalidate_graphs(c);
    c->h_tris.assign(tris, tris + 20 * n);
    CK(c->d_tris.alloc(5 * (size_t)n));
    CK(cudaMemcpyAsync(c->d_tris.p, tris, 80 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return upload_leaf_stream(c);
}

int pt_set_materials(pt_context* c, const pt_material* mats, int32_t n_mats, const pt_texture* tex, int32_t n_tex) {
    if (!c || !mats || n_mats <= 0 || n_tex < 0 || (n_tex > 0 && !tex)) return fail(PT_ERR_INVALID, "pt_set_materials: bad argument");
    static_assert(sizeof(pt_material) == sizeof(MaterialRec), "material record layout");
    // everything is validated and the new texture set is built beside the old one; the context changes only on success
    for (int i = 0; i < n_tex; i++)
        if (tex[i].width <= 0 || tex[i].height <= 0 || !tex[i].rgba) return fail(PT_ERR_INVALID, "pt_set_materials: texture %d invalid", i);
    int rc = check_mat_ids(c->h_tris.data(), (int64_t)(c->h_tris.size() / 20), (size_t)n_mats, "pt_set_materials");
    if (rc != PT_OK) return rc;
    PT_ENTER(c);
    std::vector<float4*> texels;
    std::vector<TextureRec> recs(n_tex);
    DevBuf<MaterialRec> mats_new; DevBuf<TextureRec> texrec_new;
    cudaError_t e = mats_new.alloc(n_mats);
    if (e == cudaSuccess) e = texrec_new.alloc(n_tex > 0 ? n_tex : 1);
    if (e == cudaSuccess) e = cudaMemcpyAsync(mats_new.p, mats, sizeof(pt_material) * n_mats, cudaMemcpyHostToDevice, c->stream);
    for (int i = 0; i < n_tex && e == cudaSuccess; i++) {
        size_t cnt = (size_t)tex[i].width * tex[i].height;
        float4* d = nullptr;
        e = cudaMalloc((void**)&d, cnt * sizeof(float4));
        if (e != cudaSuccess) break;
        texels.push_back(d);
        e = cudaMemcpyAsync(d, tex[i].rgba, cnt * sizeof(float4), cudaMemcpyHostToDevice, c->stream);
        recs[i].width = tex[i].width; recs[i].height = tex[i].height; recs[i].texels = d;
    }
    if (e == cudaSuccess && n_tex > 0) e = cudaMemcpyAsync(texrec_new.p, recs.data(), sizeof(TextureRec) * n_tex, cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) {
        cudaStreamSynchronize(c->stream);
        for (float4* t : texels) cudaFree(t);
        mats_new.release(); texrec_new.release();
        return fail(PT_ERR_CUDA, "pt_set_materials: %s (the previous materials stay in place)", cudaGetErrorString(e));
    }
    pt_invalidate_graphs(c);
    CK(cudaDeviceSynchronize());                                 // nothing in flight reads the old set
    for (float4* t : c->d_texels) cudaFree(t);
    c->d_mats.release(); c->d_texrec.release();
    c->d_texels.swap(texels);
    c->d_mats = mats_new; c->d_texrec = texrec_new;              // ownership moves (DevBuf is a plain pointer + count)
    c->h_mats.assign(mats, mats + n_mats);
    c->h_texels.clear(); c->h_texdims.clear();
    for (int i = 0; i < n_tex; i++) {
        c->h_texels.emplace_back(tex[i].rgba, tex[i].rgba + 4 * (size_t)tex[i].width * tex[i].height);
        c->h_texdims.emplace_back(tex[i].width, tex[i].height);
    }
    c->n_textures = n_tex;
    return PT_OK;
}

int pt_set_lights(pt_context* c, const pt_light* lights, int32_t n, float total_light_area) {
    if (!c || !lights || n <= 0) return fail(PT_ERR_INVALID, "pt_set_lights: a scene needs at least one emissive triangle (Iterative.comp:42-61)");
    int rc = check_light_tris(lights, n, (int64_t)(c->h_tris.size() / 20), "pt_set_lights");
    if (rc != PT_OK) return rc;
    PT_ENTER(c);
    static_assert(sizeof(pt_light) == sizeof(float2), "light record layout");
    pt_invalidate_graphs(c);
    CK(c->d_lights.alloc(n));
    CK(cudaMemcpyAsync(c->d_lights.p, lights, sizeof(pt_light) * n, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->h_lights.assign(lights, lights + n);
    c->n_lights = n; c->total_light_area = total_light_area;
    return PT_OK;
}

int pt_set_environment_color(pt_context* c, const float rgb[3]) {
    if (!c || !rgb) return fail(PT_ERR_INVALID, "pt_set_environment_color: bad argument");
    pt_invalidate_graphs(c);
    c->env[0] = rgb[0]; c->env[1] = rgb[1]; c->env[2] = rgb[2];
    if (c->env_size) { PT_ENTER(c); CK(cudaDeviceSynchronize()); c->d_env.release(); c->env_size = 0; }
    return PT_OK;
}

int pt_set_sun(pt_context* c, const float direction[3], float radius) {
    if (!c || !direction || !(radius > 0.0f)) return fail(PT_ERR_INVALID, "pt_set_sun: bad argument");
    pt_invalidate_graphs(c);
    for (int k = 0; k < 3; k++) c->sun_dir[k] = direction[k];
    c->sun_radius = radius;
    return PT_OK;
}

int pt_set_environment_cubemap(pt_context* c, const float* faces, int32_t size) {
    if (!c || !faces || size < 1 || size > 16384) return fail(PT_ERR_INVALID, "pt_set_environment_cubemap: bad argument");
    PT_ENTER(c);
    CK(cudaDeviceSynchronize());
    pt_invalidate_graphs(c);
    CK(c->d_env.alloc(6 * (size_t)size * size));
    CK(cudaMemcpy(c->d_env.p, faces, 6 * (size_t)size * size * sizeof(float4), cudaMemcpyHostToDevice));
    c->env_size = size;
    return PT_OK;
}

int pt_set_environment_equirect(pt_context* c, const float* rgba, int32_t width, int32_t height, int32_t cube_size) {
    if (!c || !rgba || width < 1 || height < 1 || cube_size < 1 || cube_size > 16384)
        return fail(PT_ERR_INVALID, "pt_set_environment_equirect: bad argument");
    PT_ENTER(c);
    CK(cudaDeviceSynchronize());
    pt_invalidate_graphs(c);
    DevBuf<float4> src;
    CK(src.alloc((size_t)width * height));
    cudaError_t e = cudaMemcpy(src.p, rgba, (size_t)width * height * sizeof(float4), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = c->d_env.alloc(6 * (size_t)cube_size * cube_size);
    if (e == cudaSuccess) {
        launch_equirect_to_cube(src.p, width, height, c->d_env.p, cube_size, c->stream);
        e = cudaGetLastError();
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    }
    src.release();
    if (e != cudaSuccess) { c->env_size = 0; return fail(PT_ERR_CUDA, "pt_set_environment_equirect: %s", cudaGetErrorString(e)); }
    c->env_size = cube_size;
    return PT_OK;
}

int32_t pt_environment_size(pt_context* c) { return c ? c->env_size : 0; }

int pt_download_environment(pt_context* c, float* faces) {
    if (!c || !faces) return fail(PT_ERR_INVALID, "pt_download_environment: bad argument");
    if (!c->env_size) return fail(PT_ERR_INVALID, "pt_download_environment: the environment is a constant colour");
    PT_ENTER(c);
    CK(cudaDeviceSynchronize());
    CK(cudaMemcpy(faces, c->d_env.p, 6 * (size_t)c->env_size * c->env_size * sizeof(float4), cudaMemcpyDeviceToHost));
    return PT_OK;
}

// LoadEnvironmnet (Renderer.cpp:236-323): "GENERATE COLOR x" | *.hdr (equirect -> 1024^2 cubemap) | text file naming six faces
static int load_environment_impl(pt_context* c, const char* args);
int pt_load_environment(pt_context* c, const char* args) {
    // image headers drive allocations: std::bad_alloc / length_error from a damaged file must not cross the C ABI
    try { return load_environment_impl(c, args); }
    catch (const std::exception& e) { return fail(PT_ERR_LIMIT, "pt_load_environment: %s", e.what()); }
}
static int load_environment_impl(pt_context* c, const char* args) {
    if (!c || !args) return fail(PT_ERR_INVALID, "pt_load_environment: bad argument");
    std::string a = args;
    std::string ext = a.substr(a.find_last_of('.') + 1);       // npos + 1 == 0: the whole string, as in the reference
    if (a.find_first_of("GENERATE") == 0) {
        float rgb[3]; int rc = pt_parse_environment(args, rgb);
        if (rc != PT_OK) return rc;
        return pt_set_environment_color(c, rgb);
    }
    if (ext == "hdr") {
        int w = 0, h = 0; std::vector<float> px; std::string err;
        if (!load_hdr_rgba32f(a, w, h, px, err)) return fail(PT_ERR_IO, "pt_load_environment: %s", err.c_str());
        return pt_set_environment_equirect(c, px.data(), w, h, 1024);
    }
    if (ext == "jpg") {                                           // stbi_loadf on an 8-bit file: gamma-2.2 expansion, then the same passes
        int w = 0, h = 0; std::vector<float> px; std::string err;
        if (!load_ldr_as_hdr_rgba32f(a, w, h, px, err)) return fail(PT_ERR_IO, "pt_load_environment: %s", err.c_str());
        return pt_set_environment_equirect(c, px.data(), w, h, 1024);
    }
    // TextureCubemap::LoadTexture (Texture.cpp:245-276): six lines (those starting with '#' skipped), paths relative
    // to the list's folder, order +X -X +Y -Y +Z -Z, each decoded to RGBA bytes -> normalised floats
    std::string path = a;
    for (char& ch : path) if (ch == '\\') ch = '/';
    std::string folder = path.substr(0, path.find_last_of('/') + 1);
    FILE* f = fopen(path.c_str(), "rb");
    if (!f) return fail(PT_ERR_IO, "pt_load_environment: cannot open %s", path.c_str());
    std::vector<std::string> lines; { std::string cur; int ch; while ((ch = fgetc(f)) != EOF) { if (ch == '\n') { lines.push_back(cur); cur.clear(); } else cur.push_back((char)ch); } if (!cur.empty()) lines.push_back(cur); }
    fclose(f);
    std::vector<float> faces; int size = 0, got = 0;
    for (size_t li = 0; li < lines.size() && got < 6; li++) {
        std::string fp = lines[li];
        if (!fp.empty() && fp.back() == '\r') fp.pop_back();
        if (fp.empty()) return fail(PT_ERR_INVALID, "pt_load_environment: %s: empty line in the face list", path.c_str());
        if (fp.front() == '#') continue;
        int w = 0, h = 0; std::vector<float> px;
        if (!load_image_rgba32f(folder + fp, w, h, px)) return fail(PT_ERR_IO, "pt_load_environment: cannot decode %s (supported: .ppm, .bmp, .tga)", (folder + fp).c_str());
        if (w != h || (got && w != size)) return fail(PT_ERR_INVALID, "pt_load_environment: cube faces must be square and of one size (%s is %dx%d)", fp.c_str(), w, h);
        size = w;
        faces.insert(faces.end(), px.begin(), px.end());
        got++;
    }
    if (got != 6) return fail(PT_ERR_INVALID, "pt_load_environment: %s names %d faces, 6 needed", path.c_str(), got);
    return pt_set_environment_cubemap(c, faces.data(), size);
}

static int install_bvh(pt_context* c) {
    pt_invalidate_graphs(c);
    c->host_bvh_stale = false;                                    // the host tree is the authority again
    const float* nodes = c->h_bvh.nodes.data();
    const int n_nodes = (int)(c->h_bvh.nodes.size() / 8);
    const int n_pairs = (n_nodes - 1) / 2;
    std::vector<BvhPair> pairs(n_pairs);
    auto link_of = [](const float* nd) {
        int32_t d0; memcpy(&d0, nd + 3, 4);
        int32_t link = d0 > 0 ? (d0 - 1) / 2 : ~(-d0);
        float f; memcpy(&f, &link, 4); return f;
    };
    for (int pi = 0; pi < n_pairs; pi++) {
        const float* c0 = nodes + 8 * (size_t)(2 * pi + 1);
        const float* c1 = c0 + 8;
        pairs[pi].q[0] = make_float4(c0[0], c0[1], c0[2], link_of(c0));
        pairs[pi].q[1] = make_float4(c0[4], c0[5], c0[6], 0.0f);
        pairs[pi].q[2] = make_float4(c1[0], c1[1], c1[2], link_of(c1));
        pairs[pi].q[3] = make_float4(c1[4], c1[5], c1[6], 0.0f);
    }
    // PT_SMEM_TOP = K: move the K pair records nearest the root to the front, breadth-first, so that a CTA can stage
    // records [0, K) in shared memory with one contiguous copy. Only indices change; what a ray visits does not.
    c->top_pairs = 0;
    if (c->dims.top_pairs > 0 && n_pairs > 1) {
        const int K = c->dims.top_pairs < n_pairs ? c->dims.top_pairs : n_pairs;
        auto link_i = [](float f) { int32_t v; memcpy(&v, &f, 4); return v; };
        std::vector<int> order; order.reserve(n_pairs);
        std::vector<char> taken(n_pairs, 0);
        order.push_back(0); taken[0] = 1;
        for (size_t head = 0; head < order.size() && (int)order.size() < K; head++) {
            const BvhPair& pr = pairs[order[head]];
            for (int k = 0; k < 2 && (int)order.size() < K; k++) {
                int32_t l = link_i(pr.q[2 * k].w);
                if (l >= 1 && !taken[l]) { order.push_back(l); taken[l] = 1; }
            }
        }
        for (int i = 0; i < n_pairs; i++) if (!taken[i]) order.push_back(i);
        std::vector<int> new_index(n_pairs);
        for (int i = 0; i < n_pairs; i++) new_index[order[i]] = i;
        std::vector<BvhPair> re(n_pairs);
        for (int i = 0; i < n_pairs; i++) {
            BvhPair pr = pairs[order[i]];
            for (int k = 0; k < 2; k++) {
                int32_t l = link_i(pr.q[2 * k].w);
                if (l >= 1) { int32_t nl = new_index[l]; memcpy(&pr.q[2 * k].w, &nl, 4); }
            }
            re[i] = pr;
        }
        pairs.swap(re);
        c->top_pairs = K;
    }
    CK(c->d_pairs.alloc(n_pairs));
    CK(cudaMemcpyAsync(c->d_pairs.p, pairs.data(), sizeof(BvhPair) * n_pairs, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return upload_leaf_stream(c);
}

int pt_set_bvh(pt_context* c, const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs) {
    if (!c || !nodes || !refs || n_nodes < 3 || n_refs < 1) return fail(PT_ERR_INVALID, "pt_set_bvh: bad argument (the root must be an inner node, BVH.glsl:627)");
    PT_ENTER(c);
    std::string msg; int n_leaves = 0;
    int64_t nt = c->h_tris.empty() ? -1 : (int64_t)(c->h_tris.size() / 20);
    int depth = validate_bvh(nodes, n_nodes, refs, n_refs, nt, msg, &n_leaves);
    if (depth < 0) return fail(PT_ERR_INVALID, "pt_set_bvh: %s", msg.c_str());
    if (depth >= PT_MAX_TREE_DEPTH) return fail(PT_ERR_LIMIT, "pt_set_bvh: tree depth %d exceeds the traversal stack (%d)", depth, PT_MAX_TREE_DEPTH);
    c->h_bvh.nodes.assign(nodes, nodes + 8 * (size_t)n_nodes);
    c->h_bvh.refs.assign(refs, refs + n_refs);
    c->h_bvh.max_depth = depth; c->h_bvh.n_leaves = n_leaves;
    return install_bvh(c);
}

// Reinsertion optimisation of whatever tree the context holds (host or device built, or injected): read back if need be, optimised on
// the host, installed again. stats: n_nodes / n_refs / n_leaves / max_depth / sah_cost of the result, build_ms = time of the pass.
int pt_optimize_bvh(pt_context* c, int32_t iterations, float batch_fraction, pt_bvh_stats* stats) {
    if (!c) return fail(PT_ERR_INVALID, "pt_optimize_bvh: null context");
    if (iterations < 1 || iterations > 64 || !(batch_fraction > 0.0f) || batch_fraction > 1.0f)
        return fail(PT_ERR_INVALID, "pt_optimize_bvh: iterations 1 .. 64, batch_fraction in (0, 1]");
    PT_ENTER(c);
    int rc = pt_check_scene(c, "pt_optimize_bvh", false);
    if (rc == PT_OK) rc = pt_ensure_host_bvh(c);
    if (rc != PT_OK) return rc;
    auto t0 = std::chrono::steady_clock::now();
    float before = 0.0f, after = 0.0f;
    HostBvh& b = c->h_bvh;
    const int moved = optimize_bvh_reinsertion(b.nodes.data(), (int)(b.nodes.size() / 8), iterations, batch_fraction, PT_MAX_TREE_DEPTH, &before, &after,
                                               b.refs.data(), (int)b.refs.size());
    if (moved < 0) return fail(PT_ERR_LIMIT, "pt_optimize_bvh: the optimised tree would exceed the traversal stack (%d); tree left as it was", PT_MAX_TREE_DEPTH);
    auto t1 = std::chrono::steady_clock::now();
    if (moved > 0) {
        std::string msg; int n_leaves = 0;
        int depth = validate_bvh(b.nodes.data(), (int)(b.nodes.size() / 8), b.refs.data(), (int)b.refs.size(), (int64_t)(c->h_tris.size() / 20), msg, &n_leaves);
        if (depth < 0) return fail(PT_ERR_INVALID, "pt_optimize_bvh: internal error, %s", msg.c_str());
        b.max_depth = depth; b.n_leaves = n_leaves; b.sah_cost = after;
        if ((rc = install_bvh(c)) != PT_OK) return rc;
    }
    if (stats) {
        stats->n_nodes = (int)(b.nodes.size() / 8); stats->n_refs = (int)b.refs.size(); stats->n_leaves = b.n_leaves; stats->max_depth = b.max_depth;
        stats->sah_cost = moved > 0 ? after : before;
        stats->build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
        stats->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return PT_OK;
}

int pt_build_bvh(pt_context* c, pt_bvh_stats* stats) {
    if (!c) return fail(PT_ERR_INVALID, "pt_build_bvh: null context");
    if (c->h_tris.empty()) return fail(PT_ERR_INVALID, "pt_build_bvh: call pt_set_triangles first");
    PT_ENTER(c);
    std::string msg;
    HostBvh b;
    cudaEvent_t dummy; (void)dummy;
    auto t0 = std::chrono::steady_clock::now();
    if (!build_sbvh(c->h_tris.data(), (int64_t)(c->h_tris.size() / 20), 0, b, msg)) return fail(PT_ERR_LIMIT, "pt_build_bvh: %s", msg.c_str());
    auto t1 = std::chrono::steady_clock::now();
    c->h_bvh = std::move(b);
    int rc = install_bvh(c);
    if (rc != PT_OK) return rc;
    if (stats) {
        stats->n_nodes = (int)(c->h_bvh.nodes.size() / 8);
        stats->n_refs = (int)c->h_bvh.refs.size();
        stats->n_leaves = c->h_bvh.n_leaves;
        stats->max_depth = c->h_bvh.max_depth;
        stats->sah_cost = c->h_bvh.sah_cost;
        stats->build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
        stats->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return PT_OK;
}

// The device builder (pt_bvh_gpu.cu): Morton LBVH, one triangle per leaf, reference serialised form
static int build_bvh_on_device(pt_context* c, pt_bvh_stats* stats, int ploc_radius, int split_grid);
int pt_build_bvh_device(pt_context* c, pt_bvh_stats* stats) { return build_bvh_on_device(c, stats, 0, 0); }
int pt_build_bvh_device_ploc(pt_context* c, int32_t radius, pt_bvh_stats* stats) {
    if (radius < 1 || radius > 256) return fail(PT_ERR_INVALID, "pt_build_bvh_device_ploc: radius 1 .. 256");
    return build_bvh_on_device(c, stats, radius, 0);
}
int pt_build_bvh_device_split(pt_context* c, int32_t ploc_radius, int32_t split_grid, pt_bvh_stats* stats) {
    if (ploc_radius < 0 || ploc_radius > 256) return fail(PT_ERR_INVALID, "pt_build_bvh_device_split: ploc_radius 0 (Morton hierarchy) .. 256");
    if (split_grid < 2 || split_grid > 1024) return fail(PT_ERR_INVALID, "pt_build_bvh_device_split: split_grid 2 .. 1024");
    return build_bvh_on_device(c, stats, ploc_radius, split_grid);
}
int pt_build_bvh_device_sah(pt_context* c, pt_bvh_stats* stats) { return build_bvh_on_device(c, stats, -1, 0); }
static int build_bvh_on_device(pt_context* c, pt_bvh_stats* stats, int ploc_radius, int split_grid) {
    if (!c) return fail(PT_ERR_INVALID, "pt_build_bvh_device: null context");
    if (c->h_tris.empty() || !c->d_tris.p) return fail(PT_ERR_INVALID, "pt_build_bvh_device: call pt_set_triangles first");
    const int64_t n = (int64_t)(c->h_tris.size() / 20);
    if (n < 2 || n > (1 << 29)) return fail(PT_ERR_LIMIT, "pt_build_bvh_device: 2 .. 2^29 triangles");
    PT_ENTER(c);
    auto t0 = std::chrono::steady_clock::now();
    pt_invalidate_graphs(c);
    // spatial splits first (pt_build_bvh_device_split): the leaves of the build are then m >= n references
    RefSet refset; float split_ms = 0.0f;
    if (split_grid > 0) {
        int l = presplit_device(c->d_tris.p, (int)n, split_grid, &refset, c->stream, &split_ms);
        if (l < 0) return fail(l == -(int)cudaErrorInvalidValue ? PT_ERR_LIMIT : PT_ERR_CUDA, "pt_build_bvh_device_split: %s",
                               l == -(int)cudaErrorInvalidValue ? "the split grid would more than double the scene; use a coarser split_grid"
                                                                 : cudaGetErrorString((cudaError_t)(-l)));
        c->counters.kernel_launches += l;
    }
    const int64_t m = split_grid > 0 ? refset.m : n;
    struct RefGuard { RefSet* r; ~RefGuard() { free_refset(r); } } guard{&refset};
    // the serialised tree stays on the device (read back only when somebody asks for it: pt_get_bvh, wide collapse);
    // the traversal layout (pair records + leaf stream) is written by the build kernels themselves
    CK(c->d_ser_nodes.alloc(2 * (size_t)(2 * m - 1)));
    CK(c->d_ser_refs.alloc((size_t)m));
    CK(c->d_pairs.alloc((size_t)(m - 1)));
    CK(c->d_leaf.alloc((size_t)m));
    float kernel_ms = 0.0f; int depth = 0, leaves = (int)m;      // the SAH builder makes leaves of several triangles: fewer leaves than references
    const RefSet* rs = split_grid > 0 ? &refset : nullptr;
    // scratch kept in the context: a rebuild per frame pays for kernels, not for ~40 cudaMalloc / cudaFree (if it cannot be had, the builder allocates)
    if (c->d_build_ws.alloc(build_lbvh_workspace_bytes(m)) != cudaSuccess) { c->d_build_ws.release(); cudaGetLastError(); }
    int launches = build_lbvh_device(c->d_tris.p, (int)n, c->d_ser_nodes.p, c->d_ser_refs.p, c->d_pairs.p, c->d_leaf.p, &depth,
                                     c->stream, &kernel_ms, ploc_radius, rs, &leaves, c->d_build_ws.p, c->d_build_ws.n);
    if (launches < 0) return fail(PT_ERR_CUDA, "pt_build_bvh_device: %s", cudaGetErrorString((cudaError_t)(-launches)));
    c->counters.kernel_launches += launches;
    kernel_ms += split_ms;
    if (depth >= PT_MAX_TREE_DEPTH && ploc_radius != 0) {
        // agglomerative clustering does not bound the depth (chains of small clusters joining a big one); the Morton
        // hierarchy does (30 key bits + index tie-break), so it takes over. Not an error: the note is left in pt_last_error().
        fail(PT_OK, "pt_build_bvh_device_%s tree depth %d exceeds the traversal stack (%d); Morton hierarchy used instead", ploc_radius > 0 ? "ploc: PLOC" : "sah: SAH", depth, PT_MAX_TREE_DEPTH);
        float ms2 = 0.0f;
        launches = build_lbvh_device(c->d_tris.p, (int)n, c->d_ser_nodes.p, c->d_ser_refs.p, c->d_pairs.p, c->d_leaf.p, &depth,
                                     c->stream, &ms2, 0, rs, &leaves, c->d_build_ws.p, c->d_build_ws.n);
        if (launches < 0) return fail(PT_ERR_CUDA, "pt_build_bvh_device: %s", cudaGetErrorString((cudaError_t)(-launches)));
        c->counters.kernel_launches += launches;
        kernel_ms += ms2;
    }
    if (depth >= PT_MAX_TREE_DEPTH) {
        c->d_pairs.release(); c->d_leaf.release();
        return fail(PT_ERR_LIMIT, "pt_build_bvh_device: tree depth %d exceeds the traversal stack (%d); use pt_build_bvh", depth, PT_MAX_TREE_DEPTH);
    }
    if (c->d_build_ws.n > ((size_t)512 << 20)) c->d_build_ws.release();      // big one-off builds do not keep gigabytes of scratch
    c->n_leaf = (int)m;
    c->top_pairs = 0;                                             // Karras order: the top levels are not contiguous
    c->h_bvh = HostBvh();
    c->h_bvh.max_depth = depth; c->h_bvh.n_leaves = leaves;
    c->host_bvh_stale = true; c->ser_nodes = 2 * (int64_t)leaves - 1; c->ser_refs = m;
    if (c->cfg.bvh_layout != PT_BVH_REFERENCE) {                 // the wide collapse runs on the host
        int rc = pt_ensure_host_bvh(c);
        if (rc != PT_OK) return rc;
        if ((rc = upload_wide(c)) != PT_OK) return rc;
    }
    if (stats) {
        stats->n_nodes = 2 * leaves - 1; stats->n_refs = (int)m; stats->n_leaves = leaves; stats->max_depth = depth;
        stats->sah_cost = c->host_bvh_stale ? 0.0f : c->h_bvh.sah_cost;     // known once the tree has been read back
        stats->build_ms = kernel_ms;        // CUDA-event time of the build kernels (the tree is usable at that point)
        stats->total_ms = std::chrono::duration<float, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return PT_OK;
}

// Renderer::Initialize's scene half (Renderer.cpp:519-523 -> Scene::LoadScene, Scene.cpp:233-357) from files
static int load_scene_impl(pt_context* c, const char* scene_path, const char* environment, pt_bvh_stats* stats);
int pt_load_scene(pt_context* c, const char* scene_path, const char* environment, pt_bvh_stats* stats) {
    try { return load_scene_impl(c, scene_path, environment, stats); }
    catch (const std::exception& e) { return fail(PT_ERR_LIMIT, "pt_load_scene: %s", e.what()); }
}
static int load_scene_impl(pt_context* c, const char* scene_path, const char* environment, pt_bvh_stats* stats) {
    if (!c || !scene_path) return fail(PT_ERR_INVALID, "pt_load_scene: bad argument");
    std::string path(scene_path);
    size_t dot = path.find_last_of('.');
    if (dot == std::string::npos || path.substr(dot + 1) != "obj")
        return fail(PT_ERR_UNSUPPORTED, "Unsupported file type: %s (the reference's only live loader is .obj, Scene.cpp:248-259)", scene_path);
    {
        int rc = pt_load_environment(c, environment ? environment : "GENERATE COLOR WHITE");
        if (rc != PT_OK) return rc;
    }
    pt_scene_file* sf = nullptr;
    int rc = pt_scene_file_load_obj(scene_path, &sf);
    if (rc != PT_OK) return rc;
    const int64_t n = pt_scene_file_num_triangles(sf);
    std::vector<pt_material> mats(pt_scene_file_materials(sf), pt_scene_file_materials(sf) + pt_scene_file_num_materials(sf));
    // map_Kd images -> RGBA32F textures; a file that cannot be decoded leaves the material on its Kd colour
    std::vector<std::vector<float>> texels; std::vector<pt_texture> tex;
    for (size_t m = 0; m < mats.size(); m++) {
        std::string tp = pt_scene_file_texture_path(sf, (int32_t)m);
        if (tp.empty()) continue;
        int w = 0, h = 0; std::vector<float> px;
        if (!load_image_rgba32f(tp, w, h, px)) continue;
        texels.push_back(std::move(px));
        mats[m].albedo_tex = (int32_t)texels.size() - 1;
        pt_texture t; t.width = w; t.height = h; t.rgba = nullptr; tex.push_back(t);
    }
    for (size_t i = 0; i < tex.size(); i++) tex[i].rgba = texels[i].data();
    std::vector<float> tris((size_t)n * 20);
    rc = pt_assemble_triangles(pt_scene_file_vertices(sf), pt_scene_file_material_ids(sf), n, tris.data());
    pt_scene_file_free(sf);
    if (rc != PT_OK) return fail(rc, "pt_load_scene: triangle assembly failed");
    float total = 0.0f;
    int nl = pt_build_light_table(tris.data(), n, mats.data(), (int32_t)mats.size(), nullptr, 0, &total);
    if (nl <= 0) return fail(PT_ERR_INVALID, "pt_load_scene: %s has no emissive triangle (a material needs Ke > 0; the reference samples garbage then, Iterative.comp:42-61)", scene_path);
    std::vector<pt_light> lights(nl);
    pt_build_light_table(tris.data(), n, mats.data(), (int32_t)mats.size(), lights.data(), nl, &total);
    pt_presubtract_edges(tris.data(), n);
    if ((rc = pt_set_triangles(c, tris.data(), n)) != PT_OK) return rc;
    if ((rc = pt_set_materials(c, mats.data(), (int32_t)mats.size(), tex.empty() ? nullptr : tex.data(), (int32_t)tex.size())) != PT_OK) return rc;
    if ((rc = pt_set_lights(c, lights.data(), nl, total)) != PT_OK) return rc;
    return pt_build_bvh(c, stats);
}

int pt_get_bvh(pt_context* c, float* nodes, int32_t* n_nodes, int32_t* refs, int32_t* n_refs) {
    if (!c || !n_nodes || !n_refs) return fail(PT_ERR_INVALID, "pt_get_bvh: bad argument");
    { PT_ENTER(c); int rc = pt_ensure_host_bvh(c); if (rc != PT_OK) return rc; }
    *n_nodes = (int32_t)(c->h_bvh.nodes.size() / 8);
    *n_refs = (int32_t)c->h_bvh.refs.size();
    if (nodes) memcpy(nodes, c->h_bvh.nodes.data(), c->h_bvh.nodes.size() * 4);
    if (refs) memcpy(refs, c->h_bvh.refs.data(), c->h_bvh.refs.size() * 4);
    return PT_OK;
}

int pt_set_camera(pt_context* c, const pt_camera* cam) {
    if (!c || !cam) return fail(PT_ERR_INVALID, "pt_set_camera: bad argument");
    static_assert(sizeof(pt_camera) == sizeof(CameraRec), "camera record layout");
    memcpy(&c->cam, cam, sizeof(CameraRec));
    CK(cudaSetDevice(c->cfg.device));
    // 76 bytes of uniforms (Shader::LoadCamera). Frames already queued keep their camera: stream order when every frame is
    // joined into `stream`; with chained frames (lazy join) the records live in a ring and a slot is reused only after a join.
    if (c->lazy_join) {
        c->cam_slot = (c->cam_slot + 1) % PT_CAM_RING;
        if (c->cam_slot == 0) { int rc = pt_join_slices(c); if (rc != PT_OK) return rc; }
    }
    CK(cudaMemcpyAsync(c->d_cam.p + c->cam_slot, &c->cam, sizeof(CameraRec), cudaMemcpyHostToDevice, c->stream));
    c->have_cam = true;
    return PT_OK;
}

int pt_set_rng(pt_context* c, const uint32_t* state4) {
    if (!c || !state4) return fail(PT_ERR_INVALID, "pt_set_rng: bad argument");
    PT_ENTER(c);
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    CK(cudaMemcpyAsync(c->d_rng.p, state4, npx * 16, cudaMemcpyHostToDevice, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    c->have_rng = true;
    return PT_OK;
}
int pt_seed_rng(pt_context* c, uint64_t seed) {
    if (!c) return fail(PT_ERR_INVALID, "pt_seed_rng: null context");
    PT_ENTER(c);
    launch_seed_rng(c->d_rng.p, (long long)c->cfg.width * c->cfg.height, seed, c->stream);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(c->stream));
    c->have_rng = true;
    return PT_OK;
}
int pt_get_rng(pt_context* c, uint32_t* state4) {
    if (!c || !state4) return fail(PT_ERR_INVALID, "pt_get_rng: bad argument");
    PT_ENTER(c);
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    CK(cudaMemcpyAsync(state4, c->d_rng.p, npx * 16, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PT_OK;
}

// The persistent kernels hand out ray indices from an int cursor that every resident warp bumps by up to 32 once more after
// the batch is exhausted (148 SMs x 9 CTAs x 4 warps x 32 = 170 k on a B200): batches stop well short of INT_MAX.
static const int64_t kMaxBatchRays = 1ll << 30;
static const int64_t kHostChunk = 1 << 21;      // host batches are traced in chunks of this many rays (pipelined copies)
static const int64_t kRayOrderMin = 1 << 16;    // smaller batches are not worth three extra launches
// Decide whether a batch is traced in binned order and, if so, enqueue the binning; nullptr = natural order.
static int order_rays(pt_context* c, const float4* d_rays, int64_t n, const int** perm) {
    *perm = nullptr;
    if (c->ray_order < 0 || n < (c->ray_order > 0 ? 2 : kRayOrderMin)) return PT_OK;
    if (c->ray_order == 0) {
        // auto: the binary layout once it outgrows L2 (every node step a DRAM round trip). Measured: +26 % on
        // the 10 M-triangle batch; -4 % on an L2-resident tree (binning costs what the trace gains); the wide layouts
        // are issue-bound, the halved DRAM traffic does not pay for the binning there (-4 %).
        if (c->dims.wide) return PT_OK;
        if (c->d_pairs.n * sizeof(BvhPair) + c->d_leaf.n * sizeof(LeafTri) <= c->l2_bytes) return PT_OK;
    }
    CK(c->d_ray_bins.alloc((1 << 21) + 2048));                  // counters + scan tile totals (launch_ray_order)
    CK(c->d_ray_bin.alloc((size_t)n));
    CK(c->d_ray_perm.alloc((size_t)n));
    c->counters.kernel_launches += launch_ray_order(pt_scene_dev(c), d_rays, n, c->d_ray_bins.p, c->d_ray_bin.p, c->d_ray_perm.p, c->stream);
    *perm = c->d_ray_perm.p;
    return PT_OK;
}

int pt_check_scene(pt_context* c, const char* who, bool need_shading) {
    if (!c->d_pairs.p || !c->d_leaf.p || !c->d_tris.p) return fail(PT_ERR_INVALID, "%s: triangles and BVH must be set first", who);
    if (need_shading) {
        if (!c->d_mats.p) return fail(PT_ERR_INVALID, "%s: materials not set", who);
        if (!c->d_lights.p || c->n_lights < 1) return fail(PT_ERR_INVALID, "%s: light table not set", who);
        if (!c->have_cam) return fail(PT_ERR_INVALID, "%s: camera not set", who);
        if (!c->have_rng) return fail(PT_ERR_INVALID, "%s: RNG table not set (pt_set_rng / pt_seed_rng)", who);
    }
    return PT_OK;
}

// parameters of the launch sequence that owns ray indices [first, first + count); counter block `block`
static FrameParams frame_params(pt_context* c, int first = 0, int count = -1, int block = 0) {
    FrameParams p{};
    p.scene = pt_scene_dev(c);
    p.cam = c->d_cam.p + c->cam_slot;
    p.width = c->cfg.width; p.height = c->cfg.height;
    p.slice_first = first; p.slice_count = count < 0 ? c->cfg.width * c->cfg.height : count;
    p.max_depth = c->cfg.max_depth; p.model = c->cfg.microfacet_model; p.sun = c->cfg.sun_light ? 1 : 0; p.trace_null = c->cfg.trace_null_shadow ? 1 : 0;
    p.plibm = (c->cfg.parity_flags & PT_PARITY_LIBM) ? 1 : 0;
    p.rng_table = c->d_rng.p; p.accum = c->d_accum;
    const size_t o = (size_t)first;                              // the slice's sub-range of every per-path array
    for (int q = 0; q < 2; q++) {
        p.queue[q].q0 = c->d_q[q][0].p + o; p.queue[q].q1 = c->d_q[q][1].p + o; p.queue[q].q2 = c->d_q[q][2].p + o;
        p.queue[q].q3 = c->d_q[q][3].p + o; p.queue[q].q4 = c->d_qrng[q].p + o;
    }
    p.shadow.s0 = c->d_shadow[0].p + o; p.shadow.s1 = c->d_shadow[1].p + o; p.shadow.s2 = c->d_shadow[2].p + o; p.shadow.s3 = c->d_shadow[3].p + o;
    p.hits = c->d_hits.p + o;
    p.counters = c->d_counters.p + (size_t)block * (c->cfg.max_depth + 2);
    p.totals = c->d_totals.p;
    return p;
}



// `g` consecutive frames of every slice between a fork from and a join into the main stream. Slices advance through
// their frames independently; stream-capturable (every cross-stream edge is an event).
static int join_impl(pt_context* c, bool mark_time);
static int enqueue_frames_impl(pt_context* c, int g, bool join, cudaEvent_t accum_free = nullptr, bool clear_accum = false) {
    const int npx = c->cfg.width * c->cfg.height;
    int launches = 0;
    const bool chained = c->pending_join;          // the previous frame of every slice is still on its streams, not joined
    cudaEventRecord(c->ev_fork, c->stream);
    for (int s = 0; s < c->n_slices; s++) {
        cudaStream_t a = c->slice_streams[s].a;
        cudaStreamWaitEvent(a, c->ev_fork, 0);
        if (accum_free) cudaStreamWaitEvent(a, accum_free, 0);
        if (clear_accum) {
            // a banded slice's pixels are the accumulator range of its ray indices. The caller alternates between two accumulators
            // (pt_comm.cu): the frames still in flight on stream b write the OTHER one, and stream a is behind everything that
            // wrote this one (its own kernels; the shadow kernels of two batches ago through the ev_shadow wait of the last
            // batch's first shade).
            const int first = s * c->slice_size;
            const int count = npx - first < c->slice_size ? npx - first : c->slice_size;
            cudaMemsetAsync(c->d_accum + first, 0, (size_t)count * sizeof(float4), a);
        }
    }
    for (int f = 0; f < g; f++, c->frame_seq++)
        for (int s = 0; s < c->n_slices; s++) {
            int first = s * c->slice_size;
            int count = npx - first < c->slice_size ? npx - first : c->slice_size;
            FrameParams p = frame_params(c, first, count, 2 * s + (c->frame_seq & 1));
            launches += enqueue_frame_dag(p, c->dims, c->slice_streams[s], f > 0 || chained);
        }
    for (int s = 0; s < c->n_slices; s++) {
        cudaEventRecord(c->slice_done[2 * s], c->slice_streams[s].a);
        cudaEventRecord(c->slice_done[2 * s + 1], c->slice_streams[s].b);
    }
    c->pending_join = true;
    if (join) join_impl(c, false);
    return launches;
}
int pt_enqueue_frames(pt_context* c, int g) { return enqueue_frames_impl(c, g, true); }
int pt_enqueue_frames_chained(pt_context* c, int g, cudaEvent_t accum_free, bool clear_accum) {
    return enqueue_frames_impl(c, g, false, accum_free, clear_accum && c->slices_banded);
}

static int join_impl(pt_context* c, bool mark_time) {
    if (!c->pending_join) return PT_OK;
    for (int s = 0; s < c->n_slices; s++) {
        CK(cudaStreamWaitEvent(c->stream, c->slice_done[2 * s], 0));
        CK(cudaStreamWaitEvent(c->stream, c->slice_done[2 * s + 1], 0));
        if (c->slices_presenting) CK(cudaStreamWaitEvent(c->stream, c->slice_presented[s], 0));
    }
    c->pending_join = false; c->slices_presenting = false;
    if (mark_time) CK(cudaEventRecord(c->ev1, c->stream));      // device time of chained frames runs up to their join
    return PT_OK;
}
int pt_join_slices(pt_context* c) { return join_impl(c, true); }

static int ensure_graph(pt_context* c) {
    if (c->graph.exec) return PT_OK;
    cudaGraph_t graph = nullptr;
    CK(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    int launches = pt_enqueue_frames(c, 1);
    cudaError_t e = cudaStreamEndCapture(c->stream, &graph);
    if (e != cudaSuccess || !graph) { cudaGetLastError(); return fail(PT_ERR_CUDA, "frame graph capture failed: %s", cudaGetErrorString(e)); }
    e = cudaGraphInstantiate(&c->graph.exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) { c->graph.exec = nullptr; return fail(PT_ERR_CUDA, "cudaGraphInstantiate failed: %s", cudaGetErrorString(e)); }
    c->graph.launches = launches;
    return PT_OK;
}

int pt_render_frames(pt_context* c, int32_t n_frames) {
    if (!c || n_frames < 0) return fail(PT_ERR_INVALID, "pt_render_frames: bad argument");
    CK(cudaSetDevice(c->cfg.device));
    int rc = pt_check_scene(c, "pt_render_frames", true);
    if (rc != PT_OK) return rc;
    const int epf = 3 * c->cfg.max_depth + 3;
    if (c->lazy_join && n_frames == 1 && !c->stage_timing) {
        // The interactive case (one RenderFrame per Present): the frame is chained behind the previous one slice by slice and
        // nothing is joined -- the next entry point that needs the result joins (pt_join_slices). Device time (ev0 .. ev1)
        // spans the whole chain.
        if (!c->pending_join) CK(cudaEventRecord(c->ev0, c->stream));
        c->counters.kernel_launches += enqueue_frames_impl(c, 1, false);
        CK(cudaGetLastError());
        c->num_samples += 1;
        c->counters.frames += 1;
        return PT_OK;
    }
    if ((rc = pt_join_slices(c)) != PT_OK) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    if (c->stage_timing) {
        // one launch sequence for the whole frame on the main stream (per-stage events need kernels that do not overlap)
        FrameParams p = frame_params(c);
        while ((int)c->stage_events.size() < epf * n_frames) {
            cudaEvent_t e; CK(cudaEventCreate(&e)); c->stage_events.push_back(e);
        }
        c->stage_frames = n_frames; c->stage_events_per_frame = epf;
        for (int f = 0; f < n_frames; f++)
            c->counters.kernel_launches += launch_frame(p, c->dims, c->stream, c->stage_events.data() + (size_t)f * epf);
    } else if (c->use_graph && n_frames == 1) {
        // PT_LAZY_JOIN=0: replay the captured fork/join DAG of one frame
        int rc2 = ensure_graph(c);
        if (rc2 != PT_OK) return rc2;
        CK(cudaGraphLaunch(c->graph.exec, c->stream));
        c->counters.kernel_launches += c->graph.launches;
    } else {
        // batches: slices run through all their frames without a join in between (measured faster than replaying graphs)
        c->counters.kernel_launches += pt_enqueue_frames(c, n_frames);
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaGetLastError());
    c->num_samples += n_frames;
    c->counters.frames += n_frames;
    return PT_OK;
}

int pt_synchronize(pt_context* c) {
    if (!c) return fail(PT_ERR_INVALID, "pt_synchronize: null context");
    PT_ENTER(c);
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaStreamSynchronize(c->copy_stream));
    return PT_OK;
}

int pt_reset_samples(pt_context* c) {
    if (!c) return fail(PT_ERR_INVALID, "pt_reset_samples: null context");
    PT_ENTER(c);
    // glClearTexSubImage(accum, {0,0,0,1}) + numSamples = 0, Renderer.cpp:687-691
    CK(cudaMemsetAsync(c->d_accum, 0, (size_t)c->cfg.width * c->cfg.height * sizeof(float4), c->stream));
    c->num_samples = 0;
    return PT_OK;
}
uint32_t pt_num_samples(pt_context* c) { return c ? c->num_samples : 0; }

int pt_download_accum(pt_context* c, float* rgba) {
    if (!c || !rgba) return fail(PT_ERR_INVALID, "pt_download_accum: bad argument");
    PT_ENTER(c);
    CK(cudaMemcpyAsync(rgba, c->d_accum, (size_t)c->cfg.width * c->cfg.height * 16, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PT_OK;
}
void* pt_accum_device_ptr(pt_context* c) { return c ? (void*)c->d_accum : nullptr; }
int pt_set_accum_device_ptr(pt_context* c, void* dev_rgba) {
    if (!c) return fail(PT_ERR_INVALID, "pt_set_accum_device_ptr: null context");
    pt_invalidate_graphs(c);
    c->d_accum = dev_rgba ? (float4*)dev_rgba : c->d_accum_own.p;
    return PT_OK;
}

// ---- checkpoint / resume. The reference keeps its running sum in a GL image and its RNG table in
// an SSBO and can save neither; a 4096-spp 4K render here survives a restart: the file carries exactly the
// state RenderFrame reads and writes -- per-pixel HybridTaus states, the RGBA32F sum, numSamples -- so frames rendered
// after pt_checkpoint_load are bit-identical to the frames an uninterrupted run would have rendered.
struct CheckpointHeader {
    char magic[8];                 // "B2PTCKP1"
    uint32_t width, height, num_samples, max_depth, microfacet_model, reserved;
    uint64_t payload_bytes, payload_fnv1a;
};
static uint64_t fnv1a(const void* data, size_t n, uint64_t h = 1469598103934665603ull) {
    const unsigned char* p = (const unsigned char*)data;
    for (size_t i = 0; i < n; i++) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}
int pt_checkpoint_save(pt_context* c, const char* path) {
    if (!c || !path) return fail(PT_ERR_INVALID, "pt_checkpoint_save: bad argument");
    if (!c->have_rng) return fail(PT_ERR_INVALID, "pt_checkpoint_save: no RNG table yet (pt_set_rng / pt_seed_rng)");
    PT_ENTER(c);
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    std::vector<unsigned char> payload(npx * 32);
    CK(cudaMemcpyAsync(payload.data(), c->d_rng.p, npx * 16, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaMemcpyAsync(payload.data() + npx * 16, c->d_accum, npx * 16, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    CheckpointHeader h{};
    memcpy(h.magic, "B2PTCKP1", 8);
    h.width = (uint32_t)c->cfg.width; h.height = (uint32_t)c->cfg.height; h.num_samples = c->num_samples;
    h.max_depth = (uint32_t)c->cfg.max_depth; h.microfacet_model = (uint32_t)c->cfg.microfacet_model;
    h.payload_bytes = payload.size(); h.payload_fnv1a = fnv1a(payload.data(), payload.size());
    // write beside the target and rename: a crash while saving never destroys the previous checkpoint
    std::string tmp = std::string(path) + ".part";
    FILE* f = fopen(tmp.c_str(), "wb");
    if (!f) return fail(PT_ERR_IO, "pt_checkpoint_save: cannot create %s", tmp.c_str());
    bool ok = fwrite(&h, sizeof(h), 1, f) == 1 && fwrite(payload.data(), 1, payload.size(), f) == payload.size();
    ok = (fclose(f) == 0) && ok;
    if (!ok || rename(tmp.c_str(), path) != 0) { remove(tmp.c_str()); return fail(PT_ERR_IO, "pt_checkpoint_save: cannot write %s", path); }
    return PT_OK;
}
int pt_checkpoint_load(pt_context* c, const char* path) {
    if (!c || !path) return fail(PT_ERR_INVALID, "pt_checkpoint_load: bad argument");
    FILE* f = fopen(path, "rb");
    if (!f) return fail(PT_ERR_IO, "pt_checkpoint_load: cannot open %s", path);
    CheckpointHeader h{};
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    std::vector<unsigned char> payload;
    bool ok = fread(&h, sizeof(h), 1, f) == 1 && memcmp(h.magic, "B2PTCKP1", 8) == 0;
    if (ok && (h.width != (uint32_t)c->cfg.width || h.height != (uint32_t)c->cfg.height)) {
        fclose(f);
        return fail(PT_ERR_INVALID, "pt_checkpoint_load: %s is %ux%u, the context is %dx%d", path, h.width, h.height, c->cfg.width, c->cfg.height);
    }
    if (ok) { payload.resize(npx * 32); ok = h.payload_bytes == payload.size() && fread(payload.data(), 1, payload.size(), f) == payload.size(); }
    fclose(f);
    if (!ok || fnv1a(payload.data(), payload.size()) != h.payload_fnv1a) return fail(PT_ERR_IO, "pt_checkpoint_load: %s is not a complete checkpoint", path);
    PT_ENTER(c);
    CK(cudaStreamSynchronize(c->stream));
    CK(c->d_rng.alloc(npx));
    CK(cudaMemcpy(c->d_rng.p, payload.data(), npx * 16, cudaMemcpyHostToDevice));
    CK(cudaMemcpy(c->d_accum, payload.data() + npx * 16, npx * 16, cudaMemcpyHostToDevice));
    c->have_rng = true;
    c->num_samples = h.num_samples;
    return PT_OK;
}

int pt_resolve(pt_context* c, uint32_t num_samples, float* rgb_f32, uint8_t* rgb_u8) {
    if (!c || num_samples == 0) return fail(PT_ERR_INVALID, "pt_resolve: bad argument");
    PT_ENTER(c);
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    float* d_f = nullptr; unsigned char* d_u = nullptr;
    if (rgb_f32) { CK(c->d_resolve_f32.alloc(npx * 3)); d_f = c->d_resolve_f32.p; }
    if (rgb_u8) { CK(c->d_resolve_u8.alloc(npx * 3)); d_u = c->d_resolve_u8.p; }
    launch_resolve(c->d_accum, (long long)npx, num_samples, d_f, d_u, c->stream, (c->cfg.parity_flags & PT_PARITY_LIBM) != 0);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    if (rgb_f32) CK(cudaMemcpyAsync(rgb_f32, d_f, npx * 12, cudaMemcpyDeviceToHost, c->stream));
    if (rgb_u8) CK(cudaMemcpyAsync(rgb_u8, d_u, npx * 3, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    return PT_OK;
}

int pt_present_async(pt_context* c, uint32_t num_samples, uint8_t* host_rgb_u8) {
    if (!c || num_samples == 0 || !host_rgb_u8) return fail(PT_ERR_INVALID, "pt_present_async: bad argument");
    CK(cudaSetDevice(c->cfg.device));
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int k = c->present_turn; c->present_turn ^= 1;
    const bool plibm = (c->cfg.parity_flags & PT_PARITY_LIBM) != 0;
    CK(c->d_present[k].alloc(npx * 3));
    if (c->pending_join) {
        // Chained frames (pt_render_frames(1) without a join): every slice resolves ITS OWN pixels on its own streams right
        // behind its last kernel, so slice s's next frame waits for slice s only; just the copy waits for all slices.
        // (Present.frag runs per fragment: which stream shades a pixel changes nothing.)
        for (int s = 0; s < c->n_slices; s++) {
            const SliceStreams& ss = c->slice_streams[s];
            const int first = s * c->slice_size;
            const int count = (int)npx - first < c->slice_size ? (int)npx - first : c->slice_size;
            // On stream b, right behind the slice's last shadow kernel (the last writer of its pixels). Stream a stays free
            // for the next frame's raygen + extend(0); its shade(0) -- the next writer of these pixels -- waits for
            // ev_shadow, which is therefore recorded again behind the resolve.
            CK(cudaStreamWaitEvent(ss.b, c->ev_copied[k], 0));                  // the copy that last read this buffer
            launch_resolve_slice(c->d_accum, first, count, c->cfg.width, num_samples, c->d_present[k].p, ss.b, plibm);
            c->counters.kernel_launches++;
            CK(cudaEventRecord(c->slice_presented[s], ss.b));
            CK(cudaEventRecord(ss.ev_shadow, ss.b));
            CK(cudaStreamWaitEvent(c->copy_stream, c->slice_presented[s], 0));
        }
        c->slices_presenting = true;
        CK(cudaGetLastError());
    } else {
        CK(cudaStreamWaitEvent(c->stream, c->ev_copied[k], 0));          // the copy that last used this buffer (no-op the first time)
        launch_resolve(c->d_accum, (long long)npx, num_samples, nullptr, c->d_present[k].p, c->stream, plibm);
        c->counters.kernel_launches++;
        CK(cudaGetLastError());
        CK(cudaEventRecord(c->ev_resolved, c->stream));
        CK(cudaStreamWaitEvent(c->copy_stream, c->ev_resolved, 0));
    }
    CK(cudaMemcpyAsync(host_rgb_u8, c->d_present[k].p, npx * 3, cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaEventRecord(c->ev_copied[k], c->copy_stream));
    return PT_OK;
}

// One primary ray per pixel exactly as the first bounce of RenderFrame traces it (InitRay Iterative.comp:174-210 with the
// pixel's RNG state, IfIfClosestHit), hit records in pixel order; the copy runs on the copy stream, double-buffered
int pt_trace_primary_async(pt_context* c, float* host_hits) {
    if (!c || !host_hits) return fail(PT_ERR_INVALID, "pt_trace_primary_async: bad argument");
    PT_ENTER(c);
    int rc = pt_check_scene(c, "pt_trace_primary_async", false);
    if (rc != PT_OK) return rc;
    if (!c->have_cam || !c->have_rng) return fail(PT_ERR_INVALID, "pt_trace_primary_async: camera and RNG table must be set");
    const size_t npx = (size_t)c->cfg.width * c->cfg.height;
    const int k = c->primary_turn; c->primary_turn ^= 1;
    CK(c->d_primary[k].alloc(npx));
    CK(cudaEventRecord(c->ev0, c->stream));
    CK(cudaStreamWaitEvent(c->stream, c->ev_primary_copied[k], 0));
    FrameParams p = frame_params(c);
    c->counters.kernel_launches += launch_primary(p, c->dims, c->d_primary[k].p, c->stream);
    c->counters.closest_rays += (uint64_t)npx;
    CK(cudaGetLastError());
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaEventRecord(c->ev_primary, c->stream));
    CK(cudaStreamWaitEvent(c->copy_stream, c->ev_primary, 0));
    CK(cudaMemcpyAsync(host_hits, c->d_primary[k].p, npx * sizeof(float4), cudaMemcpyDeviceToHost, c->copy_stream));
    CK(cudaEventRecord(c->ev_primary_copied[k], c->copy_stream));
    return PT_OK;
}

int pt_save_screenshot(pt_context* c, const char* path) {
    if (!c || !path) return fail(PT_ERR_INVALID, "pt_save_screenshot: bad argument");
    if (c->num_samples == 0) return fail(PT_ERR_INVALID, "pt_save_screenshot: no samples rendered");
    size_t npx = (size_t)c->cfg.width * c->cfg.height;
    std::string p(path);
    bool pfm = p.size() > 4 && p.substr(p.size() - 4) == ".pfm";
    if (pfm) {
        std::vector<float> img(npx * 3);
        int rc = pt_resolve(c, c->num_samples, img.data(), nullptr);
        if (rc != PT_OK) return rc;
        if (!write_pfm(path, c->cfg.width, c->cfg.height, img.data())) return fail(PT_ERR_IO, "cannot write %s", path);
    } else {
        std::vector<uint8_t> img(npx * 3);
        int rc = pt_resolve(c, c->num_samples, nullptr, img.data());
        if (rc != PT_OK) return rc;
        if (p.size() > 4 && p.substr(p.size() - 4) == ".png") {
            // the reference's format (SOIL_SAVE_TYPE_PNG); image row 0 is the bottom row of its GL framebuffer, PNG is top-down
            std::vector<uint8_t> flip(npx * 3);
            const size_t row = 3 * (size_t)c->cfg.width;
            for (int y = 0; y < c->cfg.height; y++) memcpy(&flip[(size_t)y * row], &img[(size_t)(c->cfg.height - 1 - y) * row], row);
            if (!write_png_rgb8(path, c->cfg.width, c->cfg.height, flip.data())) return fail(PT_ERR_IO, "cannot write %s", path);
        } else if (!write_ppm(path, c->cfg.width, c->cfg.height, img.data())) return fail(PT_ERR_IO, "cannot write %s", path);
    }
    return PT_OK;
}

int pt_trace_closest_device(pt_context* c, const void* d_rays, int64_t n, void* d_hits) {
    if (!c || !d_rays || !d_hits || n < 0 || n > kMaxBatchRays) return fail(PT_ERR_INVALID, "pt_trace_closest_device: bad argument");
    PT_ENTER(c);
    int rc = pt_check_scene(c, "pt_trace_closest", false);
    if (rc != PT_OK) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    const int* perm = nullptr;
    if ((rc = order_rays(c, (const float4*)d_rays, n, &perm)) != PT_OK) return rc;
    launch_trace_closest(pt_scene_dev(c), (const float4*)d_rays, n, (float4*)d_hits, c->d_cursor.p, c->dims, c->stream, perm);
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaGetLastError());
    c->counters.kernel_launches++;
    c->counters.closest_rays += (uint64_t)n;
    return PT_OK;
}
int pt_trace_any_device(pt_context* c, const void* d_rays, int64_t n, void* d_flags) {
    if (!c || !d_rays || !d_flags || n < 0 || n > kMaxBatchRays) return fail(PT_ERR_INVALID, "pt_trace_any_device: bad argument");
    PT_ENTER(c);
    int rc = pt_check_scene(c, "pt_trace_any", false);
    if (rc != PT_OK) return rc;
    CK(cudaEventRecord(c->ev0, c->stream));
    const int* perm = nullptr;
    if ((rc = order_rays(c, (const float4*)d_rays, n, &perm)) != PT_OK) return rc;
    launch_trace_any(pt_scene_dev(c), (const float4*)d_rays, n, (unsigned char*)d_flags, c->d_cursor.p, c->dims, c->stream, perm);
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaGetLastError());
    c->counters.kernel_launches++;
    c->counters.shadow_rays += (uint64_t)n;
    return PT_OK;
}

// Host batches: chunks of kHostChunk rays flow through a three-stage pipeline -- H2D on `in_stream`, the traversal
// kernel on the render stream, D2H on `copy_stream` -- with double-buffered staging, so a 16 M-ray batch costs about
// max(PCIe in, kernel, PCIe out) instead of their sum (pinned host memory makes the copies truly asynchronous).


static int trace_host(pt_context* c, const float* rays, int64_t n, void* out, size_t out_bytes, bool any) {
    PT_ENTER(c);
    int rc = pt_check_scene(c, any ? "pt_trace_any" : "pt_trace_closest", false);
    if (rc != PT_OK) return rc;
    const int64_t chunk = n < kHostChunk ? n : kHostChunk;
    CK(c->d_stage_rays.alloc(2 * 2 * (size_t)chunk));                 // 2 buffers x 2 float4 per ray
    if (any) CK(c->d_stage_flags.alloc(2 * (size_t)chunk)); else CK(c->d_stage_hits.alloc(2 * (size_t)chunk));
    if (!c->in_stream) {
        CK(cudaStreamCreateWithFlags(&c->in_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 2; k++) {
            CK(cudaEventCreateWithFlags(&c->ev_h_in[k], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->ev_h_done[k], cudaEventDisableTiming));
            CK(cudaEventCreateWithFlags(&c->ev_h_out[k], cudaEventDisableTiming));
        }
    }
    CK(cudaEventRecord(c->ev0, c->stream));
    int k = 0;
    for (int64_t first = 0; first < n; first += chunk, k ^= 1) {
        const int64_t m = n - first < chunk ? n - first : chunk;
        float4* d_rays = c->d_stage_rays.p + (size_t)k * 2 * chunk;
        CK(cudaStreamWaitEvent(c->in_stream, c->ev_h_done[k], 0));    // the kernel that last read this ray buffer
        CK(cudaMemcpyAsync(d_rays, rays + 8 * first, 32 * (size_t)m, cudaMemcpyHostToDevice, c->in_stream));
        CK(cudaEventRecord(c->ev_h_in[k], c->in_stream));
        CK(cudaStreamWaitEvent(c->stream, c->ev_h_in[k], 0));
        CK(cudaStreamWaitEvent(c->stream, c->ev_h_out[k], 0));        // the copy that last read this result buffer
        const int* perm = nullptr;
        if ((rc = order_rays(c, d_rays, m, &perm)) != PT_OK) return rc;
        if (any) {
            unsigned char* d_out = c->d_stage_flags.p + (size_t)k * chunk;
            launch_trace_any(pt_scene_dev(c), d_rays, m, d_out, c->d_cursor.p, c->dims, c->stream, perm);
            CK(cudaEventRecord(c->ev_h_done[k], c->stream));
            CK(cudaStreamWaitEvent(c->copy_stream, c->ev_h_done[k], 0));
            CK(cudaMemcpyAsync((unsigned char*)out + first, d_out, (size_t)m, cudaMemcpyDeviceToHost, c->copy_stream));
            c->counters.shadow_rays += (uint64_t)m;
        } else {
            float4* d_out = c->d_stage_hits.p + (size_t)k * chunk;
            launch_trace_closest(pt_scene_dev(c), d_rays, m, d_out, c->d_cursor.p, c->dims, c->stream, perm);
            CK(cudaEventRecord(c->ev_h_done[k], c->stream));
            CK(cudaStreamWaitEvent(c->copy_stream, c->ev_h_done[k], 0));
            CK(cudaMemcpyAsync((char*)out + out_bytes * first, d_out, out_bytes * (size_t)m, cudaMemcpyDeviceToHost, c->copy_stream));
            c->counters.closest_rays += (uint64_t)m;
        }
        CK(cudaEventRecord(c->ev_h_out[k], c->copy_stream));
        c->counters.kernel_launches++;
    }
    CK(cudaEventRecord(c->ev1, c->stream));
    CK(cudaGetLastError());
    CK(cudaStreamSynchronize(c->stream));
    CK(cudaStreamSynchronize(c->copy_stream));
    return PT_OK;
}

int pt_trace_closest(pt_context* c, const float* rays, int64_t n, float* hits) {
    if (!c || !rays || !hits || n < 0 || n > kMaxBatchRays) return fail(PT_ERR_INVALID, "pt_trace_closest: bad argument");
    if (n == 0) return PT_OK;
    return trace_host(c, rays, n, hits, 16, false);
}
int pt_trace_any(pt_context* c, const float* rays, int64_t n, uint8_t* flags) {
    if (!c || !rays || !flags || n < 0 || n > kMaxBatchRays) return fail(PT_ERR_INVALID, "pt_trace_any: bad argument");
    if (n == 0) return PT_OK;
    return trace_host(c, rays, n, flags, 1, true);
}

int pt_raygen_probe(pt_context* c, int64_t first, int64_t n, float* out) {
    if (!c || !out || first < 0 || n < 0 || first + n > (int64_t)c->cfg.width * c->cfg.height) return fail(PT_ERR_INVALID, "pt_raygen_probe: bad range");
    if (!c->have_cam || !c->have_rng) return fail(PT_ERR_INVALID, "pt_raygen_probe: camera and RNG table must be set");
    if (n == 0) return PT_OK;
    PT_ENTER(c);
    float4* d = nullptr;
    CK(cudaMalloc((void**)&d, 48 * (size_t)n));
    FrameParams p = frame_params(c);
    launch_raygen_probe(p, first, n, d, c->stream);
    c->counters.kernel_launches++;
    CK(cudaGetLastError());
    CK(cudaMemcpyAsync(out, d, 48 * (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    cudaFree(d);
    return PT_OK;
}

int pt_parity_libm_probe(pt_context* c, int32_t fn, const float* x, const float* y, int64_t n, float* out) {
    if (!c || !x || !out || n < 0 || fn < 0 || fn > 4 || (fn == 4 && !y)) return fail(PT_ERR_INVALID, "pt_parity_libm_probe: bad argument");
    if (n == 0) return PT_OK;
    PT_ENTER(c);
    DevBuf<float> dx, dy, dout;
    CK(dx.alloc((size_t)n)); CK(dy.alloc((size_t)n)); CK(dout.alloc((size_t)n));
    CK(cudaMemcpyAsync(dx.p, x, 4 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    if (y) CK(cudaMemcpyAsync(dy.p, y, 4 * (size_t)n, cudaMemcpyHostToDevice, c->stream));
    launch_plibm_probe(fn, dx.p, dy.p, n, dout.p, c->stream);
    c->counters.kernel_launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, dout.p, 4 * (size_t)n, cudaMemcpyDeviceToHost, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    dx.release(); dy.release(); dout.release();
    if (e != cudaSuccess) return fail(PT_ERR_CUDA, "pt_parity_libm_probe: %s", cudaGetErrorString(e));
    return PT_OK;
}

int pt_get_counters(pt_context* c, pt_counters* out) {
    if (!c || !out) return fail(PT_ERR_INVALID, "pt_get_counters: bad argument");
    PT_ENTER(c);
    unsigned long long t[4];
    CK(cudaMemcpyAsync(t, c->d_totals.p, sizeof(t), cudaMemcpyDeviceToHost, c->stream));
    CK(cudaStreamSynchronize(c->stream));
    *out = c->counters;
    out->closest_rays += t[0]; out->shadow_rays += t[1]; out->paths = t[2]; out->null_shadow_rays = t[3];
    if (c->d_visits.p) {
        unsigned long long v[4];
        CK(cudaMemcpy(v, c->d_visits.p, sizeof(v), cudaMemcpyDeviceToHost));
        out->closest_pair_tests = v[0]; out->closest_tri_tests = v[1]; out->shadow_pair_tests = v[2]; out->shadow_tri_tests = v[3];
        out->counted_closest_rays = c->counted_rays[0] + (c->count_visits ? out->closest_rays - c->counted_base[0] : 0);
        out->counted_shadow_rays = c->counted_rays[1] + (c->count_visits ? out->shadow_rays - c->counted_base[1] : 0);
    }
    return PT_OK;
}

int pt_set_visit_counting(pt_context* c, int on) {
    if (!c) return fail(PT_ERR_INVALID, "pt_set_visit_counting: null context");
    if (c->dims.wide && on) return fail(PT_ERR_UNSUPPORTED, "pt_set_visit_counting: binary layout only");
    pt_counters now;
    int rc = pt_get_counters(c, &now);                                  // joins and synchronises
    if (rc != PT_OK) return rc;
    if (on && !c->d_visits.p) { CK(c->d_visits.alloc(4)); CK(cudaMemset(c->d_visits.p, 0, 4 * sizeof(unsigned long long))); }
    if (on && !c->count_visits) { c->counted_base[0] = now.closest_rays; c->counted_base[1] = now.shadow_rays; }
    if (!on && c->count_visits) { c->counted_rays[0] += now.closest_rays - c->counted_base[0]; c->counted_rays[1] += now.shadow_rays - c->counted_base[1]; }
    c->count_visits = on != 0;
    c->dims.count_visits = on ? 1 : 0;
    pt_invalidate_graphs(c);
    return PT_OK;
}

float pt_last_device_ms(pt_context* c) {
    if (!c) return -1.0f;
    cudaSetDevice(c->cfg.device);
    if (pt_join_slices(c) != PT_OK) return -1.0f;
    if (cudaEventSynchronize(c->ev1) != cudaSuccess) return -1.0f;
    float ms = -1.0f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) != cudaSuccess) return -1.0f;
    return ms;
}
int pt_set_stage_timing(pt_context* c, int on) {
    if (!c) return fail(PT_ERR_INVALID, "pt_set_stage_timing: null context");
    c->stage_timing = on != 0;
    c->stage_frames = 0;
    return PT_OK;
}
int pt_get_stage_ms(pt_context* c, float out[5]) {
    if (!c || !out) return fail(PT_ERR_INVALID, "pt_get_stage_ms: bad argument");
    PT_ENTER(c);
    CK(cudaStreamSynchronize(c->stream));
    for (int i = 0; i < 5; i++) out[i] = 0.0f;
    const int epf = c->stage_events_per_frame;
    for (int f = 0; f < c->stage_frames; f++) {
        cudaEvent_t* ev = c->stage_events.data() + (size_t)f * epf;
        for (int k = 0; k + 1 < epf; k++) {
            float ms = 0.0f;
            CK(cudaEventElapsedTime(&ms, ev[k], ev[k + 1]));
            int stage = (k == 0) ? 0 : (k == epf - 2) ? 4 : 1 + ((k - 1) % 3);
            out[stage] += ms;
        }
    }
    return PT_OK;
}
void* pt_stream(pt_context* c) { return c ? (void*)c->stream : nullptr; }
}  // extern "C"
