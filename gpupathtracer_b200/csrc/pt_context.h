// pt_context.h -- the context behind the C ABI (private to the library): device buffers, streams, host mirrors.
// Shared by pt_api.cu (single-GPU entry points) and pt_comm.cu (sample split across GPUs over NCCL).
#pragma once
#include "../../include/b200pt.h"
#include "pt_kernels.h"
#include "pt_host.h"
#include <cuda_runtime.h>
#include <string>
#include <vector>

int pt_fail(int code, const char* fmt, ...);
#define fail pt_fail
#define CK(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return fail(PT_ERR_CUDA, "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e__), __FILE__, __LINE__); \
    } while (0)

template <class T>
struct DevBuf {
    T* p = nullptr; size_t n = 0;
    cudaError_t alloc(size_t count) {
        if (count <= n && p) return cudaSuccess;
        release();
        cudaError_t e = cudaMalloc((void**)&p, (count ? count : 1) * sizeof(T));
        if (e == cudaSuccess) n = count;
        return e;
    }
    void release() { if (p) cudaFree(p); p = nullptr; n = 0; }
};

using namespace pt;

struct pt_context {
    pt_config cfg;
    cudaStream_t stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    // A frame is cut into n_slices ranges of ray indices (pixel-disjoint, so RNG table and accum need no
    // synchronisation); every slice runs its own raygen -> (extend, shade, shadow)* sequence on its own stream,
    // so the drain phase of one slice's persistent kernel is filled by another slice's kernels.
    // Inside a slice, shadow(b) and extend(b+1) run concurrently on two streams (enqueue_frame_dag). The whole
    // fork/join structure of `g` consecutive frames is captured once into a CUDA graph and replayed.
    int n_slices = 1, slice_size = 0;
    bool slices_banded = false;                    // every slice is a whole number of 8-row tile bands: a contiguous range of the accumulator
    std::vector<SliceStreams> slice_streams;
    std::vector<cudaEvent_t> slice_done;      // 2 per slice (streams a and b)
    cudaEvent_t ev_fork = nullptr;
    bool use_graph = true;
    // Interactive loop (RenderFrame(1) + Present per frame): frames are enqueued eagerly and the slices are NOT joined into
    // `stream` after each one; slice s's next frame follows slice s's own present, so the GPU never drains between frames.
    // Every other entry point joins first (pt_join_slices). PT_LAZY_JOIN=0 restores the per-frame graph replay + join.
    bool lazy_join = true, pending_join = false, slices_presenting = false;
    unsigned frame_seq = 0;                    // counter-block parity of eagerly chained single frames
    std::vector<cudaEvent_t> slice_presented;  // per slice: its part of the last pt_present_async
    int cam_slot = 0;                          // ring of camera records: frames in flight keep the camera they were enqueued with
    struct FrameGraph { cudaGraphExec_t exec = nullptr; int launches = 0; };
    FrameGraph graph;                          // one frame, all slices
    DevBuf<CameraRec> d_cam;
    // pt_present_async: double-buffered resolve target + copy stream
    cudaStream_t copy_stream = nullptr;
    // host-batch pipeline (trace_host)
    cudaStream_t in_stream = nullptr;
    cudaEvent_t ev_h_in[2] = {nullptr, nullptr}, ev_h_done[2] = {nullptr, nullptr}, ev_h_out[2] = {nullptr, nullptr};
    DevBuf<unsigned char> d_present[2];
    DevBuf<float4> d_primary[2];
    cudaEvent_t ev_primary = nullptr, ev_primary_copied[2] = {nullptr, nullptr};
    int primary_turn = 0;
    cudaEvent_t ev_resolved = nullptr, ev_copied[2] = {nullptr, nullptr};
    int present_turn = 0;
    LaunchDims dims{};
    float last_ms = 0.0f;
    pt_counters counters{};
    uint32_t num_samples = 0;

    // host mirrors (what the reference keeps in triangleVec / nodesVec / referenceVec, Scene.h:23, BVH.h:97-98)
    std::vector<float> h_tris;
    HostBvh h_bvh;
    std::vector<pt_material> h_mats;
    std::vector<pt_light> h_lights;
    std::vector<std::vector<float>> h_texels; std::vector<std::pair<int, int>> h_texdims;   // texture mirrors (pt_group_share_scene)

    DevBuf<float4> d_tris;
    DevBuf<BvhPair> d_pairs;
    DevBuf<LeafTri> d_leaf;
    DevBuf<unsigned char> d_wide;          // PT_BVH_WIDE / PT_BVH_WIDE4 node bytes
    DevBuf<LeafTri> d_wide_tris;
    WideBvh h_wide;
    // device builder: serialised tree kept on the device, host mirror fetched on demand
    DevBuf<float4> d_ser_nodes; DevBuf<int> d_ser_refs;
    DevBuf<unsigned char> d_build_ws;          // scratch of the device builders, kept between rebuilds
    bool host_bvh_stale = false; int64_t ser_nodes = 0, ser_refs = 0;
    float wide_ms = 0.0f;
    DevBuf<MaterialRec> d_mats;
    DevBuf<TextureRec> d_texrec;
    std::vector<float4*> d_texels;
    DevBuf<float2> d_lights;
    int n_lights = 0, n_textures = 0, n_leaf = 0;
    int top_pairs = 0;                     // pair records [0, top_pairs) are the breadth-first top of the tree
    float total_light_area = 0.0f;
    float env[3] = {1.0f, 1.0f, 1.0f};
    DevBuf<float4> d_env; int env_size = 0;      // cubemap faces, 0: constant colour
    float sun_dir[3] = {0.0f, 1.0f, 0.0f}; float sun_radius = 0.0f;   // set in pt_create

    CameraRec cam{};
    bool have_cam = false, have_rng = false;

    DevBuf<uint4> d_rng;
    DevBuf<float4> d_accum_own;
    float4* d_accum = nullptr;
    DevBuf<float4> d_q[2][4];
    DevBuf<uint4> d_qrng[2];
    DevBuf<float4> d_shadow[4];
    DevBuf<float4> d_hits;
    DevBuf<BounceCounters> d_counters;
    DevBuf<unsigned long long> d_totals;
    DevBuf<int> d_cursor;
    DevBuf<unsigned long long> d_visits;   // pt_set_visit_counting: 4 visit counters written by the instrumented traversal kernels
    bool count_visits = false;
    uint64_t counted_rays[2] = {0, 0}, counted_base[2] = {0, 0};     // closest / shadow rays traced while counting was on
    // staging for host-pointer batch entry points
    DevBuf<float4> d_stage_rays, d_stage_hits;
    DevBuf<unsigned char> d_stage_flags;
    // processing order of big ray batches (launch_ray_order)
    DevBuf<unsigned> d_ray_bins, d_ray_bin; DevBuf<int> d_ray_perm;
    int ray_order = 0;                     // 0 auto, 1 always, -1 never (PT_RAY_ORDER)
    size_t l2_bytes = 0;
    DevBuf<float> d_resolve_f32;
    DevBuf<unsigned char> d_resolve_u8;
    struct pt_comm_state* comm = nullptr;  // pt_comm.cu: NCCL communicator + double-buffered batch accumulators (null: single GPU)
    // per-stage timing (pt_set_stage_timing)
    bool stage_timing = false;
    std::vector<cudaEvent_t> stage_events;
    int stage_frames = 0, stage_events_per_frame = 0;
};


// helpers of pt_api.cu that pt_comm.cu builds on (library-internal; C linkage only because pt_api.cu defines them inside its
// extern "C" block)
void pt_invalidate_graphs(pt_context* c);
pt::SceneDev pt_scene_dev(pt_context* c);
extern "C" int pt_check_scene(pt_context* c, const char* who, bool need_shading);
extern "C" int pt_join_slices(pt_context* c);   // c->stream waits for everything the slices have been given (device-side wait)
#define PT_CAM_RING 64
// n frames of every slice chained behind whatever the slice streams hold, no join (pending_join stays set). accum_free (may be
// null): waited for by every slice before it touches c->d_accum; clear_accum: every slice zeroes its own band first (needs
// c->slices_banded). Slice s is complete when c->slice_done[2 * s + 1] fires.
extern "C" int pt_enqueue_frames_chained(pt_context* c, int n_frames, cudaEvent_t accum_free, bool clear_accum);
extern "C" int pt_enqueue_frames(pt_context* c, int n_frames);            // n frames of every slice between a fork from and a join into c->stream
extern "C" int pt_ensure_host_bvh(pt_context* c);
void pt_comm_release(pt_context* c);                          // pt_comm.cu; called by pt_destroy
