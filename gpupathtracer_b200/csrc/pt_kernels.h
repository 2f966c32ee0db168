// pt_kernels.h -- launch wrappers of the sm_100a kernels (implemented in pt_kernels.cu).
#pragma once
#include "pt_layout.h"

namespace pt {

// Per-frame device counters: one row per bounce. Zeroed once per frame, so no kernel ever has to
// reset a counter another kernel is still reading.
struct BounceCounters {
    int n_paths;      // size of the path queue entering this bounce (written by raygen / previous shade)
    int n_shadow;     // size of the shadow queue produced by this bounce's shade
    int fetch_extend; // persistent-kernel work cursors
    int fetch_shadow;
    int n_null;       // NEE samples of this bounce whose radiance is exactly zero: not queued for a shadow ray
};

struct FrameParams {
    SceneDev scene;
    const CameraRec* cam;    // device copy of the camera uniforms (so a captured graph sees camera updates)
    int width, height;
    int slice_first, slice_count;   // this launch sequence owns ray indices [slice_first, slice_first + slice_count)
    int max_depth;
    int model;
    int sun;              // SUN_LIGHT (Iterative.comp:116-126), off in the reference as shipped
    int plibm;            // PT_PARITY_LIBM: transcendentals through pt_parity_libm.h (bit-identical to the CPU checkers)
    int trace_null;       // 1: queue a shadow ray even when its radiance is exactly 0 (pt_config.trace_null_shadow)
    uint4* rng_table;     // per pixel
    float4* accum;        // per pixel, running sum
    PathQueue queue[2];
    ShadowQueue shadow;
    float4* hits;         // per queue slot (queue, shadow and hit pointers are already offset to the slice's sub-range)
    BounceCounters* counters;   // [max_depth + 2], one block per slice
    unsigned long long* totals; // [0] closest rays, [1] shadow rays traced, [2] paths, [3] shadow rays not needed (radiance 0)
};

struct LaunchDims { int count_visits; int sm_count; int extend_blocks; int frame_blocks; int stream_blocks; int shade_blocks; int leaf_thresh; int fetch_thresh; int wide; int top_pairs; };

// fixed ray batches (AoS rays: n x {o.xyz, -, d.xyz, tmax})
void launch_trace_closest(const SceneDev& s, const float4* rays, long long n, float4* hits, int* cursor,
                          const LaunchDims& dims, cudaStream_t st, const int* perm = nullptr);
void launch_trace_any(const SceneDev& s, const float4* rays, long long n, unsigned char* flags, int* cursor,
                      const LaunchDims& dims, cudaStream_t st, const int* perm = nullptr);
// processing order for a batch: counting sort of the rays by origin cell + direction octant. bins: 2097152 counters + 2048 tile totals,
// ray_bin / perm: n entries each. Returns the number of kernels launched.
int launch_ray_order(const SceneDev& s, const float4* rays, long long n, unsigned* bins, unsigned* ray_bin, int* perm, cudaStream_t st);

// one frame = raygen + max_depth x (extend, shade, shadow) + tally; returns number of kernels launched
// kernel k of a frame: 0 = raygen, 1+3b = extend, 2+3b = shade, 3+3b = shadow (bounce b), last = tally
int launch_frame(const FrameParams& p, const LaunchDims& dims, cudaStream_t st, cudaEvent_t* evs);

// The same frame as a dependency graph on two streams: shadow(b) (NEE visibility of bounce b) and extend(b+1)
// (closest hit of the next bounce) only depend on shade(b), so they run concurrently, shadow on `b`, everything
// else on `a`. Safe to capture into a CUDA graph (cross-stream edges are events). `wait_prev`: the previous frame
// of this slice was enqueued on the same streams without an intervening join (its last shadow may still run).
struct SliceStreams { cudaStream_t a, b; cudaEvent_t ev_shade, ev_shadow; };
int enqueue_frame_dag(const FrameParams& p, const LaunchDims& dims, const SliceStreams& ss, bool wait_prev);

// pt_bvh_gpu.cu: LBVH on the device, output in the reference's serialised form
// A reference = a triangle restricted to a box (the unit of spatial splits). presplit_device cuts every triangle whose box is
// long against the scene (more than 1.5 cells of a `grid`^3 grid over the scene box) into pieces along a regular subdivision of
// its OWN box, clips the triangle to each piece (Sutherland-Hodgman) and keeps the non-empty ones; all other triangles become
// one reference each. Arrays are device memory owned by the caller afterwards (free_refset).
struct RefSet { int m = 0; int* tri = nullptr; float4* lo = nullptr; float4* hi = nullptr; };
int presplit_device(const float4* d_tris, int n, int grid, RefSet* out, cudaStream_t st, float* kernel_ms);   // launches, or -(cudaError)
void free_refset(RefSet* r);
int build_lbvh_device(const float4* d_tris, int n, float4* d_nodes, int* d_refs, BvhPair* d_pairs, LeafTri* d_leaf, int* depth,
                      cudaStream_t st, float* kernel_ms, int ploc_radius = 0, const RefSet* refset = nullptr, int* n_leaves = nullptr,
                      void* workspace = nullptr, size_t workspace_bytes = 0);
// bytes of scratch a build over n leaves takes at most (the workspace argument above; a smaller one is used as far as it goes)
inline size_t build_lbvh_workspace_bytes(long long n) { return (size_t)n * 352 + ((size_t)4 << 20); }

int launch_primary(const FrameParams& p, const LaunchDims& dims, float4* by_pixel, cudaStream_t st);
void launch_raygen_probe(const FrameParams& p, long long first, long long n, float4* out3, cudaStream_t st);
void launch_seed_rng(uint4* table, long long n, unsigned long long seed, cudaStream_t st);
void query_launch_dims(int device, int wide, LaunchDims* out);
void launch_resolve(const float4* accum, long long n, unsigned num_samples, float* rgb_f32, unsigned char* rgb_u8,
                    cudaStream_t st, bool plibm = false);
// the same for the pixels of ray indices [first, first + count) (8x8 Morton tiles, Iterative.comp:163-172): one slice's part of a present
void launch_resolve_slice(const float4* accum, int first, int count, int width, unsigned num_samples, unsigned char* rgb_u8,
                          cudaStream_t st, bool plibm = false);
void launch_plibm_probe(int fn, const float* x, const float* y, long long n, float* out, cudaStream_t st);
void launch_equirect_to_cube(const float4* src, int w, int h, float4* faces, int n, cudaStream_t st);

}  // namespace pt
