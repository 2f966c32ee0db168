// pt_kernels.cu -- the sm_100a kernels of the wavefront path tracer.
//
// Stage map to the reference megakernel (src/shaders/Iterative.comp:214-307):
//   k_raygen   InitRay                      :174-210
//   k_extend   IfIfClosestHit               common/BVH.glsl:634-767
//   k_shade    vertex + emission/MIS + NEE sample + BSDF sample + Russian roulette   :236-301
//   k_shadow   IfIfAnyHit + NEE add + store common/BVH.glsl:769-854, Iterative.comp:278-280,295
//   k_resolve  Present.frag:13-37
// Extend and shadow are persistent kernels: one resident grid, warps pull 32-ray packets from a
// device-side cursor. Shade is a streaming kernel that compacts survivors and shadow items with
// warp ballots + one atomic per warp.
#ifndef PT_TRACE_THREADS
#define PT_TRACE_THREADS 128   // measured, one box, same warps per SM (profiles/r07_ab_cta_size.log): 64 threads x 18 CTAs 3.26, 128 x 9 3.27,
#endif                         // 256 x 4 (64 registers) 3.18 Grays/s on the atrium
// forcing 12 CTAs/SM (40 registers, spills) was measured: atrium +0.4 %, 10M-triangle microbench -37 %
#ifndef PT_TRACE_MIN_BLOCKS
#define PT_TRACE_MIN_BLOCKS 9   // caps the traversal kernels at 56 registers (k_extend takes 64 when left alone = 8 CTAs/SM). Measured, one box,
                                // 8 / 9 / 10 (48 registers, 16-32 B of spills): atrium 3.20 / 3.22 / 3.19 Grays/s, bathroom 2.87 / 2.93 / 2.85,
                                // 10M-triangle batch 3.12 / 3.23 / 3.24, bunny primary rays 4.49 / 4.52 / 4.31
#endif
#ifndef PT_DEFAULT_CARVEOUT
#define PT_DEFAULT_CARVEOUT -1
#define PT_DEFAULT_CARVEOUT_SHADE -1
#endif
#include "pt_kernels.h"
#include "pt_trace.cuh"
#include "pt_trace_wide.cuh"
#include "pt_shade.cuh"
#include <cstdlib>

namespace pt {

#define PT_STREAM_THREADS 256
#define PT_SHADE_THREADS 128
#ifndef PT_SHADE_MIN_BLOCKS
#define PT_SHADE_MIN_BLOCKS 6        // register cap of k_shade: 85 (80 used)
#endif
#define PT_RAY_ORDER_MAX_BITS 6      // 64^3 cells x 8 octants = 2097152 bins = 2048 scan tiles
#define PT_RAY_ORDER_PER_BIN 64

// ------------------------------------------------------------------ fixed ray batches
struct BatchIO {
    const float4* rays; float4* hits; unsigned char* flags; int n; int* cur;
    const int* perm;      // optional processing order (launch_ray_order); results still land at the ray's own index
    __device__ __forceinline__ int count() const { return n; }
    __device__ __forceinline__ int* cursor() const { return cur; }
    __device__ __forceinline__ void load(int i, V3& o, V3& d, float& tmax) const {
        if (perm) i = __ldg(perm + i);
        float4 ro = __ldg(rays + 2 * (size_t)i), rd = __ldg(rays + 2 * (size_t)i + 1);
        o = xyz(ro); d = xyz(rd); tmax = rd.w;
    }
    __device__ __forceinline__ void store(int i, const HitRec& h, bool result) const {
        if (perm) i = __ldg(perm + i);
        if (flags) flags[i] = result ? 1 : 0;
        else st_stream(hits + i, make_float4(h.t, h.u, h.v, __int_as_float(h.tri)));
    }
};

// ---- processing order for big incoherent batches: rays binned by origin cell (Morton, `bits` per axis inside the root
// box) and direction octant with a counting sort over global atomics -- two passes over the rays, no stability needed.
// Neighbouring lanes then start in the same part of the tree and head the same way, which is what the caches (and on a
// tree that outgrows L2, the DRAM row buffers) reward. A ray's own traversal and result do not depend on the order.
__device__ __forceinline__ unsigned spread3(unsigned v) {      // 10 bits -> every third bit
    v = (v | (v << 16)) & 0x030000ffu;
    v = (v | (v << 8)) & 0x0300f00fu;
    v = (v | (v << 4)) & 0x030c30c3u;
    v = (v | (v << 2)) & 0x09249249u;
    return v;
}
__global__ void k_ray_bin(SceneDev s, const float4* __restrict__ rays, int n, int bits, int dbits /* > 0: octahedral direction cells of 2^dbits per axis
                          instead of the octant; < 0: the same, direction-major */, unsigned* counts, unsigned* ray_bin) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4* q = s.pairs[0].q;
    float4 a0 = __ldg(q), a1 = __ldg(q + 1), b0 = __ldg(q + 2), b1 = __ldg(q + 3);
    float lo[3] = {fminf(a0.x, b0.x), fminf(a0.y, b0.y), fminf(a0.z, b0.z)};
    float hi[3] = {fmaxf(a1.x, b1.x), fmaxf(a1.y, b1.y), fmaxf(a1.z, b1.z)};
    float4 ro = __ldg(rays + 2 * (size_t)i), rd = __ldg(rays + 2 * (size_t)i + 1);
    const int cells = 1 << bits;
    float o[3] = {ro.x, ro.y, ro.z};
    unsigned c[3];
    for (int k = 0; k < 3; k++) {
        float ext = hi[k] - lo[k];
        float f = ext > 0.0f ? (o[k] - lo[k]) / ext * (float)cells : 0.0f;
        int v = f >= 0.0f ? (int)fminf(f, (float)(cells - 1)) : 0;            // NaN and outside origins clamp
        c[k] = (unsigned)v;
    }
    unsigned morton = spread3(c[0]) | (spread3(c[1]) << 1) | (spread3(c[2]) << 2);
    unsigned oct = (rd.x < 0.0f ? 1u : 0u) | (rd.y < 0.0f ? 2u : 0u) | (rd.z < 0.0f ? 4u : 0u);
    unsigned bin = (morton << 3) | oct;
    if (dbits != 0) {
        // experiment (PT_RAY_ORDER_DIR_BITS): octahedral map of the direction, 2^|dbits| cells per axis
        const int db = dbits < 0 ? -dbits : dbits;
        const float l1 = fabsf(rd.x) + fabsf(rd.y) + fabsf(rd.z);
        float u = l1 > 0.0f ? rd.x / l1 : 0.0f, v = l1 > 0.0f ? rd.y / l1 : 0.0f;
        if (rd.z < 0.0f) { const float uu = (1.0f - fabsf(v)) * (u >= 0.0f ? 1.0f : -1.0f), vv = (1.0f - fabsf(u)) * (v >= 0.0f ? 1.0f : -1.0f); u = uu; v = vv; }
        const int dc = 1 << db;
        int iu = (int)((u * 0.5f + 0.5f) * (float)dc), iv = (int)((v * 0.5f + 0.5f) * (float)dc);
        iu = iu < 0 ? 0 : (iu >= dc ? dc - 1 : iu); iv = iv < 0 ? 0 : (iv >= dc ? dc - 1 : iv);
        unsigned dkey = 0;                                       // the two cell indices interleaved: 2 * db bits
        for (int b = 0; b < db; b++) dkey |= (((unsigned)iu >> b) & 1u) << (2 * b) | (((unsigned)iv >> b) & 1u) << (2 * b + 1);
        bin = dbits > 0 ? (morton << (2 * db)) | dkey : (dkey << (3 * bits)) | morton;
    }
    ray_bin[i] = bin;
    atomicAdd(counts + bin, 1u);
}
// exclusive scan of n_bins counters in place (n_bins a multiple of 1024, at most 1024 tiles): tile totals, then every
// tile adds the totals of the tiles before it to its own scan -- two small coalesced launches
__global__ void __launch_bounds__(1024) k_ray_bin_tile_sums(const unsigned* __restrict__ counts, unsigned* tile_sums) {
    __shared__ unsigned warp_sums[32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned x = counts[(size_t)blockIdx.x * 1024 + threadIdx.x];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) x += __shfl_down_sync(0xffffffffu, x, o);
    if (lane == 0) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
        unsigned z = warp_sums[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z += __shfl_down_sync(0xffffffffu, z, o);
        if (lane == 0) tile_sums[blockIdx.x] = z;
    }
}
__global__ void __launch_bounds__(1024) k_ray_bin_scan(unsigned* counts, const unsigned* __restrict__ tile_sums) {
    __shared__ unsigned warp_sums[32];
    __shared__ unsigned tile_base;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    // totals of the tiles before this one
    unsigned before = 0u;
    for (int t = threadIdx.x; t < (int)blockIdx.x; t += 1024) before += tile_sums[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) before += __shfl_down_sync(0xffffffffu, before, o);
    if (lane == 0) warp_sums[warp] = before;
    __syncthreads();
    if (warp == 0) {
        unsigned z = warp_sums[lane];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) z += __shfl_down_sync(0xffffffffu, z, o);
        if (lane == 0) tile_base = z;
    }
    __syncthreads();
    unsigned* mine = counts + (size_t)blockIdx.x * 1024 + threadIdx.x;
    unsigned v = *mine, x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) warp_sums[warp] = x;
    __syncthreads();
    if (warp == 0) {
        unsigned w = warp_sums[lane], z = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, z, o); if (lane >= o) z += y; }
        warp_sums[lane] = z - w;
    }
    __syncthreads();
    *mine = tile_base + warp_sums[warp] + (x - v);
}
__global__ void k_ray_bin_scatter(const unsigned* __restrict__ ray_bin, int n, unsigned* cursors, int* perm) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned pos = atomicAdd(cursors + ray_bin[i], 1u);
    perm[pos] = i;
}
int launch_ray_order(const SceneDev& s, const float4* rays, long long n, unsigned* bins, unsigned* ray_bin, int* perm, cudaStream_t st) {
    // about PT_RAY_ORDER_PER_BIN rays per bin, 2..PT_RAY_ORDER_MAX_BITS bits per axis (+3 octant bits)
    int bits = 2, max_bits = PT_RAY_ORDER_MAX_BITS, per_bin = PT_RAY_ORDER_PER_BIN;
    if (const char* e = getenv("PT_RAY_ORDER_BITS")) max_bits = atoi(e);
    if (const char* e = getenv("PT_RAY_ORDER_PER_BIN")) per_bin = atoi(e);
    if (max_bits > PT_RAY_ORDER_MAX_BITS) max_bits = PT_RAY_ORDER_MAX_BITS;
    if (per_bin < 1) per_bin = 1;
    while (bits < max_bits && (n >> (3 * (bits + 1) + 3)) >= per_bin) bits++;
    int n_bins = 1 << (3 * bits + 3);
    int dbits = 0;
    if (const char* e = getenv("PT_RAY_ORDER_DIR_BITS")) dbits = atoi(e);
    if (dbits != 0) {
        // same number of bins, split differently between origin cells and direction cells
        const int total = 3 * bits + 3, db = dbits < 0 ? -dbits : dbits;
        if (2 * db + 3 <= total) { bits = (total - 2 * db) / 3; n_bins = 1 << (3 * bits + 2 * db); } else dbits = 0;
    }
    if (n_bins < 1024) n_bins = 1024;
    const int tiles = n_bins / 1024;
    unsigned* tile_sums = bins + n_bins;
    cudaMemsetAsync(bins, 0, sizeof(unsigned) * n_bins, st);
    const int blocks = (int)((n + 255) / 256);
    k_ray_bin<<<blocks, 256, 0, st>>>(s, rays, (int)n, bits, dbits, bins, ray_bin);
    k_ray_bin_tile_sums<<<tiles, 1024, 0, st>>>(bins, tile_sums);
    k_ray_bin_scan<<<tiles, 1024, 0, st>>>(bins, tile_sums);
    k_ray_bin_scatter<<<blocks, 256, 0, st>>>(ray_bin, (int)n, bins, perm);
    return 4;
}

template <bool ANY, class TR>
__global__ void __launch_bounds__(PT_TRACE_THREADS, PT_TRACE_MIN_BLOCKS)
k_trace_batch(SceneDev s, BatchIO io, int leaf_thresh, int fetch_thresh) {
    trace_persistent<ANY, TR>(s, io, leaf_thresh, fetch_thresh);
}

void launch_trace_closest(const SceneDev& s, const float4* rays, long long n, float4* hits, int* cursor,
                          const LaunchDims& dims, cudaStream_t st, const int* perm) {
    cudaMemsetAsync(cursor, 0, sizeof(int), st);
    BatchIO io{rays, hits, nullptr, (int)n, cursor, perm};
    if (dims.wide == 8) k_trace_batch<false, WideTraverser<8>><<<dims.extend_blocks, PT_TRACE_THREADS, 0, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.wide == 4) k_trace_batch<false, WideTraverser<4>><<<dims.extend_blocks, PT_TRACE_THREADS, 0, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.count_visits) k_trace_batch<false, CountingTraverser><<<dims.extend_blocks, PT_TRACE_THREADS, 64 * s.top_pairs, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else k_trace_batch<false, Traverser><<<dims.extend_blocks, PT_TRACE_THREADS, 64 * s.top_pairs, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
}
void launch_trace_any(const SceneDev& s, const float4* rays, long long n, unsigned char* flags, int* cursor,
                      const LaunchDims& dims, cudaStream_t st, const int* perm) {
    cudaMemsetAsync(cursor, 0, sizeof(int), st);
    BatchIO io{rays, nullptr, flags, (int)n, cursor, perm};
    if (dims.wide == 8) k_trace_batch<true, WideTraverser<8>><<<dims.extend_blocks, PT_TRACE_THREADS, 0, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.wide == 4) k_trace_batch<true, WideTraverser<4>><<<dims.extend_blocks, PT_TRACE_THREADS, 0, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.count_visits) k_trace_batch<true, CountingTraverser><<<dims.extend_blocks, PT_TRACE_THREADS, 64 * s.top_pairs, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
    else k_trace_batch<true, Traverser><<<dims.extend_blocks, PT_TRACE_THREADS, 64 * s.top_pairs, st>>>(s, io, dims.leaf_thresh, dims.fetch_thresh);
}

// ------------------------------------------------------------------ ray generation
template <bool PL>
__global__ void __launch_bounds__(PT_STREAM_THREADS)
k_raygen(FrameParams p) {
    const int n = p.slice_count;
    if (blockIdx.x == 0 && threadIdx.x == 0) p.counters[0].n_paths = n;
    const PathQueue q = p.queue[0];
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        int px, py;
        ray_index_to_morton(p.slice_first + i, p.width, px, py);
        int pix = py * p.width + px;
        uint4 st = p.rng_table[pix];                         // initRNG, Random.glsl:232-237
        Rng rng; rng.x = st.x; rng.y = st.y; rng.z = st.z; rng.w = st.w;
        V3 o, d;
        init_ray<PL>(*p.cam, p.width, p.height, px, py, rng, o, d);
        q.q0[i] = f4(o, 1.0f);
        q.q1[i] = f4(d, 1.0f);
        q.q2[i] = make_float4(1.0f, 1.0f, 1.0f, __int_as_float(pix));
        q.q3[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        q.q4[i] = make_uint4(rng.x, rng.y, rng.z, rng.w);
    }
}

template <bool PL>
__global__ void k_raygen_probe(FrameParams p, long long first, long long n, float4* out3) {
    long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= n) return;
    int px, py;
    ray_index_to_morton((int)(first + k), p.width, px, py);
    int pix = py * p.width + px;
    uint4 st = p.rng_table[pix];
    Rng rng; rng.x = st.x; rng.y = st.y; rng.z = st.z; rng.w = st.w;
    V3 o, d;
    init_ray<PL>(*p.cam, p.width, p.height, px, py, rng, o, d);
    p.rng_table[pix] = make_uint4(rng.x, rng.y, rng.z, rng.w);
    out3[3 * k + 0] = make_float4(__int_as_float(px), __int_as_float(py), o.x, o.y);
    out3[3 * k + 1] = make_float4(o.z, d.x, d.y, d.z);
    out3[3 * k + 2] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

// ------------------------------------------------------------------ extend (closest hit)
struct ExtendIO {
    PathQueue q; float4* hits; BounceCounters* c;
    __device__ __forceinline__ int count() const { return c->n_paths; }
    __device__ __forceinline__ int* cursor() const { return &c->fetch_extend; }
    __device__ __forceinline__ void load(int i, V3& o, V3& d, float& tmax) const {
        float4 ro = ld_stream(q.q0 + i), rd = ld_stream(q.q1 + i);
        o = xyz(ro); d = xyz(rd); tmax = 1e20f;                      // Iterative.comp:232
    }
    __device__ __forceinline__ void store(int i, const HitRec& h, bool) const {
        st_stream(hits + i, make_float4(h.t, h.u, h.v, __int_as_float(h.tri)));
    }
};

template <class TR>
__global__ void __launch_bounds__(PT_TRACE_THREADS, PT_TRACE_MIN_BLOCKS)
k_extend(FrameParams p, int bounce, int qi, int leaf_thresh, int fetch_thresh) {
    ExtendIO io{p.queue[qi], p.hits, p.counters + bounce};
    trace_persistent<false, TR>(p.scene, io, leaf_thresh, fetch_thresh);
}
static void launch_extend(const FrameParams& p, int b, int qi, const LaunchDims& dims, int blocks, cudaStream_t st) {
    if (dims.wide == 8) k_extend<WideTraverser<8>><<<blocks, PT_TRACE_THREADS, 0, st>>>(p, b, qi, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.wide == 4) k_extend<WideTraverser<4>><<<blocks, PT_TRACE_THREADS, 0, st>>>(p, b, qi, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.count_visits) k_extend<CountingTraverser><<<blocks, PT_TRACE_THREADS, 64 * p.scene.top_pairs, st>>>(p, b, qi, dims.leaf_thresh, dims.fetch_thresh);
    else k_extend<Traverser><<<blocks, PT_TRACE_THREADS, 64 * p.scene.top_pairs, st>>>(p, b, qi, dims.leaf_thresh, dims.fetch_thresh);
}

// ------------------------------------------------------------------ shade
template <int MODEL, bool SUN>
__global__ void __launch_bounds__(PT_SHADE_THREADS, PT_SHADE_MIN_BLOCKS)
k_shade(FrameParams p, int bounce, int qi) {
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    BounceCounters* c = p.counters + bounce;
    const int n = c->n_paths;
    const PathQueue qin = p.queue[qi], qout = p.queue[qi ^ 1];
    const ShadowQueue sq = p.shadow;
    const SceneDev& s = p.scene;
    const int warps_total = (gridDim.x * blockDim.x) >> 5;
    const int warp_id = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const float neePdf = 0.5f / s.total_light_area;             // Iterative.comp:203

    for (int base = warp_id * 32; base < n; base += warps_total * 32) {
        const int i = base + lane;
        bool alive = false, has_shadow = false, null_shadow = false;
        // outputs
        V3 vpos = v3(0.0f), newdir = v3(0.0f), thr = v3(0.0f), contrib = v3(0.0f), L = v3(0.0f), ldir = v3(0.0f), lorg = v3(0.0f);
        float pdf0 = 0.0f, pdf1 = 0.0f, tmax = 0.0f;
        int pix = 0;
        Rng rng; rng.x = rng.y = rng.z = rng.w = 0;
        if (i < n) {
            float4 a0 = qin.q0[i], a1 = qin.q1[i], a2 = qin.q2[i], a3 = qin.q3[i];
            uint4 a4 = qin.q4[i];
            float4 hit = p.hits[i];
            V3 ro = xyz(a0), rd = xyz(a1);
            pdf0 = a0.w; pdf1 = a1.w;
            thr = xyz(a2); pix = __float_as_int(a2.w);
            contrib = xyz(a3);
            rng.x = a4.x; rng.y = a4.y; rng.z = a4.z; rng.w = a4.w;
            int tri = __float_as_int(hit.w);
            if (tri < 0) {
                // miss: sky radiance, NOT MIS weighted (Iterative.comp:251-255), store + freeRNG (:262-264)
                contrib = contrib + thr * sample_environment(s, rd);
                float4 acc = p.accum[pix];
                p.accum[pix] = make_float4(acc.x + contrib.x, acc.y + contrib.y, acc.z + contrib.z, 1.0f);
                p.rng_table[pix] = make_uint4(rng.x, rng.y, rng.z, rng.w);
            } else {
                // GetInterpolatedVertex, Geometry.glsl:315-337 + Iterative.comp:236-240
                const float4* tp = s.tris + 5 * (size_t)tri;
                float4 t2 = __ldg(tp + 2), t3 = __ldg(tp + 3), t4 = __ldg(tp + 4);
                float hu = hit.y, hv = hit.z, hw = 1.0f - hit.y - hit.z;
                V3 vnormal = v3(t3.w, t4.x, t4.y);
                float tu = t2.w * hu + t3.y * hv + t2.y * hw;
                float tv = t3.x * hu + t3.z * hv + t2.z * hw;
                uint32_t matId = __float_as_uint(t4.z);
                vpos = ro + rd * hit.x;
                vpos = vpos + 0.003f * vnormal;
                V3 viewDir = -rd;

                Material mat = make_material(s, matId, tu, tv);
                Interaction it = make_interaction(vnormal, viewDir);

                V3 idt = thr;
                if (bounce != 0) {                              // Iterative.comp:256-261
                    float dist = distance(ro, vpos);
                    float oldMis = mis_weight(pdf0, pdf1);
                    float g = 0.5f * avdot(vnormal, viewDir) / (dist * dist);
                    pdf0 = pdf0 * g;
                    pdf1 = pdf1 * g;
                    idt = (thr / oldMis) * pdf0 / (pdf0 + pdf1 + neePdf);
                }
                contrib = contrib + idt * mat.emission;

                // GenerateLightSample, Iterative.comp:113-147
                V3 lpos, lnrm; uint32_t lmat;
                float lightPdf, radius;
                if (SUN) {
                    // #ifdef SUN_LIGHT, Iterative.comp:116-126: a coin flip between the emitter table and the sun disk
                    float lightPick = rng.next();
                    if (lightPick > 0.5f) {
                        random_light_vertex(s, rng, lpos, lnrm, lmat);
                        lightPdf = 0.5f / s.total_light_area;
                        radius = distance(lpos, vpos);
                        tmax = radius - 0.005f;
                    } else {
                        random_sun_vertex<PT_MODEL_PLIBM(MODEL)>(s, rng, vpos, lpos, lnrm, lmat);
                        float sunArea = PT_PI * s.sun_radius * s.sun_radius;              // Iterative.comp:92
                        lightPdf = 0.5f / sunArea;
                        radius = distance(lpos, vpos);
                        tmax = 1e20f;
                    }
                } else {
                    random_light_vertex(s, rng, lpos, lnrm, lmat);
                    lightPdf = 1.0f / s.total_light_area;
                    radius = distance(lpos, vpos);
                    tmax = radius - 0.005f;
                }
                ldir = normalize(lpos - vpos);
                {
                    Interaction li = it;
                    set_incoming(li, ldir);
                    float bouncePdf = pdf_direction<MODEL>(mat, li) * avdot(lnrm, -ldir) / (radius * radius);
                    V3 lightThroughput = thr * bsdf<MODEL>(mat, li) * li.ndi * avdot(lnrm, -ldir) *
                                         mis_weight(lightPdf, bouncePdf) / (lightPdf * radius * radius);
                    float4 le = __ldg(&s.materials[lmat >> 1].emission_flag);
                    L = lightThroughput * xyz(le);              // Iterative.comp:279
                }
                lorg = vpos + 0.001f * vnormal;
                // The reference traces this ray unconditionally (Iterative.comp:276-280). When L is exactly zero (light below
                // the surface's horizon: n.i = 0) the answer cannot change the sum -- contribution + 0 -- so the sample is
                // settled here and never queued. NaN / Inf radiance is not zero and takes the normal route.
                has_shadow = p.trace_null || !(L.x == 0.0f && L.y == 0.0f && L.z == 0.0f);
                null_shadow = !has_shadow;

                // BSDF sample, Iterative.comp:282-288
                newdir = importance_sample<MODEL>(mat, it, rng, pdf0, pdf1);
                thr = thr * (bsdf<MODEL>(mat, it) * it.ndi * mis_weight(pdf0, pdf1) / pdf0);

                // Russian roulette, Iterative.comp:290-301
                float continuation = gmin(gmax(thr.x, gmax(thr.y, thr.z)), 1.0f);
                thr = thr / continuation;
                bool terminate = (rng.next() > continuation) || (bounce > p.max_depth - 2);
                alive = !terminate;
                if (terminate) p.rng_table[pix] = make_uint4(rng.x, rng.y, rng.z, rng.w);   // freeRNG
            }
        }
        // ---- compaction: one atomic per warp per queue
        unsigned m_alive = __ballot_sync(0xffffffffu, alive);
        unsigned m_shadow = __ballot_sync(0xffffffffu, has_shadow);
        unsigned m_null = __ballot_sync(0xffffffffu, null_shadow);
        int base_alive = 0, base_shadow = 0;
        if (lane == 0) {
            if (m_alive) base_alive = atomicAdd(&c[1].n_paths, __popc(m_alive));
            if (m_shadow) base_shadow = atomicAdd(&c->n_shadow, __popc(m_shadow));
            if (m_null) atomicAdd(&c->n_null, __popc(m_null));
        }
        base_alive = __shfl_sync(0xffffffffu, base_alive, 0);
        base_shadow = __shfl_sync(0xffffffffu, base_shadow, 0);
        int dest = ~pix;
        if (alive) {
            int j = base_alive + __popc(m_alive & lt_mask);
            dest = j;
            qout.q0[j] = f4(vpos, pdf0);
            qout.q1[j] = f4(newdir, pdf1);
            qout.q2[j] = f4(thr, __int_as_float(pix));
            qout.q4[j] = make_uint4(rng.x, rng.y, rng.z, rng.w);
            // q3 (contribution) is written by k_shadow once visibility is known -- or here when no ray is needed
            if (null_shadow) qout.q3[j] = f4(contrib, 0.0f);
        } else if (null_shadow) {
            float4 acc = p.accum[pix];                               // imageStore, Iterative.comp:295
            p.accum[pix] = make_float4(acc.x + contrib.x, acc.y + contrib.y, acc.z + contrib.z, 1.0f);
        }
        if (has_shadow) {
            int k = base_shadow + __popc(m_shadow & lt_mask);
            sq.s0[k] = f4(lorg, tmax);
            sq.s1[k] = f4(ldir, __int_as_float(dest));
            sq.s2[k] = f4(L, 0.0f);
            sq.s3[k] = f4(contrib, 0.0f);
        }
    }
}

// ------------------------------------------------------------------ shadow (any hit) + NEE add + store
struct ShadowIO {
    ShadowQueue sq; PathQueue qn; float4* accum; BounceCounters* c;
    __device__ __forceinline__ int count() const { return c->n_shadow; }
    __device__ __forceinline__ int* cursor() const { return &c->fetch_shadow; }
    __device__ __forceinline__ void load(int i, V3& o, V3& d, float& tmax) const {
        float4 s0 = ld_stream(sq.s0 + i), s1 = ld_stream(sq.s1 + i);
        o = xyz(s0); d = xyz(s1); tmax = s0.w;
    }
    __device__ __forceinline__ void store(int i, const HitRec&, bool occluded) const {
        float4 s2 = ld_stream(sq.s2 + i), s3 = ld_stream(sq.s3 + i);
        int dest = __float_as_int(sq.s1[i].w);
        V3 contrib = xyz(s3);
        if (!occluded) contrib = contrib + xyz(s2);                  // Iterative.comp:278-280
        if (dest >= 0) {
            st_stream(qn.q3 + dest, f4(contrib, 0.0f));
        } else {
            int pix = ~dest;                                         // imageStore, Iterative.comp:295
            float4 acc = accum[pix];
            accum[pix] = make_float4(acc.x + contrib.x, acc.y + contrib.y, acc.z + contrib.z, 1.0f);
        }
    }
};

template <class TR>
__global__ void __launch_bounds__(PT_TRACE_THREADS, PT_TRACE_MIN_BLOCKS)
k_shadow(FrameParams p, int bounce, int qi_next, int leaf_thresh, int fetch_thresh) {
    ShadowIO io{p.shadow, p.queue[qi_next], p.accum, p.counters + bounce};
    trace_persistent<true, TR>(p.scene, io, leaf_thresh, fetch_thresh);
}
static void launch_shadow(const FrameParams& p, int b, int qi_next, const LaunchDims& dims, int blocks, cudaStream_t st) {
    if (dims.wide == 8) k_shadow<WideTraverser<8>><<<blocks, PT_TRACE_THREADS, 0, st>>>(p, b, qi_next, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.wide == 4) k_shadow<WideTraverser<4>><<<blocks, PT_TRACE_THREADS, 0, st>>>(p, b, qi_next, dims.leaf_thresh, dims.fetch_thresh);
    else if (dims.count_visits) k_shadow<CountingTraverser><<<blocks, PT_TRACE_THREADS, 64 * p.scene.top_pairs, st>>>(p, b, qi_next, dims.leaf_thresh, dims.fetch_thresh);
    else k_shadow<Traverser><<<blocks, PT_TRACE_THREADS, 64 * p.scene.top_pairs, st>>>(p, b, qi_next, dims.leaf_thresh, dims.fetch_thresh);
}

__global__ void k_tally(FrameParams p) {
    if (threadIdx.x != 0 || blockIdx.x != 0) return;
    unsigned long long closest = 0, shadow = 0, nulls = 0;
    for (int b = 0; b < p.max_depth; b++) { closest += p.counters[b].n_paths; shadow += p.counters[b].n_shadow; nulls += p.counters[b].n_null; }
    // slices of one frame tally concurrently from different streams
    atomicAdd(&p.totals[0], closest); atomicAdd(&p.totals[1], shadow); atomicAdd(&p.totals[2], (unsigned long long)p.slice_count);
    atomicAdd(&p.totals[3], nulls);
}

template <bool SUN>
static void launch_shade_model(const FrameParams& p, int b, int qi, const LaunchDims& dims, cudaStream_t st) {
    switch ((p.model & 1) | (p.plibm ? 2 : 0)) {     // bit 0: Beckmann, bit 1: PT_PARITY_LIBM (pt_shade.cuh)
    case 0: k_shade<0, SUN><<<dims.shade_blocks, PT_SHADE_THREADS, 0, st>>>(p, b, qi); break;
    case 1: k_shade<1, SUN><<<dims.shade_blocks, PT_SHADE_THREADS, 0, st>>>(p, b, qi); break;
    case 2: k_shade<2, SUN><<<dims.shade_blocks, PT_SHADE_THREADS, 0, st>>>(p, b, qi); break;
    default: k_shade<3, SUN><<<dims.shade_blocks, PT_SHADE_THREADS, 0, st>>>(p, b, qi); break;
    }
}
static void launch_shade(const FrameParams& p, int b, int qi, const LaunchDims& dims, cudaStream_t st) {
    if (p.sun) launch_shade_model<true>(p, b, qi, dims, st); else launch_shade_model<false>(p, b, qi, dims, st);
}
static void launch_raygen(const FrameParams& p, const LaunchDims& dims, cudaStream_t st) {
    if (p.plibm) k_raygen<true><<<dims.stream_blocks, PT_STREAM_THREADS, 0, st>>>(p);
    else k_raygen<false><<<dims.stream_blocks, PT_STREAM_THREADS, 0, st>>>(p);
}

int launch_frame(const FrameParams& p, const LaunchDims& dims, cudaStream_t st, cudaEvent_t* evs) {
    // evs (optional): one event recorded before every kernel and one after the last, for per-stage timing
    int launches = 0;
    cudaMemsetAsync(p.counters, 0, sizeof(BounceCounters) * (p.max_depth + 2), st);
    if (evs) cudaEventRecord(evs[launches], st);
    launch_raygen(p, dims, st); launches++;
    int qi = 0;
    for (int b = 0; b < p.max_depth; b++) {
        if (evs) cudaEventRecord(evs[launches], st);
        launch_extend(p, b, qi, dims, dims.extend_blocks, st); launches++;
        if (evs) cudaEventRecord(evs[launches], st);
        launch_shade(p, b, qi, dims, st); launches++;
        if (evs) cudaEventRecord(evs[launches], st);
        launch_shadow(p, b, qi ^ 1, dims, dims.extend_blocks, st); launches++;
        qi ^= 1;
    }
    if (evs) cudaEventRecord(evs[launches], st);
    k_tally<<<1, 32, 0, st>>>(p); launches++;
    if (evs) cudaEventRecord(evs[launches], st);
    return launches;
}

int enqueue_frame_dag(const FrameParams& p, const LaunchDims& dims, const SliceStreams& ss, bool wait_prev) {
    int launches = 0;
    cudaStream_t A = ss.a, B = ss.b;
    cudaMemsetAsync(p.counters, 0, sizeof(BounceCounters) * (p.max_depth + 2), A);
    launch_raygen(p, dims, A); launches++;
    launch_extend(p, 0, 0, dims, dims.frame_blocks, A); launches++;
    // shade(0) rewrites the shadow queue and accum pixels the previous frame's last shadow kernel may still be using
    if (wait_prev) cudaStreamWaitEvent(A, ss.ev_shadow, 0);
    int qi = 0;
    for (int b = 0; b < p.max_depth; b++) {
        launch_shade(p, b, qi, dims, A); launches++;
        cudaEventRecord(ss.ev_shade, A);
        cudaStreamWaitEvent(B, ss.ev_shade, 0);
        launch_shadow(p, b, qi ^ 1, dims, dims.frame_blocks, B); launches++;
        if (b + 1 == p.max_depth) { k_tally<<<1, 32, 0, B>>>(p); launches++; }
        cudaEventRecord(ss.ev_shadow, B);
        qi ^= 1;
        if (b + 1 < p.max_depth) {
            launch_extend(p, b + 1, qi, dims, dims.frame_blocks, A); launches++;
            cudaStreamWaitEvent(A, ss.ev_shadow, 0);         // shade(b+1) reuses the shadow queue and reads q3 written by shadow(b)
        }
    }
    return launches;
}

// ---- primary rays only (pt_trace_primary_async): the first two kernels of a frame -- InitRay for every pixel, closest hit --
// and the hit records put in pixel order (the queue is in 8x8 Morton tile order, Iterative.comp:163-172)
__global__ void k_primary_gather(FrameParams p, float4* by_pixel) {
    const int n = p.slice_count;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x)
        by_pixel[__float_as_int(p.queue[0].q2[i].w)] = p.hits[i];
}
int launch_primary(const FrameParams& p, const LaunchDims& dims, float4* by_pixel, cudaStream_t st) {
    cudaMemsetAsync(p.counters, 0, sizeof(BounceCounters) * (p.max_depth + 2), st);
    launch_raygen(p, dims, st);
    launch_extend(p, 0, 0, dims, dims.extend_blocks, st);
    k_primary_gather<<<dims.stream_blocks, PT_STREAM_THREADS, 0, st>>>(p, by_pixel);
    return 3;
}

void launch_raygen_probe(const FrameParams& p, long long first, long long n, float4* out3, cudaStream_t st) {
    int blocks = (int)((n + 255) / 256);
    if (blocks > 0) {
        if (p.plibm) k_raygen_probe<true><<<blocks, 256, 0, st>>>(p, first, n, out3);
        else k_raygen_probe<false><<<blocks, 256, 0, st>>>(p, first, n, out3);
    }
}

// ------------------------------------------------------------------ RNG table seeding
// Renderer.cpp:545-553 draws every lane from uniform_int_distribution(129, UINT32_MAX) (HybridTaus
// needs its Tausworthe lanes > 128). Here: splitmix64 of (seed, pixel, lane), folded into [129, 2^32).
__device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__global__ void k_seed_rng(uint4* table, long long n, unsigned long long seed) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    unsigned long long a = splitmix64(seed ^ splitmix64((unsigned long long)i * 2 + 0));
    unsigned long long b = splitmix64(seed ^ splitmix64((unsigned long long)i * 2 + 1));
    auto lanefix = [](unsigned v) { return v < 129u ? v + 129u : v; };
    table[i] = make_uint4(lanefix((unsigned)a), lanefix((unsigned)(a >> 32)), lanefix((unsigned)b), lanefix((unsigned)(b >> 32)));
}
void launch_seed_rng(uint4* table, long long n, unsigned long long seed, cudaStream_t st) {
    int blocks = (int)((n + 255) / 256);
    if (blocks > 0) k_seed_rng<<<blocks, 256, 0, st>>>(table, n, seed);
}

// ------------------------------------------------------------------ resolve (Present.frag:13-37)
template <bool PL>
__global__ void k_resolve(const float4* __restrict__ accum, long long n, unsigned num_samples, float* rgb_f32,
                          unsigned char* rgb_u8, int first, int width) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (width > 0) {                   // slice form: i counts ray indices from `first`; the pixel is their Morton-tile position
        int px, py;
        ray_index_to_morton(first + (int)i, width, px, py);
        i = (long long)py * width + px;
    }
    const float A = 0.15f, B = 0.50f, C = 0.10f, D = 0.20f, E = 0.02f, F = 0.30f, W = 11.2f;
    const float white = ((W * (A * W + C * B) + D * E) / (W * (A * W + B) + D * F)) - E / F;
    float4 a = accum[i];
    float in[3] = {a.x, a.y, a.z};
    for (int c = 0; c < 3; c++) {
        float color = in[c] / (float)num_samples;
        color = 1.68f * color;                                 // kExposure, Renderer.cpp:18
        color *= 2.0f;
        color = ((color * (A * color + C * B) + D * E) / (color * (A * color + B) + D * F)) - E / F;
        color /= white;
        color = pow_sel<PL>(color, 1.0f / 2.2f);
        if (rgb_f32) rgb_f32[3 * i + c] = color;
        if (rgb_u8) rgb_u8[3 * i + c] = (unsigned char)(gmin(gmax(color, 0.0f), 1.0f) * 255.0f + 0.5f);
    }
}
void launch_resolve(const float4* accum, long long n, unsigned num_samples, float* rgb_f32, unsigned char* rgb_u8,
                    cudaStream_t st, bool plibm) {
    int blocks = (int)((n + 255) / 256);
    if (blocks <= 0) return;
    if (plibm) k_resolve<true><<<blocks, 256, 0, st>>>(accum, n, num_samples, rgb_f32, rgb_u8, 0, 0);
    else k_resolve<false><<<blocks, 256, 0, st>>>(accum, n, num_samples, rgb_f32, rgb_u8, 0, 0);
}
void launch_resolve_slice(const float4* accum, int first, int count, int width, unsigned num_samples, unsigned char* rgb_u8,
                          cudaStream_t st, bool plibm) {
    int blocks = (count + 255) / 256;
    if (blocks <= 0) return;
    if (plibm) k_resolve<true><<<blocks, 256, 0, st>>>(accum, count, num_samples, nullptr, rgb_u8, first, width);
    else k_resolve<false><<<blocks, 256, 0, st>>>(accum, count, num_samples, nullptr, rgb_u8, first, width);
}

// ------------------------------------------------------------------ pt_parity_libm.h on the device, argument by argument
// fn: 0 sin, 1 cos, 2 log, 3 exp, 4 pow(x, y). The CPU checkers run the same header; tests require equal bits.
__global__ void k_plibm_probe(int fn, const float* __restrict__ x, const float* __restrict__ y, long long n, float* out) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    float a = x[i], r;
    switch (fn) {
    case 0: r = ptpl::sinf_(a); break;
    case 1: r = ptpl::cosf_(a); break;
    case 2: r = ptpl::logf_(a); break;
    case 3: r = ptpl::expf_(a); break;
    default: r = ptpl::powf_(a, y[i]); break;
    }
    out[i] = r;
}
void launch_plibm_probe(int fn, const float* x, const float* y, long long n, float* out, cudaStream_t st) {
    int blocks = (int)((n + 255) / 256);
    if (blocks > 0) k_plibm_probe<<<blocks, 256, 0, st>>>(fn, x, y, n, out);
}

// ------------------------------------------------------------------ equirectangular -> cubemap
// LoadEnvironmnet, .hdr branch (Renderer.cpp:262-312): the reference draws a unit cube six times with a 90-degree
// projection (glm::perspective(radians(90), 1, ..): focal length 1) and the six lookAt views of :286-294 into the faces of
// an RGB32F cubemap; EquirectangularConverter.vert passes sampleCoords = -position, EquirectangularConverter.frag:9-19
// looks the equirect image up at (atan(v.z, v.x), asin(v.y)) * (0.1591, 0.3183) + 0.5 of the normalised vector
// (GL_LINEAR, CLAMP_TO_EDGE, row 0 of the file = t 0). Per face texel that is: position on the cube face
// p = f + nx * s + ny * u with (nx, ny) the NDC of the texel centre and (s, u, f) the lookAt basis; colour =
// equirect(normalize(-p)); alpha reads back as 1 (RGB32F storage).
__global__ void k_equirect_to_cube(const float4* __restrict__ src, int w, int h, float4* __restrict__ faces, int n) {
    long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= 6ll * n * n) return;
    int face = (int)(i / ((long long)n * n));
    int py = (int)((i / n) % n), px = (int)(i % n);
    const float centre[6][3] = {{1, 0, 0}, {-1, 0, 0}, {0, 1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}};
    const float upv[6][3] = {{0, -1, 0}, {0, -1, 0}, {0, 0, 1}, {0, 0, -1}, {0, -1, 0}, {0, -1, 0}};
    V3 f = v3(centre[face][0], centre[face][1], centre[face][2]);
    V3 up = v3(upv[face][0], upv[face][1], upv[face][2]);
    V3 sv = cross(f, up), uv = cross(sv, f);                       // glm::lookAt, axis-aligned: already unit length
    float nx = ((float)px + 0.5f) / (float)n * 2.0f - 1.0f, ny = ((float)py + 0.5f) / (float)n * 2.0f - 1.0f;
    V3 p = f + sv * nx + uv * ny;
    V3 d = normalize(-p);
    if (d.z == 0.0f) d.z = 0.0f;      // -0 -> +0: the seam column of odd-sized faces must not depend on atan2f(-0, x < 0)
    float tu = atan2f(d.z, d.x) * 0.1591f + 0.5f, tv = asinf(d.y) * 0.3183f + 0.5f;
    float x = tu * (float)w - 0.5f, y = tv * (float)h - 0.5f;
    float fx = floorf(x), fy = floorf(y);
    float wx = x - fx, wy = y - fy;
    int x0 = min(max((int)fx, 0), w - 1), x1 = min(max((int)fx + 1, 0), w - 1);
    int y0 = min(max((int)fy, 0), h - 1), y1 = min(max((int)fy + 1, 0), h - 1);
    float4 a = __ldg(src + (size_t)y0 * w + x0), b = __ldg(src + (size_t)y0 * w + x1);
    float4 c = __ldg(src + (size_t)y1 * w + x0), e = __ldg(src + (size_t)y1 * w + x1);
    faces[i] = make_float4(mixf(mixf(a.x, b.x, wx), mixf(c.x, e.x, wx), wy), mixf(mixf(a.y, b.y, wx), mixf(c.y, e.y, wx), wy),
                           mixf(mixf(a.z, b.z, wx), mixf(c.z, e.z, wx), wy), 1.0f);
}
void launch_equirect_to_cube(const float4* src, int w, int h, float4* faces, int n, cudaStream_t st) {
    long long total = 6ll * n * n;
    int blocks = (int)((total + 255) / 256);
    if (blocks > 0) k_equirect_to_cube<<<blocks, 256, 0, st>>>(src, w, h, faces, n);
}

// Steps a lane takes per vote round of trace_persistent (binary traverser). One step per round (the original scheme)
// spends ~28 of every ~90 instructions on ballots and refill checks; measured on the headline config, one box:
// burst 1: 3.02 Grays/s, 2: 3.06, 4: 3.23, 8: 3.26, 12: 3.28, 32: 3.19, unbounded (while-while): 3.03;
// + 2 leaf steps per round: 3.35; + leaf_thresh 8 -> 4: 3.38. Results are unaffected (a lane's own step sequence is
// the same whatever the schedule).
#ifndef PT_INNER_BURST
#define PT_INNER_BURST 12
#endif
#ifndef PT_LEAF_BURST
#define PT_LEAF_BURST 2
#endif
// occupancy-derived persistent grid sizes
void query_launch_dims(int device, int wide, LaunchDims* out) {
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, device);
    int per_sm_ext = 1, per_sm_shadow = 1, per_sm_shade = 1;
    out->wide = wide;
    out->count_visits = 0;
    out->top_pairs = 0;
    if (wide == 8) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_ext, k_extend<WideTraverser<8>>, PT_TRACE_THREADS, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_shadow, k_shadow<WideTraverser<8>>, PT_TRACE_THREADS, 0);
    } else if (wide == 4) {
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_ext, k_extend<WideTraverser<4>>, PT_TRACE_THREADS, 0);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_shadow, k_shadow<WideTraverser<4>>, PT_TRACE_THREADS, 0);
    } else {
        out->top_pairs = 0;
#ifdef PT_SMEM_TOP   // only when the staging code is compiled in (pt_trace.cuh)
        if (const char* e = getenv("PT_SMEM_TOP")) out->top_pairs = atoi(e);
#endif
        if (out->top_pairs < 0) out->top_pairs = 0;
        if (out->top_pairs > 700) out->top_pairs = 700;                      // 44.8 KB: below the 48-KB dynamic limit without opt-in
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_ext, k_extend<Traverser>, PT_TRACE_THREADS, 64 * out->top_pairs);
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_shadow, k_shadow<Traverser>, PT_TRACE_THREADS, 64 * out->top_pairs);
    }
    // Shared-memory carve-out of the frame kernels (percent of the SM's 228 KB; -1 = the driver's choice). The driver gives
    // the traversal kernels 100 KB although 9 CTAs x (4 KB stack + 1 KB) need 45 KB, and k_shade 32 KB: unequal preferences
    // and less L1 than possible for kernels that live on their L1 hit rate.
    {
        int carve = PT_DEFAULT_CARVEOUT, carve_shade = PT_DEFAULT_CARVEOUT_SHADE;
        if (const char* e = getenv("PT_CARVEOUT")) carve = atoi(e);
        if (const char* e = getenv("PT_CARVEOUT_SHADE")) carve_shade = atoi(e);
        if (carve >= 0) {
            cudaFuncSetAttribute(k_extend<Traverser>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_shadow<Traverser>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_trace_batch<false, Traverser>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
            cudaFuncSetAttribute(k_trace_batch<true, Traverser>, cudaFuncAttributePreferredSharedMemoryCarveout, carve);
        }
        if (carve_shade >= 0) {
            cudaFuncSetAttribute(k_shade<0, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve_shade);
            cudaFuncSetAttribute(k_shade<1, false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve_shade);
            cudaFuncSetAttribute(k_raygen<false>, cudaFuncAttributePreferredSharedMemoryCarveout, carve_shade);
        }
    }
    cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_shade, k_shade<0, false>, PT_SHADE_THREADS, 0);
    int per = per_sm_ext < per_sm_shadow ? per_sm_ext : per_sm_shadow;
    if (per < 1) per = 1;
    if (per_sm_shade < 1) per_sm_shade = 1;
    out->sm_count = prop.multiProcessorCount;
    out->extend_blocks = prop.multiProcessorCount * per;
    out->shade_blocks = prop.multiProcessorCount * per_sm_shade;
    out->stream_blocks = prop.multiProcessorCount * 4;
    // warp-scheduling thresholds (see pt_trace.cuh); env overrides exist for tuning runs only
    out->leaf_thresh = wide ? 8 : 4; out->fetch_thresh = 12;
    if (const char* e = getenv("PT_LEAF_THRESH")) out->leaf_thresh = atoi(e);
    {
        int burst = wide ? 2 : PT_INNER_BURST;             // wide nodes: 2 / 2 measured best (10M-triangle batch 2.95 -> 3.11 Grays/s)
        if (const char* e = getenv("PT_INNER_BURST")) burst = atoi(e);
        if (burst < 1) burst = 1;
        int leaf_burst = wide ? 2 : PT_LEAF_BURST;
        if (const char* e = getenv("PT_LEAF_BURST")) leaf_burst = atoi(e);
        if (leaf_burst < 1) leaf_burst = 1;
        if (leaf_burst > 255) leaf_burst = 255;
        out->leaf_thresh = (out->leaf_thresh & 0xff) | (leaf_burst << 8) | (burst << 16);      // decoded in trace_persistent
    }
    if (const char* e = getenv("PT_FETCH_THRESH")) out->fetch_thresh = atoi(e);
    if (const char* e = getenv("PT_BLOCKS_PER_SM")) out->extend_blocks = prop.multiProcessorCount * atoi(e);
    // Grid of k_extend / k_shadow inside the frame DAG (enqueue_frame_dag), where shadow(b), extend(b+1) and the other
    // slices' kernels are in flight together: a persistent grid that fills the machine keeps the next kernel's CTAs out
    // until its own drain phase, 4 CTAs/SM let two or three of them share every SM. Measured on the headline config
    // (one box, 3 slices): 9 CTAs/SM 2.070 ms/frame, 6: 2.034, 5: 2.017, 4: 2.009, 3: 2.005, 2: 2.110 -- while the same
    // kernel ALONE gets slower (extend 1.03 -> 1.30 ms at 4), so sequential launches (launch_frame, ray batches) keep the
    // full grid.
    out->frame_blocks = prop.multiProcessorCount * (per < 4 ? per : 4);
    if (const char* e = getenv("PT_FRAME_BLOCKS_PER_SM")) out->frame_blocks = prop.multiProcessorCount * atoi(e);
    if (const char* e = getenv("PT_SHADE_BLOCKS_PER_SM")) out->shade_blocks = prop.multiProcessorCount * atoi(e);
}

}  // namespace pt
