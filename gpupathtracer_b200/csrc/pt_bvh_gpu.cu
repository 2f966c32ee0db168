// pt_bvh_gpu.cu -- BVH construction on the device (pt_build_bvh_device): the fast-build alternative to the host SBVH
// builder (pt_bvh_build.cpp) for BASELINE.json configs[0] ("SAH BVH build + 1 spp primary").
//
// Replaces BoundingVolumeHierarchy::BuildBinnedSpatial (src/core/BVH.cpp:2197-2293) by a different algorithm with the
// SAME output contract: the reference's serialised tree (32-B NewLayout nodes, BVH.cpp:2261-2277: sibling pairs at
// (firstChild, firstChild + 1), firstChild odd, larger-area child first, root = 0) plus the reference list with the
// negated end-of-leaf entry (ConvertIntoLeaf, :2107-2121) -- so the oracle, the reference-layout kernels and the wide
// collapse consume it unchanged. Algorithm: 30-bit Morton codes of the triangle centroids, LSD radix sort (own
// kernels, 4 x 8 bits), the fully parallel hierarchy of Karras (HPG 2012) with the triangle index as tie-break,
// bottom-up box fitting with one atomic flag per inner node, one triangle per leaf. No spatial splits, no SAH
// search: the tree is built in ~1 ms instead of tens of ms and costs more to traverse (bench.py reports both).
#include "pt_kernels.h"
#include <cuda_runtime.h>
#include <cfloat>

namespace pt {

#define BVH_THREADS 256

// ---------------------------------------------------------------- scene bounds + Morton keys
__device__ __forceinline__ unsigned f2ord(float f) {           // order-preserving float -> uint (for atomicMin/Max)
    unsigned u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(unsigned u) {
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7fffffffu) : ~u);
}

// tris: n x 5 float4 (pre-subtracted edges): {p0, e1.x} {e1.yz, e2.xy} {e2.z, ...}
__device__ __forceinline__ void tri_box(const float4* t, float lo[3], float hi[3]) {
    float4 a = t[0], b = t[1], c = t[2];
    float p0[3] = {a.x, a.y, a.z}, e1[3] = {a.w, b.x, b.y}, e2[3] = {b.z, b.w, c.x};
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float p1 = p0[k] + e1[k], p2 = p0[k] + e2[k];
        lo[k] = fminf(p0[k], fminf(p1, p2));
        hi[k] = fmaxf(p0[k], fmaxf(p1, p2));
    }
}

__global__ void k_scene_bounds(const float4* __restrict__ tris, int n, unsigned* bounds /* 3 min, 3 max (ordered uints) */) {
    __shared__ unsigned s[6];
    if (threadIdx.x < 3) s[threadIdx.x] = 0xffffffffu; else if (threadIdx.x < 6) s[threadIdx.x] = 0u;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float lo[3], hi[3];
        tri_box(tris + 5 * (size_t)i, lo, hi);
#pragma unroll
        for (int k = 0; k < 3; k++) {
            // centroid bounds are what the Morton grid needs; triangle boxes are fitted later
            float c = 0.5f * (lo[k] + hi[k]);
            atomicMin(&s[k], f2ord(c)); atomicMax(&s[3 + k], f2ord(c));
        }
    }
    __syncthreads();
    if (threadIdx.x < 3) atomicMin(&bounds[threadIdx.x], s[threadIdx.x]);
    else if (threadIdx.x < 6) atomicMax(&bounds[threadIdx.x], s[threadIdx.x]);
}

__device__ __forceinline__ unsigned expand10(unsigned v) {      // 10 bits -> every third bit
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void k_morton(const float4* __restrict__ tris, int n, const unsigned* __restrict__ bounds, unsigned* keys, int* vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo[3], hi[3];
    tri_box(tris + 5 * (size_t)i, lo, hi);
    unsigned q[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float bl = ord2f(bounds[k]), bh = ord2f(bounds[3 + k]);
        float ext = bh - bl;
        float c = 0.5f * (lo[k] + hi[k]);
        float u = ext > 0.0f ? (c - bl) / ext : 0.0f;
        q[k] = (unsigned)fminf(fmaxf(u * 1024.0f, 0.0f), 1023.0f);
    }
    keys[i] = (expand10(q[0]) << 2) | (expand10(q[1]) << 1) | expand10(q[2]);
    vals[i] = i;
}

// ---------------------------------------------------------------- LSD radix sort, 8 bits per pass, stable
// Every block owns one contiguous tile of SORT_TILE elements. Pass = histogram (digit-major matrix hist[256][blocks]),
// exclusive scan of that matrix, stable scatter (ranks inside a tile by warp match + per-warp digit counts).
#define SORT_ITEMS 8
#define SORT_TILE (BVH_THREADS * SORT_ITEMS)

__global__ void __launch_bounds__(BVH_THREADS)
k_sort_hist(const unsigned* __restrict__ keys, int n, int shift, unsigned* hist, int n_blocks) {
    __shared__ unsigned h[256];
    h[threadIdx.x] = 0;
    __syncthreads();
    const int base = blockIdx.x * SORT_TILE;
#pragma unroll
    for (int k = 0; k < SORT_ITEMS; k++) {
        int i = base + k * BVH_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&h[(keys[i] >> shift) & 255u], 1u);
    }
    __syncthreads();
    hist[(size_t)threadIdx.x * n_blocks + blockIdx.x] = h[threadIdx.x];
}

// exclusive scan of `count` values in place, one block
__global__ void __launch_bounds__(1024)
k_sort_scan(unsigned* data, int count) {
    __shared__ unsigned warp_sums[32];
    __shared__ unsigned carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int base = 0; base < count; base += 1024) {
        int i = base + threadIdx.x;
        unsigned v = i < count ? data[i] : 0u, x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        if (lane == 31) warp_sums[warp] = x;
        __syncthreads();
        if (warp == 0) {
            unsigned w = warp_sums[lane], z = w;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, z, o); if (lane >= o) z += y; }
            warp_sums[lane] = z - w;                           // exclusive prefix of the warp totals
        }
        __syncthreads();
        unsigned excl = carry + warp_sums[warp] + (x - v);
        if (i < count) data[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
}

__global__ void __launch_bounds__(BVH_THREADS)
k_sort_scatter(const unsigned* __restrict__ keys_in, const int* __restrict__ vals_in, unsigned* keys_out, int* vals_out,
               int n, int shift, const unsigned* __restrict__ hist, int n_blocks) {
    __shared__ unsigned base[256];
    __shared__ unsigned warp_hist[BVH_THREADS / 32][256];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned lt_mask = (1u << lane) - 1u;
    base[threadIdx.x] = hist[(size_t)threadIdx.x * n_blocks + blockIdx.x];
    for (int w = 0; w < BVH_THREADS / 32; w++) warp_hist[w][threadIdx.x] = 0;
    __syncthreads();
    const int tile = blockIdx.x * SORT_TILE;
    for (int k = 0; k < SORT_ITEMS; k++) {
        const int i = tile + k * BVH_THREADS + threadIdx.x;
        const bool valid = i < n;
        unsigned key = valid ? keys_in[i] : 0u;
        unsigned d = valid ? ((key >> shift) & 255u) : 256u + (unsigned)lane;   // invalid lanes match nobody
        unsigned peers = __match_any_sync(0xffffffffu, d);
        unsigned rank = __popc(peers & lt_mask);
        if (valid && rank == 0) warp_hist[warp][d] = __popc(peers);
        __syncthreads();
        if (valid) {
            unsigned off = base[d] + rank;
            for (int w = 0; w < warp; w++) off += warp_hist[w][d];
            keys_out[off] = key;
            vals_out[off] = vals_in[i];
        }
        __syncthreads();
        unsigned add = 0;
        for (int w = 0; w < BVH_THREADS / 32; w++) { add += warp_hist[w][threadIdx.x]; warp_hist[w][threadIdx.x] = 0; }
        base[threadIdx.x] += add;
        __syncthreads();
    }
}

// ---------------------------------------------------------------- hierarchy (Karras 2012)
// common-prefix length of sorted keys i and j; ties broken by position so that every key is unique
__device__ __forceinline__ int delta(const unsigned* __restrict__ keys, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    unsigned a = keys[i], b = keys[j];
    if (a == b) return 32 + __clz(i ^ j);
    return __clz(a ^ b);
}

// child links: >= 0 inner node index, < 0 leaf ~index
__global__ void k_hierarchy(const unsigned* __restrict__ keys, int n, int2* children, int* parent_inner, int* parent_leaf) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    int j = i + l * d;
    int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    int gamma = i + s * d + min(d, 0);
    int lo = min(i, j), hi = max(i, j);
    int left = (lo == gamma) ? ~gamma : gamma;
    int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    if (left < 0) parent_leaf[~left] = i; else parent_inner[left] = i;
    if (right < 0) parent_leaf[~right] = i; else parent_inner[right] = i;
    if (i == 0) parent_inner[0] = -1;
}

// bottom-up boxes: the second thread to arrive at an inner node fits it and climbs on
__global__ void k_fit(const float4* __restrict__ tris, const int* __restrict__ vals, int n, const int2* __restrict__ children,
                      const int* __restrict__ parent_inner, const int* __restrict__ parent_leaf, float4* leaf_lo, float4* leaf_hi,
                      float4* inner_lo, float4* inner_hi, unsigned* flags) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    float lo[3], hi[3];
    tri_box(tris + 5 * (size_t)vals[l], lo, hi);
    leaf_lo[l] = make_float4(lo[0], lo[1], lo[2], 0.0f);
    leaf_hi[l] = make_float4(hi[0], hi[1], hi[2], 0.0f);
    int cur = parent_leaf[l];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&flags[cur], 1u) == 0u) return;          // first arrival: the sibling subtree is not ready
        int2 c = children[cur];
        // volatile reads: the sibling's boxes were written by another thread before its fence + atomic
        const volatile float4* alo = c.x < 0 ? leaf_lo + ~c.x : inner_lo + c.x;
        const volatile float4* ahi = c.x < 0 ? leaf_hi + ~c.x : inner_hi + c.x;
        const volatile float4* blo = c.y < 0 ? leaf_lo + ~c.y : inner_lo + c.y;
        const volatile float4* bhi = c.y < 0 ? leaf_hi + ~c.y : inner_hi + c.y;
        inner_lo[cur] = make_float4(fminf(alo->x, blo->x), fminf(alo->y, blo->y), fminf(alo->z, blo->z), 0.0f);
        inner_hi[cur] = make_float4(fmaxf(ahi->x, bhi->x), fmaxf(ahi->y, bhi->y), fmaxf(ahi->z, bhi->z), 0.0f);
        cur = parent_inner[cur];
    }
}


// ---------------------------------------------------------------- PLOC (pt_build_bvh_device_ploc)
// Parallel locally-ordered clustering (Meister & Bittner 2018) on the Morton-sorted leaves: every cluster looks for the
// neighbour within `radius` positions whose union with it has the smallest surface area, mutual nearest neighbours merge,
// the cluster array is compacted in order, repeat until one cluster is left. Bottom-up agglomeration with a surface-area
// criterion gives trees close to a SAH sweep (the Karras hierarchy above splits by Morton bits only), at the price of
// ~4 small kernels per round and one 4-byte read-back of the cluster count. Output goes through the same k_emit.
__global__ void k_ploc_leaves(const float4* __restrict__ tris, const int* __restrict__ vals, int n, float4* leaf_lo, float4* leaf_hi,
                              int* cid, float4* clo, float4* chi) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    float lo[3], hi[3];
    tri_box(tris + 5 * (size_t)vals[l], lo, hi);
    float4 a = make_float4(lo[0], lo[1], lo[2], 0.0f), b = make_float4(hi[0], hi[1], hi[2], 0.0f);
    leaf_lo[l] = a; leaf_hi[l] = b;
    cid[l] = ~l; clo[l] = a; chi[l] = b;
}

__device__ __forceinline__ float union_area(float4 alo, float4 ahi, float4 blo, float4 bhi) {
    float x = fmaxf(ahi.x, bhi.x) - fminf(alo.x, blo.x), y = fmaxf(ahi.y, bhi.y) - fminf(alo.y, blo.y), z = fmaxf(ahi.z, bhi.z) - fminf(alo.z, blo.z);
    return x * (y + z) + y * z;
}

// nearest neighbour by union area inside [i - radius, i + radius]; ties go to the lower index (the measure is symmetric, so
// the pair with the smallest measure that involves the lowest index is always mutual: every round merges at least once)
__global__ void k_ploc_nn(const float4* __restrict__ clo, const float4* __restrict__ chi, int m, int radius, int* nn) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    float4 lo = clo[i], hi = chi[i];
    float best = FLT_MAX; int bj = -1;
    int j0 = max(0, i - radius), j1 = min(m - 1, i + radius);
    for (int j = j0; j <= j1; j++) {
        if (j == i) continue;
        float a = union_area(lo, hi, clo[j], chi[j]);
        if (a < best || bj < 0) { best = a; bj = j; }
    }
    nn[i] = bj;
}

// mutual pairs merge: the lower position creates the inner node and keeps the merged cluster, the higher one is dropped.
// Inner node ids are handed out downwards from n - 2, so the last merge (the root) is inner node 0, as k_emit expects.
__global__ void __launch_bounds__(BVH_THREADS)
k_ploc_merge(int* cid, float4* clo, float4* chi, const int* __restrict__ nn, int m, int n, int* counter, int2* children,
             float4* inner_lo, float4* inner_hi, int* parent_inner, int* parent_leaf, int* valid, int* block_counts, int* subtree) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int v = 0;
    if (i < m) {
        v = 1;
        int j = nn[i];
        if (j >= 0 && nn[j] == i) {
            if (i < j) {
                int id = n - 2 - atomicAdd(counter, 1);
                int a = cid[i], b = cid[j];
                float4 alo = clo[i], ahi = chi[i], blo = clo[j], bhi = chi[j];
                float4 lo = make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.0f);
                float4 hi = make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f);
                children[id] = make_int2(a, b);
                inner_lo[id] = lo; inner_hi[id] = hi;
                subtree[id] = 1 + (a < 0 ? 0 : subtree[a]) + (b < 0 ? 0 : subtree[b]);      // inner nodes below and including id
                if (a < 0) parent_leaf[~a] = id; else parent_inner[a] = id;
                if (b < 0) parent_leaf[~b] = id; else parent_inner[b] = id;
                cid[i] = id; clo[i] = lo; chi[i] = hi;
            } else v = 0;
        }
        valid[i] = v;
    }
    int cnt = __syncthreads_count(v);
    if (threadIdx.x == 0) block_counts[blockIdx.x] = cnt;
}

// exclusive scan of the per-block counts (one block; nb <= 2^29 / 256), total -> *total
__global__ void __launch_bounds__(1024)
k_ploc_scan(int* counts, int nb, int* total) {
    __shared__ int part[1024];
    const int t = threadIdx.x;
    const int per = (nb + 1023) / 1024;
    const int b0 = min(nb, t * per), b1 = min(nb, b0 + per);
    int s = 0;
    for (int k = b0; k < b1; k++) s += counts[k];
    part[t] = s;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        int v = t >= off ? part[t - off] : 0;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    int run = part[t] - s;                                   // exclusive prefix of this thread's chunk
    for (int k = b0; k < b1; k++) { int c = counts[k]; counts[k] = run; run += c; }
    if (t == 1023) *total = part[1023];
}

__global__ void __launch_bounds__(BVH_THREADS)
k_ploc_compact(const int* __restrict__ cid, const float4* __restrict__ clo, const float4* __restrict__ chi, const int* __restrict__ valid,
               const int* __restrict__ block_off, int m, int* cid2, float4* clo2, float4* chi2) {
    __shared__ int wsum[BVH_THREADS / 32];
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    int v = i < m ? valid[i] : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    unsigned b = __ballot_sync(0xffffffffu, v);
    if (lane == 0) wsum[warp] = __popc(b);
    __syncthreads();
    if (!v) return;
    int off = block_off[blockIdx.x] + __popc(b & ((1u << lane) - 1u));
    for (int w = 0; w < warp; w++) off += wsum[w];
    cid2[off] = cid[i]; clo2[off] = clo[i]; chi2[off] = chi[i];
}

// PLOC hands out node ids in creation order (level by level, bottom-up), which scatters a subtree over the whole array.
// Relabel depth-first (node, first child's subtree, second child's subtree): a node's index is its depth plus the sizes of
// the first-child subtrees it passes on the right, found by one climb to the root.
__global__ void k_ploc_relabel(const int2* __restrict__ children, const int* __restrict__ parent_inner, const int* __restrict__ subtree,
                               int n_inner, int* remap) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_inner) return;
    int idx = 0;
    for (int cur = i, p = parent_inner[i]; p >= 0; cur = p, p = parent_inner[p]) {
        int2 c = children[p];
        idx += 1 + ((c.y == cur && c.x >= 0) ? subtree[c.x] : 0);
    }
    remap[i] = idx;
}

__device__ __forceinline__ float box_area(float4 lo, float4 hi) {
    float x = hi.x - lo.x, y = hi.y - lo.y, z = hi.z - lo.z;
    return 2.0f * (x * (y + z) + y * z);
}

// the reference's serialised form: inner node i keeps its children at (2i+1, 2i+2), larger surface area first
// (BVH.cpp:2240-2247); leaf l owns reference-list entry l = -(triangle index), the negated "last of leaf" entry (:2117)
__global__ void k_emit(const int* __restrict__ vals, int n, const int2* __restrict__ children, const float4* __restrict__ leaf_lo,
                       const float4* __restrict__ leaf_hi, const float4* __restrict__ inner_lo, const float4* __restrict__ inner_hi,
                       float4* nodes /* 2 float4 per node */, int* refs, const float4* __restrict__ tris, BvhPair* pairs, LeafTri* leaf,
                       const int* __restrict__ remap /* optional: inner node id -> position */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    refs[i] = -vals[i];
    if (leaf) {
        // what upload_leaf_stream does on the host: intersection record of the leaf's triangle. The stream marks the end of a
        // leaf with ~index, which triangle 0 can carry too -- unlike the reference list's -0 == 0 (BVH.cpp:2117), whose
        // run-on into the next leaf costs triangle tests and changes no result. (A tree read back and re-installed through
        // pt_set_bvh gets the copied-marker behaviour of the reference again.)
        const int tri = vals[i];
        const float4* t = tris + 5 * (size_t)tri;
        float4 t0 = t[0], t1 = t[1], t2 = t[2];
        leaf[i].r[0] = t0; leaf[i].r[1] = t1;
        leaf[i].r[2] = make_float4(t2.x, __int_as_float(~tri), 0.0f, 0.0f);
        leaf[i].r[3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    if (i >= n - 1) return;
    int2 c = children[i];
    float4 lo0 = c.x < 0 ? leaf_lo[~c.x] : inner_lo[c.x], hi0 = c.x < 0 ? leaf_hi[~c.x] : inner_hi[c.x];
    float4 lo1 = c.y < 0 ? leaf_lo[~c.y] : inner_lo[c.y], hi1 = c.y < 0 ? leaf_hi[~c.y] : inner_hi[c.y];
    const int me = remap ? remap[i] : i;
    int link0 = c.x < 0 ? -(~c.x) : 2 * (remap ? remap[c.x] : c.x) + 1, link1 = c.y < 0 ? -(~c.y) : 2 * (remap ? remap[c.y] : c.y) + 1;
    if (box_area(lo0, hi0) < box_area(lo1, hi1)) {
        float4 t = lo0; lo0 = lo1; lo1 = t; t = hi0; hi0 = hi1; hi1 = t;
        int k = link0; link0 = link1; link1 = k;
    }
    size_t at = 2 * (size_t)(2 * me + 1);
    nodes[at + 0] = make_float4(lo0.x, lo0.y, lo0.z, __int_as_float(link0));
    nodes[at + 1] = make_float4(hi0.x, hi0.y, hi0.z, 0.0f);
    nodes[at + 2] = make_float4(lo1.x, lo1.y, lo1.z, __int_as_float(link1));
    nodes[at + 3] = make_float4(hi1.x, hi1.y, hi1.z, 0.0f);
    if (pairs) {
        // install_bvh's re-layout: one 64-B record per sibling pair; link = pair index of an inner child, ~leaf-stream offset of a leaf
        int p0 = link0 > 0 ? (link0 - 1) / 2 : ~(-link0), p1 = link1 > 0 ? (link1 - 1) / 2 : ~(-link1);
        pairs[me].q[0] = make_float4(lo0.x, lo0.y, lo0.z, __int_as_float(p0));
        pairs[me].q[1] = make_float4(hi0.x, hi0.y, hi0.z, 0.0f);
        pairs[me].q[2] = make_float4(lo1.x, lo1.y, lo1.z, __int_as_float(p1));
        pairs[me].q[3] = make_float4(hi1.x, hi1.y, hi1.z, 0.0f);
    }
    if (i == 0) {
        nodes[0] = make_float4(inner_lo[0].x, inner_lo[0].y, inner_lo[0].z, __int_as_float(1));
        nodes[1] = make_float4(inner_hi[0].x, inner_hi[0].y, inner_hi[0].z, 0.0f);
    }
}

// depth of the tree = longest leaf-to-root climb
__global__ void k_depth(const int* __restrict__ parent_inner, const int* __restrict__ parent_leaf, int n, int* max_depth) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    int d = 0;
    for (int cur = parent_leaf[l]; cur >= 0; cur = parent_inner[cur]) d++;
    d = max(d, __shfl_xor_sync(0xffffffffu, d, 16)); d = max(d, __shfl_xor_sync(0xffffffffu, d, 8));
    d = max(d, __shfl_xor_sync(0xffffffffu, d, 4)); d = max(d, __shfl_xor_sync(0xffffffffu, d, 2));
    d = max(d, __shfl_xor_sync(0xffffffffu, d, 1));
    if ((threadIdx.x & 31) == 0) atomicMax(max_depth, d);
}

// ---------------------------------------------------------------- driver
#define BCK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { err = e__; goto done; } } while (0)

// d_tris: n x 5 float4 (device). d_nodes: 2 * (2n - 1) float4, d_refs: n ints (device, caller-owned). d_pairs (n - 1) /
// d_leaf (n): optional, the traversal layout of pt_layout.h written directly. *depth: depth of the tree.
// Returns the number of kernels launched (> 0) or -(cudaError) on failure; *kernel_ms = device time of the build.
// ploc_radius > 0: the hierarchy is built by PLOC with that search radius instead of Karras' Morton-bit splits.
int build_lbvh_device(const float4* d_tris, int n, float4* d_nodes, int* d_refs, BvhPair* d_pairs, LeafTri* d_leaf, int* depth,
                      cudaStream_t st, float* kernel_ms, int ploc_radius) {
    cudaError_t err = cudaSuccess;
    int launches = 0;
    const int n_blocks = (n + SORT_TILE - 1) / SORT_TILE;
    const int grid_n = (n + BVH_THREADS - 1) / BVH_THREADS;
    unsigned *bounds = nullptr, *keys[2] = {nullptr, nullptr}, *hist = nullptr, *flags = nullptr;
    int *vals[2] = {nullptr, nullptr}, *parent_inner = nullptr, *parent_leaf = nullptr, *d_depth = nullptr;
    int2* children = nullptr;
    float4 *leaf_lo = nullptr, *leaf_hi = nullptr, *inner_lo = nullptr, *inner_hi = nullptr;
    int *cid[2] = {nullptr, nullptr}, *nn = nullptr, *valid = nullptr, *block_counts = nullptr, *d_m = nullptr, *subtree = nullptr, *remap = nullptr;
    float4 *clo[2] = {nullptr, nullptr}, *chi[2] = {nullptr, nullptr};
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    const unsigned init_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    BCK(cudaEventCreate(&e0)); BCK(cudaEventCreate(&e1));
    BCK(cudaMalloc((void**)&bounds, 6 * sizeof(unsigned)));
    for (int k = 0; k < 2; k++) { BCK(cudaMalloc((void**)&keys[k], (size_t)n * 4)); BCK(cudaMalloc((void**)&vals[k], (size_t)n * 4)); }
    BCK(cudaMalloc((void**)&hist, (size_t)256 * n_blocks * 4));
    BCK(cudaMalloc((void**)&flags, (size_t)n * 4));
    BCK(cudaMalloc((void**)&parent_inner, (size_t)n * 4)); BCK(cudaMalloc((void**)&parent_leaf, (size_t)n * 4));
    BCK(cudaMalloc((void**)&children, (size_t)n * sizeof(int2)));
    BCK(cudaMalloc((void**)&leaf_lo, (size_t)n * 16)); BCK(cudaMalloc((void**)&leaf_hi, (size_t)n * 16));
    BCK(cudaMalloc((void**)&inner_lo, (size_t)n * 16)); BCK(cudaMalloc((void**)&inner_hi, (size_t)n * 16));
    BCK(cudaMemcpyAsync(bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
    BCK(cudaMemsetAsync(flags, 0, (size_t)n * 4, st));
    BCK(cudaMalloc((void**)&d_depth, 4));
    BCK(cudaMemsetAsync(d_depth, 0, 4, st));
    if (ploc_radius > 0) {
        for (int k = 0; k < 2; k++) {
            BCK(cudaMalloc((void**)&cid[k], (size_t)n * 4)); BCK(cudaMalloc((void**)&clo[k], (size_t)n * 16)); BCK(cudaMalloc((void**)&chi[k], (size_t)n * 16));
        }
        BCK(cudaMalloc((void**)&nn, (size_t)n * 4)); BCK(cudaMalloc((void**)&valid, (size_t)n * 4));
        BCK(cudaMalloc((void**)&block_counts, (size_t)grid_n * 4));
        BCK(cudaMalloc((void**)&subtree, (size_t)n * 4)); BCK(cudaMalloc((void**)&remap, (size_t)n * 4));
        BCK(cudaMalloc((void**)&d_m, 8));
        BCK(cudaMemsetAsync(d_m, 0, 8, st));                     // [0] = cluster count of the round, [1] = inner nodes created
    }
    BCK(cudaEventRecord(e0, st));
    k_scene_bounds<<<min(grid_n, 148 * 8), BVH_THREADS, 0, st>>>(d_tris, n, bounds); launches++;
    k_morton<<<grid_n, BVH_THREADS, 0, st>>>(d_tris, n, bounds, keys[0], vals[0]); launches++;
    for (int pass = 0; pass < 4; pass++) {
        const int src = pass & 1, dst = src ^ 1, shift = 8 * pass;
        k_sort_hist<<<n_blocks, BVH_THREADS, 0, st>>>(keys[src], n, shift, hist, n_blocks); launches++;
        k_sort_scan<<<1, 1024, 0, st>>>(hist, 256 * n_blocks); launches++;
        k_sort_scatter<<<n_blocks, BVH_THREADS, 0, st>>>(keys[src], vals[src], keys[dst], vals[dst], n, shift, hist, n_blocks); launches++;
    }
    // 4 passes: the sorted arrays are back in buffer 0
    if (ploc_radius > 0) {
        k_ploc_leaves<<<grid_n, BVH_THREADS, 0, st>>>(d_tris, vals[0], n, leaf_lo, leaf_hi, cid[0], clo[0], chi[0]); launches++;
        int m = n, cur = 0;
        while (m > 1) {
            const int nb = (m + BVH_THREADS - 1) / BVH_THREADS;
            k_ploc_nn<<<nb, BVH_THREADS, 0, st>>>(clo[cur], chi[cur], m, ploc_radius, nn); launches++;
            k_ploc_merge<<<nb, BVH_THREADS, 0, st>>>(cid[cur], clo[cur], chi[cur], nn, m, n, d_m + 1, children, inner_lo, inner_hi,
                                                     parent_inner, parent_leaf, valid, block_counts, subtree); launches++;
            k_ploc_scan<<<1, 1024, 0, st>>>(block_counts, nb, d_m); launches++;
            k_ploc_compact<<<nb, BVH_THREADS, 0, st>>>(cid[cur], clo[cur], chi[cur], valid, block_counts, m, cid[cur ^ 1], clo[cur ^ 1], chi[cur ^ 1]); launches++;
            int m_new = 0;
            BCK(cudaMemcpyAsync(&m_new, d_m, 4, cudaMemcpyDeviceToHost, st));
            BCK(cudaStreamSynchronize(st));
            if (m_new < 1 || m_new >= m) { err = cudaErrorUnknown; goto done; }     // cannot happen: every round merges at least one pair
            m = m_new; cur ^= 1;
        }
        BCK(cudaMemsetAsync(parent_inner, 0xff, 4, st));          // the root (inner node 0) has no parent
        k_ploc_relabel<<<grid_n, BVH_THREADS, 0, st>>>(children, parent_inner, subtree, n - 1, remap); launches++;
    } else {
        k_hierarchy<<<grid_n, BVH_THREADS, 0, st>>>(keys[0], n, children, parent_inner, parent_leaf); launches++;
        k_fit<<<grid_n, BVH_THREADS, 0, st>>>(d_tris, vals[0], n, children, parent_inner, parent_leaf, leaf_lo, leaf_hi, inner_lo, inner_hi, flags); launches++;
    }
    k_emit<<<grid_n, BVH_THREADS, 0, st>>>(vals[0], n, children, leaf_lo, leaf_hi, inner_lo, inner_hi, d_nodes, d_refs, d_tris, d_pairs, d_leaf, ploc_radius > 0 ? remap : nullptr); launches++;
    k_depth<<<grid_n, BVH_THREADS, 0, st>>>(parent_inner, parent_leaf, n, d_depth); launches++;
    BCK(cudaEventRecord(e1, st));
    if (depth) BCK(cudaMemcpyAsync(depth, d_depth, 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaGetLastError());
    BCK(cudaStreamSynchronize(st));
    if (kernel_ms) BCK(cudaEventElapsedTime(kernel_ms, e0, e1));
done:
    cudaFree(bounds); cudaFree(keys[0]); cudaFree(keys[1]); cudaFree(vals[0]); cudaFree(vals[1]); cudaFree(hist); cudaFree(flags);
    cudaFree(parent_inner); cudaFree(parent_leaf); cudaFree(children); cudaFree(d_depth);
    cudaFree(leaf_lo); cudaFree(leaf_hi); cudaFree(inner_lo); cudaFree(inner_hi);
    for (int k = 0; k < 2; k++) { cudaFree(cid[k]); cudaFree(clo[k]); cudaFree(chi[k]); }
    cudaFree(nn); cudaFree(valid); cudaFree(block_counts); cudaFree(d_m); cudaFree(subtree); cudaFree(remap);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return err == cudaSuccess ? launches : -(int)err;
}

}  // namespace pt
