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

// What the n leaves of a build are: the triangles themselves (ref_* null: leaf i = triangle i, its box = the triangle's), or a
// set of REFERENCES from the pre-split pass below (leaf i = triangle ref_tri[i] restricted to the box [ref_lo[i], ref_hi[i]]).
struct LeafSrc { const float4* tris; const int* ref_tri; const float4* ref_lo; const float4* ref_hi; };
__device__ __forceinline__ void leaf_box(const LeafSrc& s, int id, float lo[3], float hi[3]) {
    if (s.ref_lo) {
        float4 a = s.ref_lo[id], b = s.ref_hi[id];
        lo[0] = a.x; lo[1] = a.y; lo[2] = a.z; hi[0] = b.x; hi[1] = b.y; hi[2] = b.z;
    } else tri_box(s.tris + 5 * (size_t)id, lo, hi);
}
__device__ __forceinline__ int leaf_tri(const LeafSrc& s, int id) { return s.ref_tri ? s.ref_tri[id] : id; }

__global__ void k_scene_bounds(LeafSrc src, int n, unsigned* bounds /* 3 min, 3 max (ordered uints) */) {
    __shared__ unsigned s[6];
    if (threadIdx.x < 3) s[threadIdx.x] = 0xffffffffu; else if (threadIdx.x < 6) s[threadIdx.x] = 0u;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float lo[3], hi[3];
        leaf_box(src, i, lo, hi);
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

__global__ void k_morton(LeafSrc src, int n, const unsigned* __restrict__ bounds, unsigned* keys, int* vals) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float lo[3], hi[3];
    leaf_box(src, i, lo, hi);
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
__global__ void k_fit(LeafSrc src, const int* __restrict__ vals, int n, const int2* __restrict__ children,
                      const int* __restrict__ parent_inner, const int* __restrict__ parent_leaf, float4* leaf_lo, float4* leaf_hi,
                      float4* inner_lo, float4* inner_hi, unsigned* flags) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    float lo[3], hi[3];
    leaf_box(src, vals[l], lo, hi);
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
__global__ void k_ploc_leaves(LeafSrc src, const int* __restrict__ vals, int n, float4* leaf_lo, float4* leaf_hi,
                              int* cid, float4* clo, float4* chi) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n) return;
    float lo[3], hi[3];
    leaf_box(src, vals[l], lo, hi);
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
                       const int* __restrict__ remap /* optional: inner node id -> position */, const int* __restrict__ ref_tri,
                       const unsigned char* __restrict__ leaf_end = nullptr /* optional: 1 where a position ends its leaf (null: everywhere) */,
                       const int* __restrict__ live = nullptr /* optional: parent of every inner id, < -1 for ids that were never used */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int tri_of_leaf = ref_tri ? ref_tri[vals[i]] : vals[i];
    const bool ends = leaf_end ? leaf_end[i] != 0 : true;
    refs[i] = ends ? -tri_of_leaf : tri_of_leaf;
    if (leaf) {
        // what upload_leaf_stream does on the host: intersection record of the leaf's triangle. The stream marks the end of a
        // leaf with ~index, which triangle 0 can carry too -- unlike the reference list's -0 == 0 (BVH.cpp:2117), whose
        // run-on into the next leaf costs triangle tests and changes no result. (A tree read back and re-installed through
        // pt_set_bvh gets the copied-marker behaviour of the reference again.)
        const int tri = tri_of_leaf;
        const float4* t = tris + 5 * (size_t)tri;
        float4 t0 = t[0], t1 = t[1], t2 = t[2];
        leaf[i].r[0] = t0; leaf[i].r[1] = t1;
        leaf[i].r[2] = make_float4(t2.x, __int_as_float(ends ? ~tri : tri), 0.0f, 0.0f);
        leaf[i].r[3] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    }
    if (i >= n - 1) return;
    if (live && live[i] < -1) return;
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

// ---------------------------------------------------------------- top-down binned SAH on the device (pt_build_bvh_device_sah)
// The reference's builder is a top-down SAH search (FindBestObjectSplit, BVH.cpp:1619-1763: 8 bins per axis, sweep below 8
// references); its spatial splits are worth 2 % of SAH cost on the atrium and nothing on the bunny scene, the
// top-down search and the multi-triangle leaves are worth the 8-10 % that separate it from Morton / PLOC trees. Here the same search runs on the device,
// breadth first: every level bins the primitives of all nodes larger than SAH_SMALL at once (SAH_BINS bins per axis over the
// node's CENTROID box, counts + boxes accumulated with atomics: block-private in shared memory for the node a block starts in, global for
// the rest), one thread per node picks the cheapest of the 3 x (SAH_BINS - 1) planes and creates the children, and the primitives
// are partitioned into the other index buffer; subtrees of <= SAH_SMALL primitives are finished by ONE thread each with an exact
// sweep, which is also where leaves of SEVERAL triangles are made by the reference's SAH leaf rule (k_sah_small; the LBVH / PLOC builders
// keep one primitive per leaf). Inner ids are handed out by an atomic counter (root = 0; ids of nodes that turned into leaves stay unused) and made
// depth-first by k_ploc_relabel, so the emitted tree does not depend on the order atomics happened to land in.
#ifndef SAH_BINS
#define SAH_BINS 16
#endif
#ifndef SAH_SMALL
#define SAH_SMALL 16              // <= 32: one lane per primitive in k_sah_small
#endif
struct SahNode { int begin, end; };                            // position range of an inner node in the index buffer
struct SahSplit { int axis, plane, nl; int left, right; };     // axis < 0: median by position; children: inner id, or ~leaf position

__device__ __forceinline__ int sah_bin(float c, float lo, float hi) {
    const float e = hi - lo;
    if (!(e > 0.0f)) return 0;
    int b = (int)((c - lo) / e * (float)SAH_BINS);
    return b < 0 ? 0 : (b >= SAH_BINS ? SAH_BINS - 1 : b);
}
__device__ __forceinline__ float half_area(const float lo[3], const float hi[3]) {
    const float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    return x * (y + z) + y * z;
}

__global__ void k_sah_init(LeafSrc src, int n, float4* prim_lo, float4* prim_hi, int* idx, int* node_of, unsigned* root_bounds, unsigned* cb /* root's centroid box */) {
    __shared__ unsigned s[12];
    if (threadIdx.x < 12) s[threadIdx.x] = (threadIdx.x % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float lo[3], hi[3];
        leaf_box(src, i, lo, hi);
        prim_lo[i] = make_float4(lo[0], lo[1], lo[2], 0.0f); prim_hi[i] = make_float4(hi[0], hi[1], hi[2], 0.0f);
        idx[i] = i; node_of[i] = 0;
#pragma unroll
        for (int k = 0; k < 3; k++) {
            atomicMin(&s[k], f2ord(lo[k])); atomicMax(&s[3 + k], f2ord(hi[k]));
            const unsigned c = f2ord(0.5f * (lo[k] + hi[k]));
            atomicMin(&s[6 + k], c); atomicMax(&s[9 + k], c);
        }
    }
    __syncthreads();
    if (threadIdx.x < 3) atomicMin(&root_bounds[threadIdx.x], s[threadIdx.x]);
    else if (threadIdx.x < 6) atomicMax(&root_bounds[threadIdx.x], s[threadIdx.x]);
    else if (threadIdx.x < 9) atomicMin(&cb[threadIdx.x - 6], s[threadIdx.x]);
    else if (threadIdx.x < 12) atomicMax(&cb[threadIdx.x - 6], s[threadIdx.x]);
}
__global__ void k_sah_root(const unsigned* __restrict__ root_bounds, int n, SahNode* nodes, float4* inner_lo, float4* inner_hi, int* parent_inner,
                           int* subtree, int* slot_of, int* lists /* [0] large count, [1] small count */, int* large, int* small) {
    nodes[0].begin = 0; nodes[0].end = n;
    inner_lo[0] = make_float4(ord2f(root_bounds[0]), ord2f(root_bounds[1]), ord2f(root_bounds[2]), 0.0f);
    inner_hi[0] = make_float4(ord2f(root_bounds[3]), ord2f(root_bounds[4]), ord2f(root_bounds[5]), 0.0f);
    parent_inner[0] = -1; subtree[0] = n - 1;
    if (n > SAH_SMALL) { large[0] = 0; slot_of[0] = 0; lists[0] = 1; lists[1] = 0; }
    else { small[0] = 0; slot_of[0] = -1; lists[0] = 0; lists[1] = 1; }
}

// bins of every large node of this level: cnt[slot][axis][bin], box[slot][axis][bin][6] (ordered uints)
__global__ void __launch_bounds__(BVH_THREADS)
k_sah_bin(const int* __restrict__ idx, const int* __restrict__ node_of, int n, const int* __restrict__ slot_of, const float4* __restrict__ prim_lo,
          const float4* __restrict__ prim_hi, const unsigned* __restrict__ cb, unsigned* cnt, unsigned* box) {
    __shared__ unsigned s_cnt[3 * SAH_BINS];
    __shared__ unsigned s_box[3 * SAH_BINS * 6];
    __shared__ int s_node;
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (threadIdx.x == 0) { int p0 = blockIdx.x * blockDim.x; s_node = p0 < n ? node_of[p0] : -1; }
    for (int k = threadIdx.x; k < 3 * SAH_BINS; k += blockDim.x) s_cnt[k] = 0u;
    for (int k = threadIdx.x; k < 3 * SAH_BINS * 6; k += blockDim.x) s_box[k] = (k % 6) < 3 ? 0xffffffffu : 0u;
    __syncthreads();
    const int first = s_node;
    const int first_slot = first >= 0 ? slot_of[first] : -1;
    if (pos < n) {
        const int nd = node_of[pos];
        const int slot = nd >= 0 ? slot_of[nd] : -1;
        if (slot >= 0) {
            const int prim = idx[pos];
            const float4 plo = prim_lo[prim], phi = prim_hi[prim];
            const float l[3] = {plo.x, plo.y, plo.z}, h[3] = {phi.x, phi.y, phi.z};
            const unsigned* nb = cb + (size_t)nd * 6;
            const float nl[3] = {ord2f(nb[0]), ord2f(nb[1]), ord2f(nb[2])}, nh[3] = {ord2f(nb[3]), ord2f(nb[4]), ord2f(nb[5])};
            const bool priv = nd == first;
#pragma unroll
            for (int ax = 0; ax < 3; ax++) {
                const int b = sah_bin(0.5f * (l[ax] + h[ax]), nl[ax], nh[ax]);
                unsigned* c = priv ? &s_cnt[ax * SAH_BINS + b] : &cnt[((size_t)slot * 3 + ax) * SAH_BINS + b];
                unsigned* x = priv ? &s_box[(ax * SAH_BINS + b) * 6] : &box[(((size_t)slot * 3 + ax) * SAH_BINS + b) * 6];
                atomicAdd(c, 1u);
#pragma unroll
                for (int k = 0; k < 3; k++) { atomicMin(&x[k], f2ord(l[k])); atomicMax(&x[3 + k], f2ord(h[k])); }
            }
        }
    }
    __syncthreads();
    if (first_slot >= 0) {
        for (int k = threadIdx.x; k < 3 * SAH_BINS; k += blockDim.x) {
            const unsigned c = s_cnt[k];
            if (!c) continue;
            atomicAdd(&cnt[(size_t)first_slot * 3 * SAH_BINS + k], c);
            unsigned* x = &box[((size_t)first_slot * 3 * SAH_BINS + k) * 6];
#pragma unroll
            for (int q = 0; q < 3; q++) { atomicMin(&x[q], s_box[k * 6 + q]); atomicMax(&x[3 + q], s_box[k * 6 + 3 + q]); }
        }
    }
}

__global__ void k_sah_box_init(unsigned* box, size_t count) {
    size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
    if (i < count) box[i] = (i % 6) < 3 ? 0xffffffffu : 0u;
}

// children of node `nd` after a split that puts nl primitives on the left: ids, ranges, boxes, parents; appended to the next lists
__device__ void sah_make_children(int nd, int begin, int end, int nl, const float llo[3], const float lhi[3], const float rlo[3], const float rhi[3],
                                  SahNode* nodes, float4* inner_lo, float4* inner_hi, int2* children, int* parent_inner, int* parent_leaf,
                                  int* subtree, int* slot_of, int* next_lists, int* next_large, int* next_small, int* alloc, unsigned* cb,
                                  float4* leaf_lo, float4* leaf_hi, unsigned char* leaf_end, int& left, int& right) {
    const int cnts[2] = {nl, end - begin - nl}, begins[2] = {begin, begin + nl};
    int link[2];
    for (int k = 0; k < 2; k++) {
        if (cnts[k] == 1) {
            link[k] = ~begins[k]; parent_leaf[begins[k]] = nd; leaf_end[begins[k]] = 1;
            const float* lo = k ? rlo : llo; const float* hi = k ? rhi : lhi;
            leaf_lo[begins[k]] = make_float4(lo[0], lo[1], lo[2], 0.0f); leaf_hi[begins[k]] = make_float4(hi[0], hi[1], hi[2], 0.0f);
            continue;
        }
        const int id = atomicAdd(alloc, 1);
        link[k] = id;
        nodes[id].begin = begins[k]; nodes[id].end = begins[k] + cnts[k];
        const float* lo = k ? rlo : llo; const float* hi = k ? rhi : lhi;
        inner_lo[id] = make_float4(lo[0], lo[1], lo[2], 0.0f); inner_hi[id] = make_float4(hi[0], hi[1], hi[2], 0.0f);
        parent_inner[id] = nd; subtree[id] = cnts[k] - 1;
        for (int q = 0; q < 6; q++) cb[(size_t)id * 6 + q] = q < 3 ? 0xffffffffu : 0u;      // filled by k_sah_partition
        if (cnts[k] > SAH_SMALL) { const int s = atomicAdd(&next_lists[0], 1); next_large[s] = id; slot_of[id] = s; }
        else { const int s = atomicAdd(&next_lists[1], 1); next_small[s] = id; slot_of[id] = -1; }
    }
    left = link[0]; right = link[1];
    // the child with the larger surface goes first, as in the reference's serialised form (BVH.cpp:2240-2247): k_emit orders the pair
    // that way whatever stands here, but the depth-first relabel follows THIS order, and the host builder's layout has the larger
    // child's subtree right behind its parent
    if (half_area(rlo, rhi) > half_area(llo, lhi)) { const int t = link[0]; link[0] = link[1]; link[1] = t; }
    children[nd] = make_int2(link[0], link[1]);
}

// one WARP per large node: the 3 x (SAH_BINS - 1) planes are spread over the lanes, each summing its two sides from the node's bins
// (staged in shared memory); the cheapest wins, the lowest (axis, plane) among equals
#define SAH_SPLIT_WARPS 4
#define SAH_PLANES (3 * (SAH_BINS - 1))
__device__ __forceinline__ void sah_side(const unsigned* c, const unsigned* x, int b0, int b1, float lo[3], float hi[3], int& count) {
    lo[0] = lo[1] = lo[2] = 3.4e38f; hi[0] = hi[1] = hi[2] = -3.4e38f; count = 0;
    for (int b = b0; b < b1; b++) {
        if (!c[b]) continue;
        count += (int)c[b];
        for (int q = 0; q < 3; q++) { lo[q] = fminf(lo[q], ord2f(x[b * 6 + q])); hi[q] = fmaxf(hi[q], ord2f(x[b * 6 + 3 + q])); }
    }
}
__global__ void __launch_bounds__(32 * SAH_SPLIT_WARPS)
k_sah_split(const int* __restrict__ large, int n_large, const unsigned* __restrict__ cnt, const unsigned* __restrict__ box,
            SahNode* nodes, float4* inner_lo, float4* inner_hi, int2* children, int* parent_inner, int* parent_leaf, int* subtree,
            int* slot_of, int* next_lists, int* next_large, int* next_small, int* alloc, SahSplit* splits, unsigned* part_cnt /* 2 per slot */, unsigned* cb,
            float4* leaf_lo, float4* leaf_hi, unsigned char* leaf_end) {
    __shared__ unsigned s_cnt[SAH_SPLIT_WARPS][3 * SAH_BINS], s_box[SAH_SPLIT_WARPS][3 * SAH_BINS * 6];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int slot = blockIdx.x * SAH_SPLIT_WARPS + w;
    if (slot >= n_large) return;                                  // whole warps leave
    const unsigned full = 0xffffffffu;
    for (int k = lane; k < 3 * SAH_BINS; k += 32) s_cnt[w][k] = cnt[(size_t)slot * 3 * SAH_BINS + k];
    for (int k = lane; k < 3 * SAH_BINS * 6; k += 32) s_box[w][k] = box[(size_t)slot * 3 * SAH_BINS * 6 + k];
    __syncwarp();
    float best = 3.4e38f; int bc = 0x7fffffff;
    for (int cand = lane; cand < SAH_PLANES; cand += 32) {
        const int ax = cand / (SAH_BINS - 1), k = cand % (SAH_BINS - 1) + 1;     // plane k: bins < k on the left
        float llo[3], lhi[3], rlo[3], rhi[3]; int lc, rc;
        sah_side(s_cnt[w] + ax * SAH_BINS, s_box[w] + ax * SAH_BINS * 6, 0, k, llo, lhi, lc);
        sah_side(s_cnt[w] + ax * SAH_BINS, s_box[w] + ax * SAH_BINS * 6, k, SAH_BINS, rlo, rhi, rc);
        if (lc == 0 || rc == 0) continue;
        const float cost = half_area(llo, lhi) * (float)lc + half_area(rlo, rhi) * (float)rc;
        if (cost < best) { best = cost; bc = cand; }
    }
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        const float ob = __shfl_xor_sync(full, best, off); const int oc = __shfl_xor_sync(full, bc, off);
        if (ob < best || (ob == best && oc < bc)) { best = ob; bc = oc; }
    }
    if (lane != 0) return;
    const int nd = large[slot];
    const int begin = nodes[nd].begin, end = nodes[nd].end, count = end - begin;
    float bl[6], br[6];
    SahSplit sp;
    if (best < 3.4e38f) {
        sp.axis = bc / (SAH_BINS - 1); sp.plane = bc % (SAH_BINS - 1) + 1;
        int rc;
        sah_side(s_cnt[w] + sp.axis * SAH_BINS, s_box[w] + sp.axis * SAH_BINS * 6, 0, sp.plane, bl, bl + 3, sp.nl);
        sah_side(s_cnt[w] + sp.axis * SAH_BINS, s_box[w] + sp.axis * SAH_BINS * 6, sp.plane, SAH_BINS, br, br + 3, rc);
    } else {
        // every centroid in one bin on all three axes: halve by position; children keep the parent's box (conservative)
        sp.axis = -1; sp.plane = 0; sp.nl = count / 2;
        const float4 a = inner_lo[nd], b = inner_hi[nd];
        bl[0] = br[0] = a.x; bl[1] = br[1] = a.y; bl[2] = br[2] = a.z; bl[3] = br[3] = b.x; bl[4] = br[4] = b.y; bl[5] = br[5] = b.z;
    }
    sah_make_children(nd, begin, end, sp.nl, bl, bl + 3, br, br + 3, nodes, inner_lo, inner_hi, children, parent_inner, parent_leaf, subtree, slot_of,
                      next_lists, next_large, next_small, alloc, cb, leaf_lo, leaf_hi, leaf_end, sp.left, sp.right);
    splits[slot] = sp;
    part_cnt[2 * slot] = 0u; part_cnt[2 * slot + 1] = 0u;
}

// primitives of large nodes move to their side in the other buffer; everything else is copied through
__global__ void __launch_bounds__(BVH_THREADS)
k_sah_partition(const int* __restrict__ idx, const int* __restrict__ node_of, int n, const int* __restrict__ slot_of_level /* slots as of THIS level */,
                const SahNode* __restrict__ nodes_level /* ranges as of this level (begin/end of the node being split) */, const SahSplit* __restrict__ splits,
                const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi, unsigned* cb, unsigned* part_cnt, int* idx_out, int* node_out) {
    const int pos = blockIdx.x * blockDim.x + threadIdx.x;
    if (pos >= n) return;
    const int nd = node_of[pos];
    const int slot = nd >= 0 ? slot_of_level[nd] : -1;
    const int prim = idx[pos];
    if (slot < 0) { if (nd < 0) { idx_out[pos] = prim; node_out[pos] = nd; } return; }   // small nodes are rewritten by k_sah_small
    const SahSplit sp = splits[slot];
    const int begin = nodes_level[nd].begin;
    bool left;
    const float4 plo = prim_lo[prim], phi = prim_hi[prim];
    const float cen[3] = {0.5f * (plo.x + phi.x), 0.5f * (plo.y + phi.y), 0.5f * (plo.z + phi.z)};
    if (sp.axis < 0) left = pos - begin < sp.nl;
    else {
        const float c = sp.axis == 0 ? cen[0] : sp.axis == 1 ? cen[1] : cen[2];
        left = sah_bin(c, ord2f(cb[(size_t)nd * 6 + sp.axis]), ord2f(cb[(size_t)nd * 6 + 3 + sp.axis])) < sp.plane;
    }
    // one atomic per (node, side) group of the warp
    const unsigned key = ((unsigned)slot << 1) | (left ? 0u : 1u);
    const unsigned peers = __match_any_sync(__activemask(), key);
    const int leader = __ffs(peers) - 1, lane = threadIdx.x & 31;
    unsigned base = 0;
    if (lane == leader) base = atomicAdd(&part_cnt[key], (unsigned)__popc(peers));
    base = __shfl_sync(peers, base, leader);
    const int at = begin + (left ? 0 : sp.nl) + (int)base + __popc(peers & ((1u << lane) - 1u));
    idx_out[at] = prim;
    const int child = left ? sp.left : sp.right;
    node_out[at] = child >= 0 ? child : -1;
    // the child's centroid box (what the next level bins over): reduced over the warp's group first
    const int child_count = left ? sp.nl : nodes_level[nd].end - begin - sp.nl;
    if (child_count > SAH_SMALL) {                                // small subtrees are swept exactly, no bins
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const unsigned o = f2ord(cen[q]);
            const unsigned mn = __reduce_min_sync(peers, o), mx = __reduce_max_sync(peers, o);
            if (lane == leader) { atomicMin(&cb[(size_t)child * 6 + q], mn); atomicMax(&cb[(size_t)child * 6 + 3 + q], mx); }
        }
    }
}

// one WARP finishes the whole subtree of a node with <= SAH_SMALL primitives: exact SAH over every centroid plane of every axis
// (3c candidates, up to two per lane, each an O(c) pass over the node's primitives in shared memory), stable partition by ballot.
// This is also where LEAVES of several triangles are made, by the reference's rule (AdjustSAH + the leaf test, BVH.cpp:2123-2131, as
// pt_bvh_build.cpp applies it): a node becomes a leaf unless  Ct + Ci * (A_l n_l + A_r n_r) / A  <  Ci * n  -- triangles that share a
// box (the two halves of a quad) stay together instead of hanging under an inner node of the very same box.
#define SAH_SMALL_WARPS 4
#define SAH_CT 1.23f
#define SAH_CI 5.33f
#define SAH_DEAD (-2)             // parent_inner of an inner id that was handed out but not used
__global__ void __launch_bounds__(32 * SAH_SMALL_WARPS)
k_sah_small(const int* __restrict__ small, int n_small, const int* __restrict__ idx, const SahNode* __restrict__ nodes_in,
            const float4* __restrict__ prim_lo, const float4* __restrict__ prim_hi, float4* inner_lo, float4* inner_hi, int2* children,
            int* parent_inner, int* parent_leaf, int* alloc, int* idx_out, int* node_out, float4* leaf_lo, float4* leaf_hi,
            unsigned char* leaf_end, int multi_leaf) {
    __shared__ float s_lo[SAH_SMALL_WARPS][SAH_SMALL][3], s_hi[SAH_SMALL_WARPS][SAH_SMALL][3];
    __shared__ int s_id[SAH_SMALL_WARPS][SAH_SMALL], s_stack[SAH_SMALL_WARPS][SAH_SMALL][4];
    const int w = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int t = blockIdx.x * SAH_SMALL_WARPS + w;
    if (t >= n_small) return;                                     // whole warps leave
    const unsigned full = 0xffffffffu, lt = (1u << lane) - 1u;
    const int root = small[t];
    const int rb = nodes_in[root].begin, cnt0 = nodes_in[root].end - rb;
    float (*lo)[3] = s_lo[w]; float (*hi)[3] = s_hi[w]; int* ids = s_id[w];
    if (lane < cnt0) {
        const int prim = idx[rb + lane];
        const float4 a = prim_lo[prim], b = prim_hi[prim];
        ids[lane] = prim; lo[lane][0] = a.x; lo[lane][1] = a.y; lo[lane][2] = a.z; hi[lane][0] = b.x; hi[lane][1] = b.y; hi[lane][2] = b.z;
    }
    int next_id = 0;                                              // the subtree has at most cnt0 - 1 inner nodes; `root` is one of them
    if (cnt0 > 2) { if (lane == 0) next_id = atomicAdd(alloc, cnt0 - 2); next_id = __shfl_sync(full, next_id, 0); }
    // stack of open nodes: (parent id, side of the parent's child pair, first, count); parent -1 = the subtree's root, whose id exists
    if (lane == 0) { s_stack[w][0][0] = -1; s_stack[w][0][1] = 0; s_stack[w][0][2] = 0; s_stack[w][0][3] = cnt0; }
    int sp = 1;
    __syncwarp();
    while (sp > 0) {
        sp--;
        const int par = s_stack[w][sp][0], side = s_stack[w][sp][1], f = s_stack[w][sp][2], c = s_stack[w][sp][3];
        __syncwarp();
        const bool mine = lane < c;
        const int i = f + (mine ? lane : 0);
        // the node's own box
        float nlo[3], nhi[3];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            nlo[q] = ord2f(__reduce_min_sync(full, mine ? f2ord(lo[i][q]) : 0xffffffffu));
            nhi[q] = ord2f(__reduce_max_sync(full, mine ? f2ord(hi[i][q]) : 0u));
        }
        float best = 3.4e38f; int bc = 0x7fffffff;
        for (int cand = lane; cand < 3 * c; cand += 32) {
            // plane at primitive j's centroid: left = centroids below it (ties by slot), j itself goes right
            const int ax = cand / c, j = f + cand % c;
            const float cj = 0.5f * (lo[j][ax] + hi[j][ax]);
            float llo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, lhi[3] = {-3.4e38f, -3.4e38f, -3.4e38f}, rlo[3] = {3.4e38f, 3.4e38f, 3.4e38f}, rhi[3] = {-3.4e38f, -3.4e38f, -3.4e38f};
            int nl = 0;
            for (int k = f; k < f + c; k++) {
                const float ck = 0.5f * (lo[k][ax] + hi[k][ax]);
                const bool left = ck < cj || (ck == cj && k < j);
                if (left) { nl++; for (int q = 0; q < 3; q++) { llo[q] = fminf(llo[q], lo[k][q]); lhi[q] = fmaxf(lhi[q], hi[k][q]); } }
                else for (int q = 0; q < 3; q++) { rlo[q] = fminf(rlo[q], lo[k][q]); rhi[q] = fmaxf(rhi[q], hi[k][q]); }
            }
            if (nl == 0) continue;
            const float cost = half_area(llo, lhi) * (float)nl + half_area(rlo, rhi) * (float)(c - nl);
            if (cost < best) { best = cost; bc = cand; }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {                  // cheapest plane, the lowest candidate among equals
            const float ob = __shfl_xor_sync(full, best, off); const int oc = __shfl_xor_sync(full, bc, off);
            if (ob < best || (ob == best && oc < bc)) { best = ob; bc = oc; }
        }
        // there always is a plane for c >= 2 (every centroid but an axis' smallest has somebody on its left) unless boxes hold NaN
        const bool have = best < 3.4e38f;
        const float node_area = half_area(nlo, nhi);
        const bool whole_tree_root = par < 0 && root == 0;       // inner node 0 must stay an inner node (k_emit: nodes[0])
        const bool as_leaf = !whole_tree_root && (!have || (multi_leaf && (!(node_area > 0.0f) || !(SAH_CT + SAH_CI * best / node_area < SAH_CI * (float)c))));
        if (as_leaf) {
            if (lane == 0) {
                const int first = rb + f;
                if (par < 0) {                                    // the subtree's root: its parent (a binned node) already links to it as an inner node
                    const int p = parent_inner[root];
                    const int2 ch = children[p];
                    reinterpret_cast<int*>(&children[p])[ch.y == root ? 1 : 0] = ~first;
                    parent_leaf[first] = p; parent_inner[root] = SAH_DEAD;
                } else { reinterpret_cast<int*>(&children[par])[side] = ~first; parent_leaf[first] = par; }
                leaf_lo[first] = make_float4(nlo[0], nlo[1], nlo[2], 0.0f); leaf_hi[first] = make_float4(nhi[0], nhi[1], nhi[2], 0.0f);
                leaf_end[first + c - 1] = 1;
                // the reference list marks the end of a leaf by the sign of its last entry, which triangle 0 cannot carry
                if (c > 1 && ids[f + c - 1] == 0) { ids[f + c - 1] = ids[f + c - 2]; ids[f + c - 2] = 0; }
            }
            __syncwarp();
            continue;
        }
        const int nd = par < 0 ? root : next_id++;
        const int ax = have ? bc / c : 0, bj = have ? f + bc % c : f;
        const float pivot = 0.5f * (lo[bj][ax] + hi[bj][ax]);
        const float ci = 0.5f * (lo[i][ax] + hi[i][ax]);
        const bool left = mine && (have ? (ci < pivot || (ci == pivot && i < bj)) : lane < c / 2);
        const unsigned bl = __ballot_sync(full, left), br = __ballot_sync(full, mine && !left);
        const int nl = __popc(bl);
        const int at = f + (left ? __popc(bl & lt) : nl + __popc(br & lt));
        const int my_id = ids[i]; float ml[3], mh[3];
        for (int q = 0; q < 3; q++) { ml[q] = lo[i][q]; mh[q] = hi[i][q]; }
        float cl[3], chh[3], cr[3], crh[3];
#pragma unroll
        for (int q = 0; q < 3; q++) {
            const unsigned ol = f2ord(ml[q]), oh = f2ord(mh[q]);
            cl[q] = ord2f(__reduce_min_sync(full, left ? ol : 0xffffffffu)); chh[q] = ord2f(__reduce_max_sync(full, left ? oh : 0u));
            cr[q] = ord2f(__reduce_min_sync(full, mine && !left ? ol : 0xffffffffu)); crh[q] = ord2f(__reduce_max_sync(full, mine && !left ? oh : 0u));
        }
        __syncwarp();
        if (mine) { ids[at] = my_id; for (int q = 0; q < 3; q++) { lo[at][q] = ml[q]; hi[at][q] = mh[q]; } }
        const int cnts[2] = {nl, c - nl}, firsts[2] = {f, f + nl};
        // the child with the larger surface takes side 0 (see sah_make_children)
        const int swap = half_area(cr, crh) > half_area(cl, chh) ? 1 : 0;
        if (lane == 0) {
            if (par >= 0) {
                reinterpret_cast<int*>(&children[par])[side] = nd; parent_inner[nd] = par;
                inner_lo[nd] = make_float4(nlo[0], nlo[1], nlo[2], 0.0f); inner_hi[nd] = make_float4(nhi[0], nhi[1], nhi[2], 0.0f);
            }
            int push = sp;
            for (int k = 0; k < 2; k++) {
                const int sd = k ^ swap;
                if (cnts[k] == 1) {
                    const int pos = rb + firsts[k];
                    reinterpret_cast<int*>(&children[nd])[sd] = ~pos; parent_leaf[pos] = nd; leaf_end[pos] = 1;
                    leaf_lo[pos] = k ? make_float4(cr[0], cr[1], cr[2], 0.0f) : make_float4(cl[0], cl[1], cl[2], 0.0f);
                    leaf_hi[pos] = k ? make_float4(crh[0], crh[1], crh[2], 0.0f) : make_float4(chh[0], chh[1], chh[2], 0.0f);
                    continue;
                }
                s_stack[w][push][0] = nd; s_stack[w][push][1] = sd; s_stack[w][push][2] = firsts[k]; s_stack[w][push][3] = cnts[k]; push++;
            }
        }
        sp += (cnts[0] > 1 ? 1 : 0) + (cnts[1] > 1 ? 1 : 0);
        __syncwarp();
    }
    if (lane < cnt0) { idx_out[rb + lane] = ids[lane]; node_out[rb + lane] = -1; }
}
// leaves below every inner node (subtree[] starts at -1: leaves - 1 = inner nodes of the subtree). Every leaf climbs to the root; lanes
// of a warp that meet at the same ancestor merge into one climber carrying their sum, so the top of the tree sees one atomic per warp
__global__ void k_sah_count(const unsigned char* __restrict__ leaf_end, int n, const int* __restrict__ parent_inner, const int* __restrict__ parent_leaf,
                            int* subtree, int* n_leaves) {
    const unsigned full = 0xffffffffu;
    const int l = blockIdx.x * blockDim.x + threadIdx.x, lane = threadIdx.x & 31;
    const bool first = l < n && (l == 0 || leaf_end[l - 1] != 0);
    const unsigned m = __ballot_sync(full, first);
    if (lane == 0 && m) atomicAdd(n_leaves, __popc(m));
    int cur = first ? parent_leaf[l] : -1, w = first ? 1 : 0;
    while (__any_sync(full, cur >= 0)) {
        const unsigned peers = __match_any_sync(full, cur);
        const int total = __reduce_add_sync(peers, w);
        if (cur >= 0) {
            if (lane == __ffs(peers) - 1) { atomicAdd(&subtree[cur], total); w = total; cur = parent_inner[cur]; }
            else { cur = -1; w = 0; }
        }
    }
}
__global__ void k_sah_leaves(const int* __restrict__ idx, int n, int* vals) {
    int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l < n) vals[l] = idx[l];
}

// ---------------------------------------------------------------- triangle pre-splitting (spatial splits for the device builders)
// FindBestSpatialSplit (BVH.cpp:1793-1984) lets the reference cut a triangle that straddles a good split plane; Morton /
// PLOC builders cluster whole triangles, so a wall quad that spans the room drags a room-sized box into every subtree it
// lands in. Here long triangles are cut BEFORE the sort (Karras & Aila 2013, simplified): a triangle whose box is longer than
// 1.5 cells of a grid^3 grid over the scene box along some axis is cut into ceil(extent / cell) slabs along that axis (at
// most PRESPLIT_MAX_PIECES boxes per triangle), the triangle is clipped to every piece and empty pieces are dropped. Pieces
// are work items of their own (one thread per piece, owner found by binary search in the scanned piece counts), so the two
// ground triangles of a scene that need 4096 pieces each cost what 8192 small triangles cost.
#define PRESPLIT_MAX_PIECES 4096
#define SCAN_TILE 2048           // elements per block of the 3-kernel scan (256 threads x 8)

__global__ void k_tri_scene_box(const float4* __restrict__ tris, int n, unsigned* bounds) {
    __shared__ unsigned s[6];
    if (threadIdx.x < 3) s[threadIdx.x] = 0xffffffffu; else if (threadIdx.x < 6) s[threadIdx.x] = 0u;
    __syncthreads();
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float lo[3], hi[3];
        tri_box(tris + 5 * (size_t)i, lo, hi);
#pragma unroll
        for (int k = 0; k < 3; k++) { atomicMin(&s[k], f2ord(lo[k])); atomicMax(&s[3 + k], f2ord(hi[k])); }
    }
    __syncthreads();
    if (threadIdx.x < 3) atomicMin(&bounds[threadIdx.x], s[threadIdx.x]);
    else if (threadIdx.x < 6) atomicMax(&bounds[threadIdx.x], s[threadIdx.x]);
}

// pieces per axis of triangle i, packed 12 bits each (<= 4096 in total)
__device__ __forceinline__ unsigned presplit_plan(const float lo[3], const float hi[3], float cell) {
    int nk[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {
        float e = hi[k] - lo[k];
        nk[k] = e > 1.5f * cell ? (int)fminf(ceilf(e / cell), 4095.0f) : 1;
    }
    while ((long long)nk[0] * nk[1] * nk[2] > PRESPLIT_MAX_PIECES) {          // coarsen the axis with the most pieces
        int k = nk[0] >= nk[1] ? (nk[0] >= nk[2] ? 0 : 2) : (nk[1] >= nk[2] ? 1 : 2);
        nk[k] = (nk[k] + 1) / 2;
    }
    return (unsigned)nk[0] | ((unsigned)nk[1] << 12) | ((unsigned)nk[2] << 24);
}
__global__ void k_presplit_count(const float4* __restrict__ tris, int n, const unsigned* __restrict__ bounds, int grid, unsigned* pieces /* n + 1 */) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i > n) return;
    if (i == n) { pieces[n] = 0u; return; }
    float ext = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; k++) ext = fmaxf(ext, ord2f(bounds[3 + k]) - ord2f(bounds[k]));
    float lo[3], hi[3];
    tri_box(tris + 5 * (size_t)i, lo, hi);
    unsigned plan = presplit_plan(lo, hi, ext / (float)grid);
    pieces[i] = (plan & 4095u) * ((plan >> 12) & 4095u) * (plan >> 24);
}

// ---- exclusive scan of `count` unsigned values in place: tile sums, scan of the tile sums (one block), tile-local scan + base
__global__ void __launch_bounds__(BVH_THREADS) k_scan_tile_sums(const unsigned* __restrict__ data, int count, int* sums) {
    __shared__ unsigned w[BVH_THREADS / 32];
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * 8;
    unsigned t = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) if (base + k < count) t += data[base + k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) w[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) { unsigned z = 0; for (int k = 0; k < BVH_THREADS / 32; k++) z += w[k]; sums[blockIdx.x] = (int)z; }
}
__global__ void __launch_bounds__(BVH_THREADS) k_scan_apply(unsigned* data, int count, const int* __restrict__ tile_base) {
    __shared__ unsigned w[BVH_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int base = blockIdx.x * SCAN_TILE + threadIdx.x * 8;
    unsigned v[8], t = 0;
#pragma unroll
    for (int k = 0; k < 8; k++) { v[k] = base + k < count ? data[base + k] : 0u; t += v[k]; }
    unsigned x = t;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { unsigned y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
    if (lane == 31) w[warp] = x;
    __syncthreads();
    unsigned run = (unsigned)tile_base[blockIdx.x] + (x - t);
    for (int k = 0; k < warp; k++) run += w[k];
#pragma unroll
    for (int k = 0; k < 8; k++) { if (base + k < count) data[base + k] = run; run += v[k]; }
}

// clip polygon `in` (n vertices) against the half space x[axis] >= value (keep_greater) or <= value
__device__ int clip_halfspace(const float (*in)[3], int n, int axis, float value, bool keep_greater, float (*out)[3]) {
    int m = 0;
    for (int i = 0; i < n; i++) {
        const float* a = in[i]; const float* b = in[(i + 1) % n];
        const bool ia = keep_greater ? a[axis] >= value : a[axis] <= value;
        const bool ib = keep_greater ? b[axis] >= value : b[axis] <= value;
        if (ia) { out[m][0] = a[0]; out[m][1] = a[1]; out[m][2] = a[2]; m++; }
        if (ia != ib) {
            const float t = (value - a[axis]) / (b[axis] - a[axis]);
            out[m][0] = a[0] + t * (b[0] - a[0]); out[m][1] = a[1] + t * (b[1] - a[1]); out[m][2] = a[2] + t * (b[2] - a[2]);
            out[m][axis] = value;
            m++;
        }
    }
    return m;
}

// one thread per piece: owner triangle by binary search in the scanned piece counts, the piece's box, the clipped triangle's box
__global__ void k_presplit_pieces(const float4* __restrict__ tris, int n, const unsigned* __restrict__ bounds, int grid,
                                  const unsigned* __restrict__ offsets /* n + 1, scanned */, unsigned total, int* p_tri, float4* p_lo,
                                  float4* p_hi, unsigned* p_valid /* total + 1 */) {
    unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p > total) return;
    if (p == total) { p_valid[total] = 0u; return; }
    int a = 0, b = n;                                          // last i with offsets[i] <= p
    while (b - a > 1) { int mid = (a + b) >> 1; if (offsets[mid] <= p) a = mid; else b = mid; }
    const int i = a;
    const unsigned j = p - offsets[i];
    float ext = 0.0f;
#pragma unroll
    for (int k = 0; k < 3; k++) ext = fmaxf(ext, ord2f(bounds[3 + k]) - ord2f(bounds[k]));
    float lo[3], hi[3];
    const float4* t = tris + 5 * (size_t)i;
    tri_box(t, lo, hi);
    const unsigned plan = presplit_plan(lo, hi, ext / (float)grid);
    const unsigned nk[3] = {plan & 4095u, (plan >> 12) & 4095u, plan >> 24};
    p_tri[p] = i;
    if (nk[0] * nk[1] * nk[2] == 1u) {
        p_lo[p] = make_float4(lo[0], lo[1], lo[2], 0.0f); p_hi[p] = make_float4(hi[0], hi[1], hi[2], 0.0f);
        p_valid[p] = 1u;
        return;
    }
    const unsigned ix[3] = {j % nk[0], (j / nk[0]) % nk[1], j / (nk[0] * nk[1])};
    float clo[3], chi[3];
#pragma unroll
    for (int k = 0; k < 3; k++) {                              // piece ix of nk equal slabs of the triangle's own box; end planes exact
        const float e = hi[k] - lo[k];
        clo[k] = ix[k] == 0 ? lo[k] : lo[k] + e * ((float)ix[k] / (float)nk[k]);
        chi[k] = ix[k] + 1 == nk[k] ? hi[k] : lo[k] + e * ((float)(ix[k] + 1) / (float)nk[k]);
    }
    float4 r0 = t[0], r1 = t[1], r2 = t[2];
    float A[10][3], B[10][3];
    A[0][0] = r0.x; A[0][1] = r0.y; A[0][2] = r0.z;
    A[1][0] = r0.x + r0.w; A[1][1] = r0.y + r1.x; A[1][2] = r0.z + r1.y;
    A[2][0] = r0.x + r1.z; A[2][1] = r0.y + r1.w; A[2][2] = r0.z + r2.x;
    int m = 3;
    for (int k = 0; k < 3 && m > 0; k++) {
        if (nk[k] == 1u) continue;                             // the triangle lies inside its own box along this axis
        m = clip_halfspace(A, m, k, clo[k], true, B);
        m = m > 0 ? clip_halfspace(B, m, k, chi[k], false, A) : 0;
    }
    if (m < 3) { p_valid[p] = 0u; p_lo[p] = make_float4(0, 0, 0, 0); p_hi[p] = make_float4(0, 0, 0, 0); return; }
    float blo[3] = {A[0][0], A[0][1], A[0][2]}, bhi[3] = {A[0][0], A[0][1], A[0][2]};
    for (int q = 1; q < m; q++)
#pragma unroll
        for (int k = 0; k < 3; k++) { blo[k] = fminf(blo[k], A[q][k]); bhi[k] = fmaxf(bhi[k], A[q][k]); }
#pragma unroll
    for (int k = 0; k < 3; k++) {
        // interpolated vertices carry rounding error in the two coordinates that were not cut: widen by a few ulps of the
        // coordinates involved, then never beyond the piece or the triangle's own box
        const float pad = 4.0f * 1.1920929e-7f * fmaxf(fabsf(lo[k]), fabsf(hi[k]));
        blo[k] = fmaxf(fmaxf(blo[k] - pad, clo[k]), lo[k]);
        bhi[k] = fminf(fminf(bhi[k] + pad, chi[k]), hi[k]);
    }
    p_lo[p] = make_float4(blo[0], blo[1], blo[2], 0.0f); p_hi[p] = make_float4(bhi[0], bhi[1], bhi[2], 0.0f);
    p_valid[p] = 1u;
}
__global__ void k_presplit_compact(const int* __restrict__ p_tri, const float4* __restrict__ p_lo, const float4* __restrict__ p_hi,
                                   const unsigned* __restrict__ offsets /* scanned validity, total + 1 */, unsigned total, int* r_tri,
                                   float4* r_lo, float4* r_hi) {
    unsigned p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= total) return;
    const unsigned at = offsets[p];
    if (offsets[p + 1] == at) return;                          // dropped piece
    r_tri[at] = p_tri[p]; r_lo[at] = p_lo[p]; r_hi[at] = p_hi[p];
}

// ---------------------------------------------------------------- driver
#define BCK(call) do { cudaError_t e__ = (call); if (e__ != cudaSuccess) { err = e__; goto done; } } while (0)

void free_refset(RefSet* r) {
    if (!r) return;
    cudaFree(r->tri); cudaFree(r->lo); cudaFree(r->hi);
    r->tri = nullptr; r->lo = r->hi = nullptr; r->m = 0;
}

// See pt_kernels.h. Two 4-byte read-backs (piece count, reference count) size the arrays.
int presplit_device(const float4* d_tris, int n, int grid, RefSet* out, cudaStream_t st, float* kernel_ms) {
    cudaError_t err = cudaSuccess;
    int launches = 0;
    unsigned *bounds = nullptr, *pieces = nullptr, *valid = nullptr;
    int *tile_sums = nullptr, *d_total = nullptr, *p_tri = nullptr;
    float4 *p_lo = nullptr, *p_hi = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    const unsigned init_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    int total = 0, m = 0;
    *out = RefSet();
    auto scan = [&](unsigned* data, int count) {
        const int nb = (count + SCAN_TILE - 1) / SCAN_TILE;
        k_scan_tile_sums<<<nb, BVH_THREADS, 0, st>>>(data, count, tile_sums);
        k_ploc_scan<<<1, 1024, 0, st>>>(tile_sums, nb, d_total);
        k_scan_apply<<<nb, BVH_THREADS, 0, st>>>(data, count, tile_sums);
        launches += 3;
    };
    BCK(cudaEventCreate(&e0)); BCK(cudaEventCreate(&e1));
    BCK(cudaMalloc((void**)&bounds, sizeof(init_bounds)));
    BCK(cudaMalloc((void**)&pieces, ((size_t)n + 1) * 4));
    BCK(cudaMalloc((void**)&d_total, 4));
    BCK(cudaMemcpyAsync(bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
    BCK(cudaEventRecord(e0, st));
    k_tri_scene_box<<<min((n + BVH_THREADS - 1) / BVH_THREADS, 148 * 8), BVH_THREADS, 0, st>>>(d_tris, n, bounds); launches++;
    k_presplit_count<<<(n + 1 + BVH_THREADS - 1) / BVH_THREADS, BVH_THREADS, 0, st>>>(d_tris, n, bounds, grid, pieces); launches++;
    BCK(cudaMalloc((void**)&tile_sums, ((size_t)(n + 1) * 2 / SCAN_TILE + 64) * 4));      // pieces <= ~2 n in practice; checked below
    scan(pieces, n + 1);
    BCK(cudaMemcpyAsync(&total, d_total, 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    if (total < n || (long long)total > 2ll * n + (1 << 20)) { err = cudaErrorInvalidValue; goto done; }   // more than doubling the scene: refuse
    BCK(cudaMalloc((void**)&p_tri, (size_t)total * 4)); BCK(cudaMalloc((void**)&p_lo, (size_t)total * 16)); BCK(cudaMalloc((void**)&p_hi, (size_t)total * 16));
    BCK(cudaMalloc((void**)&valid, ((size_t)total + 1) * 4));
    cudaFree(tile_sums); tile_sums = nullptr;
    BCK(cudaMalloc((void**)&tile_sums, ((size_t)(total + 1) / SCAN_TILE + 64) * 4));
    k_presplit_pieces<<<(total + 1 + BVH_THREADS - 1) / BVH_THREADS, BVH_THREADS, 0, st>>>(d_tris, n, bounds, grid, pieces, (unsigned)total, p_tri, p_lo, p_hi, valid); launches++;
    scan(valid, total + 1);
    BCK(cudaMemcpyAsync(&m, d_total, 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaStreamSynchronize(st));
    if (m < 1 || m > total) { err = cudaErrorUnknown; goto done; }
    BCK(cudaMalloc((void**)&out->tri, (size_t)m * 4)); BCK(cudaMalloc((void**)&out->lo, (size_t)m * 16)); BCK(cudaMalloc((void**)&out->hi, (size_t)m * 16));
    k_presplit_compact<<<(total + BVH_THREADS - 1) / BVH_THREADS, BVH_THREADS, 0, st>>>(p_tri, p_lo, p_hi, valid, (unsigned)total, out->tri, out->lo, out->hi); launches++;
    out->m = m;
    BCK(cudaEventRecord(e1, st));
    BCK(cudaGetLastError());
    BCK(cudaStreamSynchronize(st));
    if (kernel_ms) BCK(cudaEventElapsedTime(kernel_ms, e0, e1));
done:
    cudaFree(bounds); cudaFree(pieces); cudaFree(valid); cudaFree(tile_sums); cudaFree(d_total); cudaFree(p_tri); cudaFree(p_lo); cudaFree(p_hi);
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    if (err != cudaSuccess) free_refset(out);
    return err == cudaSuccess ? launches : -(int)err;
}

// d_tris: n x 5 float4 (device). d_nodes: 2 * (2n - 1) float4, d_refs: n ints (device, caller-owned). d_pairs (n - 1) /
// d_leaf (n): optional, the traversal layout of pt_layout.h written directly. *depth: depth of the tree.
// Returns the number of kernels launched (> 0) or -(cudaError) on failure; *kernel_ms = device time of the build.
// ploc_radius > 0: the hierarchy is built by PLOC with that search radius instead of Karras' Morton-bit splits.
int build_lbvh_device(const float4* d_tris, int n, float4* d_nodes, int* d_refs, BvhPair* d_pairs, LeafTri* d_leaf, int* depth,
                      cudaStream_t st, float* kernel_ms, int ploc_radius, const RefSet* refset, int* n_leaves_out, void* workspace, size_t workspace_bytes) {
    // scratch comes out of the caller's workspace while it lasts (a rebuild per frame should not pay ~40 cudaMalloc / cudaFree), else from cudaMalloc
    struct Arena {
        char* base; size_t cap, off;
        cudaError_t get(void** p, size_t bytes) {
            const size_t need = (bytes + 255) & ~(size_t)255;
            if (base && off + need <= cap) { *p = base + off; off += need; return cudaSuccess; }
            return cudaMalloc(p, bytes);
        }
        void put(void* p) { if (p && !(base && (char*)p >= base && (char*)p < base + cap)) cudaFree(p); }
    } ws{(char*)workspace, workspace ? workspace_bytes : 0, 0};
    cudaError_t err = cudaSuccess;
    int launches = 0;
    // with a reference set the leaves are its m references (n becomes m below); d_nodes / d_refs / d_pairs / d_leaf are sized by m
    LeafSrc src{d_tris, refset ? refset->tri : nullptr, refset ? refset->lo : nullptr, refset ? refset->hi : nullptr};
    if (refset) n = refset->m;
    const int n_blocks = (n + SORT_TILE - 1) / SORT_TILE;
    const int grid_n = (n + BVH_THREADS - 1) / BVH_THREADS;
    unsigned *bounds = nullptr, *keys[2] = {nullptr, nullptr}, *hist = nullptr, *flags = nullptr;
    int *vals[2] = {nullptr, nullptr}, *parent_inner = nullptr, *parent_leaf = nullptr, *d_depth = nullptr;
    int2* children = nullptr;
    float4 *leaf_lo = nullptr, *leaf_hi = nullptr, *inner_lo = nullptr, *inner_hi = nullptr;
    int *cid[2] = {nullptr, nullptr}, *nn = nullptr, *valid = nullptr, *block_counts = nullptr, *d_m = nullptr, *subtree = nullptr, *remap = nullptr;
    float4 *clo[2] = {nullptr, nullptr}, *chi[2] = {nullptr, nullptr};
    float4 *sah_prim_lo = nullptr, *sah_prim_hi = nullptr;
    int *sah_idx[2] = {nullptr, nullptr}, *sah_node_of[2] = {nullptr, nullptr}, *sah_slot_of = nullptr, *sah_large[2] = {nullptr, nullptr}, *sah_small[2] = {nullptr, nullptr};
    int *sah_lists = nullptr, *sah_alloc = nullptr;
    unsigned *sah_cnt = nullptr, *sah_box = nullptr, *sah_part = nullptr, *sah_cb = nullptr;
    unsigned char* sah_leaf_end = nullptr;
    int n_leaves = 0, sah_multi_leaf = 1;
    SahNode* sah_nodes = nullptr; SahSplit* sah_splits = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    const unsigned init_bounds[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
    BCK(cudaEventCreate(&e0)); BCK(cudaEventCreate(&e1));
    BCK(ws.get((void**)&bounds, 6 * sizeof(unsigned)));
    for (int k = 0; k < 2; k++) { BCK(ws.get((void**)&keys[k], (size_t)n * 4)); BCK(ws.get((void**)&vals[k], (size_t)n * 4)); }
    BCK(ws.get((void**)&hist, (size_t)256 * n_blocks * 4));
    BCK(ws.get((void**)&flags, (size_t)n * 4));
    BCK(ws.get((void**)&parent_inner, (size_t)n * 4)); BCK(ws.get((void**)&parent_leaf, (size_t)n * 4));
    BCK(ws.get((void**)&children, (size_t)n * sizeof(int2)));
    BCK(ws.get((void**)&leaf_lo, (size_t)n * 16)); BCK(ws.get((void**)&leaf_hi, (size_t)n * 16));
    BCK(ws.get((void**)&inner_lo, (size_t)n * 16)); BCK(ws.get((void**)&inner_hi, (size_t)n * 16));
    BCK(cudaMemcpyAsync(bounds, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
    BCK(cudaMemsetAsync(flags, 0, (size_t)n * 4, st));
    BCK(ws.get((void**)&d_depth, 4));
    BCK(cudaMemsetAsync(d_depth, 0, 4, st));
    if (ploc_radius > 0) {
        for (int k = 0; k < 2; k++) {
            BCK(ws.get((void**)&cid[k], (size_t)n * 4)); BCK(ws.get((void**)&clo[k], (size_t)n * 16)); BCK(ws.get((void**)&chi[k], (size_t)n * 16));
        }
        BCK(ws.get((void**)&nn, (size_t)n * 4)); BCK(ws.get((void**)&valid, (size_t)n * 4));
        BCK(ws.get((void**)&block_counts, (size_t)grid_n * 4));
        BCK(ws.get((void**)&subtree, (size_t)n * 4)); BCK(ws.get((void**)&remap, (size_t)n * 4));
        BCK(ws.get((void**)&d_m, 8));
        BCK(cudaMemsetAsync(d_m, 0, 8, st));                     // [0] = cluster count of the round, [1] = inner nodes created
    }
    if (ploc_radius < 0) {
        // ---- top-down binned SAH (kernels above). Level loop on the host: one 8-byte read-back per level sizes the launches.
        const int max_slots = n / (SAH_SMALL + 1) + 2;
        BCK(ws.get((void**)&sah_prim_lo, (size_t)n * 16)); BCK(ws.get((void**)&sah_prim_hi, (size_t)n * 16));
        for (int k = 0; k < 2; k++) { BCK(ws.get((void**)&sah_idx[k], (size_t)n * 4)); BCK(ws.get((void**)&sah_node_of[k], (size_t)n * 4)); }
        BCK(ws.get((void**)&sah_nodes, (size_t)n * sizeof(SahNode))); BCK(ws.get((void**)&sah_slot_of, (size_t)n * 4));
        BCK(ws.get((void**)&subtree, (size_t)n * 4)); BCK(ws.get((void**)&remap, (size_t)n * 4));
        for (int k = 0; k < 2; k++) { BCK(ws.get((void**)&sah_large[k], (size_t)max_slots * 4)); BCK(ws.get((void**)&sah_small[k], (size_t)(n / 2 + 2) * 4)); }
        BCK(ws.get((void**)&sah_lists, 4 * 4)); BCK(ws.get((void**)&sah_alloc, 4));
        BCK(ws.get((void**)&sah_cb, (size_t)n * 6 * 4));
        BCK(ws.get((void**)&sah_leaf_end, (size_t)n));
        BCK(cudaMemsetAsync(sah_leaf_end, 0, (size_t)n, st));
        BCK(cudaMemsetAsync(parent_inner, 0xfe, (size_t)n * 4, st));       // every inner id starts out unused (< -1)
        BCK(cudaMemsetAsync(parent_leaf, 0xff, (size_t)n * 4, st));        // -1: only the first position of a leaf gets a parent
        if (const char* e = getenv("PT_SAH_SINGLE_LEAF")) sah_multi_leaf = atoi(e) ? 0 : 1;
        BCK(cudaMemcpyAsync(sah_cb, init_bounds, sizeof(init_bounds), cudaMemcpyHostToDevice, st));
        BCK(ws.get((void**)&sah_cnt, (size_t)max_slots * 3 * SAH_BINS * 4)); BCK(ws.get((void**)&sah_box, (size_t)max_slots * 3 * SAH_BINS * 6 * 4));
        BCK(ws.get((void**)&sah_splits, (size_t)max_slots * sizeof(SahSplit))); BCK(ws.get((void**)&sah_part, (size_t)max_slots * 2 * 4));
        const int one = 1;
        BCK(cudaMemcpyAsync(sah_alloc, &one, 4, cudaMemcpyHostToDevice, st));             // inner id 0 is the root
    }
    BCK(cudaEventRecord(e0, st));
    if (ploc_radius < 0) {
        const int max_slots = n / (SAH_SMALL + 1) + 2;
        k_sah_init<<<min(grid_n, 148 * 8), BVH_THREADS, 0, st>>>(src, n, sah_prim_lo, sah_prim_hi, sah_idx[0], sah_node_of[0], bounds, sah_cb); launches++;
        k_sah_root<<<1, 1, 0, st>>>(bounds, n, sah_nodes, inner_lo, inner_hi, parent_inner, subtree, sah_slot_of, sah_lists, sah_large[0], sah_small[0]); launches++;
        int cur = 0, lvl = 0;
        for (;; lvl++) {
            int counts[2] = {0, 0};
            BCK(cudaMemcpyAsync(counts, sah_lists + 2 * (lvl & 1), 8, cudaMemcpyDeviceToHost, st));
            BCK(cudaStreamSynchronize(st));
            if (counts[0] == 0 && counts[1] == 0) break;
            if (counts[0] > max_slots || counts[1] > n / 2 + 2 || lvl > 4 * PT_MAX_TREE_DEPTH) { err = cudaErrorUnknown; goto done; }
            int* next_lists = sah_lists + 2 * ((lvl + 1) & 1);
            BCK(cudaMemsetAsync(next_lists, 0, 8, st));
            const int lv = lvl & 1, nx = lv ^ 1;
            if (counts[0] > 0) {
                const size_t bins = (size_t)counts[0] * 3 * SAH_BINS;
                BCK(cudaMemsetAsync(sah_cnt, 0, bins * 4, st));
                k_sah_box_init<<<(unsigned)((bins * 6 + BVH_THREADS - 1) / BVH_THREADS), BVH_THREADS, 0, st>>>(sah_box, bins * 6); launches++;
                k_sah_bin<<<grid_n, BVH_THREADS, 0, st>>>(sah_idx[cur], sah_node_of[cur], n, sah_slot_of, sah_prim_lo, sah_prim_hi, sah_cb, sah_cnt, sah_box); launches++;
                k_sah_split<<<(counts[0] + SAH_SPLIT_WARPS - 1) / SAH_SPLIT_WARPS, 32 * SAH_SPLIT_WARPS, 0, st>>>(sah_large[lv], counts[0], sah_cnt, sah_box, sah_nodes, inner_lo, inner_hi, children, parent_inner,
                                                                      parent_leaf, subtree, sah_slot_of, next_lists, sah_large[nx], sah_small[nx], sah_alloc, sah_splits, sah_part, sah_cb, leaf_lo, leaf_hi, sah_leaf_end); launches++;
            }
            // copy-through + partition of the large nodes, then the small subtrees (disjoint positions of the other buffer)
            k_sah_partition<<<grid_n, BVH_THREADS, 0, st>>>(sah_idx[cur], sah_node_of[cur], n, sah_slot_of, sah_nodes, sah_splits, sah_prim_lo, sah_prim_hi, sah_cb,
                                                            sah_part, sah_idx[cur ^ 1], sah_node_of[cur ^ 1]); launches++;
            if (counts[1] > 0) {
                k_sah_small<<<(counts[1] + SAH_SMALL_WARPS - 1) / SAH_SMALL_WARPS, 32 * SAH_SMALL_WARPS, 0, st>>>(sah_small[lv], counts[1], sah_idx[cur], sah_nodes, sah_prim_lo, sah_prim_hi, inner_lo, inner_hi, children,
                                                                  parent_inner, parent_leaf, sah_alloc, sah_idx[cur ^ 1], sah_node_of[cur ^ 1], leaf_lo, leaf_hi, sah_leaf_end,
                                                                  sah_multi_leaf); launches++;
            }
            cur ^= 1;
        }
        k_sah_leaves<<<grid_n, BVH_THREADS, 0, st>>>(sah_idx[cur], n, vals[0]); launches++;
        BCK(cudaMemsetAsync(subtree, 0xff, (size_t)n * 4, st));            // -1 + the leaves below = inner nodes of the subtree
        BCK(cudaMemsetAsync(sah_lists, 0, 4, st));
        k_sah_count<<<grid_n, BVH_THREADS, 0, st>>>(sah_leaf_end, n, parent_inner, parent_leaf, subtree, sah_lists); launches++;
        BCK(cudaMemcpyAsync(&n_leaves, sah_lists, 4, cudaMemcpyDeviceToHost, st));
        k_ploc_relabel<<<grid_n, BVH_THREADS, 0, st>>>(children, parent_inner, subtree, n - 1, remap); launches++;
        goto emit;
    }
    k_scene_bounds<<<min(grid_n, 148 * 8), BVH_THREADS, 0, st>>>(src, n, bounds); launches++;
    k_morton<<<grid_n, BVH_THREADS, 0, st>>>(src, n, bounds, keys[0], vals[0]); launches++;
    for (int pass = 0; pass < 4; pass++) {
        const int src = pass & 1, dst = src ^ 1, shift = 8 * pass;
        k_sort_hist<<<n_blocks, BVH_THREADS, 0, st>>>(keys[src], n, shift, hist, n_blocks); launches++;
        k_sort_scan<<<1, 1024, 0, st>>>(hist, 256 * n_blocks); launches++;
        k_sort_scatter<<<n_blocks, BVH_THREADS, 0, st>>>(keys[src], vals[src], keys[dst], vals[dst], n, shift, hist, n_blocks); launches++;
    }
    // 4 passes: the sorted arrays are back in buffer 0
    if (ploc_radius > 0) {
        k_ploc_leaves<<<grid_n, BVH_THREADS, 0, st>>>(src, vals[0], n, leaf_lo, leaf_hi, cid[0], clo[0], chi[0]); launches++;
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
        k_fit<<<grid_n, BVH_THREADS, 0, st>>>(src, vals[0], n, children, parent_inner, parent_leaf, leaf_lo, leaf_hi, inner_lo, inner_hi, flags); launches++;
    }
emit:
    k_emit<<<grid_n, BVH_THREADS, 0, st>>>(vals[0], n, children, leaf_lo, leaf_hi, inner_lo, inner_hi, d_nodes, d_refs, d_tris, d_pairs, d_leaf, ploc_radius != 0 ? remap : nullptr, src.ref_tri,
                                            sah_leaf_end, ploc_radius < 0 ? parent_inner : nullptr); launches++;
    k_depth<<<grid_n, BVH_THREADS, 0, st>>>(parent_inner, parent_leaf, n, d_depth); launches++;
    BCK(cudaEventRecord(e1, st));
    if (depth) BCK(cudaMemcpyAsync(depth, d_depth, 4, cudaMemcpyDeviceToHost, st));
    BCK(cudaGetLastError());
    BCK(cudaStreamSynchronize(st));
    if (kernel_ms) BCK(cudaEventElapsedTime(kernel_ms, e0, e1));
    if (n_leaves_out) *n_leaves_out = ploc_radius < 0 ? n_leaves : n;
done:
    ws.put(bounds); ws.put(keys[0]); ws.put(keys[1]); ws.put(vals[0]); ws.put(vals[1]); ws.put(hist); ws.put(flags);
    ws.put(parent_inner); ws.put(parent_leaf); ws.put(children); ws.put(d_depth);
    ws.put(leaf_lo); ws.put(leaf_hi); ws.put(inner_lo); ws.put(inner_hi);
    for (int k = 0; k < 2; k++) { ws.put(cid[k]); ws.put(clo[k]); ws.put(chi[k]); }
    ws.put(nn); ws.put(valid); ws.put(block_counts); ws.put(d_m); ws.put(subtree); ws.put(remap);
    ws.put(sah_prim_lo); ws.put(sah_prim_hi); ws.put(sah_slot_of); ws.put(sah_lists); ws.put(sah_alloc); ws.put(sah_cnt); ws.put(sah_box);
    ws.put(sah_part); ws.put(sah_nodes); ws.put(sah_splits); ws.put(sah_cb); ws.put(sah_leaf_end);
    for (int k = 0; k < 2; k++) { ws.put(sah_idx[k]); ws.put(sah_node_of[k]); ws.put(sah_large[k]); ws.put(sah_small[k]); }
    if (e0) cudaEventDestroy(e0);
    if (e1) cudaEventDestroy(e1);
    return err == cudaSuccess ? launches : -(int)err;
}

}  // namespace pt
