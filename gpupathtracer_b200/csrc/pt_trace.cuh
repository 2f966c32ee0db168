// pt_trace.cuh -- closest-hit / any-hit traversal of the PT_BVH_REFERENCE layout.
// Replaces IfIfClosestHit (common/BVH.glsl:634-767), IfIfAnyHit (:769-854), IntersectNodeFast (:379-391),
// IntersectLeaf (:420-437) and IntersectTriangleMT (common/Geometry.glsl:198-225) of the reference.
//
// Per RAY, the sequence of node visits, triangle tests and arithmetic is exactly the reference's, hence
// the same (triId, t, u, v) bit for bit. What is different is how a WARP schedules its 32 rays:
//   * a ray's next operation is either an INNER step (one sibling-pair test) or a LEAF step (one
//     triangle test). Instead of running "inner block; leaf block" every iteration (the shader's if-if
//     loop, which executes the expensive leaf block at ~5 % lane utilisation almost every time), the warp
//     votes: leaf steps run only when at least `leaf_thresh` lanes wait at a leaf or no lane has an inner
//     step left. Waiting does not change a ray's own sequence, so results cannot change;
//   * persistent threads with dynamic fetch: lanes whose ray has finished pull new rays from a device
//     cursor (one atomic per warp, ballot + popc ranks) once `fetch_thresh` lanes are idle.
#pragma once
#include "pt_math.cuh"
#include "pt_layout.h"

namespace pt {

#define PT_STACK_SIZE PT_MAX_TREE_DEPTH   // reference: 27 (BVH.glsl:489); upload rejects deeper trees with PT_ERR_LIMIT

struct HitRec { float t, u, v; int tri; };

// Per-lane traversal stack. Default: a local-memory array (L1-resident; lanes at the same depth share a wavefront).
// Compile-time experiment -DPT_SMEM_STACK=S: the S entries nearest the bottom live in shared memory, transposed
// ([entry][thread]: conflict-free, one wavefront per push/pop whatever the lanes' depths), deeper ones stay local.
#ifndef PT_SMEM_STACK
#define PT_SMEM_STACK 8
#endif
#if PT_SMEM_STACK > 0
__device__ __forceinline__ int* smem_stack() {
    __shared__ int s_stack[PT_SMEM_STACK * PT_TRACE_THREADS];
    return s_stack + threadIdx.x;
}
__device__ __forceinline__ void stack_push(int* stack, int sp, int v) {
    if (sp < PT_SMEM_STACK) smem_stack()[sp * PT_TRACE_THREADS] = v; else stack[sp] = v;
}
__device__ __forceinline__ int stack_at(const int* stack, int sp) {
    return sp < PT_SMEM_STACK ? smem_stack()[sp * PT_TRACE_THREADS] : stack[sp];
}
#else
__device__ __forceinline__ void stack_push(int* stack, int sp, int v) { stack[sp] = v; }
__device__ __forceinline__ int stack_at(const int* stack, int sp) { return stack[sp]; }
#endif

// Compile-time experiment -DPT_POP_CULL (closest hit only): the far child's entry distance is kept beside its link
// and an INNER entry whose distance exceeds the current hit distance is dropped at pop time without fetching its pair
// record. Exactly equivalent: child boxes lie inside the parent's box, the slab arithmetic is monotonic in the box
// planes, so both children of such a node fail `t_entry <= min(t_exit, t)` and the reference pops again. Leaf entries are
// never dropped (a triangle's Moeller-Trumbore t is not bounded by its box's t_entry in floating point).
#ifdef PT_POP_CULL
__device__ __forceinline__ float* smem_tstack() {
    __shared__ float s_tstack[(PT_SMEM_STACK > 0 ? PT_SMEM_STACK : 1) * PT_TRACE_THREADS];
    return s_tstack + threadIdx.x;
}
#endif

// slab test of one child (IntersectNodeFast); returns t_entry
__device__ __forceinline__ float slab_test(float4 bmin, float4 bmax, V3 inv, V3 oinv, float tcur, bool& hit) {
    float ax = bmin.x * inv.x + oinv.x, ay = bmin.y * inv.y + oinv.y, az = bmin.z * inv.z + oinv.z;
    float bx = bmax.x * inv.x + oinv.x, by = bmax.y * inv.y + oinv.y, bz = bmax.z * inv.z + oinv.z;
    float t_entry = gmax(gmin(ax, bx), gmax(gmin(ay, by), gmin(az, bz)));
    float t_exit = gmin(gmax(ax, bx), gmin(gmax(ay, by), gmin(gmax(az, bz), tcur)));
    hit = (t_entry <= t_exit) && (t_exit > 0.0f);
    return t_entry;
}

// Moeller-Trumbore with pre-subtracted edges; strict bounds as in the shader
__device__ __forceinline__ bool tri_test(const LeafTri* __restrict__ rec, V3 o, V3 d, HitRec& h, bool& last) {
    F8 lo = ldg256(&rec->r[0]);
    float2 tail = __ldg(reinterpret_cast<const float2*>(&rec->r[2]));
    float4 r0 = lo.lo, r1 = lo.hi;
    float4 r2 = make_float4(tail.x, tail.y, 0.0f, 0.0f);
    V3 p0 = v3(r0.x, r0.y, r0.z);
    V3 v01 = v3(r0.w, r1.x, r1.y);
    V3 v02 = v3(r1.z, r1.w, r2.x);
    int id = __float_as_int(r2.y);
    last = id < 0;
    id = last ? ~id : id;
    V3 p = cross(d, v02);
    float det = dot(v01, p);
    V3 tv = o - p0;
#ifdef PT_TRI_EARLY_REJECT
    // Experiment: u, v and t are numerator * (1 / det), so each has the sign of numerator x det (or is 0 / NaN / inf when one of
    // them is); a numerator that is zero or of the other sign than det cannot pass `u > 0`, `v > 0`, `t > 0`. Rejecting on the
    // signs skips the IEEE division (a ~20-instruction subroutine) when every lane of the leaf round rejects; everything that
    // is not rejected here goes through the unchanged arithmetic below, so accepted hits keep their bits.
    const float un = dot(tv, p);
    if (un == 0.0f || ((un < 0.0f) != (det < 0.0f))) return false;
    V3 q = cross(tv, v01);
    const float vn = dot(d, q), tn = dot(v02, q);
    if (vn == 0.0f || ((vn < 0.0f) != (det < 0.0f)) || tn == 0.0f || ((tn < 0.0f) != (det < 0.0f))) return false;
    float idet = 1.0f / det;
    float u = un * idet;
    float v = vn * idet;
    float t = tn * idet;
#else
    float idet = 1.0f / det;
    float u = dot(tv, p) * idet;
    V3 q = cross(tv, v01);
    float v = dot(d, q) * idet;
    float t = dot(v02, q) * idet;
#endif
    if (u > 0.0f && u < 1.0f && v > 0.0f && u + v < 1.0f && t < h.t && t > 0.0f) {
        h.t = t; h.u = u; h.v = v; h.tri = id;
        return true;
    }
    return false;
}

#ifndef PT_PAIR_FETCH_MIN
#define PT_PAIR_FETCH_MIN 12          // lanes that want a pair record before the warp fetches cooperatively (below: solo loads)
#endif
// Per-lane traversal state machine. `current` >= 0: pair record to test next (inner step);
// `current` < 0: ~current = next record of the leaf triangle stream to test (leaf step).
struct Traverser {
    typedef int StackEntry;
#ifdef PT_POP_CULL
    static const int kStack = 2 * PT_STACK_SIZE;
#else
    static const int kStack = PT_STACK_SIZE;
#endif
    V3 o, d, inv, oinv;
    HitRec h;
    int current, sp;
    bool result;          // "some triangle of the current leaf was accepted"
#ifdef PT_POP_CULL   // distances live in the upper half of the lane's stack array
    static __device__ __forceinline__ void tpush(int* stack, int at, float t) { if (at < PT_SMEM_STACK) smem_tstack()[at * PT_TRACE_THREADS] = t; else stack[PT_STACK_SIZE + at] = __float_as_int(t); }
    static __device__ __forceinline__ float tat(const int* stack, int at) { return at < PT_SMEM_STACK ? smem_tstack()[at * PT_TRACE_THREADS] : __int_as_float(stack[PT_STACK_SIZE + at]); }
#endif

    __device__ __forceinline__ void begin(V3 org, V3 dir, float tmax) {
        o = org; d = dir;
        inv = v3(1.0f) / d;
        oinv = -o * inv;
        h.t = tmax; h.u = 0.0f; h.v = 0.0f; h.tri = -1;
        current = 0;      // pair record of the root's children (RootFirstChild, BVH.glsl:627)
        sp = 0;
        result = false;
    }
    // Stage the top of the tree (pair records [0, top_pairs), breadth-first) in shared memory, once per CTA.
    // Compile-time experiment (build with PT_NVCC_EXTRA=-DPT_SMEM_TOP, run with PT_SMEM_TOP=<pairs>): measured slower than
    // L1 on both workloads, and even switched off at run time the extra compare/branch in inner_step cost 3.8 % on the
    // atrium (3.02 -> 2.90 Grays/s), so it is not compiled in by default.
    static __device__ __forceinline__ void prologue(const SceneDev& s) {
#ifdef PT_SMEM_TOP
        extern __shared__ float4 s_top[];
        const float4* src = reinterpret_cast<const float4*>(s.pairs);
        for (int i = threadIdx.x; i < 4 * s.top_pairs; i += blockDim.x) s_top[i] = __ldg(src + i);
        __syncthreads();
#else
        (void)s;
#endif
    }
    __device__ __forceinline__ void idle() { current = 0; sp = 0; }
    __device__ __forceinline__ void epilogue(unsigned long long*) {}          // see CountingTraverser
    __device__ __forceinline__ bool at_leaf() const { return current < 0; }
    // returns true when the ray is finished
    template <bool ANY>
    __device__ __forceinline__ bool pop(const int* stack) {
#ifdef PT_POP_CULL
        if (!ANY) {
            while (sp > 0) {
                current = stack_at(stack, --sp);
                if (current >= 0 && tat(stack, sp) > h.t) continue;
                return false;
            }
            return true;
        }
#endif
        if (sp == 0) return true;
        current = stack_at(stack, --sp);
        return false;
    }
    // ---- pair-record fetch. Solo: the lane's own record with two 256-bit loads. Cooperative (kPairFetch, called by a CONVERGED
    // warp): lanes 2k / 2k+1 fetch the record of lane 2k by halves with one load instruction, then the record of lane 2k+1
    // with a second one, and swap the halves they do not need with 7 SHFL.BFLY. The L1 cost of a scattered fetch is one tag
    // look-up per (instruction, 128-B line), so a record costs one look-up instead of two: tools/l1_microbench.cu measures
    // 1.41 instead of 2.33 SM-cycles per record at 32 lanes, 1.68 / 2.32 at 20, 2.34 / 2.18 at 10 (profiles/r08_l1_microbench.log).
    // MEASURED AND NOT ADOPTED (profiles/r08_pair_fetch.md): in the traversal kernels the tag accesses do drop by 30 %, but SHFL runs on
    // the same LSU data pipe (counted there as shared-memory wavefronts), so the binding pipe carries as many wavefronts as before,
    // and the warp-uniform loop + selects + shuffles double the warp instructions: 3284 -> 2438 Mrays/s on the headline (lane pairs
    // in lockstep instead of the whole warp, -DPT_PAIR_FETCH_MIN=0: 2222). Build with PT_NVCC_EXTRA=-DPT_PAIR_FETCH to reproduce.
#ifdef PT_PAIR_FETCH
    static const bool kPairFetch = true;
#else
    static const bool kPairFetch = false;
#endif
    static __device__ __forceinline__ void fetch_solo(const SceneDev& s, int cur, float4& a0, float4& a1, float4& b0, float4& b1) {
        const float4* q = s.pairs[cur].q;
        F8 ca = ldg256(q), cb = ldg256(q + 2);
        a0 = ca.lo; a1 = ca.hi; b0 = cb.lo; b1 = cb.hi;
    }
    static __device__ __forceinline__ void fetch_coop(const SceneDev& s, unsigned FULL /* lanes that call together: the warp, or the lane pair */,
                                                      int cur /* < 0: this lane wants nothing */, int other /* the partner's cur */,
                                                      float4& a0, float4& a1, float4& b0, float4& b1) {
        const int half = threadIdx.x & 1;
        const int even_cur = half ? other : cur, odd_cur = half ? cur : other;
        F8 e, o;
        e.lo = e.hi = o.lo = o.hi = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        if (even_cur >= 0) e = ldg256(s.pairs[even_cur].q + 2 * half);     // lane 2k: child 0 of its record, lane 2k+1: child 1 of it
        if (odd_cur >= 0) o = ldg256(s.pairs[odd_cur].q + 2 * half);       // the record of lane 2k+1 likewise
        // the even lane keeps e and needs the partner's e (its child 1); the odd lane keeps o and needs the partner's o (its child 0)
        const float4 g0 = half ? e.lo : o.lo, g1 = half ? e.hi : o.hi;
        float4 r0, r1;
        r0.x = __shfl_xor_sync(FULL, g0.x, 1); r0.y = __shfl_xor_sync(FULL, g0.y, 1); r0.z = __shfl_xor_sync(FULL, g0.z, 1);
        r0.w = __shfl_xor_sync(FULL, g0.w, 1);
        r1.x = __shfl_xor_sync(FULL, g1.x, 1); r1.y = __shfl_xor_sync(FULL, g1.y, 1); r1.z = __shfl_xor_sync(FULL, g1.z, 1);
        r1.w = 0.0f;                                                        // padding word of the record
        if (half) { a0 = r0; a1 = r1; b0 = o.lo; b1 = o.hi; } else { a0 = e.lo; a1 = e.hi; b0 = r0; b1 = r1; }
    }
    template <bool ANY>
    __device__ __forceinline__ bool inner_step(const SceneDev& s, int* stack) {
        float4 a0, a1, b0, b1;
#ifdef PT_SMEM_TOP
        if (current < s.top_pairs) {
            extern __shared__ float4 s_top[];
            const float4* q = s_top + 4 * current;
            a0 = q[0]; a1 = q[1]; b0 = q[2]; b1 = q[3];
        } else
#endif
        fetch_solo(s, current, a0, a1, b0, b1);
        return inner_apply<ANY>(s, stack, a0, a1, b0, b1);
    }
    // the pair test proper on a fetched record: both boxes, near child next, far child pushed
    template <bool ANY>
    __device__ __forceinline__ bool inner_apply(const SceneDev& s, int* stack, float4 a0, float4 a1, float4 b0, float4 b1) {
        bool hit0, hit1;
        float d0 = slab_test(a0, a1, inv, oinv, h.t, hit0);
        float d1 = slab_test(b0, b1, inv, oinv, h.t, hit1);
        int sub0 = __float_as_int(a0.w), sub1 = __float_as_int(b0.w);
        if (hit0 && hit1) {
            // Near child first. For ANY this differs from IfIfAnyHit's fixed order (BVH.glsl:806-812) but cannot
            // change the flag: until the first accepted triangle h.t stays at tmax, so which leaves are reachable
            // does not depend on the visiting order, and the answer is "some reachable leaf holds a hit".
#ifdef PT_ANY_FIXED_ORDER      // experiment: any-hit rays keep the reference's child order (larger child first, BVH.glsl:806-812)
            bool swap = ANY ? false : d0 > d1;
#else
            bool swap = d0 > d1;
#endif
            current = swap ? sub1 : sub0;
            const int far_child = swap ? sub0 : sub1;
#ifdef PT_POP_CULL
            if (!ANY) tpush(stack, sp, swap ? d0 : d1);
#endif
            stack_push(stack, sp++, far_child);
#ifdef PT_PREFETCH_FAR   // measured (build with PT_NVCC_EXTRA=-DPT_PREFETCH_FAR): atrium 3.03 -> 2.85 Grays/s, 10M-triangle batch 2.40 -> 1.91:
                         // the prefetch costs an L1 tag look-up on a pipe that is already the bottleneck, and half the pushed nodes are culled
            if (far_child >= 0) asm volatile("prefetch.global.L1 [%0];" :: "l"(s.pairs + far_child));
            else asm volatile("prefetch.global.L1 [%0];" :: "l"(s.leaf_tris + ~far_child));
#endif
            return false;
        }
        if (hit0 != hit1) { current = hit0 ? sub0 : sub1; return false; }
        return pop<ANY>(stack);
    }
    // one triangle of the current leaf; the leaf is always tested to its end marker (IntersectLeaf)
    template <bool ANY>
    __device__ __forceinline__ bool leaf_step(const SceneDev& s, const int* stack) {
        int ri = ~current;
        bool last;
        bool hit = tri_test(s.leaf_tris + ri, o, d, h, last);
        result = result || hit;
        if (!last && ri + 1 < s.n_leaf_tris) { current = ~(ri + 1); return false; }
        if (ANY && result) return true;                // IfIfAnyHit returns after the first leaf that hit
        return pop<ANY>(stack);
    }
};

// The same walk with visit counters (pt_set_visit_counting): pair tests and triangle tests of every ray this lane traced,
// summed per warp and added to SceneDev::visits at the end of the kernel -- the N_inner / N_tri of SURVEY 8d's algorithmic
// bytes, measured on the rays the kernel actually traced. An instrumented variant: the timed kernels are `Traverser`.
struct CountingTraverser : Traverser {
    unsigned long long n_pairs = 0, n_tris = 0;
    template <bool ANY>
    __device__ __forceinline__ bool inner_step(const SceneDev& s, int* stack) { n_pairs++; return Traverser::inner_step<ANY>(s, stack); }
    template <bool ANY>
    __device__ __forceinline__ bool inner_apply(const SceneDev& s, int* stack, float4 a0, float4 a1, float4 b0, float4 b1) {
        n_pairs++; return Traverser::inner_apply<ANY>(s, stack, a0, a1, b0, b1);
    }
    template <bool ANY>
    __device__ __forceinline__ bool leaf_step(const SceneDev& s, const int* stack) { n_tris++; return Traverser::leaf_step<ANY>(s, stack); }
    __device__ __forceinline__ void epilogue(unsigned long long* out) {
        unsigned long long a = n_pairs, b = n_tris;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { a += __shfl_down_sync(0xffffffffu, a, o); b += __shfl_down_sync(0xffffffffu, b, o); }
        if ((threadIdx.x & 31) == 0 && out) { atomicAdd(out, a); atomicAdd(out + 1, b); }
    }
};

// The warp-scheduled persistent traversal loop. IO supplies:
//   int  count() const;  int* cursor() const;
//   void load(int i, V3& o, V3& d, float& tmax) const;
//   void store(int i, const HitRec& h, bool result) const;
template <bool ANY, class TR, class IO>
__device__ __forceinline__ void trace_persistent(const SceneDev& s, const IO& io, int leaf_thresh, int fetch_thresh) {
    const unsigned FULL = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n = io.count();
    int* cursor = io.cursor();
    // bits 16.. of leaf_thresh: inner steps a lane may take per vote round (the vote/refill bookkeeping is ~28
    // instructions against ~60 for a pair test, so re-voting after every single step is expensive)
    const int burst = leaf_thresh >> 16, leaf_burst = (leaf_thresh >> 8) & 0xff;
    leaf_thresh &= 0xff;
    TR::prologue(s);
    TR tr;
    typename TR::StackEntry stack[TR::kStack];   // per-lane traversal stack (local memory, L1-resident)
    bool active = false, exhausted = false;
    int ridx = -1;
    tr.idle();
    while (true) {
        // ---- refill idle lanes
        unsigned m_idle = __ballot_sync(FULL, !active);
        if (!exhausted && (m_idle == FULL || __popc(m_idle) >= fetch_thresh)) {
            int cnt = __popc(m_idle), base = 0;
            int leader = __ffs(m_idle) - 1;
            if (lane == leader) base = atomicAdd(cursor, cnt);
            base = __shfl_sync(FULL, base, leader);
            if (!active) {
                int my = base + __popc(m_idle & lt_mask);
                if (my < n) {
                    V3 o, d; float tmax;
                    io.load(my, o, d, tmax);
                    tr.begin(o, d, tmax);
                    ridx = my;
                    active = true;
                }
            }
            if (base + cnt >= n) exhausted = true;
        }
        // ---- vote on the phase
        bool at_leaf = active && tr.at_leaf();
        unsigned m_leaf = __ballot_sync(FULL, at_leaf);
        unsigned m_active = __ballot_sync(FULL, active);
        if (m_active == 0) { if (exhausted) break; else continue; }
        unsigned m_inner = m_active & ~m_leaf;
        bool done = false;
        if (m_inner != 0 && __popc(m_leaf) < leaf_thresh) {
            if constexpr (TR::kPairFetch) {
#if PT_PAIR_FETCH_MIN > 0
                // the burst as a warp-uniform loop, so that lane pairs can share the loads of a step (fetch_coop)
                bool go = active && !at_leaf;
                for (int k = 0; k < burst; k++) {
                    const unsigned m_go = __ballot_sync(FULL, go);
                    if (m_go == 0) break;
                    float4 a0, a1, b0, b1;
                    if (__popc(m_go) >= PT_PAIR_FETCH_MIN) {
                        const int cur = go ? tr.current : -1;
                        TR::fetch_coop(s, FULL, cur, __shfl_xor_sync(FULL, cur, 1), a0, a1, b0, b1);
                    } else if (go) TR::fetch_solo(s, tr.current, a0, a1, b0, b1);
                    if (go) {
                        done = tr.template inner_apply<ANY>(s, stack, a0, a1, b0, b1);
                        go = !done && !tr.at_leaf();
                    }
                }
#else
                // lane PAIRS in lockstep, the warp is not: a lane whose partner still has steps to take stays as its helper
                bool go = active && !at_leaf;
                const unsigned pm = 3u << (lane & ~1);
                for (int k = 0; k < burst; k++) {
                    const int cur = go ? tr.current : -1;
                    const int other = __shfl_xor_sync(pm, cur, 1);
                    if (cur < 0 && other < 0) break;
                    float4 a0, a1, b0, b1;
                    TR::fetch_coop(s, pm, cur, other, a0, a1, b0, b1);
                    if (go) {
                        done = tr.template inner_apply<ANY>(s, stack, a0, a1, b0, b1);
                        go = !done && !tr.at_leaf();
                    }
                }
#endif
            } else if (active && !at_leaf) {
                done = tr.template inner_step<ANY>(s, stack);
                for (int k = 1; k < burst && !done && !tr.at_leaf(); k++) done = tr.template inner_step<ANY>(s, stack);
            }
        } else {
            if (at_leaf) {
                done = tr.template leaf_step<ANY>(s, stack);
                for (int k = 1; k < leaf_burst && !done && tr.at_leaf(); k++) done = tr.template leaf_step<ANY>(s, stack);
            }
        }
        if (done) {
            io.store(ridx, tr.h, tr.result);
            active = false;
        }
    }
    tr.epilogue(s.visits ? s.visits + (ANY ? 2 : 0) : nullptr);
}

}  // namespace pt
