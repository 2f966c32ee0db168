// pt_trace_wide.cuh -- closest-hit / any-hit traversal of the PT_BVH_WIDE layout (8-wide, 8-bit quantised child
// boxes; pt_layout.h WideNode, built by pt_bvh_wide.cpp).
//
// Same contract as pt_trace.cuh (IfIfClosestHit / IfIfAnyHit of common/BVH.glsl:634-854) with a different tree walk:
// one step decodes and tests 8 child boxes, so a ray needs ~3x fewer dependent memory round trips. Triangles are
// tested with the SAME IntersectTriangleMT arithmetic (tri_test), so (t, u, v) of a hit are bit-identical to the
// reference-layout path whenever both report the same triangle. What can differ is WHICH triangles are reached:
// quantised boxes are supersets of the fp32 boxes, so this walk reaches every triangle the reference walk reaches,
// plus the few the reference's non-conservative fp32 slab test culls although the ray does hit them.
//
// Traversal state follows Ylitie et al. 2017: a "node group" G = {child_base, hit bits << 24 | imask} and a
// "triangle group" T = {tri_base, triangle hit bits}; children are visited in descending (slot ^ octinv) order, the
// builder having placed children in slots by their position inside the parent.
#pragma once
#include "pt_trace.cuh"

namespace pt {

// the words of one node, whichever the width
template <int N> struct NodeWords;
template <> struct NodeWords<8> {
    float px, py, pz; unsigned hdr, child_base, tri_base;
    unsigned meta[2], lo[3][2], hi[3][2];
    __device__ __forceinline__ void load(const void* nodes, unsigned idx) {
        const uint4* nd = reinterpret_cast<const uint4*>(nodes) + 6 * (size_t)idx;        // 96 B
        F8 w0 = ldg256(nd), w1 = ldg256(nd + 2), w2 = ldg256(nd + 4);
        px = w0.lo.x; py = w0.lo.y; pz = w0.lo.z; hdr = __float_as_uint(w0.lo.w);
        child_base = __float_as_uint(w0.hi.x); tri_base = __float_as_uint(w0.hi.y);
        meta[0] = __float_as_uint(w0.hi.z); meta[1] = __float_as_uint(w0.hi.w);
        lo[0][0] = __float_as_uint(w1.lo.x); lo[0][1] = __float_as_uint(w1.lo.y);
        lo[1][0] = __float_as_uint(w1.lo.z); lo[1][1] = __float_as_uint(w1.lo.w);
        lo[2][0] = __float_as_uint(w1.hi.x); lo[2][1] = __float_as_uint(w1.hi.y);
        hi[0][0] = __float_as_uint(w1.hi.z); hi[0][1] = __float_as_uint(w1.hi.w);
        hi[1][0] = __float_as_uint(w2.lo.x); hi[1][1] = __float_as_uint(w2.lo.y);
        hi[2][0] = __float_as_uint(w2.lo.z); hi[2][1] = __float_as_uint(w2.lo.w);
    }
};
template <> struct NodeWords<4> {
    float px, py, pz; unsigned hdr, child_base, tri_base;
    unsigned meta[1], lo[3][1], hi[3][1];
    __device__ __forceinline__ void load(const void* nodes, unsigned idx) {
        const uint4* nd = reinterpret_cast<const uint4*>(nodes) + 4 * (size_t)idx;        // 64 B
        F8 w0 = ldg256(nd), w1 = ldg256(nd + 2);
        px = w0.lo.x; py = w0.lo.y; pz = w0.lo.z; hdr = __float_as_uint(w0.lo.w);
        child_base = __float_as_uint(w0.hi.x); tri_base = __float_as_uint(w0.hi.y);
        meta[0] = __float_as_uint(w0.hi.z);
        lo[0][0] = __float_as_uint(w0.hi.w); lo[1][0] = __float_as_uint(w1.lo.x); lo[2][0] = __float_as_uint(w1.lo.y);
        hi[0][0] = __float_as_uint(w1.lo.z); hi[1][0] = __float_as_uint(w1.lo.w); hi[2][0] = __float_as_uint(w1.hi.x);
    }
};

template <int N>
struct WideTraverser {
    typedef uint2 StackEntry;
    static const int kStack = 32;
    static const bool kPairFetch = false;      // the cooperative record fetch of pt_trace.cuh is for the 64-B pair records
    __device__ __forceinline__ void epilogue(unsigned long long*) {}
    V3 o, d, inv;
    HitRec h;
    uint2 G, T;
    int sp;
    unsigned octinv;
    bool result;

    static __device__ __forceinline__ void prologue(const SceneDev&) {}
    __device__ __forceinline__ void idle() { G = make_uint2(0, 0); T = make_uint2(0, 0); sp = 0; }
    __device__ __forceinline__ void begin(V3 org, V3 dir, float tmax) {
        o = org; d = dir;
        inv = v3(1.0f) / d;
        h.t = tmax; h.u = 0.0f; h.v = 0.0f; h.tri = -1;
        unsigned oct = (d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u);
        octinv = 7u - oct;
        G = make_uint2(0u, 0x80000000u);          // the root as a one-child group
        T = make_uint2(0u, 0u);
        sp = 0;
        result = false;
    }
    __device__ __forceinline__ bool at_leaf() const { return T.y != 0u; }
    __device__ __forceinline__ bool next_group(const uint2* stack) {
        if (T.y != 0u || (G.y & 0xff000000u) != 0u) return false;
        if (sp == 0) return true;
        G = stack[--sp];
        return false;
    }

    // one wide node: pick the nearest pending child of G, test its N children
    template <bool ANY>
    __device__ __forceinline__ bool inner_step(const SceneDev& s, uint2* stack) {
        const unsigned bit = 31u - (unsigned)__clz((int)(G.y & 0xff000000u));
        G.y &= ~(1u << bit);
        const unsigned slot = (bit - 24u) ^ octinv;
        const unsigned rel = __popc(G.y & 0xffu & ~(0xffffffffu << slot));
        NodeWords<N> w;
        w.load(s.wide_nodes, G.x + rel);
        if (G.y & 0xff000000u) stack[sp++] = G;
        const unsigned hdr = w.hdr;
        // A quantised plane q (0..255) lies at t = q * A + B with A = 2^e / d, B = (p - o) / d. Instead of 6N integer->float
        // conversions per node (I2F runs at a quarter of the FFMA rate), PRMT drops the byte into the mantissa of
        // 1.0f: v = 1 + q / 2^15, and t = v * (2^15 A) + (B - 2^15 A) is one FFMA. The offset costs |A| / 512 of
        // absolute accuracy; `dx` below widens every slab by a few ulps of the operands so the test stays conservative.
        const float ax = __uint_as_float(((hdr & 0xffu) + 15u) << 23) * inv.x;
        const float ay = __uint_as_float((((hdr >> 8) & 0xffu) + 15u) << 23) * inv.y;
        const float az = __uint_as_float((((hdr >> 16) & 0xffu) + 15u) << 23) * inv.z;
        const float b0x = (w.px - o.x) * inv.x, b0y = (w.py - o.y) * inv.y, b0z = (w.pz - o.z) * inv.z;
        const float dx = (fabsf(ax) + fabsf(b0x)) * 4.76837158e-7f, dy = (fabsf(ay) + fabsf(b0y)) * 4.76837158e-7f,
                    dz = (fabsf(az) + fabsf(b0z)) * 4.76837158e-7f;
        const float bnx = (b0x - ax) - dx, bfx = (b0x - ax) + dx;
        const float bny = (b0y - ay) - dy, bfy = (b0y - ay) + dy;
        const float bnz = (b0z - az) - dz, bfz = (b0z - az) + dz;
        // planes met first / last along the ray: swap lo and hi where the direction is negative
        const bool nx = inv.x < 0.0f, ny = inv.y < 0.0f, nz = inv.z < 0.0f;
        const float tcur = h.t;
        unsigned hitmask = 0u;
#pragma unroll
        for (int half = 0; half < N / 4; half++) {
            const unsigned nxw = nx ? w.hi[0][half] : w.lo[0][half], fxw = nx ? w.lo[0][half] : w.hi[0][half];
            const unsigned nyw = ny ? w.hi[1][half] : w.lo[1][half], fyw = ny ? w.lo[1][half] : w.hi[1][half];
            const unsigned nzw = nz ? w.hi[2][half] : w.lo[2][half], fzw = nz ? w.lo[2][half] : w.hi[2][half];
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const unsigned sel = 0x7604u + 0x10u * j;        // {0x3F, 0x80, byte j, 0x00}
                const float tlx = __fmaf_rn(__uint_as_float(__byte_perm(nxw, 0x3F800000u, sel)), ax, bnx);
                const float thx = __fmaf_rn(__uint_as_float(__byte_perm(fxw, 0x3F800000u, sel)), ax, bfx);
                const float tly = __fmaf_rn(__uint_as_float(__byte_perm(nyw, 0x3F800000u, sel)), ay, bny);
                const float thy = __fmaf_rn(__uint_as_float(__byte_perm(fyw, 0x3F800000u, sel)), ay, bfy);
                const float tlz = __fmaf_rn(__uint_as_float(__byte_perm(nzw, 0x3F800000u, sel)), az, bnz);
                const float thz = __fmaf_rn(__uint_as_float(__byte_perm(fzw, 0x3F800000u, sel)), az, bfz);
                const float tmin = fmaxf(fmaxf(tlx, tly), fmaxf(tlz, 0.0f));
                const float tmax = fminf(fminf(thx, thy), fminf(thz, tcur));
                if (tmin <= tmax) {
                    const unsigned m = (w.meta[half] >> (8 * j)) & 0xffu;
                    unsigned idx = m & 31u;
                    if (idx >= 24u) idx ^= octinv;               // inner child: priority by octant
                    hitmask |= (m >> 5) << idx;
                }
            }
        }
        G = make_uint2(w.child_base, (hitmask & 0xff000000u) | (hdr >> 24));
        T = make_uint2(w.tri_base, hitmask & 0x00ffffffu);
        return next_group(stack);
    }

    // one triangle of the current triangle group
    template <bool ANY>
    __device__ __forceinline__ bool leaf_step(const SceneDev& s, const uint2* stack) {
        const unsigned bit = (unsigned)__ffs((int)T.y) - 1u;
        T.y &= T.y - 1u;
        bool last;
        bool hit = tri_test(s.wide_tris + (T.x + bit), o, d, h, last);
        result = result || hit;
        if (ANY && result) return true;
        return next_group(stack);
    }
};

}  // namespace pt
