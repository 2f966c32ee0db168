// pt_trace.cuh -- closest-hit / any-hit traversal of the PT_BVH_REFERENCE layout.
// Replaces IfIfClosestHit (common/BVH.glsl:634-767), IfIfAnyHit (:769-854), IntersectNodeFast (:379-391),
// IntersectLeaf (:420-437) and IntersectTriangleMT (common/Geometry.glsl:198-225) of the reference.
// Same node visit order and the same arithmetic (see pt_math.cuh), hence the same (triId, t, u, v).
#pragma once
#include "pt_math.cuh"
#include "pt_layout.h"

namespace pt {

#define PT_STACK_SIZE 64   // reference: 27 (BVH.glsl:489); upload rejects deeper trees with PT_ERR_LIMIT

struct HitRec { float t, u, v; int tri; };

// slab test of one child; returns t_entry
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
    float4 r0 = __ldg(&rec->r[0]);
    float4 r1 = __ldg(&rec->r[1]);
    float4 r2 = __ldg(&rec->r[2]);
    V3 p0 = v3(r0.x, r0.y, r0.z);
    V3 v01 = v3(r0.w, r1.x, r1.y);
    V3 v02 = v3(r1.z, r1.w, r2.x);
    int id = __float_as_int(r2.y);
    last = id < 0;
    id = last ? ~id : id;
    V3 p = cross(d, v02);
    float det = dot(v01, p);
    float idet = 1.0f / det;
    V3 tv = o - p0;
    float u = dot(tv, p) * idet;
    V3 q = cross(tv, v01);
    float v = dot(d, q) * idet;
    float t = dot(v02, q) * idet;
    if (u > 0.0f && u < 1.0f && v > 0.0f && u + v < 1.0f && t < h.t && t > 0.0f) {
        h.t = t; h.u = u; h.v = v; h.tri = id;
        return true;
    }
    return false;
}

// ANY = false: closest hit (near child first, far child pushed).
// ANY = true : stops after the first LEAF that produced a hit (the leaf itself is fully tested).
template <bool ANY>
__device__ __forceinline__ bool traverse(const SceneDev& s, V3 o, V3 d, HitRec& h) {
    V3 inv = v3(1.0f) / d;
    V3 oinv = -o * inv;
    int stack[PT_STACK_SIZE];
    int sp = 0;
    int current = 0;          // pair record of the root's children (RootFirstChild, BVH.glsl:627)
    bool result = false;
    while (true) {
        if (current >= 0) {
            const float4* q = s.pairs[current].q;
            float4 a0 = __ldg(q + 0), a1 = __ldg(q + 1), b0 = __ldg(q + 2), b1 = __ldg(q + 3);
            bool hit0, hit1;
            float d0 = slab_test(a0, a1, inv, oinv, h.t, hit0);
            float d1 = slab_test(b0, b1, inv, oinv, h.t, hit1);
            int sub0 = __float_as_int(a0.w), sub1 = __float_as_int(b0.w);
            if (hit0 && hit1) {
                bool swap = ANY ? false : (d0 > d1);
                current = swap ? sub1 : sub0;
                stack[sp++] = swap ? sub0 : sub1;
            } else if (hit0 != hit1) {
                current = hit0 ? sub0 : sub1;
            } else {
                if (sp == 0) break;
                current = stack[--sp];
            }
        }
        if (current < 0) {
            const LeafTri* rec = s.leaf_tris + (~current);
            const LeafTri* end = s.leaf_tris + s.n_leaf_tris;
            bool last = false;
            while (!last && rec < end) {
                bool hit = tri_test(rec, o, d, h, last);
                result = result || hit;
                ++rec;
            }
            if (ANY && result) return true;
            if (sp == 0) break;
            current = stack[--sp];
        }
    }
    return result;
}

}  // namespace pt
