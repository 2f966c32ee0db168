// pt_layout.h -- HBM data layout of libb200pt (see DESIGN.md "Data layout in HBM").
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

// Deepest tree the binary traversal kernels accept (entries of the per-lane stack). The reference's shader has 27
// (BVH.glsl:489) and is undefined beyond; uploads of deeper trees are refused with PT_ERR_LIMIT. 96 rather than 64 since
// the PLOC builder came out 64-66 deep on the 10 M-triangle nested-shell mesh; entries beyond the few in shared memory
// are local memory that costs nothing until touched.
#ifndef PT_MAX_TREE_DEPTH
#define PT_MAX_TREE_DEPTH 96
#endif

namespace pt {

// ---- BVH, PT_BVH_REFERENCE layout -------------------------------------------------------------
// The reference stores one 32-B node per child and addresses sibling pairs at (firstChild,
// firstChild+1) with firstChild ODD (root = 0, its children = 1,2; BVH.cpp:2224-2255), so every
// pair straddles a 64-B boundary. Here one 64-B aligned record holds the pair = 2 x 256-bit loads
// from 2 sectors of one line:
//   q[0] = {c0.min.xyz, link0}   q[1] = {c0.max.xyz, 0}
//   q[2] = {c1.min.xyz, link1}   q[3] = {c1.max.xyz, 0}
// link >= 1 : index of the pair record holding that child's children
// link <  0 : leaf, ~link = index of its first record in the leaf triangle stream
// Topology, fp32 boxes, child order and leaf order are those of the reference tree, so traversal
// visits the same nodes in the same order.
struct BvhPair { float4 q[4]; };

// ---- leaf triangle stream ----------------------------------------------------------------------
// The reference walks leaf -> reference list (4 B) -> 80-B triangle. Here the reference list is
// resolved at upload time into a stream of 64-B (2 x 256-bit loads, 32-B aligned) intersection records in
// reference-list order; 40 B are used, the rest is padding that keeps every record inside one 128-B line:
//   r[0] = {p0.xyz, e1.x}  r[1] = {e1.y, e1.z, e2.x, e2.y}  r[2] = {e2.z, id, 0, 0}  r[3] = 0
// id = triangle index, stored as ~index when the reference-list entry carried the end-of-leaf
// marker (negative entry, BVH.cpp:2117). Entry "0" cannot carry the marker in the reference
// (-0 == 0): that leaf runs on into the next one; the stream reproduces this because the marker
// is copied, not recomputed.
struct LeafTri { float4 r[4]; };

// ---- BVH, PT_BVH_WIDE layout ----------------------------------------------------------------------
// 8-wide node, child boxes quantised to 8 bits per plane on a per-node power-of-two grid; 96 B, 32-B aligned =
// 3 x 256-bit loads (an 80-B record would need 5 x 128-bit loads: the traversal kernels are bound by L1 tag
// lookups per lane, not by bytes). Field-for-field WideNodeHost (pt_host.h):
//   w0 = {p.x, p.y, p.z, ex|ey<<8|ez<<16|imask<<24, child_base, tri_base, meta[0..3], meta[4..7]}
//   w1 = {qlo_x[0..3], qlo_x[4..7], qlo_y.., qlo_y.., qlo_z.., qlo_z.., qhi_x.., qhi_x..}
//   w2 = {qhi_y.., qhi_y.., qhi_z.., qhi_z.., pad x 4}
// Leaves (1-3 triangles per slot, at most 24 per node) index a second LeafTri stream in node order.
// PT_BVH_WIDE4: the same scheme with 4 children in 64 B = 2 x 256-bit loads (WideNode4Host).
struct WideNode { uint4 w[6]; };

// ---- shading tables ------------------------------------------------------------------------------
// Full reference triangle (5 x vec4, Scene.cpp:339-347 upload) is kept for shading: uv, normal, matId.
struct MaterialRec { float4 albedo_tex, emission_flag, props_tex; };   // == pt_material
struct TextureRec { int width, height; const float4* texels; };

struct CameraRec {
    float lower_left[3], horizontal[3], vertical[3], position[3], u[3], v[3];
    float lens_radius;
};

// ---- wavefront path state (SoA of float4 arrays, compacted every bounce, ping-pong) -----------
//   q0 = {ray.o.xyz, bxdfPdf0}      (ray.o doubles as lastPosition, Iterative.comp:286-288)
//   q1 = {ray.d.xyz, bxdfPdf1}      (viewDir == -ray.d)
//   q2 = {throughput.xyz, pixel index}
//   q3 = {contribution.xyz, bounce i}
//   q4 = HybridTaus state (uvec4)
struct PathQueue {
    float4* q0; float4* q1; float4* q2; float4* q3; uint4* q4;
};

// One NEE shadow item (64 B): ray + what to do with the answer.
//   s0 = {o.xyz, tmax}  s1 = {d.xyz, dest}  s2 = {L.xyz (already times Le), 0}  s3 = {contribution.xyz, 0}
// dest >= 0 : path continues in slot `dest` of the next queue; q3.xyz = contribution (+ L if visible)
// dest <  0 : path ended by Russian roulette; accum[~dest] += contribution (+ L if visible)
struct ShadowQueue { float4* s0; float4* s1; float4* s2; float4* s3; };

struct SceneDev {
    const BvhPair* pairs;
    const LeafTri* leaf_tris;
    int n_leaf_tris;
    int top_pairs;               // pair records [0, top_pairs) = the top levels in breadth-first order, staged in shared memory
    const void* wide_nodes;      // PT_BVH_WIDE (96-B nodes) / PT_BVH_WIDE4 (64-B nodes)
    const LeafTri* wide_tris;
    const float4* tris;          // n_tris x 5
    const MaterialRec* materials;
    const TextureRec* textures;
    int n_textures;
    const float2* lights;        // {cumulative area, tri index bits}
    int n_lights;
    float total_light_area;
    float env[3];                // constant environment ("GENERATE COLOR"), used when env_size == 0
    float sun_dir[3];            // uniform sunDir / sunRadius (Iterative.comp:89-90, Renderer.cpp:23-26,626-627)
    float sun_radius;
    unsigned long long* visits;  // optional (pt_set_visit_counting): {closest pair tests, closest triangle tests, shadow pair tests, shadow triangle tests}
    const float4* env_faces;     // cubemap: 6 x env_size x env_size RGBA32F, faces +X -X +Y -Y +Z -Z, rows as uploaded
    int env_size;
};

}  // namespace pt
