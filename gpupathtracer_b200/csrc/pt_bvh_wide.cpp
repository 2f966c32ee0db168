// pt_bvh_wide.cpp -- collapse of the binary SBVH into the PT_BVH_WIDE layout: 8-wide nodes with child boxes
// quantised to 8 bits per plane on a per-node power-of-two grid (the "compressed wide BVH" of Ylitie, Karras,
// Laine, HPG 2017), 96 B per node = 3 x 256-bit loads (pt_layout.h).
//
// The reference has no counterpart: its device tree is the 32-B binary NewLayout node array (BVH.cpp:2261-2277)
// walked one sibling pair per step (BVH.glsl:634-767). BASELINE.json's north_star asks for "a compact 8-wide or
// quantised node format"; this is it. Input is the reference-form tree (whoever built it: the reference's
// BuildBinnedSpatial or pt_bvh_build.cpp), so triangle-to-leaf assignment and spatial splits are inherited.
//
// Conservative by construction: every quantised child box contains the fp32 box it replaces, hence every
// triangle the binary traversal can reach is reachable here (DESIGN.md discusses the residual differences).
#include "pt_host.h"
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <deque>
#include <cstdlib>

namespace pt {
namespace {

struct BNode {
    float lo[3], hi[3];
    int left = -1, right = -1;         // both -1: leaf
    std::vector<int> tris;
};

inline float box_area(const float* lo, const float* hi) {
    float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
    if (x < 0 || y < 0 || z < 0) return 0.0f;
    return 2.0f * (x * (y + z) + y * z);
}

struct Collapser {
    const HostBvh& in;
    const float* tris;                  // n x 20, pre-subtracted edges
    std::vector<BNode> b;

    int32_t data0(int i) const { int32_t d; std::memcpy(&d, in.nodes.data() + 8 * (size_t)i + 3, 4); return d; }

    // bounds of triangle `t` clipped to [lo, hi] (SBVH leaves hold references whose boxes were clipped)
    void tri_bounds(int t, const float* clo, const float* chi, float* lo, float* hi) const {
        const float* v = tris + 20 * (size_t)t;
        for (int k = 0; k < 3; k++) {
            float a = v[k], bb = v[k] + v[3 + k], c = v[k] + v[6 + k];
            lo[k] = std::max(std::min(a, std::min(bb, c)), clo[k]);
            hi[k] = std::min(std::max(a, std::max(bb, c)), chi[k]);
            if (lo[k] > hi[k]) { lo[k] = clo[k]; hi[k] = chi[k]; }     // numerically outside: keep the leaf box
        }
    }

    // reference-form tree -> pointer-free binary nodes; leaves keep their triangle lists
    int import(int i) {
        int me = (int)b.size();
        b.emplace_back();
        const float* nd = in.nodes.data() + 8 * (size_t)i;
        for (int k = 0; k < 3; k++) { b[me].lo[k] = nd[k]; b[me].hi[k] = nd[4 + k]; }
        int32_t d0 = data0(i);
        if (d0 > 0) {
            int l = import(d0); b[me].left = l;
            int r = import(d0 + 1); b[me].right = r;
        } else {
            // IntersectLeaf (BVH.glsl:420-437): walk until the negated entry; entry 0 cannot carry the marker, so
            // that leaf runs on into the next one -- reproduced, those triangles are tested by the reference too
            std::vector<int> t;
            for (size_t k = (size_t)(-d0); k < in.refs.size(); k++) {
                int32_t r = in.refs[k];
                t.push_back(r < 0 ? -r : r);
                if (r < 0) break;
            }
            std::sort(t.begin(), t.end());
            t.erase(std::unique(t.begin(), t.end()), t.end());
            b[me].tris = std::move(t);
        }
        return me;
    }

    // a wide leaf slot holds at most 3 triangles: larger leaves become small median-split subtrees
    void split_big_leaf(int me) {
        if (b[me].left >= 0 || b[me].tris.size() <= 3) return;
        std::vector<int> t = std::move(b[me].tris);
        b[me].tris.clear();
        int ax = 0; float ext = -1.0f;
        for (int k = 0; k < 3; k++) if (b[me].hi[k] - b[me].lo[k] > ext) { ext = b[me].hi[k] - b[me].lo[k]; ax = k; }
        std::vector<std::pair<float, int>> key;
        for (int id : t) {
            float lo[3], hi[3]; tri_bounds(id, b[me].lo, b[me].hi, lo, hi);
            key.push_back({0.5f * (lo[ax] + hi[ax]), id});
        }
        std::sort(key.begin(), key.end());
        size_t half = key.size() / 2;
        int kids[2];
        for (int side = 0; side < 2; side++) {
            int c = (int)b.size(); b.emplace_back(); kids[side] = c;
            for (int k = 0; k < 3; k++) { b[c].lo[k] = FLT_MAX; b[c].hi[k] = -FLT_MAX; }
            size_t from = side ? half : 0, to = side ? key.size() : half;
            for (size_t q = from; q < to; q++) {
                float lo[3], hi[3]; tri_bounds(key[q].second, b[me].lo, b[me].hi, lo, hi);
                for (int k = 0; k < 3; k++) { b[c].lo[k] = std::min(b[c].lo[k], lo[k]); b[c].hi[k] = std::max(b[c].hi[k], hi[k]); }
                b[c].tris.push_back(key[q].second);
            }
        }
        b[me].left = kids[0]; b[me].right = kids[1];
        split_big_leaf(kids[0]); split_big_leaf(kids[1]);
    }
};

// exponent byte e (IEEE-754 biased) of the smallest power of two with 2^e * 255 >= extent
inline uint8_t grid_exponent(float extent) {
    if (!(extent > 0.0f)) return 1;
    int e = (int)std::ceil(std::log2((double)extent / 255.0));
    e = std::min(std::max(e + 127, 1), 230);
    while (e < 230 && std::ldexp(255.0, e - 127) < (double)extent) e++;
    return (uint8_t)e;
}

}  // namespace

// One wide node before serialisation: children in ascending slot order (slot = one of 8 octant-ordered positions
// even when the node holds at most 4 children).
struct NodeRec {
    float p[3]; uint8_t e[3]; uint8_t imask; uint32_t child_base, tri_base;
    int n = 0;
    uint8_t meta[8], qlo[3][8], qhi[3][8];
};

static void serialise(const NodeRec& r, int width, uint8_t* dst) {
    if (width == 8) {
        WideNodeHost w; std::memset(&w, 0, sizeof(w));
        std::memcpy(w.p, r.p, 12); std::memcpy(w.e, r.e, 3); w.imask = r.imask; w.child_base = r.child_base; w.tri_base = r.tri_base;
        for (int k = 0; k < 8; k++) {
            w.meta[k] = k < r.n ? r.meta[k] : 0;
            for (int a = 0; a < 3; a++) { w.qlo[a][k] = k < r.n ? r.qlo[a][k] : 255; w.qhi[a][k] = k < r.n ? r.qhi[a][k] : 0; }
        }
        std::memcpy(dst, &w, sizeof(w));
    } else {
        WideNode4Host w; std::memset(&w, 0, sizeof(w));
        std::memcpy(w.p, r.p, 12); std::memcpy(w.e, r.e, 3); w.imask = r.imask; w.child_base = r.child_base; w.tri_base = r.tri_base;
        for (int k = 0; k < 4; k++) {
            w.meta[k] = k < r.n ? r.meta[k] : 0;
            for (int a = 0; a < 3; a++) { w.qlo[a][k] = k < r.n ? r.qlo[a][k] : 255; w.qhi[a][k] = k < r.n ? r.qhi[a][k] : 0; }
        }
        std::memcpy(dst, &w, sizeof(w));
    }
}

bool build_wide_bvh(const HostBvh& in, const float* tris_pre, int64_t n_tris, int width, WideBvh& out, std::string& msg) {
    (void)n_tris;
    if (width != 4 && width != 8) { msg = "wide BVH width must be 4 or 8"; return false; }
    if (in.nodes.size() < 24) { msg = "wide collapse needs a tree whose root is an inner node"; return false; }
    Collapser c{in, tris_pre, {}};
    c.b.reserve(in.nodes.size() / 8 + 16);
    int root = c.import(0);
    for (size_t i = 0, n0 = c.b.size(); i < n0; i++) c.split_big_leaf((int)i);
    const std::vector<BNode>& b = c.b;

    std::vector<NodeRec> recs;
    out.tri_ids.clear(); out.depth = 0; out.width = width; out.node_size = width == 8 ? 96 : 64;
    struct Item { int wide, bin, depth; };
    std::deque<Item> queue;
    recs.emplace_back();
    queue.push_back({0, root, 0});
    while (!queue.empty()) {
        Item it = queue.front(); queue.pop_front();
        out.depth = std::max(out.depth, it.depth);
        // ---- gather up to `width` children: repeatedly open the inner child with the largest surface area
        int kids[8]; int nk = 0;
        kids[nk++] = b[it.bin].left; kids[nk++] = b[it.bin].right;
        while (nk < width) {
            int best = -1; float best_area = -1.0f;
            for (int k = 0; k < nk; k++) {
                const BNode& x = b[kids[k]];
                if (x.left < 0) continue;
                float a = box_area(x.lo, x.hi);
                if (a > best_area) { best_area = a; best = k; }
            }
            if (best < 0) break;
            int open = kids[best];
            kids[best] = b[open].left; kids[nk++] = b[open].right;
        }
        // ---- node frame: origin = min corner of the children, one power-of-two grid per axis
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int k = 0; k < nk; k++)
            for (int a = 0; a < 3; a++) { lo[a] = std::min(lo[a], b[kids[k]].lo[a]); hi[a] = std::max(hi[a], b[kids[k]].hi[a]); }
        // ---- slot assignment: slot s stands for the octant (s&1 ? +x : -x, s&2 ? +y : -y, s&4 ? +z : -z) of the node;
        //      traversal visits slots in descending (s ^ octinv), i.e. a ray travelling towards +x meets the -x slots
        //      first. Greedy maximum of <child centre - node centre, octant direction> over the 8 x nk table.
        int slot_of[8]; bool slot_used[8] = {false, false, false, false, false, false, false, false};
        bool kid_done[8] = {false, false, false, false, false, false, false, false};
        float cost[8][8];
        for (int k = 0; k < nk; k++)
            for (int s = 0; s < 8; s++) {
                float d = 0.0f;
                for (int a = 0; a < 3; a++) {
                    float cc = 0.5f * (b[kids[k]].lo[a] + b[kids[k]].hi[a]) - 0.5f * (lo[a] + hi[a]);
                    d += ((s >> a) & 1) ? cc : -cc;
                }
                cost[k][s] = d;
            }
        for (int round = 0; round < nk; round++) {
            int bk = -1, bs = -1; float bc = -FLT_MAX;
            for (int k = 0; k < nk; k++) if (!kid_done[k])
                for (int s = 0; s < 8; s++) if (!slot_used[s] && cost[k][s] > bc) { bc = cost[k][s]; bk = k; bs = s; }
            kid_done[bk] = true; slot_used[bs] = true; slot_of[bk] = bs;
        }
        int kid_in_slot[8] = {-1, -1, -1, -1, -1, -1, -1, -1};
        for (int k = 0; k < nk; k++) kid_in_slot[slot_of[k]] = kids[k];

        NodeRec w; std::memset(&w, 0, sizeof(w));
        for (int a = 0; a < 3; a++) { w.p[a] = lo[a]; w.e[a] = grid_exponent(hi[a] - lo[a]); }
        w.child_base = (uint32_t)recs.size();
        w.tri_base = (uint32_t)out.tri_ids.size();
        int tri_off = 0;
        for (int s = 0; s < 8; s++) {
            int kid = kid_in_slot[s];
            if (kid < 0) continue;
            const int k = w.n++;                              // position in the node's child arrays
            const BNode& x = b[kid];
            for (int a = 0; a < 3; a++) {
                double scale = std::ldexp(1.0, (int)w.e[a] - 127);
                double ql = std::floor(((double)x.lo[a] - (double)w.p[a]) / scale);
                double qh = std::ceil(((double)x.hi[a] - (double)w.p[a]) / scale);
                ql = std::min(std::max(ql, 0.0), 255.0); qh = std::min(std::max(qh, 0.0), 255.0);
                // the kernel evaluates p + scale * q in fp32: make sure that still encloses the fp32 box
                while (ql > 0.0 && w.p[a] + (float)scale * (float)ql > x.lo[a]) ql -= 1.0;
                while (qh < 255.0 && w.p[a] + (float)scale * (float)qh < x.hi[a]) qh += 1.0;
                w.qlo[a][k] = (uint8_t)ql; w.qhi[a][k] = (uint8_t)qh;
            }
            if (x.left >= 0) {
                w.imask |= (uint8_t)(1u << s);
                w.meta[k] = (uint8_t)((1u << 5) | (24 + s));
                queue.push_back({(int)recs.size(), kid, it.depth + 1});     // inner children: contiguous, slot order
                recs.emplace_back();
            } else {
                int cnt = (int)x.tris.size();                // 1..3 after split_big_leaf
                if (cnt < 1 || cnt > 3) { msg = "internal: leaf with an unsupported triangle count"; return false; }
                w.meta[k] = (uint8_t)((((1u << cnt) - 1u) << 5) | (unsigned)tri_off);
                for (int id : x.tris) out.tri_ids.push_back(id);
                tri_off += cnt;
            }
        }
        recs[it.wide] = w;
    }
    if (out.depth >= 30) { msg = "wide tree too deep for the traversal stack"; return false; }
    out.n_nodes = (int64_t)recs.size();
    out.bytes.assign((size_t)out.n_nodes * out.node_size, 0);
    for (int64_t i = 0; i < out.n_nodes; i++) serialise(recs[i], width, out.bytes.data() + (size_t)i * out.node_size);
    return true;
}

}  // namespace pt

extern "C" int pt_bvh_collapse_host(const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs, const float* tris,
                                    int64_t n_tris, int32_t width, uint8_t** node_bytes, int64_t* n_wide, int32_t* node_size,
                                    int32_t** tri_ids, int64_t* n_tri_ids, int32_t* depth) {
    if (!nodes || !refs || !tris || !node_bytes || !n_wide || !node_size || !tri_ids || !n_tri_ids) return PT_ERR_INVALID;
    std::string msg; int n_leaves = 0;
    if (pt::validate_bvh(nodes, n_nodes, refs, n_refs, n_tris, msg, &n_leaves) < 0) return PT_ERR_INVALID;
    pt::HostBvh in;
    in.nodes.assign(nodes, nodes + 8 * (size_t)n_nodes);
    in.refs.assign(refs, refs + n_refs);
    pt::WideBvh w;
    if (!pt::build_wide_bvh(in, tris, n_tris, width, w, msg)) return PT_ERR_LIMIT;
    *node_bytes = (uint8_t*)std::malloc(w.bytes.size() ? w.bytes.size() : 1);
    std::memcpy(*node_bytes, w.bytes.data(), w.bytes.size());
    *tri_ids = (int32_t*)std::malloc((w.tri_ids.size() ? w.tri_ids.size() : 1) * 4);
    std::memcpy(*tri_ids, w.tri_ids.data(), w.tri_ids.size() * 4);
    *n_wide = w.n_nodes; *node_size = w.node_size; *n_tri_ids = (int64_t)w.tri_ids.size();
    if (depth) *depth = w.depth;
    return PT_OK;
}
