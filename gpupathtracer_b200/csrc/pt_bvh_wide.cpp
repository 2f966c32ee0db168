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
#include <atomic>
#include <thread>
#include <cstdlib>

namespace pt {
namespace {

struct BNode {
    float lo[3], hi[3];
    int left = -1, right = -1;         // both -1: leaf
    int tri_first = 0, tri_count = 0;  // leaf: range in Collapser::leaf_tris
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
    std::vector<int> leaf_tris;         // triangle ids of all leaves, one contiguous range per leaf
    std::vector<int> big_leaves;        // leaves with more than 3 triangles (split_big_leaf)

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

    // reference-form tree -> pointer-free binary nodes (pre-order, explicit stack: the serialised tree may be deep);
    // leaves keep their triangle lists
    int import_tree() {
        struct Todo { int ser, parent; bool right; };
        std::vector<Todo> stack;
        stack.push_back({0, -1, false});
        while (!stack.empty()) {
            Todo td = stack.back(); stack.pop_back();
            int me = (int)b.size();
            b.emplace_back();
            BNode& x = b[me];
            if (td.parent >= 0) { if (td.right) b[td.parent].right = me; else b[td.parent].left = me; }
            const float* nd = in.nodes.data() + 8 * (size_t)td.ser;
            for (int k = 0; k < 3; k++) { x.lo[k] = nd[k]; x.hi[k] = nd[4 + k]; }
            int32_t d0 = data0(td.ser);
            if (d0 > 0) {
                stack.push_back({d0 + 1, me, true});
                stack.push_back({d0, me, false});
            } else {
                // IntersectLeaf (BVH.glsl:420-437): walk until the negated entry; entry 0 cannot carry the marker, so
                // that leaf runs on into the next one -- reproduced, those triangles are tested by the reference too
                size_t first = leaf_tris.size();
                for (size_t k = (size_t)(-d0); k < in.refs.size(); k++) {
                    int32_t r = in.refs[k];
                    leaf_tris.push_back(r < 0 ? -r : r);
                    if (r < 0) break;
                }
                if (leaf_tris.size() - first > 1) {
                    std::sort(leaf_tris.begin() + first, leaf_tris.end());
                    leaf_tris.erase(std::unique(leaf_tris.begin() + first, leaf_tris.end()), leaf_tris.end());
                }
                x.tri_first = (int)first; x.tri_count = (int)(leaf_tris.size() - first);
                if (x.tri_count > 3) big_leaves.push_back(me);
            }
        }
        return 0;
    }

    // a wide leaf slot holds at most 3 triangles: larger leaves become small median-split subtrees
    void split_big_leaf(int me) {
        if (b[me].left >= 0 || b[me].tri_count <= 3) return;
        std::vector<int> t(leaf_tris.begin() + b[me].tri_first, leaf_tris.begin() + b[me].tri_first + b[me].tri_count);
        b[me].tri_count = 0;
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
            b[c].tri_first = (int)leaf_tris.size();
            for (size_t q = from; q < to; q++) {
                float lo[3], hi[3]; tri_bounds(key[q].second, b[me].lo, b[me].hi, lo, hi);
                for (int k = 0; k < 3; k++) { b[c].lo[k] = std::min(b[c].lo[k], lo[k]); b[c].hi[k] = std::max(b[c].hi[k], hi[k]); }
                leaf_tris.push_back(key[q].second);
            }
            b[c].tri_count = (int)(to - from);
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

// Breadth-first emission of wide nodes below a set of binary subtree roots. One Emitter is private to one thread:
// node and triangle indices are local and rebased when the pieces are concatenated.
struct alignas(128) Emitter {
    struct Item { int wide, bin, depth; };
    const std::vector<BNode>& b;
    const std::vector<int>& leaf_tris;
    int width;
    std::vector<NodeRec> recs;
    std::vector<int> tri_ids;
    std::vector<Item> queue; size_t head = 0;
    int depth = 0;
    bool bad_leaf = false;

    Emitter(const std::vector<BNode>& b_, const std::vector<int>& lt, int w) : b(b_), leaf_tris(lt), width(w) {}
    bool empty() const { return head == queue.size(); }
    size_t pending() const { return queue.size() - head; }
    void seed(int bin, int d) { queue.push_back({(int)recs.size(), bin, d}); recs.emplace_back(); }

    void step() {
        Item it = queue[head++];
        depth = std::max(depth, it.depth);
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
        double scale[3];
        for (int a = 0; a < 3; a++) { w.p[a] = lo[a]; w.e[a] = grid_exponent(hi[a] - lo[a]); scale[a] = std::ldexp(1.0, (int)w.e[a] - 127); }
        w.child_base = (uint32_t)recs.size();
        w.tri_base = (uint32_t)tri_ids.size();
        int tri_off = 0;
        for (int s = 0; s < 8; s++) {
            int kid = kid_in_slot[s];
            if (kid < 0) continue;
            const int k = w.n++;                              // position in the node's child arrays
            const BNode& x = b[kid];
            for (int a = 0; a < 3; a++) {
                double ql = std::floor(((double)x.lo[a] - (double)w.p[a]) / scale[a]);
                double qh = std::ceil(((double)x.hi[a] - (double)w.p[a]) / scale[a]);
                ql = std::min(std::max(ql, 0.0), 255.0); qh = std::min(std::max(qh, 0.0), 255.0);
                // the kernel evaluates p + scale * q in fp32: make sure that still encloses the fp32 box
                while (ql > 0.0 && w.p[a] + (float)scale[a] * (float)ql > x.lo[a]) ql -= 1.0;
                while (qh < 255.0 && w.p[a] + (float)scale[a] * (float)qh < x.hi[a]) qh += 1.0;
                w.qlo[a][k] = (uint8_t)ql; w.qhi[a][k] = (uint8_t)qh;
            }
            if (x.left >= 0) {
                w.imask |= (uint8_t)(1u << s);
                w.meta[k] = (uint8_t)((1u << 5) | (24 + s));
                seed(kid, it.depth + 1);                      // inner children: contiguous, slot order
            } else {
                int cnt = x.tri_count;                        // 1..3 after split_big_leaf
                if (cnt < 1 || cnt > 3) { bad_leaf = true; cnt = std::min(std::max(cnt, 1), 3); }
                w.meta[k] = (uint8_t)((((1u << cnt) - 1u) << 5) | (unsigned)tri_off);
                for (int q = 0; q < cnt; q++) tri_ids.push_back(leaf_tris[(size_t)x.tri_first + q]);
                tri_off += cnt;
            }
        }
        recs[it.wide] = w;
    }
};

// Node order: the top of the tree breadth-first until kFrontier subtrees are open, then each of those subtrees
// breadth-first in its own block (emitted by a pool of threads; the order does not depend on the thread count).
bool build_wide_bvh(const HostBvh& in, const float* tris_pre, int64_t n_tris, int width, WideBvh& out, std::string& msg) {
    (void)n_tris;
    if (width != 4 && width != 8) { msg = "wide BVH width must be 4 or 8"; return false; }
    if (in.nodes.size() < 24) { msg = "wide collapse needs a tree whose root is an inner node"; return false; }
    Collapser c{in, tris_pre, {}, {}, {}};
    c.b.reserve(in.nodes.size() / 8 + 16);
    c.leaf_tris.reserve(in.refs.size());
    int root = c.import_tree();
    for (int leaf : c.big_leaves) c.split_big_leaf(leaf);

    out.tri_ids.clear(); out.depth = 0; out.width = width; out.node_size = width == 8 ? 96 : 64;
    constexpr size_t kFrontier = 1024;
    Emitter top(c.b, c.leaf_tris, width);
    top.seed(root, 0);
    while (!top.empty() && top.pending() < kFrontier) top.step();
    const size_t n_sub = top.pending();
    std::vector<Emitter> sub;
    sub.reserve(n_sub);
    for (size_t k = 0; k < n_sub; k++) {
        sub.emplace_back(c.b, c.leaf_tris, width);
        sub.back().seed(top.queue[top.head + k].bin, top.queue[top.head + k].depth);
    }
    {
        std::atomic<size_t> next{0};
        auto work = [&]() { for (size_t k; (k = next.fetch_add(1)) < n_sub;) while (!sub[k].empty()) sub[k].step(); };
        int n_threads = (int)std::min<size_t>(std::max(1u, std::thread::hardware_concurrency()), std::max<size_t>(n_sub / 4, 1));
        std::vector<std::thread> pool;
        for (int t = 1; t < n_threads; t++) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
    }
    // ---- concatenate: sub[k].recs[0] is the node the top already allocated; the rest forms block k
    std::vector<size_t> node_off(n_sub + 1), tri_off(n_sub + 1);
    node_off[0] = top.recs.size(); tri_off[0] = top.tri_ids.size();
    bool bad = top.bad_leaf; out.depth = top.depth;
    for (size_t k = 0; k < n_sub; k++) {
        node_off[k + 1] = node_off[k] + sub[k].recs.size() - 1;
        tri_off[k + 1] = tri_off[k] + sub[k].tri_ids.size();
        bad = bad || sub[k].bad_leaf; out.depth = std::max(out.depth, sub[k].depth);
    }
    if (bad) { msg = "internal: leaf with an unsupported triangle count"; return false; }
    if (out.depth >= 30) { msg = "wide tree too deep for the traversal stack"; return false; }
    if (node_off[n_sub] >= (size_t)1 << 31 || tri_off[n_sub] >= (size_t)1 << 31) { msg = "wide tree exceeds 2^31 nodes or triangle slots"; return false; }
    out.n_nodes = (int64_t)node_off[n_sub];
    out.bytes.assign((size_t)out.n_nodes * out.node_size, 0);
    out.tri_ids.resize(tri_off[n_sub]);
    std::copy(top.tri_ids.begin(), top.tri_ids.end(), out.tri_ids.begin());
    {
        const size_t n_top = top.recs.size();
        std::atomic<size_t> next{0};
        auto work = [&]() {
            for (size_t k; (k = next.fetch_add(1)) < n_sub + 1;) {
                if (k == n_sub) {        // the top block; its last n_sub nodes (the open subtrees) are written below
                    for (size_t i = 0; i + n_sub < n_top; i++) serialise(top.recs[i], width, out.bytes.data() + i * out.node_size);
                    continue;
                }
                Emitter& e = sub[k];
                for (size_t i = 0; i < e.recs.size(); i++) {
                    NodeRec r = e.recs[i];
                    r.child_base = (uint32_t)(r.child_base - 1 + node_off[k]);
                    r.tri_base = (uint32_t)(r.tri_base + tri_off[k]);
                    size_t at = i == 0 ? (size_t)top.queue[top.head + k].wide : node_off[k] + i - 1;
                    serialise(r, width, out.bytes.data() + at * out.node_size);
                }
                std::copy(e.tri_ids.begin(), e.tri_ids.end(), out.tri_ids.begin() + tri_off[k]);
            }
        };
        int n_threads = (int)std::max(1u, std::thread::hardware_concurrency());
        std::vector<std::thread> pool;
        for (int t = 1; t < n_threads; t++) pool.emplace_back(work);
        work();
        for (auto& t : pool) t.join();
    }
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
