// pt_bvh_build.cpp -- this library's own SBVH builder (host, multi-threaded).
//
// Replaces BoundingVolumeHierarchy::BuildBinnedSpatial (src/core/BVH.cpp:2197-2293) and its helpers
// (FindBestObjectSplit :1619-1763, FindBestSpatialSplit :1793-1984, BuildSBVH :2128-2169, the BFS
// serialisation + BlockingOptimizedCache :2224-2258, :2443-2490). Same contract, different program:
//   * same SAH model and constants (traversal 1.23, intersection 5.33, BVH.cpp:1592-1593), same leaf
//     criterion (split iff 1.23 + SAH/area < 5.33 n, :2123-2126,:2158), same spatial-split gate
//     (overlap area > 1e-5 root area, :2135, :2018-2021) and reference un-splitting (Stich et al. 4.4);
//   * 16 bins over the centroid range instead of 8 over the node box, exact triangle/plane clipping for
//     spatial splits instead of chopped AABBs, subtree-parallel over host threads (the reference is
//     single-threaded), depth-bounded instead of exit(-1) at depth 48;
//   * output in the reference's serialised form (32-B nodes, sibling pairs adjacent, larger-area child
//     first, root = 0 / children = 1,2; reference list with negated end-of-leaf marker), so the oracle and
//     the reference-layout kernels traverse it unchanged. Triangle 0 cannot carry the marker (-0 == 0,
//     BVH.cpp:2117): leaves are emitted in ascending index order so it only ends a leaf when it is alone in
//     it, in which case that leaf runs on into the next one exactly as in the reference (harmless).
#include "pt_host.h"
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <mutex>
#include <thread>
#include <memory>
#include <chrono>
#include <cstdlib>

namespace pt {
namespace {

constexpr float kCostTraversal = 1.23f;
constexpr int kMaxBins = 64;
// tuning-run overrides (read per build): PT_BVH_CI = intersection cost of the SAH model, PT_BVH_BINS = bins per axis
static float kCostIntersection = 5.33f;
static int kBins = 16;
constexpr int kMaxDepth = 56;          // the traversal stack holds PT_MAX_TREE_DEPTH (96) entries
constexpr float kAlpha = 1e-5f;
static bool g_no_spatial = false;     // PT_BVH_NO_SPATIAL=1 (experiments): object splits only, to tell what the spatial splits are worth
constexpr int kParallelThreshold = 8192;

struct Box {
    float lo[3], hi[3];
    void reset() { for (int k = 0; k < 3; k++) { lo[k] = FLT_MAX; hi[k] = -FLT_MAX; } }
    void grow(const float* p) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], p[k]); hi[k] = std::max(hi[k], p[k]); } }
    void grow(const Box& b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
    void clip(const Box& b) { for (int k = 0; k < 3; k++) { lo[k] = std::max(lo[k], b.lo[k]); hi[k] = std::min(hi[k], b.hi[k]); } }
    bool valid() const { return lo[0] <= hi[0] && lo[1] <= hi[1] && lo[2] <= hi[2]; }
    float area() const {               // AABB::SurfaceArea, math/AABB.cpp:24-70
        if (!valid()) return 0.0f;
        float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2];
        return 2.0f * (x * (y + z) + y * z);
    }
};

struct Ref { Box box; int idx; };

struct Node {
    Box box;
    std::unique_ptr<Node> child[2];
    std::vector<int> leaf;             // triangle indices when leaf
};

struct Split {
    float sah = FLT_MAX;
    int axis = -1, bin = 0;
    bool spatial = false, sweep = false;
    float lo = 0.0f, k1 = 0.0f, width = 0.0f;   // binning transform of the winning axis
    Box lbox, rbox;
    int nl = 0, nr = 0;
};

inline int bin_of(float c, float lo, float k1) {
    int b = (int)(k1 * (c - lo));
    return std::min(std::max(b, 0), kBins - 1);
}
inline float centroid(const Ref& r, int ax) { return 0.5f * (r.box.lo[ax] + r.box.hi[ax]); }

struct Builder {
    const float* tris;                 // n x 20, pre-subtracted edges
    float root_area = 0.0f;
    int n_threads = 1;
    std::atomic<int> active_threads{1};

    void verts(int idx, float v[3][3]) const {
        const float* t = tris + 20 * (size_t)idx;
        for (int k = 0; k < 3; k++) { v[0][k] = t[k]; v[1][k] = t[k] + t[3 + k]; v[2][k] = t[k] + t[6 + k]; }
    }

    // exact split of a reference by the plane x[dim] = pos: clip the triangle, then intersect with the old box
    void split_ref(const Ref& r, int dim, float pos, Ref& L, Ref& R) const {
        float v[3][3]; verts(r.idx, v);
        L.idx = R.idx = r.idx;
        L.box.reset(); R.box.reset();
        for (int e = 0; e < 3; e++) {
            const float* a = v[e]; const float* b = v[(e + 1) % 3];
            float pa = a[dim], pb = b[dim];
            if (pa <= pos) L.box.grow(a);
            if (pa >= pos) R.box.grow(a);
            if ((pa < pos && pb > pos) || (pa > pos && pb < pos)) {
                float t = std::min(std::max((pos - pa) / (pb - pa), 0.0f), 1.0f);
                float q[3];
                for (int k = 0; k < 3; k++) q[k] = a[k] + (b[k] - a[k]) * t;
                q[dim] = pos;
                L.box.grow(q); R.box.grow(q);
            }
        }
        L.box.hi[dim] = std::min(L.box.hi[dim], pos);
        R.box.lo[dim] = std::max(R.box.lo[dim], pos);
        L.box.clip(r.box); R.box.clip(r.box);
    }

    void object_split_binned(const std::vector<Ref>& refs, Split& best) const {
        const int n = (int)refs.size();
        for (int ax = 0; ax < 3; ax++) {
            float clo = FLT_MAX, chi = -FLT_MAX;
            for (const Ref& r : refs) { float c = centroid(r, ax); clo = std::min(clo, c); chi = std::max(chi, c); }
            if (!(chi > clo)) continue;
            float k1 = kBins / (chi - clo);
            Box bb[kMaxBins]; int cnt[kMaxBins];
            for (int b = 0; b < kBins; b++) { bb[b].reset(); cnt[b] = 0; }
            for (const Ref& r : refs) { int b = bin_of(centroid(r, ax), clo, k1); bb[b].grow(r.box); cnt[b]++; }
            Box rb[kMaxBins]; Box acc; acc.reset();
            for (int b = kBins - 1; b >= 1; b--) { acc.grow(bb[b]); rb[b] = acc; }
            Box lb; lb.reset(); int nl = 0;
            for (int b = 1; b < kBins; b++) {
                lb.grow(bb[b - 1]); nl += cnt[b - 1];
                int nr = n - nl;
                if (nl == 0 || nr == 0) continue;
                float sah = kCostIntersection * (lb.area() * nl + rb[b].area() * nr);
                if (sah < best.sah) {
                    best.sah = sah; best.axis = ax; best.bin = b; best.spatial = false; best.sweep = false;
                    best.lo = clo; best.k1 = k1; best.lbox = lb; best.rbox = rb[b]; best.nl = nl; best.nr = nr;
                }
            }
        }
    }

    // full sweep for small nodes; on success refs are left sorted along best.axis and best.bin = split position
    void object_split_sweep(std::vector<Ref>& refs, Split& best) const {
        const int n = (int)refs.size();
        std::vector<Box> rb(n);
        for (int ax = 0; ax < 3; ax++) {
            std::sort(refs.begin(), refs.end(), [ax](const Ref& a, const Ref& b) {
                float ca = centroid(a, ax), cb = centroid(b, ax);
                return ca < cb || (ca == cb && a.idx < b.idx);
            });
            Box acc; acc.reset();
            for (int i = n - 1; i >= 1; i--) { acc.grow(refs[i].box); rb[i] = acc; }
            Box lb; lb.reset();
            for (int i = 1; i < n; i++) {
                lb.grow(refs[i - 1].box);
                float sah = kCostIntersection * (lb.area() * i + rb[i].area() * (n - i));
                if (sah < best.sah) {
                    best.sah = sah; best.axis = ax; best.bin = i; best.spatial = false; best.sweep = true;
                    best.lbox = lb; best.rbox = rb[i]; best.nl = i; best.nr = n - i;
                }
            }
        }
        if (best.sweep && best.axis != 2) {
            int ax = best.axis;
            std::sort(refs.begin(), refs.end(), [ax](const Ref& a, const Ref& b) {
                float ca = centroid(a, ax), cb = centroid(b, ax);
                return ca < cb || (ca == cb && a.idx < b.idx);
            });
        }
    }

    void spatial_split(const std::vector<Ref>& refs, const Box& nb, Split& best) const {
        for (int ax = 0; ax < 3; ax++) {
            float lo = nb.lo[ax], hi = nb.hi[ax];
            if (!(hi > lo)) continue;
            float k1 = kBins / (hi - lo), width = (hi - lo) / kBins;
            Box bb[kMaxBins]; int enter[kMaxBins], leave[kMaxBins];
            for (int b = 0; b < kBins; b++) { bb[b].reset(); enter[b] = leave[b] = 0; }
            for (const Ref& r : refs) {
                int b0 = bin_of(r.box.lo[ax], lo, k1), b1 = bin_of(r.box.hi[ax], lo, k1);
                enter[b0]++; leave[b1]++;
                Ref cur = r;
                for (int b = b0; b < b1; b++) {
                    Ref l, rr;
                    split_ref(cur, ax, lo + width * (float)(b + 1), l, rr);
                    if (l.box.valid()) bb[b].grow(l.box);
                    cur = rr;
                    if (!cur.box.valid()) break;
                }
                if (cur.box.valid()) bb[b1].grow(cur.box);
            }
            Box rb[kMaxBins]; int rc[kMaxBins]; Box acc; acc.reset(); int c = 0;
            for (int b = kBins - 1; b >= 1; b--) { acc.grow(bb[b]); c += leave[b]; rb[b] = acc; rc[b] = c; }
            Box lb; lb.reset(); int nl = 0;
            for (int b = 1; b < kBins; b++) {
                lb.grow(bb[b - 1]); nl += enter[b - 1];
                int nr = rc[b];
                if (nl == 0 || nr == 0) continue;
                float sah = kCostIntersection * (lb.area() * nl + rb[b].area() * nr);
                if (sah < best.sah) {
                    best.sah = sah; best.axis = ax; best.bin = b; best.spatial = true; best.sweep = false;
                    best.lo = lo; best.k1 = k1; best.width = width; best.lbox = lb; best.rbox = rb[b]; best.nl = nl; best.nr = nr;
                }
            }
        }
    }
    // distribute refs according to `s`; spatial splits may duplicate or un-split references
    void partition(std::vector<Ref>& refs, const Split& s, std::vector<Ref>& L, std::vector<Ref>& R) const {
        L.clear(); R.clear();
        if (!s.spatial) {
            L.reserve(s.nl); R.reserve(s.nr);
            if (s.sweep) {
                L.assign(refs.begin(), refs.begin() + s.bin);
                R.assign(refs.begin() + s.bin, refs.end());
            } else {
                for (const Ref& r : refs) (bin_of(centroid(r, s.axis), s.lo, s.k1) < s.bin ? L : R).push_back(r);
            }
            return;
        }
        const int ax = s.axis;
        const float lo = s.lo;
        const float pos = lo + s.width * (float)s.bin;
        Box bl = s.lbox, br = s.rbox;
        int nl = s.nl, nr = s.nr;
        for (const Ref& r : refs) {
            int b0 = bin_of(r.box.lo[ax], lo, s.k1), b1 = bin_of(r.box.hi[ax], lo, s.k1);
            if (b0 < s.bin && b1 >= s.bin) {
                // straddles the plane: split, or un-split when cheaper (Stich et al. 2009, section 4.4)
                Ref l, rr;
                split_ref(r, ax, pos, l, rr);
                Box ul = bl; ul.grow(r.box);
                Box ur = br; ur.grow(r.box);
                float csplit = bl.area() * nl + br.area() * nr;
                float cl = ul.area() * nl + br.area() * (nr - 1);
                float cr = bl.area() * (nl - 1) + ur.area() * nr;
                bool lv = l.box.valid(), rv = rr.box.valid();
                if ((!rv && lv) || (lv && cl < csplit && cl <= cr)) { L.push_back(r); bl = ul; nr--; }
                else if ((!lv && rv) || (rv && cr < csplit)) { R.push_back(r); br = ur; nl--; }
                else if (lv && rv) { L.push_back(l); R.push_back(rr); }
                else { L.push_back(r); }
            } else if (b0 < s.bin) L.push_back(r);
            else R.push_back(r);
        }
    }

    static Box bounds_of(const std::vector<Ref>& refs) { Box b; b.reset(); for (const Ref& r : refs) b.grow(r.box); return b; }

    void make_leaf(Node* node, const std::vector<Ref>& refs) const {
        node->leaf.reserve(refs.size());
        for (const Ref& r : refs) node->leaf.push_back(r.idx);
        // duplicates of the same triangle inside one leaf (possible after un-splitting) are useless
        std::sort(node->leaf.begin(), node->leaf.end());
        node->leaf.erase(std::unique(node->leaf.begin(), node->leaf.end()), node->leaf.end());
        // sorted ascending: triangle 0, if present with others, is first and never carries the end marker
    }

    void build_node(Node* node, std::vector<Ref>& refs, int depth) {
        const int n = (int)refs.size();
        node->box = bounds_of(refs);
        if (n <= 1 || depth >= kMaxDepth) { make_leaf(node, refs); return; }
        Split best;
        if (n > kBins) object_split_binned(refs, best);
        else object_split_sweep(refs, best);
        // spatial split only where the object split leaves overlapping children (BVH.cpp:2015-2021)
        bool try_spatial = !g_no_spatial;
        if (best.axis >= 0) {
            Box ov = best.lbox; ov.clip(best.rbox);
            try_spatial = !g_no_spatial && ov.valid() && ov.area() > kAlpha * root_area;
        }
        if (try_spatial) {
            Split sp = best;
            spatial_split(refs, node->box, sp);
            if (sp.spatial) best = sp;
        }
        float area = node->box.area();
        float cost = (best.axis >= 0 && area > 0.0f) ? kCostTraversal + best.sah / area : FLT_MAX;   // AdjustSAH, BVH.cpp:2123-2126
        if (!(cost < kCostIntersection * (float)n)) { make_leaf(node, refs); return; }
        std::vector<Ref> L, R;
        partition(refs, best, L, R);
        if (L.empty() || R.empty() || (best.spatial && (int)(L.size() + R.size()) > 2 * n)) {
            // degenerate outcome (all references un-split to one side): fall back to a median object split
            int ax = 0; float ext = -1.0f;
            for (int k = 0; k < 3; k++) if (node->box.hi[k] - node->box.lo[k] > ext) { ext = node->box.hi[k] - node->box.lo[k]; ax = k; }
            std::sort(refs.begin(), refs.end(), [ax](const Ref& a, const Ref& b) {
                float ca = centroid(a, ax), cb = centroid(b, ax);
                return ca < cb || (ca == cb && a.idx < b.idx);
            });
            L.assign(refs.begin(), refs.begin() + n / 2);
            R.assign(refs.begin() + n / 2, refs.end());
        }
        std::vector<Ref>().swap(refs);
        node->child[0].reset(new Node());
        node->child[1].reset(new Node());
        Node* c0 = node->child[0].get(); Node* c1 = node->child[1].get();
        bool spawn = (int)std::min(L.size(), R.size()) >= kParallelThreshold && active_threads.load() < n_threads;
        if (spawn && active_threads.fetch_add(1) < n_threads) {
            std::thread t([this, c1, &R, depth]() { build_node(c1, R, depth + 1); active_threads.fetch_sub(1); });
            build_node(c0, L, depth + 1);
            t.join();
        } else {
            if (spawn) active_threads.fetch_sub(1);
            build_node(c0, L, depth + 1);
            build_node(c1, R, depth + 1);
        }
    }
};

}  // namespace
namespace {

struct Flat {
    std::vector<float>& nodes;
    std::vector<int32_t>& refs;
    int max_depth = 0, n_leaves = 0;
    double cost = 0.0;

    void put(int index, const Node* nd, int32_t data0) {
        float* o = nodes.data() + 8 * (size_t)index;
        o[0] = nd->box.lo[0]; o[1] = nd->box.lo[1]; o[2] = nd->box.lo[2];
        std::memcpy(o + 3, &data0, 4);
        o[4] = nd->box.hi[0]; o[5] = nd->box.hi[1]; o[6] = nd->box.hi[2];
        o[7] = 0.0f;
    }
    int32_t emit_leaf(const Node* nd) {
        int32_t off = (int32_t)refs.size();
        for (int idx : nd->leaf) refs.push_back(idx);
        // The end-of-leaf marker is the sign of the last entry (BVH.cpp:2117) and triangle 0 cannot carry it (-0 == 0): the
        // reference's traversal then runs on into the next leaf. Same results, wasted triangle tests -- so triangle 0 is kept
        // out of the last slot whenever the leaf has another entry (a one-entry leaf of triangle 0 keeps the quirk).
        if (refs.back() == 0 && nd->leaf.size() > 1) std::swap(refs[refs.size() - 1], refs[refs.size() - 2]);
        refs.back() = -refs.back();
        n_leaves++;
        cost += kCostIntersection * nd->box.area() * (double)nd->leaf.size();
        return -off;
    }
};

}  // namespace

bool build_sbvh(const float* tris_pre, int64_t n, int n_threads, HostBvh& out, std::string& msg) {
    if (n < 2) { msg = "need at least 2 triangles (the root must be an inner node, BVH.glsl:627)"; return false; }
    if (const char* e = getenv("PT_BVH_CI")) kCostIntersection = (float)atof(e);
    if (const char* e = getenv("PT_BVH_NO_SPATIAL")) g_no_spatial = atoi(e) != 0;
    if (const char* e = getenv("PT_BVH_BINS")) kBins = std::min(std::max(atoi(e), 2), kMaxBins);
    Builder b;
    b.tris = tris_pre;
    b.n_threads = n_threads > 0 ? n_threads : (int)std::max(1u, std::thread::hardware_concurrency());
    std::vector<Ref> refs((size_t)n);
    for (int64_t i = 0; i < n; i++) {
        float v[3][3]; b.verts((int)i, v);
        refs[i].idx = (int)i; refs[i].box.reset();
        for (int k = 0; k < 3; k++) refs[i].box.grow(v[k]);
    }
    Node root;
    root.box = Builder::bounds_of(refs);
    b.root_area = root.box.area();
    b.build_node(&root, refs, 0);
    if (!root.child[0]) {
        // every triangle coincides: force one split so the root is an inner node
        std::vector<int> all = root.leaf; root.leaf.clear();
        root.child[0].reset(new Node()); root.child[1].reset(new Node());
        root.child[0]->box = root.box; root.child[1]->box = root.box;
        root.child[0]->leaf.assign(all.begin(), all.begin() + all.size() / 2);
        root.child[1]->leaf.assign(all.begin() + all.size() / 2, all.end());
        if (root.child[0]->leaf.empty()) root.child[0]->leaf = root.child[1]->leaf;
    }
    // count nodes
    size_t count = 0;
    { std::vector<const Node*> st{&root}; while (!st.empty()) { const Node* x = st.back(); st.pop_back(); count++; if (x->child[0]) { st.push_back(x->child[0].get()); st.push_back(x->child[1].get()); } } }
    out.nodes.assign(8 * count, 0.0f);
    out.refs.clear(); out.refs.reserve((size_t)n + n / 4);
    Flat f{out.nodes, out.refs};
    // Depth-first emission of sibling pairs: the pair of a node's children is placed right after the node's own
    // pair has been emitted and its first (larger-area) child's subtree follows immediately.
    struct Item { const Node* node; int index; int depth; };
    std::vector<Item> st;
    int next = 1;
    f.cost += kCostTraversal * root.box.area();
    st.push_back({&root, 0, 0});
    while (!st.empty()) {
        Item it = st.back(); st.pop_back();
        const Node* nd = it.node;
        f.max_depth = std::max(f.max_depth, it.depth);
        if (!nd->child[0]) { f.put(it.index, nd, f.emit_leaf(nd)); continue; }
        const Node* a = nd->child[0].get(); const Node* c = nd->child[1].get();
        if (a->box.area() < c->box.area()) std::swap(a, c);       // larger surface area first, BVH.cpp:2240-2247
        int first = next; next += 2;
        f.put(it.index, nd, first);
        if (it.index != 0) f.cost += kCostTraversal * nd->box.area();
        st.push_back({c, first + 1, it.depth + 1});
        st.push_back({a, first, it.depth + 1});                     // popped first
    }
    out.max_depth = f.max_depth;
    out.n_leaves = f.n_leaves;
    out.sah_cost = (float)(f.cost / root.box.area());
    if (out.max_depth >= 64) { msg = "tree too deep"; return false; }
    return true;
}
}  // namespace pt

extern "C" {
int pt_bvh_build_host(const float* tris, int64_t n, int32_t n_threads, float** nodes, int32_t* n_nodes,
                      int32_t** refs, int32_t* n_refs, pt_bvh_stats* stats) {
    if (!tris || !nodes || !n_nodes || !refs || !n_refs) return PT_ERR_INVALID;
    pt::HostBvh b; std::string msg;
    auto t0 = std::chrono::steady_clock::now();
    if (!pt::build_sbvh(tris, n, n_threads, b, msg)) return PT_ERR_LIMIT;
    auto t1 = std::chrono::steady_clock::now();
    *n_nodes = (int32_t)(b.nodes.size() / 8); *n_refs = (int32_t)b.refs.size();
    *nodes = (float*)std::malloc(b.nodes.size() * 4);
    *refs = (int32_t*)std::malloc(b.refs.size() * 4);
    std::memcpy(*nodes, b.nodes.data(), b.nodes.size() * 4);
    std::memcpy(*refs, b.refs.data(), b.refs.size() * 4);
    if (stats) {
        stats->n_nodes = *n_nodes; stats->n_refs = *n_refs; stats->n_leaves = b.n_leaves; stats->max_depth = b.max_depth;
        stats->sah_cost = b.sah_cost; stats->build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();
        stats->total_ms = stats->build_ms;
    }
    return PT_OK;
}
void pt_free(void* p) { std::free(p); }
}
