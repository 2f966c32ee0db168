// pt_bvh_opt.cpp -- reinsertion optimisation of a finished BVH (host; pt_bvh_optimize_host / pt_optimize_bvh).
//
// The reference carries a ReinsertionOptimize pass that it never calls (BVH.cpp, listed as dead code in SURVEY.md 2); this is the
// idea done properly (Bittner, Hapala, Havran: "Fast insertion-based optimization of bounding volume hierarchies", CGF 2013, in
// the batched form of Meister & Bittner 2018): take the nodes with the largest surface, and for each look -- branch and bound over
// the whole tree -- for the place where hanging its subtree costs the least surface area summed over all inner nodes; move it if
// that beats where it is. Topology changes, leaves and the reference list do not: the tree keeps the reference's serialised form
// (children of a node adjacent at an odd index, larger surface first, leaf = -(offset into the reference list)) and is re-laid out
// depth first at the end. What it buys is fewer node visits per ray, which is the one thing the traversal kernels' binding limit
// (L1 look-ups per visited pair record) scales with.
#include "pt_host.h"
#include <algorithm>
#include <cfloat>
#include <cstring>
#include <numeric>
#include <vector>

namespace pt {
namespace {

struct Box3 {
    float lo[3], hi[3];
    float half_area() const { float x = hi[0] - lo[0], y = hi[1] - lo[1], z = hi[2] - lo[2]; return x * (y + z) + y * z; }
    void grow(const Box3& b) { for (int k = 0; k < 3; k++) { lo[k] = std::min(lo[k], b.lo[k]); hi[k] = std::max(hi[k], b.hi[k]); } }
};

// Working form: node i = {box, link}; link > 0: index of its first child (odd; the second is link + 1), link <= 0: leaf payload.
struct Tree {
    std::vector<Box3> box;
    std::vector<int32_t> link;
    std::vector<int32_t> parent;
    int n = 0;
    bool leaf(int i) const { return link[i] <= 0; }
    static int sibling(int i) { return (i & 1) ? i + 1 : i - 1; }
    void refit_from(int i) {                      // boxes of i and its ancestors from their children
        for (; i >= 0; i = parent[i]) {
            if (leaf(i)) continue;
            Box3 b = box[link[i]]; b.grow(box[link[i] + 1]);
            box[i] = b;
        }
    }
};

struct Candidate { int to = -1; float gain = 0.0f; };

// Best place to hang the subtree `in`: the node `to` that gets a new parent shared with `in`. Gain = surface removed from the inner
// nodes by taking `in` out (its parent disappears, the ancestors shrink) minus surface added by the new parent and the grown ancestors.
Candidate find_reinsertion(const Tree& t, int in, std::vector<std::pair<float, int>>& stack) {
    Candidate best;
    const float node_area = t.box[in].half_area();
    const int par = t.parent[in];
    float area_diff = t.box[par].half_area();     // the parent node goes away
    int sib = Tree::sibling(in);
    Box3 pivot_box = t.box[sib];
    int pivot = par;
    do {
        stack.clear();
        stack.emplace_back(area_diff, sib);
        while (!stack.empty()) {
            auto top = stack.back(); stack.pop_back();
            if (top.first - node_area <= best.gain) continue;        // even a free merge could not beat the best so far
            Box3 merged = t.box[top.second]; merged.grow(t.box[in]);
            const float reinsert = top.first - merged.half_area();
            if (reinsert > best.gain) { best.gain = reinsert; best.to = top.second; }
            if (!t.leaf(top.second)) {
                const float child = reinsert + t.box[top.second].half_area();
                stack.emplace_back(child, t.link[top.second]);
                stack.emplace_back(child, t.link[top.second] + 1);
            }
        }
        if (pivot != par) {                                            // ancestors shrink once `in` is gone
            pivot_box.grow(t.box[sib]);
            area_diff += t.box[pivot].half_area() - pivot_box.half_area();
        }
        if (pivot == 0) break;
        sib = Tree::sibling(pivot);
        pivot = t.parent[pivot];
    } while (true);
    if (best.to == Tree::sibling(in) || best.to == par) best.to = -1;   // where it already is
    return best;
}

// Move the subtree `in` next to `to`. The pair slot (in, sibling) is reused for the new pair (to, in); the sibling moves up into
// the parent's slot; `to`'s old slot becomes the new inner node.
void reinsert(Tree& t, int in, int to) {
    const int sib = Tree::sibling(in), par = t.parent[in];
    const Box3 sib_box = t.box[sib], in_box = t.box[in], to_box = t.box[to];
    const int32_t sib_link = t.link[sib], in_link = t.link[in], to_link = t.link[to];
    const int first = std::min(in, sib);                                // the pair slot that is being recycled
    // sibling -> parent's slot
    t.box[par] = sib_box; t.link[par] = sib_link;
    if (sib_link > 0) { t.parent[sib_link] = par; t.parent[sib_link + 1] = par; }
    // `to` may have been the sibling itself (then it now lives in `par`) -- excluded by the caller; but it may be a node whose index changed
    const int to_now = to;
    // new pair: (to, in) in the recycled slot
    t.box[first] = to_box; t.link[first] = to_link;
    if (to_link > 0) { t.parent[to_link] = first; t.parent[to_link + 1] = first; }
    t.box[first + 1] = in_box; t.link[first + 1] = in_link;
    if (in_link > 0) { t.parent[in_link] = first + 1; t.parent[in_link + 1] = first + 1; }
    // `to`'s old slot becomes their parent
    t.link[to_now] = first;
    t.parent[first] = to_now; t.parent[first + 1] = to_now;
    t.refit_from(par);
    t.refit_from(to_now);
}

int depth_of(const Tree& t) {
    int deepest = 0;
    std::vector<std::pair<int, int>> st{{0, 0}};
    while (!st.empty()) {
        auto [i, d] = st.back(); st.pop_back();
        deepest = std::max(deepest, d);
        if (!t.leaf(i)) { st.emplace_back(t.link[i], d + 1); st.emplace_back(t.link[i] + 1, d + 1); }
    }
    return deepest;
}

}  // namespace

// nodes: n x 8 floats in the reference's serialised form, rewritten in place. Returns the number of subtrees moved, or -1 when the
// result would be deeper than max_depth (the input is left as it was).
int optimize_bvh_reinsertion(float* nodes, int n_nodes, int iterations, float batch_fraction, int max_depth, float* sah_before, float* sah_after,
                             const int32_t* refs, int n_refs) {
    if (n_nodes < 7) return 0;
    Tree t; t.n = n_nodes;
    t.box.resize(n_nodes); t.link.resize(n_nodes); t.parent.assign(n_nodes, -1);
    for (int i = 0; i < n_nodes; i++) {
        const float* nd = nodes + 8 * (size_t)i;
        for (int k = 0; k < 3; k++) { t.box[i].lo[k] = nd[k]; t.box[i].hi[k] = nd[4 + k]; }
        std::memcpy(&t.link[i], nd + 3, 4);
    }
    for (int i = 0; i < n_nodes; i++) if (t.link[i] > 0) { t.parent[t.link[i]] = i; t.parent[t.link[i] + 1] = i; }
    if (sah_before) *sah_before = bvh_sah_cost(nodes, n_nodes, refs, n_refs);
    std::vector<int> order;
    std::vector<std::pair<float, int>> stack;
    int moved = 0;
    for (int it = 0; it < iterations; it++) {
        // the nodes with the largest surface first (not the root, not its children: they have no grandparent to inherit the sibling)
        order.clear();
        for (int i = 3; i < n_nodes; i++) order.push_back(i);
        const size_t batch = std::max<size_t>(1, (size_t)((double)order.size() * batch_fraction));
        std::partial_sort(order.begin(), order.begin() + std::min(batch, order.size()), order.end(),
                          [&](int a, int b) { float x = t.box[a].half_area(), y = t.box[b].half_area(); return x > y || (x == y && a < b); });
        order.resize(std::min(batch, order.size()));
        int moved_now = 0;
        for (int in : order) {
            if (t.parent[in] <= 0) continue;                              // became a child of the root meanwhile
            Candidate c = find_reinsertion(t, in, stack);
            if (c.to < 0 || !(c.gain > 0.0f)) continue;
            // `to` must not lie inside the subtree of `in`, nor be an ancestor slot that the move recycles
            bool inside = false;
            for (int a = c.to; a >= 0; a = t.parent[a]) if (a == in) { inside = true; break; }
            if (inside) continue;
            reinsert(t, in, c.to);
            moved_now++;
        }
        moved += moved_now;
        if (moved_now == 0) break;
    }
    if (depth_of(t) >= max_depth) return -1;
    // re-emit depth first, larger child first (BVH.cpp:2240-2247), the pair of a node's children right behind the node's own pair
    std::vector<float> out((size_t)n_nodes * 8, 0.0f);
    auto put = [&](int at, const Box3& b, int32_t link) {
        float* o = out.data() + 8 * (size_t)at;
        o[0] = b.lo[0]; o[1] = b.lo[1]; o[2] = b.lo[2]; std::memcpy(o + 3, &link, 4);
        o[4] = b.hi[0]; o[5] = b.hi[1]; o[6] = b.hi[2]; o[7] = 0.0f;
    };
    struct Item { int node, at; };
    std::vector<Item> st{{0, 0}};
    int next = 1;
    while (!st.empty()) {
        Item x = st.back(); st.pop_back();
        if (t.leaf(x.node)) { put(x.at, t.box[x.node], t.link[x.node]); continue; }
        int a = t.link[x.node], b = a + 1;
        if (t.box[a].half_area() < t.box[b].half_area()) std::swap(a, b);
        const int first = next; next += 2;
        put(x.at, t.box[x.node], first);
        st.push_back({b, first + 1});
        st.push_back({a, first});
    }
    std::memcpy(nodes, out.data(), out.size() * 4);
    if (sah_after) *sah_after = bvh_sah_cost(nodes, n_nodes, refs, n_refs);
    return moved;
}

}  // namespace pt
