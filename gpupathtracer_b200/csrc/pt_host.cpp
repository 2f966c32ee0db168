// pt_host.cpp -- host-side parts of the hot path that involve no device work:
// scene assembly (Scene.cpp:263-337), camera image plane (math/Camera.cpp:6-22), tree validation,
// SAH cost (BVH.cpp:2174-2195) and the headless image writers that stand in for
// Renderer::SaveScreenshot (Renderer.cpp:697-705).
#include "pt_host.h"
#include "pt_layout.h"
#include <cmath>
#include <cstring>
#include <cstdio>
#include <algorithm>

namespace pt {

namespace {
struct H3 { float x, y, z; };
inline H3 h3(float x, float y, float z) { return H3{x, y, z}; }
inline H3 operator-(H3 a, H3 b) { return h3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline H3 operator+(H3 a, H3 b) { return h3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline H3 operator*(H3 a, float s) { return h3(a.x * s, a.y * s, a.z * s); }
inline H3 operator/(H3 a, float s) { return h3(a.x / s, a.y / s, a.z / s); }
inline H3 operator-(H3 a) { return h3(-a.x, -a.y, -a.z); }
inline float dot(H3 a, H3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline H3 cross(H3 a, H3 b) { return h3(a.y * b.z - b.y * a.z, a.z * b.x - b.z * a.x, a.x * b.y - b.x * a.y); }
inline H3 normalize(H3 a) { return a * (1.0f / std::sqrt(dot(a, a))); }
inline float dist(H3 a, H3 b) { H3 d = a - b; return std::sqrt(dot(d, d)); }
inline float half_area(const float* n) {   // AABB::SurfaceAreaHalf, math/AABB.cpp:62-70
    float sx = n[4] - n[0], sy = n[5] - n[1], sz = n[6] - n[2];
    return sx * (sy + sz) + sy * sz;
}
}  // namespace

int validate_bvh(const float* nodes, int n_nodes, const int32_t* refs, int n_refs, int64_t n_tris, std::string& msg,
                 int* n_leaves) {
    if (n_nodes < 3 || (n_nodes % 2) == 0) { msg = "node count must be odd and >= 3 (sibling pairs behind the root)"; return -1; }
    auto data0 = [&](int i) { int32_t d; std::memcpy(&d, nodes + 8 * (size_t)i + 3, 4); return d; };
    if (data0(0) != 1) { msg = "root must be an inner node whose children are nodes 1,2 (BVH.cpp:2224-2255)"; return -1; }
    std::vector<std::pair<int, int>> stack;   // node, depth
    stack.push_back({0, 0});
    int max_depth = 0, leaves = 0;
    size_t visited = 0;
    while (!stack.empty()) {
        auto [i, d] = stack.back(); stack.pop_back();
        if (++visited > (size_t)n_nodes) { msg = "cycle in node links"; return -1; }
        max_depth = std::max(max_depth, d);
        int32_t d0 = data0(i);
        if (d0 > 0) {
            if ((d0 & 1) == 0 || d0 + 1 >= n_nodes) { msg = "inner node links must be odd and in range"; return -1; }
            stack.push_back({d0, d + 1});
            stack.push_back({d0 + 1, d + 1});
        } else {
            leaves++;
            int off = -d0;
            if (off >= n_refs) { msg = "leaf offset past the reference list"; return -1; }
        }
    }
    if (n_tris >= 0) {
        for (int i = 0; i < n_refs; i++) {
            int64_t idx = refs[i] < 0 ? -(int64_t)refs[i] : refs[i];
            if (idx >= n_tris) { msg = "reference index out of range"; return -1; }
        }
    }
    if (n_leaves) *n_leaves = leaves;
    return max_depth;
}

float bvh_sah_cost(const float* nodes, int n_nodes, const int32_t* refs, int n_refs) {
    const float costTraversal = 1.23f, costIntersection = 5.33f;   // BVH.cpp:1592-1593
    double cost = 0.0;
    for (int i = 0; i < n_nodes; i++) {
        const float* nd = nodes + 8 * (size_t)i;
        int32_t d0; std::memcpy(&d0, nd + 3, 4);
        float sa = 2.0f * half_area(nd);
        if (d0 > 0) cost += costTraversal * sa;
        else {
            int cnt = 0;
            for (int k = -d0; k < n_refs; k++) { cnt++; if (refs[k] < 0) break; }
            cost += costIntersection * sa * cnt;
        }
    }
    return (float)(cost / (2.0 * half_area(nodes)));
}

bool write_ppm(const char* path, int w, int h, const uint8_t* rgb) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    std::fprintf(f, "P6\n%d %d\n255\n", w, h);
    // image row 0 is the bottom row of the reference's GL framebuffer; PPM is top-down
    for (int y = h - 1; y >= 0; y--) std::fwrite(rgb + 3 * (size_t)y * w, 1, 3 * (size_t)w, f);
    return std::fclose(f) == 0;
}
bool write_pfm(const char* path, int w, int h, const float* rgb) {
    FILE* f = std::fopen(path, "wb");
    if (!f) return false;
    std::fprintf(f, "PF\n%d %d\n-1.0\n", w, h);    // PFM is bottom-up, little endian
    std::fwrite(rgb, 4, 3 * (size_t)w * h, f);
    return std::fclose(f) == 0;
}

}  // namespace pt

using namespace pt;

extern "C" {

int pt_assemble_triangles(const float* verts, const uint32_t* mat_ids, int64_t n_tris, float* out) {
    if (!verts || !mat_ids || !out || n_tris < 0) return PT_ERR_INVALID;
    for (int64_t i = 0; i < n_tris; i++) {
        const float* v0 = verts + 24 * i; const float* v1 = v0 + 8; const float* v2 = v0 + 16;
        H3 p0 = h3(v0[0], v0[1], v0[2]), p1 = h3(v1[0], v1[1], v1[2]), p2 = h3(v2[0], v2[1], v2[2]);
        H3 v01 = p1 - p0, v02 = p2 - p0;
        H3 n = normalize(cross(normalize(v01), normalize(v02)));               // Scene.cpp:279-281
        H3 avg = (h3(v0[3], v0[4], v0[5]) + h3(v1[3], v1[4], v1[5]) + h3(v2[3], v2[4], v2[5])) / 3.0f;
        if (dot(n, avg) < 0.0f) n = -n;                                        // Scene.cpp:284-287
        float* o = out + 20 * i;
        o[0] = p0.x; o[1] = p0.y; o[2] = p0.z; o[3] = p1.x; o[4] = p1.y; o[5] = p1.z; o[6] = p2.x; o[7] = p2.y; o[8] = p2.z;
        o[9] = v0[6]; o[10] = v0[7]; o[11] = v1[6]; o[12] = v1[7]; o[13] = v2[6]; o[14] = v2[7];
        o[15] = n.x; o[16] = n.y; o[17] = n.z;
        std::memcpy(o + 18, mat_ids + i, 4);
        o[19] = 0.0f;
    }
    return PT_OK;
}

int pt_build_light_table(const float* tris_abs, int64_t n_tris, const pt_material* mats, int32_t n_mats,
                         pt_light* out, int32_t capacity, float* total_area) {
    if (!tris_abs || !mats || !total_area) return PT_ERR_INVALID;
    std::vector<pt_light> em;
    for (int64_t i = 0; i < n_tris; i++) {
        const float* t = tris_abs + 20 * i;
        uint32_t m; std::memcpy(&m, t + 18, 4);
        if ((int32_t)(m / 2) >= n_mats) return PT_ERR_INVALID;
        if (mats[m / 2].is_emissive == 1) {                                    // Scene.cpp:303-317
            H3 p0 = h3(t[0], t[1], t[2]), p1 = h3(t[3], t[4], t[5]), p2 = h3(t[6], t[7], t[8]);
            float a = dist(p0, p2), b = dist(p0, p1), c = dist(p2, p1);
            float s = (a + b + c) / 2;
            pt_light l; l.cumulative_area = sqrtf(s * (s - a) * (s - b) * (s - c)); l.tri_index = (uint32_t)i;
            em.push_back(l);
        }
    }
    std::sort(em.begin(), em.end(), [](const pt_light& l, const pt_light& r) { return l.cumulative_area < r.cumulative_area; });
    float total = 0.0f;
    for (pt_light& l : em) { total += l.cumulative_area; l.cumulative_area = total; }   // Scene.cpp:326-330
    *total_area = total;
    if (out) {
        if ((int64_t)em.size() > capacity) return PT_ERR_LIMIT;
        std::memcpy(out, em.data(), em.size() * sizeof(pt_light));
    }
    return (int)em.size();
}

void pt_presubtract_edges(float* tris, int64_t n) {
    for (int64_t i = 0; i < n; i++) {
        float* t = tris + 20 * i;
        t[3] = t[3] - t[0]; t[4] = t[4] - t[1]; t[5] = t[5] - t[2];
        t[6] = t[6] - t[0]; t[7] = t[7] - t[1]; t[8] = t[8] - t[2];
    }
}

void pt_camera_image_plane(float aspect, float fov, float focus, float aperture, const float pos[3], const float rot[3],
                           pt_camera* out) {
    float yaw = rot[0], pitch = rot[1];
    H3 d = h3(std::cos(pitch) * std::sin(yaw), std::sin(pitch), std::cos(pitch) * -std::cos(yaw));
    d = normalize(-d);
    float image_height = 2 * std::tan(fov / 2);
    float image_width = aspect * image_height;
    H3 u = normalize(cross(h3(0, 1, 0), d));
    H3 v = cross(d, u);
    H3 horizontal = u * image_width * focus;
    H3 vertical = v * image_height * focus;
    H3 ll = -horizontal / 2.0f - vertical / 2.0f - d * focus;
    auto st = [](float* o, H3 a) { o[0] = a.x; o[1] = a.y; o[2] = a.z; };
    st(out->lower_left, ll); st(out->horizontal, horizontal); st(out->vertical, vertical);
    st(out->position, h3(pos[0], pos[1], pos[2])); st(out->u, u); st(out->v, v);
    out->lens_radius = aperture / 2;
}

}  // extern "C"

extern "C" int pt_bvh_optimize_host(float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs, int32_t iterations, float batch_fraction,
                                    float* sah_before, float* sah_after) {
    if (!nodes || !refs || n_nodes < 3 || iterations < 1 || !(batch_fraction > 0.0f) || batch_fraction > 1.0f) return PT_ERR_INVALID;
    const int moved = pt::optimize_bvh_reinsertion(nodes, n_nodes, iterations, batch_fraction, PT_MAX_TREE_DEPTH, sah_before, sah_after, refs, n_refs);
    return moved < 0 ? PT_ERR_LIMIT : moved;
}
extern "C" float pt_bvh_sah_cost(const float* nodes, int32_t n_nodes, const int32_t* refs, int32_t n_refs) {
    if (!nodes || !refs || n_nodes < 1) return -1.0f;
    return pt::bvh_sah_cost(nodes, n_nodes, refs, n_refs);
}
