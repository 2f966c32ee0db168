// pt_obj.cpp -- the file half of Scene::LoadScene: Wavefront .obj/.mtl and scene.txt (host only, no device work).
//
// Replaces LoadOBJ (src/core/Scene.cpp:120-231), CreateMatInstance (:86-118), the sky material (:236-241), the
// scene.txt reader of main() (src/Program.cpp:70-84) and the "GENERATE COLOR" branch of LoadEnvironmnet
// (src/core/Renderer.cpp:236-260). The reference delegates parsing to tinyobjloader (extern/tinyobjloader, pinned at
// a1e8bad); this is an own parser that follows tinyobjloader's PUBLISHED behaviour for the features the reference
// consumes, so that the same file gives the same triangles in the same order:
//   * v / vn / vt / f (v, v/vt, v//vn, v/vt/vn; negative = relative to the elements read so far);
//   * a new shape starts at every `g` / `o` line that follows faces; `usemtl` switches the per-face material without
//     starting a shape; the REFERENCE then gives every triangle of a shape the material of the shape's FIRST face
//     (Scene.cpp:182, `material_ids[0]`) -- reproduced;
//   * triangles pass through, quads are cut along the SHORTER diagonal (tiny_obj_loader.h:1429-1527), larger polygons
//     are ear-clipped in the plane of their dominant axes;
//   * missing normal = (0,1,0), missing texcoord = (0,0) (Scene.cpp:190-206);
//   * .mtl: newmtl, Kd, Ks, Ke, Ns, illum, map_Kd; defaults Kd = Ks = Ke = 0, Ns = 1 (tinyobj InitMaterial);
//     file looked up next to the .obj (tiny_obj_loader.h:3289-3301).
// Material rules are the reference's: matId = 2 * (1 + index in the .mtl order) (Scene.cpp:153-158), sky = 0,
// roughness = sqrt(2 / (Ns + 2)) (:160-161), metallic = max(Ks) > 0.3 (:164-170), isEmissive = sum(Ke) > 1e-5 (:109).
#include "pt_host.h"
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <map>
#include <sstream>

namespace pt {
namespace {

struct MtlRec {
    std::string name, map_kd;
    float kd[3] = {0, 0, 0}, ks[3] = {0, 0, 0}, ke[3] = {0, 0, 0};
    float ns = 1.0f;
    int illum = 0;
};

struct VIdx { int v, vt, vn; };

inline bool is_space(char c) { return c == ' ' || c == '\t'; }
inline const char* skip_ws(const char* p) { while (is_space(*p)) p++; return p; }
inline bool at_eol(const char* p) { return *p == '\0' || *p == '\r' || *p == '\n'; }

// Decimal reader with tinyobjloader's grammar AND arithmetic (tiny_obj_loader.h:866-1006, "tryParseDouble"): digits are
// accumulated in double in reading order (integer part by *10, fraction digit k by 10^-k from a table for k < 8),
// the exponent is applied as ldexp(m * 5^e, e). Not correctly rounded like strtod -- but it is what the reference's
// vertex data goes through, and the last float bit of a coordinate decides which side of a box plane a ray passes.
// "nan"/"inf"/hex are not numbers here: the word is skipped and the default is returned.
inline bool read_decimal(const char* s, const char* s_end, double& result) {
    if (s >= s_end) return false;
    double mantissa = 0.0; int exponent = 0; char sign = '+', exp_sign = '+';
    const char* c = s; int read = 0; bool leading_dot = false;
    auto digit = [](char ch) { return ch >= '0' && ch <= '9'; };
    if (*c == '+' || *c == '-') { sign = *c; c++; if (c != s_end && *c == '.') leading_dot = true; }
    else if (digit(*c)) {}
    else if (*c == '.') leading_dot = true;
    else return false;
    if (!leading_dot) {
        while (c != s_end && digit(*c)) { mantissa *= 10; mantissa += (int)(*c - '0'); c++; read++; }
        if (read == 0) return false;
    }
    if (c != s_end) {
        bool more = true;
        if (*c == '.') {
            static const double lut[] = {1.0, 0.1, 0.01, 0.001, 0.0001, 0.00001, 0.000001, 0.0000001};
            c++; read = 1;
            while (c != s_end && digit(*c)) { mantissa += (int)(*c - '0') * (read < 8 ? lut[read] : std::pow(10.0, -read)); read++; c++; }
        } else if (*c == 'e' || *c == 'E') {
        } else more = false;
        if (more && c != s_end && (*c == 'e' || *c == 'E')) {
            c++;
            if (c != s_end && (*c == '+' || *c == '-')) { exp_sign = *c; c++; }
            else if (c != s_end && digit(*c)) {}
            else return false;
            read = 0;
            while (c != s_end && digit(*c)) {
                if (exponent > 2147483647 / 10) return false;
                exponent = exponent * 10 + (int)(*c - '0'); c++; read++;
            }
            if (exp_sign == '-') exponent = -exponent;
            if (read == 0) return false;
        }
    }
    result = (sign == '+' ? 1 : -1) * (exponent ? std::ldexp(mantissa * std::pow(5.0, exponent), exponent) : mantissa);
    return true;
}
// one whitespace-delimited word as a real; the word is consumed whether or not it is a number (tinyobj parseReal)
inline float parse_float(const char*& p, float def = 0.0f) {
    p = skip_ws(p);
    const char* end = p;
    while (*end && *end != ' ' && *end != '\t' && *end != '\r') end++;
    double val = def;
    read_decimal(p, end, val);
    p = end;
    return (float)val;
}

// .obj indices are 1-based; negative = relative to the `n` elements read so far; 0 is invalid (-> -1 here)
inline bool fix_index(int idx, int n, int& out) {
    if (idx > 0) { out = idx - 1; return true; }
    if (idx < 0) { out = n + idx; return out >= 0; }
    return false;
}

bool parse_triple(const char*& p, int nv, int nvn, int nvt, VIdx& out) {
    out.v = out.vt = out.vn = -1;
    char* end = nullptr;
    long a = std::strtol(p, &end, 10);
    if (end == p) return false;
    if (!fix_index((int)a, nv, out.v)) return false;
    p = end;
    if (*p != '/') return true;
    p++;
    if (*p == '/') {                           // v//vn
        p++;
        long c = std::strtol(p, &end, 10);
        if (end == p) return true;
        if (!fix_index((int)c, nvn, out.vn)) return false;
        p = end;
        return true;
    }
    long b = std::strtol(p, &end, 10);         // v/vt[/vn]
    if (end != p) { if (!fix_index((int)b, nvt, out.vt)) return false; p = end; }
    if (*p != '/') return true;
    p++;
    long c = std::strtol(p, &end, 10);
    if (end != p) { if (!fix_index((int)c, nvn, out.vn)) return false; p = end; }
    return true;
}

std::string rest_of_line(const char* p) {
    p = skip_ws(p);
    std::string s(p);
    while (!s.empty() && (s.back() == '\r' || s.back() == '\n' || is_space(s.back()))) s.pop_back();
    return s;
}

// tinyobj ParseTextureNameAndOption (tiny_obj_loader.h:1222-1306): options take a FIXED number of words (-o/-s/-t three,
// -mm two, the rest one) and the first word that is not an option starts the file name, which runs to the end of the
// line. ("-o 0.1 tex.jpg" therefore loses its file name to -o: as published.)
std::string texture_name(const char* p) {
    static const struct { const char* opt; int words; } kOpts[] = {
        {"-blendu", 1}, {"-blendv", 1}, {"-clamp", 1}, {"-boost", 1}, {"-bm", 1}, {"-o", 3}, {"-s", 3}, {"-t", 3},
        {"-type", 1}, {"-texres", 1}, {"-imfchan", 1}, {"-mm", 2}, {"-colorspace", 1}};
    std::string name;
    while (!at_eol(p)) {
        p = skip_ws(p);
        bool is_opt = false;
        for (const auto& o : kOpts) {
            size_t n = std::strlen(o.opt);
            if (!std::strncmp(p, o.opt, n) && is_space(p[n])) {
                p += n;
                for (int k = 0; k < o.words; k++) { p = skip_ws(p); while (!at_eol(p) && !is_space(*p)) p++; }
                is_opt = true;
                break;
            }
        }
        if (is_opt) continue;
        name = p;
        break;
    }
    // LoadMtl trims trailing blanks from every line before it is interpreted
    while (!name.empty() && (name.back() == '\r' || name.back() == '\n' || is_space(name.back()))) name.pop_back();
    return name;
}

// SplitString(s, ' ', '\\') of tinyobj (:1886-1909): blanks separate names unless escaped by a backslash
std::vector<std::string> split_escaped(const std::string& s) {
    std::vector<std::string> out; std::string tok; bool esc = false;
    for (char ch : s) {
        if (esc) esc = false;
        else if (ch == '\\') { esc = true; continue; }
        else if (ch == ' ') { if (!tok.empty()) out.push_back(tok); tok.clear(); continue; }
        tok += ch;
    }
    out.push_back(tok);
    return out;
}

bool load_mtl(const std::string& path, std::vector<MtlRec>& mats, std::map<std::string, int>& by_name, std::string& warn) {
    std::ifstream f(path);
    if (!f) { warn += "material file '" + path + "' not found; faces keep the sky material\n"; return false; }
    std::string line;
    MtlRec cur;
    // tinyobj LoadMtl (tiny_obj_loader.h:1982-1987, :2306-2310): a record is flushed at the next `newmtl` when it has a
    // name, and unconditionally at the end of the file; std::map::insert keeps the FIRST record of a repeated name
    auto flush = [&]() {
        by_name.insert(std::make_pair(cur.name, (int)mats.size()));
        mats.push_back(cur);
    };
    while (std::getline(f, line)) {
        const char* p = skip_ws(line.c_str());
        if (at_eol(p) || *p == '#') continue;
        if (!std::strncmp(p, "newmtl", 6) && is_space(p[6])) {
            if (!cur.name.empty()) flush();
            cur = MtlRec();
            cur.name = rest_of_line(p + 7);
        } else if (p[0] == 'K' && p[1] == 'd' && is_space(p[2])) { p += 2; for (int k = 0; k < 3; k++) cur.kd[k] = parse_float(p); }
        else if (p[0] == 'K' && p[1] == 's' && is_space(p[2])) { p += 2; for (int k = 0; k < 3; k++) cur.ks[k] = parse_float(p); }
        else if (p[0] == 'K' && p[1] == 'e' && is_space(p[2])) { p += 2; for (int k = 0; k < 3; k++) cur.ke[k] = parse_float(p); }
        else if (p[0] == 'N' && p[1] == 's' && is_space(p[2])) { p += 2; cur.ns = parse_float(p); }
        else if (!std::strncmp(p, "illum", 5) && is_space(p[5])) { p += 5; cur.illum = (int)parse_float(p); }
        else if (!std::strncmp(p, "map_Kd", 6) && is_space(p[6])) { cur.map_kd = texture_name(p + 7); }
    }
    flush();
    return true;
}

}  // namespace

// ------------------------------------------------------------------------------------------------ the parsed file
struct SceneFile {
    std::vector<float> verts;            // n_tris x 3 x {pos.xyz, normal.xyz, uv.xy}
    std::vector<uint32_t> mat_ids;       // per triangle: matId of its first vertex (Scene.cpp:289)
    std::vector<pt_material> materials;  // [0] = sky
    std::vector<std::string> tex_paths;  // per material: folder + map_Kd, "" when none
    std::string warnings;
    int n_shapes = 0;
};

namespace {

struct Face { std::vector<VIdx> idx; };

struct ObjParser {
    std::vector<float> v, vn, vt;
    std::vector<MtlRec> mtl;
    std::map<std::string, int> mtl_by_name;
    std::string warn;
    // the shape being assembled: triangulated corner indices + per-triangle tinyobj material id
    std::vector<VIdx> shape_idx;
    std::vector<int> shape_mat;
    std::vector<Face> group;             // faces since the last flush (tinyobj PrimGroup::faceGroup)
    int material = -1;
    SceneFile* out = nullptr;

    const float* pos(int i) const { return v.data() + 3 * (size_t)i; }
    bool pos_ok(int i) const { return i >= 0 && 3 * (size_t)i + 2 < v.size(); }

    void emit_tri(const VIdx& a, const VIdx& b, const VIdx& c) {
        shape_idx.push_back(a); shape_idx.push_back(b); shape_idx.push_back(c);
        shape_mat.push_back(material);
    }

    // ear clipping of a polygon with more than 4 corners, in the plane of its dominant axes
    void clip_polygon(const Face& f) {
        const size_t n = f.idx.size();
        size_t axes[2] = {1, 2};
        for (size_t k = 0; k < n; k++) {
            int i0 = f.idx[k].v, i1 = f.idx[(k + 1) % n].v, i2 = f.idx[(k + 2) % n].v;
            if (!pos_ok(i0) || !pos_ok(i1) || !pos_ok(i2)) continue;
            const float *a = pos(i0), *b = pos(i1), *c = pos(i2);
            float e0[3] = {b[0] - a[0], b[1] - a[1], b[2] - a[2]}, e1[3] = {c[0] - b[0], c[1] - b[1], c[2] - b[2]};
            float cx = std::fabs(e0[1] * e1[2] - e0[2] * e1[1]), cy = std::fabs(e0[2] * e1[0] - e0[0] * e1[2]),
                  cz = std::fabs(e0[0] * e1[1] - e0[1] * e1[0]);
            const float eps = 1.1920929e-7f;
            if (cx > eps || cy > eps || cz > eps) {
                if (!(cx > cy && cx > cz)) { axes[0] = 0; if (cz > cx && cz > cy) axes[1] = 1; }
                break;
            }
        }
        // tinyobjloader's built-in ear clipping (tiny_obj_loader.h:1640-1795), restated: walk a "guess" corner around the
        // remaining polygon; the corner triple is an ear when its turn has the sign of tinyobj's per-triple area term
        // (v0 x v1, taken about the origin -- as published, not the polygon area) and no other remaining corner lies
        // inside it (crossing-number test); the middle corner is then removed. The iteration budget is reset whenever
        // a corner is removed.
        std::vector<VIdx> rest = f.idx;
        auto coord = [&](const VIdx& i, int ax, float& out) {
            size_t vi = (size_t)i.v;
            if (i.v < 0 || vi * 3 + axes[0] >= v.size() || vi * 3 + axes[1] >= v.size()) { out = 0.0f; return false; }
            out = v[vi * 3 + axes[ax]]; return true;
        };
        auto inside3 = [](const float* vx, const float* vy, float tx, float ty) {
            bool c = false;
            for (int i = 0, j = 2; i < 3; j = i++)
                if (((vy[i] > ty) != (vy[j] > ty)) && (tx < (vx[j] - vx[i]) * (ty - vy[i]) / (vy[j] - vy[i]) + vx[i])) c = !c;
            return c;
        };
        size_t guess = 0, budget = rest.size(), prev = rest.size();
        while (rest.size() > 3 && budget > 0) {
            size_t m = rest.size();
            if (guess >= m) guess -= m;
            if (prev != m) { prev = m; budget = m; } else budget--;
            VIdx ind[3]; float vx[3], vy[3];
            for (int k = 0; k < 3; k++) { ind[k] = rest[(guess + k) % m]; coord(ind[k], 0, vx[k]); coord(ind[k], 1, vy[k]); }
            float e0x = vx[1] - vx[0], e0y = vy[1] - vy[0], e1x = vx[2] - vx[1], e1y = vy[2] - vy[1];
            float cross = e0x * e1y - e0y * e1x;
            float area = (vx[0] * vy[1] - vy[0] * vx[1]) * 0.5f;
            if (cross * area < 0.0f) { guess += 1; continue; }
            bool overlap = false;
            for (size_t other = 3; other < m; other++) {
                size_t idx = (guess + other) % m;
                float tx, ty;
                if (!coord(rest[idx], 0, tx)) continue;
                coord(rest[idx], 1, ty);
                if (inside3(vx, vy, tx, ty)) { overlap = true; break; }
            }
            if (overlap) { guess += 1; continue; }
            emit_tri(ind[0], ind[1], ind[2]);
            rest.erase(rest.begin() + (long)((guess + 1) % m));
        }
        if (rest.size() == 3) emit_tri(rest[0], rest[1], rest[2]);
    }

    // tinyobj exportGroupsToShape: triangulate the pending faces into the current shape with the CURRENT material
    void flush_group() {
        for (const Face& f : group) {
            const size_t n = f.idx.size();
            if (n < 3) { warn += "Degenerated face found\n"; continue; }
            if (n == 3) { emit_tri(f.idx[0], f.idx[1], f.idx[2]); continue; }
            if (n == 4) {
                bool ok = true;
                for (int k = 0; k < 4; k++) ok = ok && pos_ok(f.idx[k].v);
                if (!ok) { warn += "Face with invalid vertex index found.\n"; continue; }
                const float *p0 = pos(f.idx[0].v), *p1 = pos(f.idx[1].v), *p2 = pos(f.idx[2].v), *p3 = pos(f.idx[3].v);
                float e02[3] = {p2[0] - p0[0], p2[1] - p0[1], p2[2] - p0[2]}, e13[3] = {p3[0] - p1[0], p3[1] - p1[1], p3[2] - p1[2]};
                float s02 = e02[0] * e02[0] + e02[1] * e02[1] + e02[2] * e02[2];
                float s13 = e13[0] * e13[0] + e13[1] * e13[1] + e13[2] * e13[2];
                if (s02 < s13) { emit_tri(f.idx[0], f.idx[1], f.idx[2]); emit_tri(f.idx[0], f.idx[2], f.idx[3]); }
                else { emit_tri(f.idx[0], f.idx[1], f.idx[3]); emit_tri(f.idx[1], f.idx[2], f.idx[3]); }
                continue;
            }
            clip_polygon(f);
        }
        group.clear();
    }

    // LoadOBJ's vertex loop (Scene.cpp:178-214) for one finished shape
    void finish_shape() {
        flush_group();
        if (shape_idx.empty()) return;
        // every triangle of the shape gets the material of the shape's FIRST face (Scene.cpp:182)
        int first = shape_mat[0];
        uint32_t mat_id = first < 0 ? 0u : 2u * (uint32_t)(first + 1);
        for (size_t t = 0; t < shape_idx.size() / 3; t++) {
            for (int c = 0; c < 3; c++) {
                const VIdx& i = shape_idx[3 * t + c];
                float rec[8] = {0, 0, 0, 0, 1, 0, 0, 0};
                if (pos_ok(i.v)) { rec[0] = pos(i.v)[0]; rec[1] = pos(i.v)[1]; rec[2] = pos(i.v)[2]; }
                if (i.vn >= 0 && 3 * (size_t)i.vn + 2 < vn.size()) { rec[3] = vn[3 * i.vn]; rec[4] = vn[3 * i.vn + 1]; rec[5] = vn[3 * i.vn + 2]; }
                if (i.vt >= 0 && 2 * (size_t)i.vt + 1 < vt.size()) { rec[6] = vt[2 * i.vt]; rec[7] = vt[2 * i.vt + 1]; }
                out->verts.insert(out->verts.end(), rec, rec + 8);
            }
            out->mat_ids.push_back(mat_id);
        }
        out->n_shapes++;
        shape_idx.clear(); shape_mat.clear();
    }
};

}  // namespace
}  // namespace pt

namespace pt {

static std::string folder_of(const std::string& path) {
    size_t pos = path.find_last_of("/\\");
    return pos == std::string::npos ? std::string() : path.substr(0, pos + 1);
}

bool load_obj_file(const std::string& path, SceneFile& out, std::string& err) {
    std::ifstream f(path);
    if (!f) { err = "cannot open '" + path + "'"; return false; }
    const std::string folder = folder_of(path);
    ObjParser P;
    P.out = &out;
    std::string line;
    int line_no = 0;
    while (std::getline(f, line)) {
        line_no++;
        const char* p = skip_ws(line.c_str());
        if (at_eol(p) || *p == '#') continue;
        if (p[0] == 'v' && is_space(p[1])) { p += 2; for (int k = 0; k < 3; k++) P.v.push_back(parse_float(p)); continue; }
        if (p[0] == 'v' && p[1] == 'n' && is_space(p[2])) { p += 3; for (int k = 0; k < 3; k++) P.vn.push_back(parse_float(p)); continue; }
        if (p[0] == 'v' && p[1] == 't' && is_space(p[2])) { p += 3; for (int k = 0; k < 2; k++) P.vt.push_back(parse_float(p)); continue; }
        if (p[0] == 'f' && is_space(p[1])) {
            p = skip_ws(p + 2);
            Face face;
            while (!at_eol(p)) {
                VIdx vi;
                if (!parse_triple(p, (int)(P.v.size() / 3), (int)(P.vn.size() / 3), (int)(P.vt.size() / 2), vi)) {
                    char buf[96]; std::snprintf(buf, sizeof(buf), "Failed to parse `f' line (line %d)", line_no);
                    err = buf; return false;
                }
                face.idx.push_back(vi);
                while (!at_eol(p) && !is_space(*p)) p++;
                p = skip_ws(p);
            }
            P.group.push_back(std::move(face));
            continue;
        }
        if (!std::strncmp(p, "usemtl", 6) && is_space(p[6])) {
            std::string name = rest_of_line(p + 7);
            int id = -1;
            auto it = P.mtl_by_name.find(name);
            if (it != P.mtl_by_name.end()) id = it->second;
            else P.warn += "material [ '" + name + "' ] not found in .mtl\n";
            if (id != P.material) { P.flush_group(); P.material = id; }      // per-face material, same shape
            continue;
        }
        if (!std::strncmp(p, "mtllib", 6) && is_space(p[6])) {
            // several candidate files may be named; the first that opens is used (tiny_obj_loader.h:2720-2768)
            for (const std::string& name : split_escaped(rest_of_line(p + 7)))
                if (!name.empty() && load_mtl(folder + name, P.mtl, P.mtl_by_name, P.warn)) break;
            continue;
        }
        if ((p[0] == 'g' || p[0] == 'o') && is_space(p[1])) { P.finish_shape(); continue; }   // a bare "g" is ignored, as in tinyobj
        // s, l, p, t, vp ... : not consumed by the reference
    }
    P.finish_shape();
    if (out.mat_ids.empty()) { err = "'" + path + "' holds no faces"; return false; }

    // ---- materials: sky first (Scene.cpp:236-241), then one record per .mtl material in file order (:153-175)
    out.materials.clear(); out.tex_paths.clear();
    pt_material sky; std::memset(&sky, 0, sizeof(sky));
    sky.albedo_tex = -1; sky.props_tex = -1; sky.is_emissive = 1;
    sky.emission[0] = 25.0f * 30.0f; sky.emission[1] = 25.0f * 26.0f; sky.emission[2] = 25.0f * 19.0f;
    sky.props[1] = 1.0f;
    out.materials.push_back(sky); out.tex_paths.push_back("");
    for (const MtlRec& m : P.mtl) {
        pt_material r; std::memset(&r, 0, sizeof(r));
        r.albedo_tex = -1; r.props_tex = -1;
        float tr_ggx_roughness = 2.0f / (m.ns + 2.0f);                      // Scene.cpp:160
        float beckmann_roughness = std::sqrt(tr_ggx_roughness);             // :161
        float metallic = std::max(m.ks[0], std::max(m.ks[1], m.ks[2])) > 0.3f ? 1.0f : 0.0f;   // :164-170
        for (int k = 0; k < 3; k++) { r.albedo[k] = m.kd[k]; r.emission[k] = m.ke[k]; }
        r.is_emissive = (m.ke[0] + m.ke[1] + m.ke[2] > 1e-5f) ? 1 : 0;      // :109
        r.props[0] = 0.0f; r.props[1] = beckmann_roughness; r.props[2] = metallic;   // matprop colour, :103
        out.materials.push_back(r);
        out.tex_paths.push_back(m.map_kd.empty() ? std::string() : folder + m.map_kd);
    }
    out.warnings = P.warn;
    return true;
}

// ---- small image readers for map_Kd (the reference decodes through SOIL2; no decoder library exists here):
// binary/ASCII PPM here; TGA and BMP below, JPEG and PNG in pt_image.cpp -- each held byte-identical to the stb_image the
// reference's SOIL2 carries. Texels become RGBA32F = byte / 255 (GL_RGBA8 unorm, Texture.cpp:146-156), top row first.
static bool read_file(const std::string& path, std::vector<unsigned char>& buf) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    buf.assign(std::istreambuf_iterator<char>(f), std::istreambuf_iterator<char>());
    return true;
}

// Byte cursor with stb_image's end-of-data behaviour: reads past the end give 0, a negative skip goes to the end.
struct LdrReader {
    const std::vector<unsigned char>& b; size_t pos = 0;
    int get8() { return pos < b.size() ? b[pos++] : 0; }
    int get16le() { int z = get8(); return z + (get8() << 8); }
    uint32_t get32le() { uint32_t z = (uint32_t)get16le(); return z + ((uint32_t)get16le() << 16); }
    void skip(long n) { if (n < 0) pos = b.size(); else pos = std::min(b.size(), pos + (size_t)n); }
};

// .bmp as stbi_load(.., 4) of the reference's stb_image v2.15 returns it: header sizes 12 / 40 / 56 / 108 / 124; 4- and
// 8-bit palettes, 16 / 24 / 32 bits with default or BI_BITFIELDS masks (channels widened by bit replication); no RLE, no
// 1-bit; a 32-bit file of a 40/56-byte header whose alpha is zero everywhere comes back opaque; rows bottom-up unless the
// height is negative. Quirk kept: with a 40-byte header + BI_BITFIELDS the three masks are read and THEN
// `offset - 14 - header` bytes are skipped, so the pixels start 12 bytes late when `offset` accounts for the masks.
static bool decode_bmp_rgba8(const std::vector<unsigned char>& file, int& w, int& h, std::vector<unsigned char>& out) {
    LdrReader s{file};
    if (s.get8() != 'B' || s.get8() != 'M') return false;
    s.get32le(); s.get16le(); s.get16le();
    const int offset = (int)s.get32le(), hsz = (int)s.get32le();
    uint32_t mr = 0, mg = 0, mb = 0, ma = 0, all_a = 255;
    if (hsz != 12 && hsz != 40 && hsz != 56 && hsz != 108 && hsz != 124) return false;
    int iw, ih;
    if (hsz == 12) { iw = s.get16le(); ih = s.get16le(); } else { iw = (int)s.get32le(); ih = (int)s.get32le(); }
    if (s.get16le() != 1) return false;
    const int bpp = s.get16le();
    if (bpp == 1) return false;
    if (hsz != 12) {
        const int compress = (int)s.get32le();
        if (compress == 1 || compress == 2) return false;
        for (int i = 0; i < 5; i++) s.get32le();
        if (hsz == 40 || hsz == 56) {
            if (hsz == 56) for (int i = 0; i < 4; i++) s.get32le();
            if (bpp == 16 || bpp == 32) {
                if (compress == 0) {
                    if (bpp == 32) { mr = 0xffu << 16; mg = 0xffu << 8; mb = 0xffu; ma = 0xffu << 24; all_a = 0; }
                    else { mr = 31u << 10; mg = 31u << 5; mb = 31u; }
                } else if (compress == 3) {
                    mr = s.get32le(); mg = s.get32le(); mb = s.get32le();
                    if (mr == mg && mg == mb) return false;
                } else return false;
            }
        } else {
            mr = s.get32le(); mg = s.get32le(); mb = s.get32le(); ma = s.get32le();
            s.get32le();
            for (int i = 0; i < 12; i++) s.get32le();
            if (hsz == 124) for (int i = 0; i < 4; i++) s.get32le();
        }
    }
    const bool flip = ih > 0;
    ih = ih < 0 ? -ih : ih;
    if (iw <= 0 || ih <= 0 || (double)iw * ih > 1.0e9) return false;
    w = iw; h = ih;
    int psize = 0;
    if (hsz == 12) { if (bpp < 24) psize = (offset - 14 - 24) / 3; }
    else if (bpp < 16) psize = (offset - 14 - hsz) >> 2;
    out.assign((size_t)w * h * 4, 255);
    size_t z = 0;
    if (bpp < 16) {
        if (psize <= 0 || psize > 256) return false;     // a pixel offset smaller than the headers makes psize negative: damaged file
        unsigned char pal[256][3];
        for (int i = 0; i < psize; i++) { pal[i][2] = (unsigned char)s.get8(); pal[i][1] = (unsigned char)s.get8(); pal[i][0] = (unsigned char)s.get8(); if (hsz != 12) s.get8(); }
        for (int i = psize; i < 256; i++) pal[i][0] = pal[i][1] = pal[i][2] = 0;     // stb reads its uninitialised stack here; files that index past the palette are damaged
        s.skip((long)offset - 14 - hsz - (long)psize * (hsz == 12 ? 3 : 4));
        int width;
        if (bpp == 4) width = (w + 1) >> 1; else if (bpp == 8) width = w; else return false;
        const int pad = (-width) & 3;
        for (int j = 0; j < h; j++) {
            for (int i = 0; i < w; i += 2) {
                int v = s.get8(), v2 = 0;
                if (bpp == 4) { v2 = v & 15; v >>= 4; }
                out[z] = pal[v][0]; out[z + 1] = pal[v][1]; out[z + 2] = pal[v][2]; z += 4;
                if (i + 1 == w) break;
                v = bpp == 8 ? s.get8() : v2;
                out[z] = pal[v][0]; out[z + 1] = pal[v][1]; out[z + 2] = pal[v][2]; z += 4;
            }
            s.skip(pad);
        }
    } else {
        auto high_bit = [](uint32_t v) { int n = -1; while (v) { n++; v >>= 1; } return n; };
        auto bitcount = [](uint32_t v) { int n = 0; while (v) { n += (int)(v & 1u); v >>= 1; } return n; };
        auto shiftsigned = [](int v, int shift, int bits) {
            if (shift < 0) v <<= -shift; else v >>= shift;
            int result = v;
            for (int q = bits; q < 8; q += bits) result += v >> q;
            return result;
        };
        s.skip((long)offset - 14 - hsz);
        int width = bpp == 24 ? 3 * w : bpp == 16 ? 2 * w : 0;
        const int pad = (-width) & 3;
        int easy = 0;
        if (bpp == 24) easy = 1;
        else if (bpp == 32 && mb == 0xffu && mg == 0xff00u && mr == 0x00ff0000u && ma == 0xff000000u) easy = 2;
        int rshift = 0, gshift = 0, bshift = 0, ashift = 0, rcount = 0, gcount = 0, bcount = 0, acount = 0;
        if (!easy) {
            if (!mr || !mg || !mb) return false;
            rshift = high_bit(mr) - 7; rcount = bitcount(mr); gshift = high_bit(mg) - 7; gcount = bitcount(mg);
            bshift = high_bit(mb) - 7; bcount = bitcount(mb); ashift = high_bit(ma) - 7; acount = bitcount(ma);
            if (rcount == 0 || gcount == 0 || bcount == 0) return false;
        }
        for (int j = 0; j < h; j++) {
            for (int i = 0; i < w; i++, z += 4) {
                if (easy) {
                    out[z + 2] = (unsigned char)s.get8(); out[z + 1] = (unsigned char)s.get8(); out[z] = (unsigned char)s.get8();
                    unsigned a = easy == 2 ? (unsigned)s.get8() : 255u;
                    all_a |= a; out[z + 3] = (unsigned char)a;
                } else {
                    uint32_t v = bpp == 16 ? (uint32_t)s.get16le() : s.get32le();
                    out[z] = (unsigned char)(shiftsigned((int)(v & mr), rshift, rcount) & 255);
                    out[z + 1] = (unsigned char)(shiftsigned((int)(v & mg), gshift, gcount) & 255);
                    out[z + 2] = (unsigned char)(shiftsigned((int)(v & mb), bshift, bcount) & 255);
                    int a = ma ? shiftsigned((int)(v & ma), ashift, acount) : 255;
                    all_a |= (uint32_t)a; out[z + 3] = (unsigned char)(a & 255);
                }
            }
            s.skip(pad);
        }
    }
    if (all_a == 0) for (size_t i = 3; i < out.size(); i += 4) out[i] = 255;
    if (flip)
        for (int j = 0; j < h / 2; j++)
            std::swap_ranges(out.begin() + (size_t)j * w * 4, out.begin() + (size_t)(j + 1) * w * 4, out.begin() + (size_t)(h - 1 - j) * w * 4);
    return true;
}

// .tga as stbi_load(.., 4) of stb_image v2.15 returns it: image types 1 / 2 / 3 and their RLE forms 9 / 10 / 11; 8-bit grey,
// 16-bit grey + alpha, 15/16-bit RGB (5 bits each, (c * 255) / 31, no alpha), 24/32-bit BGR(A); colour maps with 8- or
// 16-bit indices and entries of any of those formats (stb skips `first entry index` BYTES before the map and maps an index
// past the end to entry 0); origin bit 5 = top-down, the horizontal flip bit is ignored; TGA has no magic number, so the
// header must pass stb's plausibility test.
static bool decode_tga_rgba8(const std::vector<unsigned char>& file, int& w, int& h, std::vector<unsigned char>& out) {
    if (file.size() < 18) return false;
    LdrReader s{file};
    const int id_len = s.get8(), indexed = s.get8();
    int type = s.get8();
    const int pal_start = s.get16le(), pal_len = s.get16le(), pal_bits = s.get8();
    s.get16le(); s.get16le();
    const int tw = s.get16le(), th = s.get16le(), bpp = s.get8(), desc = s.get8();
    // stbi__tga_test
    if (indexed > 1) return false;
    if (indexed == 1) {
        if (type != 1 && type != 9) return false;
        if (pal_bits != 8 && pal_bits != 15 && pal_bits != 16 && pal_bits != 24 && pal_bits != 32) return false;
        if (bpp != 8 && bpp != 16) return false;
    } else {
        if (type != 2 && type != 3 && type != 10 && type != 11) return false;
        if (bpp != 8 && bpp != 15 && bpp != 16 && bpp != 24 && bpp != 32) return false;
    }
    if (tw < 1 || th < 1 || (double)tw * th > 1.0e9) return false;
    bool rle = false;
    if (type >= 8) { type -= 8; rle = true; }
    const bool inverted = ((desc >> 5) & 1) == 0;                        // bottom-up file: rows are flipped after decoding
    bool rgb16 = false;
    auto comp_of = [&](int bits, bool grey) -> int {
        switch (bits) {
            case 8: return 1;
            case 16: if (grey) return 2;                                   // fall through
            case 15: rgb16 = true; return 3;
            case 24: case 32: return bits / 8;
            default: return 0;
        }
    };
    const int comp = indexed ? comp_of(pal_bits, false) : comp_of(bpp, type == 3);
    if (!comp) return false;
    w = tw; h = th;
    const size_t n = (size_t)w * h;
    std::vector<unsigned char> data(n * comp), palette;
    auto read_rgb16 = [&](unsigned char* o) {
        unsigned px = (unsigned)s.get16le();
        o[0] = (unsigned char)((((px >> 10) & 31) * 255) / 31); o[1] = (unsigned char)((((px >> 5) & 31) * 255) / 31); o[2] = (unsigned char)(((px & 31) * 255) / 31);
    };
    s.skip(id_len);
    if (indexed) {
        s.skip(pal_start);
        palette.assign((size_t)pal_len * comp, 0);
        if (rgb16) for (int i = 0; i < pal_len; i++) read_rgb16(palette.data() + (size_t)i * comp);
        else {
            if (s.pos + palette.size() > file.size()) return false;       // "bad palette"
            std::memcpy(palette.data(), file.data() + s.pos, palette.size()); s.pos += palette.size();
        }
    }
    unsigned char raw[4] = {0, 0, 0, 0};
    int rle_count = 0; bool repeating = false, read_next = true;
    for (size_t i = 0; i < n; i++) {
        if (rle) {
            if (rle_count == 0) { int cmd = s.get8(); rle_count = 1 + (cmd & 127); repeating = (cmd >> 7) != 0; read_next = true; }
            else if (!repeating) read_next = true;
        } else read_next = true;
        if (read_next) {
            if (indexed) {
                int idx = bpp == 8 ? s.get8() : s.get16le();
                if (idx >= pal_len) idx = 0;
                for (int j = 0; j < comp; j++) raw[j] = palette.empty() ? 0 : palette[(size_t)idx * comp + j];
            } else if (rgb16) read_rgb16(raw);
            else for (int j = 0; j < comp; j++) raw[j] = (unsigned char)s.get8();
            read_next = false;
        }
        for (int j = 0; j < comp; j++) data[i * comp + j] = raw[j];
        rle_count--;
    }
    out.assign(n * 4, 255);
    for (int y = 0; y < h; y++) {
        const int src = inverted ? h - 1 - y : y;
        for (int x = 0; x < w; x++) {
            const unsigned char* q = data.data() + ((size_t)src * w + x) * comp;
            unsigned char* o = out.data() + ((size_t)y * w + x) * 4;
            if (comp == 1) { o[0] = o[1] = o[2] = q[0]; }
            else if (comp == 2) { o[0] = o[1] = o[2] = q[0]; o[3] = q[1]; }
            else if (rgb16) { o[0] = q[0]; o[1] = q[1]; o[2] = q[2]; }
            else { o[0] = q[2]; o[1] = q[1]; o[2] = q[0]; if (comp == 4) o[3] = q[3]; }
        }
    }
    return true;
}

bool load_image_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba) {
    std::vector<unsigned char> b;
    if (!read_file(path, b) || b.size() < 18) return false;
    auto put = [&](size_t px, int r, int g, int bl, int a) {
        rgba[4 * px] = r / 255.0f; rgba[4 * px + 1] = g / 255.0f; rgba[4 * px + 2] = bl / 255.0f; rgba[4 * px + 3] = a / 255.0f;
    };
    if (b[0] == 0xff && b[1] == 0xd8) {
        std::vector<unsigned char> px; std::string err;
        if (!decode_jpeg_rgba8(b, w, h, px, err)) return false;
        rgba.resize(px.size());
        for (size_t i = 0; i < px.size(); i++) rgba[i] = px[i] / 255.0f;
        return true;
    }
    if (b[0] == 137 && b[1] == 'P' && b[2] == 'N' && b[3] == 'G') {
        std::vector<unsigned char> px;
        if (!decode_png_rgba8(b, w, h, px)) return false;
        rgba.resize(px.size());
        for (size_t i = 0; i < px.size(); i++) rgba[i] = px[i] / 255.0f;
        return true;
    }
    if (b[0] == 'P' && (b[1] == '6' || b[1] == '5' || b[1] == '3')) {
        // binary .ppm / .pgm as stb_image reads them: bytes are taken as they are (no rescaling by the maximum value, which must
        // be <= 255), grey replicated to RGB. ASCII "P3" is an extension of this reader (stb refuses it); its values are scaled.
        size_t pos = 2; int vals[3], got = 0;
        while (got < 3 && pos < b.size()) {
            while (pos < b.size() && (b[pos] == ' ' || b[pos] == '\n' || b[pos] == '\r' || b[pos] == '\t' || b[pos] == '\v' || b[pos] == '\f')) pos++;
            if (pos < b.size() && b[pos] == '#') { while (pos < b.size() && b[pos] != '\n' && b[pos] != '\r') pos++; continue; }
            int v = 0; bool any = false;
            while (pos < b.size() && b[pos] >= '0' && b[pos] <= '9') { if (v > 100000000) return false; v = v * 10 + (b[pos] - '0'); pos++; any = true; }
            if (!any) return false;
            vals[got++] = v;
        }
        w = vals[0]; h = vals[1];
        if (got < 3 || w <= 0 || h <= 0 || vals[2] > 255 || (double)w * h > 1.0e9) return false;
        rgba.assign((size_t)w * h * 4, 1.0f);
        if (b[1] != '3') {
            const size_t nc = b[1] == '6' ? 3 : 1;
            pos++;                                                       // the one whitespace byte after the maximum value
            if (b.size() < pos + (size_t)w * h * nc) return false;
            for (size_t i = 0; i < (size_t)w * h; i++) {
                const unsigned char* q = b.data() + pos + nc * i;
                if (nc == 3) put(i, q[0], q[1], q[2], 255); else put(i, q[0], q[0], q[0], 255);
            }
        } else {
            if (vals[2] <= 0) return false;
            for (size_t i = 0; i < (size_t)w * h; i++) {
                int c[3];
                for (int k = 0; k < 3; k++) {
                    while (pos < b.size() && (b[pos] < '0' || b[pos] > '9')) pos++;
                    int v = 0; while (pos < b.size() && b[pos] >= '0' && b[pos] <= '9') { if (v < 100000) v = v * 10 + (b[pos] - '0'); pos++; }
                    c[k] = v * 255 / vals[2];
                }
                put(i, c[0], c[1], c[2], 255);
            }
        }
        return true;
    }
    std::vector<unsigned char> px;
    if (b[0] == 'B' && b[1] == 'M') {
        if (!decode_bmp_rgba8(b, w, h, px)) return false;
    } else if (!decode_tga_rgba8(b, w, h, px)) return false;
    rgba.resize(px.size());
    for (size_t i = 0; i < px.size(); i++) rgba[i] = px[i] / 255.0f;
    return true;
}

}  // namespace pt

// ------------------------------------------------------------------------------------------------ C ABI
struct pt_scene_file { pt::SceneFile s; };

extern "C" {

int pt_scene_file_load_obj(const char* obj_path, pt_scene_file** out) {
    if (!obj_path || !out) return PT_ERR_INVALID;
    *out = nullptr;
    pt_scene_file* sf = new pt_scene_file();
    std::string err;
    bool ok = false;
    try { ok = pt::load_obj_file(obj_path, sf->s, err); }
    catch (const std::exception& e) { err = std::string("pt_scene_file_load_obj: ") + e.what(); }
    if (!ok) {
        pt::set_last_error(err.c_str());
        delete sf;
        return PT_ERR_IO;
    }
    *out = sf;
    return PT_OK;
}
void pt_scene_file_free(pt_scene_file* sf) { delete sf; }
int64_t pt_scene_file_num_triangles(const pt_scene_file* sf) { return sf ? (int64_t)sf->s.mat_ids.size() : 0; }
const float* pt_scene_file_vertices(const pt_scene_file* sf) { return sf ? sf->s.verts.data() : nullptr; }
const uint32_t* pt_scene_file_material_ids(const pt_scene_file* sf) { return sf ? sf->s.mat_ids.data() : nullptr; }
int32_t pt_scene_file_num_materials(const pt_scene_file* sf) { return sf ? (int32_t)sf->s.materials.size() : 0; }
const pt_material* pt_scene_file_materials(const pt_scene_file* sf) { return sf ? sf->s.materials.data() : nullptr; }
const char* pt_scene_file_texture_path(const pt_scene_file* sf, int32_t material) {
    if (!sf || material < 0 || material >= (int32_t)sf->s.tex_paths.size()) return "";
    return sf->s.tex_paths[material].c_str();
}
int32_t pt_scene_file_num_shapes(const pt_scene_file* sf) { return sf ? sf->s.n_shapes : 0; }
const char* pt_scene_file_warnings(const pt_scene_file* sf) { return sf ? sf->s.warnings.c_str() : ""; }

int pt_image_load(const char* path, int32_t* width, int32_t* height, float** rgba) {
    if (!path || !width || !height || !rgba) return PT_ERR_INVALID;
    int w = 0, h = 0; std::vector<float> px;
    bool ok = false;
    try { ok = pt::load_image_rgba32f(path, w, h, px); }
    catch (const std::exception& e) {      // a header that asks for more memory than there is: std::bad_alloc / length_error must not cross the C ABI
        pt::set_last_error((std::string("cannot decode '") + path + "': " + e.what()).c_str());
        return PT_ERR_LIMIT;
    }
    if (!ok) {
        pt::set_last_error((std::string("cannot decode '") + path + "' (supported: .jpg, .png, .bmp, .tga, .ppm as stb_image reads them)").c_str());
        return PT_ERR_IO;
    }
    *width = w; *height = h;
    *rgba = (float*)std::malloc(px.size() * sizeof(float));
    std::memcpy(*rgba, px.data(), px.size() * sizeof(float));
    return PT_OK;
}

int pt_scene_txt_read(const char* path, pt_scene_txt* out) {
    if (!path || !out) return PT_ERR_INVALID;
    std::ifstream f(path);
    if (!f) { pt::set_last_error((std::string("cannot open '") + path + "'").c_str()); return PT_ERR_IO; }
    std::memset(out, 0, sizeof(*out));
    std::string obj, env;
    std::getline(f, obj); std::getline(f, env);                       // Program.cpp:74-76
    while (!obj.empty() && (obj.back() == '\r' || obj.back() == ' ')) obj.pop_back();
    while (!env.empty() && (env.back() == '\r' || env.back() == ' ')) env.pop_back();
    f >> out->camera_position[0] >> out->camera_position[1] >> out->camera_position[2];    // :77-79
    f >> out->camera_rotation[0] >> out->camera_rotation[1] >> out->camera_rotation[2];
    if (obj.empty() || obj.size() >= sizeof(out->scene_path) || env.size() >= sizeof(out->environment)) {
        pt::set_last_error("scene.txt: line 1 must be the scene path, line 2 the environment string");
        return PT_ERR_INVALID;
    }
    std::memcpy(out->scene_path, obj.c_str(), obj.size());
    std::memcpy(out->environment, env.c_str(), env.size());
    return PT_OK;
}

int pt_parse_environment(const char* args, float rgb[3]) {
    if (!args || !rgb) return PT_ERR_INVALID;
    // LoadEnvironmnet, Renderer.cpp:236-260: "GENERATE COLOR BLACK|WHITE", anything else after COLOR is red "for error"
    std::stringstream parser{std::string(args)};
    std::string cmd, arg0, arg1;
    parser >> cmd >> arg0 >> arg1;
    if (cmd != "GENERATE" || arg0 != "COLOR") {
        pt::set_last_error("environment: a string whose first character is one of \"GENERATE\" is taken as a command (find_first_of, Renderer.cpp:238) and the only command is \"GENERATE COLOR <BLACK|WHITE>\"");
        return PT_ERR_UNSUPPORTED;
    }
    if (arg1 == "BLACK") { rgb[0] = rgb[1] = rgb[2] = 0.0f; }
    else if (arg1 == "WHITE") { rgb[0] = rgb[1] = rgb[2] = 1.0f; }
    else { rgb[0] = 1.0f; rgb[1] = rgb[2] = 0.0f; }
    return PT_OK;
}

}  // extern "C"
