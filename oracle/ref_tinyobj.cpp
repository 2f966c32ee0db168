// oracle/ref_tinyobj.cpp -- TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// The reference parses scenes with tinyobjloader (extern/tinyobjloader, pinned a1e8bad; Scene.cpp:120-146). This
// harness compiles that UNMODIFIED header where it lies under /root/reference (oracle/Makefile, target `ref`) and
// hands tinyobj::ObjReader's raw result to the tests, which restate LoadOBJ's rules themselves and compare with
// libb200pt's own parser (csrc/pt_obj.cpp).
#define TINYOBJLOADER_IMPLEMENTATION
#include "tiny_obj_loader.h"
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

template <class T>
static T* dup(const std::vector<T>& v) {
    T* p = (T*)std::malloc((v.empty() ? 1 : v.size()) * sizeof(T));
    if (!v.empty()) std::memcpy(p, v.data(), v.size() * sizeof(T));
    return p;
}

extern "C" {
// corner_idx: 3 ints per corner (vertex, normal, texcoord); face_mat: tinyobj material id per triangle;
// shape_first_face: n_shapes + 1 offsets into the face list; mat_fields: 11 floats per material
// (Kd, Ks, Ke, Ns, illum); texnames: diffuse_texname per material joined by '\n'.
int ref_tinyobj_load(const char* path, int* n_corners, int** corner_idx, int* n_faces, int** face_mat, int* n_shapes,
                     int** shape_first_face, int* nv, float** v, int* nvn, float** vn, int* nvt, float** vt, int* n_mats,
                     float** mat_fields, char** texnames) {
    tinyobj::ObjReaderConfig cfg;
    cfg.mtl_search_path = "";                       // as Scene.cpp:126
    tinyobj::ObjReader reader;
    if (!reader.ParseFromFile(path, cfg)) return -1;
    const auto& attrib = reader.GetAttrib();
    const auto& shapes = reader.GetShapes();
    const auto& materials = reader.GetMaterials();
    std::vector<int> ci, fm, sf;
    for (const auto& s : shapes) {
        sf.push_back((int)fm.size());
        size_t off = 0;
        for (size_t f = 0; f < s.mesh.num_face_vertices.size(); f++) {
            size_t fv = s.mesh.num_face_vertices[f];
            for (size_t k = 0; k < fv; k++) {
                const tinyobj::index_t& i = s.mesh.indices[off + k];
                ci.push_back(i.vertex_index); ci.push_back(i.normal_index); ci.push_back(i.texcoord_index);
            }
            off += fv;
            fm.push_back(s.mesh.material_ids[f]);
        }
    }
    sf.push_back((int)fm.size());
    std::vector<float> mf; std::string names;
    for (const auto& m : materials) {
        for (int k = 0; k < 3; k++) mf.push_back(m.diffuse[k]);
        for (int k = 0; k < 3; k++) mf.push_back(m.specular[k]);
        for (int k = 0; k < 3; k++) mf.push_back(m.emission[k]);
        mf.push_back(m.shininess); mf.push_back((float)m.illum);
        names += m.diffuse_texname; names += '\n';
    }
    *n_corners = (int)ci.size() / 3; *corner_idx = dup(ci);
    *n_faces = (int)fm.size(); *face_mat = dup(fm);
    *n_shapes = (int)shapes.size(); *shape_first_face = dup(sf);
    *nv = (int)attrib.vertices.size() / 3; *v = dup(attrib.vertices);
    *nvn = (int)attrib.normals.size() / 3; *vn = dup(attrib.normals);
    *nvt = (int)attrib.texcoords.size() / 2; *vt = dup(attrib.texcoords);
    *n_mats = (int)materials.size(); *mat_fields = dup(mf);
    char* nm = (char*)std::malloc(names.size() + 1); std::memcpy(nm, names.c_str(), names.size() + 1); *texnames = nm;
    return 0;
}
void ref_tinyobj_free(void* p) { std::free(p); }
}
