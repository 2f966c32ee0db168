// pt_host.h -- host-side (CPU) parts of the hot path: scene assembly, camera, SBVH builder, image IO.
#pragma once
#include <stdint.h>
#include <vector>
#include <string>
#include "../../include/b200pt.h"

namespace pt {

struct HostBvh {
    std::vector<float> nodes;      // n x 8, reference serialised form
    std::vector<int32_t> refs;
    int max_depth = 0;
    int n_leaves = 0;
    float sah_cost = 0.0f;
};

// Validates a reference-form tree and returns its depth (or -1 with msg set).
int validate_bvh(const float* nodes, int n_nodes, const int32_t* refs, int n_refs, int64_t n_tris, std::string& msg,
                 int* n_leaves);

// SAH cost as BVH.cpp:2174-2195 (CalculateCost) evaluated on the serialised tree.
float bvh_sah_cost(const float* nodes, int n_nodes, const int32_t* refs, int n_refs);

// Reinsertion optimisation of a reference-form tree in place (pt_bvh_opt.cpp). Returns the subtrees moved, -1 if the result would
// be max_depth deep or deeper (input untouched).
int optimize_bvh_reinsertion(float* nodes, int n_nodes, int iterations, float batch_fraction, int max_depth, float* sah_before, float* sah_after,
                             const int32_t* refs, int n_refs);

// Own builder (pt_bvh_build.cpp): tris_pre = n x 20 floats with pre-subtracted edges.
bool build_sbvh(const float* tris_pre, int64_t n, int n_threads, HostBvh& out, std::string& msg);

// PT_BVH_WIDE (pt_bvh_wide.cpp): 96-B compressed 8-wide node, the exact bytes the device reads (pt_layout.h WideNode).
struct WideNodeHost {
    float p[3];                 // grid origin = min corner of the children
    uint8_t e[3];               // per-axis grid step 2^(e-127) as an IEEE-754 exponent byte
    uint8_t imask;              // bit s set: slot s is an inner node
    uint32_t child_base;        // index of the first inner child (children are contiguous, in slot order)
    uint32_t tri_base;          // index of the first leaf triangle in the wide leaf stream
    uint8_t meta[8];            // inner: 001_11sss (bit index 24 + s); leaf: unary triangle count << 5 | offset; empty: 0
    uint8_t qlo[3][8], qhi[3][8];
    uint8_t pad[16];
};
static_assert(sizeof(WideNodeHost) == 96, "wide node layout");
// 4-wide variant, 64 B = 2 x 256-bit loads; children sit in array positions 0..3 in ascending SLOT order (a slot is
// still one of 8 octant-ordered positions, carried by meta / imask), unused positions have meta 0 and an inverted box.
struct WideNode4Host {
    float p[3];
    uint8_t e[3];
    uint8_t imask;              // over the 8 slots
    uint32_t child_base, tri_base;
    uint8_t meta[4];
    uint8_t qlo[3][4], qhi[3][4];
    uint8_t pad[12];
};
static_assert(sizeof(WideNode4Host) == 64, "wide-4 node layout");
struct WideBvh {
    std::vector<uint8_t> bytes;       // n_nodes x node_size, the exact device bytes
    int64_t n_nodes = 0;
    int node_size = 0, width = 0;
    std::vector<int32_t> tri_ids;     // triangle index per entry of the wide leaf stream
    int depth = 0;
};
bool build_wide_bvh(const HostBvh& in, const float* tris_pre, int64_t n_tris, int width, WideBvh& out, std::string& msg);

// pt_obj.cpp: Wavefront .obj/.mtl with the reference's material rules (Scene.cpp:120-231)
struct SceneFile;
bool load_obj_file(const std::string& path, SceneFile& out, std::string& err);
bool load_image_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba);
bool decode_png_rgba8(const std::vector<unsigned char>& file, int& w, int& h, std::vector<unsigned char>& rgba);          // pt_image.cpp
bool decode_jpeg_rgba8(const std::vector<unsigned char>& file, int& w, int& h, std::vector<unsigned char>& rgba, std::string& err);   // pt_image.cpp
bool write_png_rgb8(const char* path, int w, int h, const unsigned char* rgb_top_down);                                    // pt_image.cpp
bool load_ldr_as_hdr_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba, std::string& err);        // pt_image.cpp
bool load_hdr_rgba32f(const std::string& path, int& w, int& h, std::vector<float>& rgba, std::string& err);   // pt_image.cpp
void set_last_error(const char* msg);       // message returned by pt_last_error() on this thread

bool write_ppm(const char* path, int w, int h, const uint8_t* rgb);
bool write_pfm(const char* path, int w, int h, const float* rgb);

}  // namespace pt
