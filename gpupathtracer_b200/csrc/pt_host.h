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

// Own builder (pt_bvh_build.cpp): tris_pre = n x 20 floats with pre-subtracted edges.
bool build_sbvh(const float* tris_pre, int64_t n, int n_threads, HostBvh& out, std::string& msg);

bool write_ppm(const char* path, int w, int h, const uint8_t* rgb);
bool write_pfm(const char* path, int w, int h, const float* rgb);

}  // namespace pt
