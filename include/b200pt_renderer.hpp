// include/b200pt_renderer.hpp -- header-only C++ facade over the C ABI (b200pt.h) with the reference's own class shape:
// the eight public methods of `Renderer` (src/core/Renderer.h:14-24) keep their names and meaning, headless.
// A maintainer of the reference swaps `#include "core/Renderer.h"` for this header in Program.cpp.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <stdexcept>
#include <string>
#include <vector>
#include "b200pt.h"

namespace b200pt {

// Stand-in for misc/Window.h: the renderer only reads Width/Height (Renderer.cpp:450-451).
struct Window { uint32_t Width = 1280, Height = 720; };

// math/Camera.h:14-46, same constructor and setters; GenerateImagePlane fills the 7 uniforms of Shader::LoadCamera.
class Camera {
public:
    Camera(float aspect, float fieldofview, float focus, float aperture)
        : aspect_(aspect), fov_(fieldofview), focus_(focus), aperture_(aperture) {}
    void SetPosition(const float p[3]) { for (int i = 0; i < 3; i++) pos_[i] = p[i]; }
    void SetRotation(const float r[3]) { for (int i = 0; i < 3; i++) rot_[i] = r[i]; }
    void GenerateImagePlane() { pt_camera_image_plane(aspect_, fov_, focus_, aperture_, pos_, rot_, &plane_); }
    const pt_camera& Plane() const { return plane_; }
private:
    float aspect_, fov_, focus_, aperture_;
    float pos_[3] = {0, 0, 0}, rot_[3] = {0, 0, 0};
    pt_camera plane_{};
};

// What Scene::LoadScene (Scene.cpp:233-357) leaves behind, as plain arrays.
struct SceneArrays {
    std::vector<float> verts;            // n_tris x 3 x {pos, normal, uv}
    std::vector<uint32_t> mat_ids;       // per triangle, = 2 * material index
    std::vector<pt_material> materials;  // [0] = sky (Scene.cpp:236-241)
    std::vector<pt_texture> textures;
    float env_color[3] = {1, 1, 1};      // "GENERATE COLOR WHITE"
};

class Renderer {
public:
    ~Renderer() { CleanUp(); }

    // Renderer::Initialize(Window*, scenePath, env) (Renderer.cpp:448-636): scene assembly, BVH build, uploads, RNG tables.
    // n_gpus > 1: the samples of every RenderFrame are split across that many GPUs (one host thread, NCCL reduce to GPU 0);
    // the scene is loaded and the tree built ONCE, on GPU 0, and copied to the others.
    void Initialize(const Window* window, const SceneArrays& scene, int max_depth = 66, uint64_t seed = 0, int n_gpus = 1,
                    int microfacet_model = PT_MICROFACET_GGX) {
        create(window, max_depth, microfacet_model, PT_BVH_REFERENCE, n_gpus);
        pt_context* ctx = pt_group_context(group_, 0);
        const int64_t n = (int64_t)scene.mat_ids.size();
        std::vector<float> tris((size_t)n * 20);
        check(pt_assemble_triangles(scene.verts.data(), scene.mat_ids.data(), n, tris.data()));           // Scene.cpp:263-292
        float total = 0.0f;
        int nl = pt_build_light_table(tris.data(), n, scene.materials.data(), (int32_t)scene.materials.size(), nullptr, 0, &total);
        if (nl < 1) throw std::runtime_error("scene has no emissive triangle (Iterative.comp:42-61 needs one)");
        std::vector<pt_light> lights((size_t)nl);
        pt_build_light_table(tris.data(), n, scene.materials.data(), (int32_t)scene.materials.size(), lights.data(), nl, &total);
        pt_presubtract_edges(tris.data(), n);                                                                // Scene.cpp:334-337
        check(pt_set_triangles(ctx, tris.data(), n));
        check(pt_set_materials(ctx, scene.materials.data(), (int32_t)scene.materials.size(), scene.textures.data(),
                               (int32_t)scene.textures.size()));
        check(pt_set_lights(ctx, lights.data(), nl, total));
        check(pt_set_environment_color(ctx, scene.env_color));
        check(pt_build_bvh(ctx, &bvh_stats_));                                                               // Scene.cpp:294
        optimize(ctx, n);
        finish(seed);
    }
    // The reference's own signature: Initialize(Window*, scenePath, env) (Renderer.cpp:448, called at Program.cpp:81).
    // scenePath is an .obj (the only live loader, Scene.cpp:248-259); env is any LoadEnvironmnet string: "GENERATE COLOR WHITE",
    // a .hdr / .jpg equirectangular image, or a text file naming six cube faces.
    void Initialize(const Window* window, const char* scenePath, const std::string& env, int max_depth = 66, uint64_t seed = 0,
                    int bvh_layout = PT_BVH_REFERENCE, int n_gpus = 1, int microfacet_model = PT_MICROFACET_GGX) {
        create(window, max_depth, microfacet_model, bvh_layout, n_gpus);
        check(pt_load_scene(pt_group_context(group_, 0), scenePath, env.c_str(), &bvh_stats_));             // Scene::LoadScene
        optimize(pt_group_context(group_, 0), bvh_stats_.n_refs);
        finish(seed);
    }
    // The reinsertion pass the reference carries and never calls (ReinsertionOptimize, BVH.cpp): +2-3 % frames for ~0.5 s per million
    // triangles of host time, same hits. Skipped for scenes where that would be felt (> 2 M triangles); a tree it cannot improve
    // within the traversal stack is left as it is. Set B200PT_OPTIMIZE_BVH=0 to skip it.
    void optimize(pt_context* ctx, int64_t n_refs) {
        const char* e = std::getenv("B200PT_OPTIMIZE_BVH");
        if ((e && e[0] == '0') || n_refs > 2000000) return;
        pt_bvh_stats st{};
        if (pt_optimize_bvh(ctx, 4, 0.1f, &st) == PT_OK) bvh_stats_ = st;
    }
    void RenderFrame(const Camera& camera) {                      // Renderer.cpp:651-662: +1 sample per pixel on every GPU
        check(pt_group_set_camera(group_, &camera.Plane()));
        check(pt_group_render_batch(group_, 1));
    }
    // Renderer::Present (Renderer.cpp:681-685): headless resolve into a host buffer (width*height*3 bytes, row 0 = bottom).
    void Present(uint8_t* rgb8) { PresentAsync(rgb8); Synchronize(); }
    // The same without stalling the render stream: the copy to (pinned) host memory overlaps the next RenderFrame;
    // the image is complete after Synchronize(). At most two presents in flight.
    void PresentAsync(uint8_t* rgb8) { check(pt_group_present_async(group_, rgb8)); }
    void Synchronize() { check(pt_group_synchronize(group_)); }
    void ResetSamples() { check(pt_group_reset_samples(group_)); }   // Renderer.cpp:687-691
    uint32_t GetNumSamples() { return pt_group_num_samples(group_); }   // Renderer.cpp:693-695 (all GPUs together)
    void SaveScreenshot(const std::string& filename) { check(pt_group_save_screenshot(group_, filename.c_str())); }   // :697-705
    // beyond the reference: a long render that survives a restart (RNG table + running sum + numSamples); single GPU
    void SaveCheckpoint(const std::string& path) { check(pt_checkpoint_save(Context(), path.c_str())); }
    void LoadCheckpoint(const std::string& path) { check(pt_checkpoint_load(Context(), path.c_str())); }
    // LoadEnvironmnet (Renderer.cpp:236-323) after Initialize, e.g. to swap the sky of a loaded scene
    void LoadEnvironment(const std::string& args) {
        check(pt_load_environment(Context(), args.c_str()));
        if (pt_group_size(group_) > 1) check(pt_group_share_scene(group_));
    }
    void CleanUp() { if (group_) { pt_group_destroy(group_); group_ = nullptr; } }                           // :638-647
    // Renderer::RenderReference (the stale CPU renderer, Renderer.cpp:1075-1187) has no counterpart: out of scope.
    const pt_bvh_stats& BvhStats() const { return bvh_stats_; }
    pt_context* Context() { return group_ ? pt_group_context(group_, 0) : nullptr; }
    int NumGpus() { return group_ ? pt_group_size(group_) : 0; }

private:
    static void check(int rc) { if (rc != PT_OK) throw std::runtime_error(pt_last_error()); }
    void create(const Window* window, int max_depth, int microfacet_model, int bvh_layout, int n_gpus) {
        pt_config cfg{};
        cfg.width = (int32_t)window->Width; cfg.height = (int32_t)window->Height;
        cfg.max_depth = max_depth; cfg.microfacet_model = microfacet_model; cfg.bvh_layout = bvh_layout;
        check(pt_create_multi(&cfg, nullptr, n_gpus, &group_));        // devices 0 .. n_gpus-1
        width_ = window->Width; height_ = window->Height;
    }
    void finish(uint64_t seed) {
        if (pt_group_size(group_) > 1) check(pt_group_share_scene(group_));
        check(pt_group_seed_rng(group_, seed));                                                              // Renderer.cpp:545-553
    }
    pt_group* group_ = nullptr;
    pt_bvh_stats bvh_stats_{};
    uint32_t width_ = 0, height_ = 0;
};

}  // namespace b200pt
