// tools/headless_main.cpp -- the reference's main() (src/Program.cpp:55-152) without the window: scene.txt in, image out.
//   g++ -std=c++17 tools/headless_main.cpp -Iinclude -Lgpupathtracer_b200 -lb200pt -Wl,-rpath,$PWD/gpupathtracer_b200 -o headless
//   ./headless [scene.txt] [spp] [out.png] [n_gpus]    (the reference saves PNG screenshots, Renderer.cpp:698; .ppm/.pfm also work;
//                                                       spp is per GPU: n_gpus GPUs deliver spp x n_gpus samples per pixel)
#include "b200pt_renderer.hpp"
#include <cstdio>
#include <cstdlib>

int main(int argc, char** argv) {
    const char* scene_txt = argc > 1 ? argv[1] : "scene.txt";
    int spp = argc > 2 ? std::atoi(argv[2]) : 64;
    const char* out = argc > 3 ? argv[3] : "out.png";
    const int n_gpus = argc > 4 ? std::atoi(argv[4]) : 1;
    pt_scene_txt st;
    if (pt_scene_txt_read(scene_txt, &st) != PT_OK) { std::fprintf(stderr, "%s\n", pt_last_error()); return 1; }
    b200pt::Window window; window.Width = 1280; window.Height = 720;               // Program.cpp:22-23
    b200pt::Camera camera(1280.0f / 720.0f, 0.7853982f, 90.0f, 0.0f);              // Program.cpp:34 (fov 45 deg, focus 90, aperture 0)
    try {
        b200pt::Renderer renderer;
        renderer.Initialize(&window, st.scene_path, st.environment, 66, 0, PT_BVH_REFERENCE, n_gpus);   // Program.cpp:81
        camera.SetPosition(st.camera_position); camera.SetRotation(st.camera_rotation);   // :82-83
        camera.GenerateImagePlane();                                               // :117
        for (int i = 0; i < spp; i++) renderer.RenderFrame(camera);                // :122
        renderer.SaveScreenshot(out);                                              // :129
        std::printf("%u samples per pixel -> %s (BVH: %d nodes, %.1f ms)\n", renderer.GetNumSamples(), out,
                    renderer.BvhStats().n_nodes, renderer.BvhStats().build_ms);
    } catch (const std::exception& e) { std::fprintf(stderr, "%s\n", e.what()); return 1; }
    return 0;
}
