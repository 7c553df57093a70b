// Runs the C++ drop-in adapter (include/mobilert_b200_renderer.hpp) the way the reference's Linux entry would
// (System_dependent/Native/C_wrapper.cpp:76-232): the reference's OWN Scene / Camera objects (Scenes.cpp built-in
// scenes, or OBJLoader + CameraFactory for an OBJ scene) are flattened by MobileRT::B200Renderer and rendered on the
// GPU through the C-ABI; the bitmap is written to a file for tests/test_gpu_round2.py to compare with the frames the
// unmodified reference rendered (tests/golden/reference_frames.npz).
//
// Built only where /root/reference exists (oracle/Makefile, target adapter), against the reference's headers and the
// compiled reference library oracle/_ref/libmobilert_ref.so (for Scenes.cpp / loaders), into oracle/_ref/.
//   b200_adapter_demo <scene 0..3 | obj mtl cam> <shader> <width> <height> <spp> <spl> <out.raw>
#include "mobilert_b200_renderer.hpp"

#include "Components/Loaders/CameraFactory.hpp"
#include "Components/Loaders/OBJLoader.hpp"
#include "Components/Samplers/StaticHaltonSeq.hpp"
#include "Scenes/Scenes.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <unordered_map>

using namespace MobileRT;

int main(int argc, char **argv) {
    if (argc != 8 && argc != 10) {
        ::std::fprintf(stderr, "usage: %s <scene 0..3 | obj mtl cam> <shader> <width> <height> <spp> <spl> <out.raw>\n", argv[0]);
        return 2;
    }
    const bool objScene {argc == 10};
    int a {objScene ? 4 : 2};
    const int shader {::std::atoi(argv[a])}, width {::std::atoi(argv[a + 1])}, height {::std::atoi(argv[a + 2])};
    const int spp {::std::atoi(argv[a + 3])}, spl {::std::atoi(argv[a + 4])};
    const char *outPath {argv[a + 5]};
    try {
        const float ratio {static_cast<float>(width) / static_cast<float>(height)};
        Scene scene {};
        ::std::unique_ptr<Camera> camera {};
        ::glm::vec3 maxDist {1.0F, 1.0F, 1.0F};
        if (objScene) {
            ::std::ifstream ifObj {argv[1]}, ifMtl {argv[2]}, ifCam {argv[3]};
            ::Components::OBJLoader objLoader {ifObj, ifMtl};
            if (!objLoader.isProcessed()) { ::std::fprintf(stderr, "OBJ not processed\n"); return 3; }
            ::std::unordered_map<::std::string, Texture> texturesCache {};
            if (!objLoader.fillScene(&scene, []() { return ::MobileRT::std::make_unique<::Components::StaticHaltonSeq>(); }, argv[1], texturesCache)) { return 3; }
            ::std::istream iCam {ifCam.rdbuf()};
            camera = ::Components::CameraFactory {}.loadFromFile(iCam, ratio);
        } else {
            switch (::std::atoi(argv[1])) {                        // C_wrapper.cpp:80-100
                case 0: scene = cornellBox_Scene(::std::move(scene)); camera = cornellBox_Cam(ratio); break;
                case 1: scene = spheres_Scene(::std::move(scene)); camera = spheres_Cam(ratio); maxDist = ::glm::vec3 {8, 8, 8}; break;
                case 2: scene = cornellBox2_Scene(::std::move(scene)); camera = cornellBox_Cam(ratio); break;
                default: scene = spheres2_Scene(::std::move(scene)); camera = spheres2_Cam(ratio); maxDist = ::glm::vec3 {8, 8, 8}; break;
            }
        }
        errno = 0;
        B200Renderer renderer {scene, *camera, width, height, spp, spl, shader, MRT_SAMPLER_AUTO, 0u, maxDist};
        ::std::vector<::std::int32_t> bitmap(static_cast<size_t>(width) * static_cast<size_t>(height), 0);
        renderer.renderFrame(bitmap.data(), 1);
        if (errno != 0) { ::std::fprintf(stderr, "errno left at %d\n", errno); return 4; }   // the front-ends poll errno right after the frame (C_wrapper.cpp:234)
        ::std::FILE *f {::std::fopen(outPath, "wb")};
        if (f == nullptr) { return 5; }
        ::std::fwrite(bitmap.data(), sizeof(::std::int32_t), bitmap.size(), f);
        ::std::fclose(f);
        ::std::printf("rays %llu sample %d\n", static_cast<unsigned long long>(renderer.getTotalCastedRays()), renderer.getSample());
    } catch (const ::std::exception &e) {
        ::std::fprintf(stderr, "exception: %s\n", e.what());
        return 1;
    }
    return 0;
}
