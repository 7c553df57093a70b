// Runs the C++ drop-in adapter (include/mobilert_b200_renderer.hpp) the way the reference's Linux entry would
// (System_dependent/Native/C_wrapper.cpp:76-232): the reference's OWN Scene / Camera objects (Scenes.cpp built-in
// scenes, or OBJLoader + CameraFactory for an OBJ scene) are flattened by MobileRT::B200Renderer and rendered on the
// GPU through the C-ABI; the bitmap is written to a file for tests/test_gpu_round2.py to compare with the frames the
// unmodified reference rendered (tests/golden/reference_frames.npz).
//
// Built only where /root/reference exists (oracle/Makefile, target adapter), against the reference's headers and the
// compiled reference library oracle/_ref/libmobilert_ref.so (for Scenes.cpp / loaders), into oracle/_ref/.
//   b200_adapter_demo <scene 0..3 | obj mtl cam> <shader> <width> <height> <spp> <spl> <out.raw>
// With MRT_DEMO_ASYNC=1 the frame is rendered by MobileRT::B200RenderThread (the render thread + State machine of
// Android_JNI/JNI_layer.cpp:406-462,744-913) with MRT_FLAG_PROGRESSIVE while this thread polls getSample() and the bitmap like
// the UI does; MRT_DEMO_STOP_AT=<sample> calls stopRender(true) once that sample has been reached.
#include "mobilert_b200_renderer.hpp"

#include "Components/Loaders/CameraFactory.hpp"
#include "Components/Loaders/OBJLoader.hpp"
#include "Components/Samplers/StaticHaltonSeq.hpp"
#include "Scenes/Scenes.hpp"

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <thread>
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
        const char *asyncEnv {::std::getenv("MRT_DEMO_ASYNC")}, *stopEnv {::std::getenv("MRT_DEMO_STOP_AT")};
        const bool async {asyncEnv != nullptr && ::std::atoi(asyncEnv) != 0};
        B200Renderer renderer {scene, *camera, width, height, spp, spl, shader, MRT_SAMPLER_AUTO, async ? static_cast<::std::uint32_t>(MRT_FLAG_PROGRESSIVE) : 0u, maxDist};
        ::std::vector<::std::int32_t> bitmap(static_cast<size_t>(width) * static_cast<size_t>(height), 0);
        if (async) {
            B200RenderThread worker {renderer};
            worker.startRender(false);
            worker.renderIntoBitmap(bitmap.data(), 1, true);
            const int stopAt {stopEnv != nullptr ? ::std::atoi(stopEnv) : -1};
            int partialFrames {0}, polls {0}, busySeen {0};
            bool stopped {false};
            while (worker.getState() == B200RenderThread::State::BUSY || worker.getState() == B200RenderThread::State::STOPPED) {
                const int sample {worker.getSample()};
                ++polls;
                if (worker.getState() == B200RenderThread::State::BUSY) { ++busySeen; }
                size_t filled {0};
                for (size_t i {0}; i < bitmap.size(); i += 97) { filled += bitmap[i] != 0 ? 1u : 0u; }
                if (sample > 0 && sample < spp && filled * 97 * 2 > bitmap.size()) { ++partialFrames; }
                if (stopAt >= 0 && sample >= stopAt && !stopped) { worker.stopRender(true); stopped = true; }
                ::std::this_thread::sleep_for(::std::chrono::milliseconds(1));
            }
            worker.finishRender();
            if (!worker.getError().empty()) { ::std::fprintf(stderr, "worker: %s\n", worker.getError().c_str()); return 6; }
            ::std::printf("async polls %d busy %d partial %d stopped %d final_sample %d state %d secs %.4f\n", polls, busySeen, partialFrames, stopped ? 1 : 0,
                          renderer.getSample(), static_cast<int>(worker.getState()), worker.getRenderingSeconds());
        } else {
            renderer.renderFrame(bitmap.data(), 1);
        }
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
