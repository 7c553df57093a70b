/*
 * mobilert_b200_renderer.hpp — C++ adapter with MobileRT::Renderer's public surface (MobileRT/Renderer.hpp:41-63) on
 * top of the C-ABI in mobilert_b200.h. Header-only; compiles inside the reference tree (it includes the reference's own
 * headers) and is what System_dependent/Native/C_wrapper.cpp:207-232 and Android_JNI/JNI_layer.cpp:691-694 would
 * construct instead of MobileRT::Renderer when Config::accelerator == ACC_B200.
 *
 * The reference keeps Sphere's fields, AreaLight::triangle_ and PointLight::position_ private without getters; the
 * adapter reads them directly, so this translation unit is compiled with -fno-access-control (as oracle/ref_shim is),
 * or the maintainer adds the four one-line getters named in INTEGRATION.md.
 */
#ifndef MOBILERT_B200_RENDERER_HPP
#define MOBILERT_B200_RENDERER_HPP

#include "mobilert_b200.h"

#include "Components/Cameras/Orthographic.hpp"
#include "Components/Cameras/Perspective.hpp"
#include "Components/Lights/AreaLight.hpp"
#include "Components/Lights/PointLight.hpp"
#include "MobileRT/Camera.hpp"
#include "MobileRT/Scene.hpp"

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstring>
#include <memory>
#include <mutex>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

namespace MobileRT {

class B200Renderer final {
    mrt_handle *handle_ {nullptr};

    static void put3(float *dst, const ::glm::vec3 &v) { dst[0] = v.x; dst[1] = v.y; dst[2] = v.z; }

    static mrt_triangle flatten(const Triangle &t) {
        mrt_triangle r;
        put3(r.AC, t.getAC()); put3(r.AB, t.getAB()); put3(r.A, t.getA());
        put3(r.nA, t.getNormalA()); put3(r.nB, t.getNormalB()); put3(r.nC, t.getNormalC());
        const ::glm::vec2 a {t.getTexCoordA()}, b {t.getTexCoordB()}, c {t.getTexCoordC()};
        r.tA[0] = a.x; r.tA[1] = a.y; r.tB[0] = b.x; r.tB[1] = b.y; r.tC[0] = c.x; r.tC[1] = c.y;
        r.material = t.getMaterialIndex();
        return r;
    }

    /* The camera object as the front-end holds it (C_wrapper.cpp:76-141 keeps the Camera, not its lookAt): its members go
     * over unchanged (MRT_CAMERA_*_STATE), so the device's generateRay sees the very numbers Camera::generateRay sees. */
    static mrt_camera_desc flatten(const Camera &camera) {
        mrt_camera_desc c;
        ::std::memset(&c, 0, sizeof(c));
        put3(c.position, camera.position_); put3(c.look_at, camera.direction_); put3(c.up, camera.up_); put3(c.right, camera.right_);
        if (const auto *p = dynamic_cast<const ::Components::Perspective *>(&camera)) {
            c.kind = MRT_CAMERA_PERSPECTIVE_STATE; c.a = p->hFov_; c.b = p->vFov_;   // the stored radians (getHFov() converts back to degrees)
        } else if (const auto *o = dynamic_cast<const ::Components::Orthographic *>(&camera)) {
            c.kind = MRT_CAMERA_ORTHOGRAPHIC_STATE; c.a = o->sizeH_; c.b = o->sizeV_;
        } else {
            throw ::std::runtime_error("B200Renderer: unknown camera type");
        }
        return c;
    }

public:
    /* Same inputs the front-ends hold right before they build Shader + Renderer (C_wrapper.cpp:76-203): the Scene
     * (before it is moved into the Shader), the camera, the Config fields. */
    B200Renderer(const Scene &scene, const Camera &camera,
                 ::std::int32_t width, ::std::int32_t height, ::std::int32_t samplesPixel, ::std::int32_t samplesLight,
                 ::std::int32_t shader,
                 // which static table the front-end's own samplers read: MRT_SAMPLER_AUTO = StaticHaltonSeq as in
                 // C_wrapper.cpp:109,145,176; the Android layer passes MRT_SAMPLER_STATIC_PCG (JNI_layer.cpp:601,616-618,636)
                 ::std::int32_t frontEndSampler = MRT_SAMPLER_AUTO,
                 ::std::uint32_t flags = 0u /* MRT_FLAG_*, e.g. MRT_FLAG_PROGRESSIVE for a UI that shows the bitmap while it renders */,
                 const ::glm::vec3 &depthMaxPoint = ::glm::vec3 {1.0F, 1.0F, 1.0F} /* DepthMap's maxPoint (C_wrapper.cpp:80-138) */,
                 ::std::int32_t accelerator = MRT_ACC_B200 /* or MRT_ACC_NAIVE / MRT_ACC_REGULAR_GRID for those accelerators' visibility rules */) {
        ::std::vector<mrt_triangle> tris; tris.reserve(scene.triangles_.size());
        for (const Triangle &t : scene.triangles_) { tris.push_back(flatten(t)); }
        ::std::vector<mrt_sphere> spheres;
        for (const Sphere &s : scene.spheres_) {
            mrt_sphere r; put3(r.center, s.center_); r.sq_radius = s.sqRadius_; r.material = s.materialIndex_; spheres.push_back(r);
        }
        ::std::vector<mrt_plane> planes;
        for (const Plane &p : scene.planes_) {
            mrt_plane r; put3(r.normal, p.getNormal()); put3(r.point, p.getPoint()); r.material = p.getMaterialIndex(); planes.push_back(r);
        }
        ::std::vector<mrt_material> mats;
        ::std::vector<mrt_texture> textures;          // Material::texture_ (Material.hpp:43), shared images listed once
        ::std::vector<int32_t> materialTexture;
        for (const Material &m : scene.materials_) {
            mrt_material r; put3(r.Le, m.Le_); put3(r.Kd, m.Kd_); put3(r.Ks, m.Ks_); put3(r.Kt, m.Kt_); r.ior = m.refractiveIndice_; mats.push_back(r);
            int32_t tex {-1};
            if (m.texture_.isValid()) {
                const uint8_t *pixels {m.texture_.image_};
                for (size_t k {0}; k < textures.size(); ++k) { if (textures[k].pixels == pixels) { tex = static_cast<int32_t>(k); } }
                if (tex < 0) {
                    tex = static_cast<int32_t>(textures.size());
                    textures.push_back(mrt_texture {pixels, m.texture_.width_, m.texture_.height_, m.texture_.channels_});
                }
            }
            materialTexture.push_back(tex);
        }
        ::std::vector<mrt_light> lights;
        for (const auto &l : scene.lights_) {
            mrt_light r;
            ::std::memset(&r, 0, sizeof(r));
            put3(r.Le, l->radiance_.Le_);
            if (const auto *a = dynamic_cast<const ::Components::AreaLight *>(l.get())) {
                r.kind = MRT_LIGHT_AREA; r.triangle = flatten(a->triangle_);
            } else if (const auto *p = dynamic_cast<const ::Components::PointLight *>(l.get())) {
                r.kind = MRT_LIGHT_POINT; put3(r.position, p->position_);
            } else {
                throw ::std::runtime_error("B200Renderer: unknown light type");
            }
            lights.push_back(r);
        }
        mrt_scene_desc sd;
        ::std::memset(&sd, 0, sizeof(sd));
        sd.textures = textures.data(); sd.n_textures = static_cast<int64_t>(textures.size()); sd.material_texture = materialTexture.data();
        sd.triangles = tris.data(); sd.n_triangles = static_cast<int64_t>(tris.size());
        sd.spheres = spheres.data(); sd.n_spheres = static_cast<int64_t>(spheres.size());
        sd.planes = planes.data(); sd.n_planes = static_cast<int64_t>(planes.size());
        sd.materials = mats.data(); sd.n_materials = static_cast<int64_t>(mats.size());
        sd.lights = lights.data(); sd.n_lights = static_cast<int64_t>(lights.size());
        const mrt_camera_desc cd = flatten(camera);
        mrt_render_desc rd;
        ::std::memset(&rd, 0, sizeof(rd));
        rd.width = width; rd.height = height; rd.samples_pixel = samplesPixel; rd.samples_light = samplesLight;
        rd.shader = shader; rd.accelerator = accelerator; rd.pixel_sampler = MRT_SAMPLER_AUTO; rd.seed = 1u;
        put3(rd.depth_max_point, depthMaxPoint);
        rd.device = -1; rd.flags = flags;
        if (frontEndSampler == MRT_SAMPLER_STATIC_PCG) {
            rd.light_sampler = MRT_SAMPLER_STATIC_PCG;
            rd.pixel_sampler = samplesPixel > 1 ? MRT_SAMPLER_STATIC_PCG : MRT_SAMPLER_CONSTANT;
        }
        if (mrt_create(&sd, &cd, &rd, nullptr, &handle_) != MRT_OK) {
            // the front-ends catch ::std::exception around construction (C_wrapper.cpp:257-263, JNI_layer.cpp:706-715)
            throw ::std::runtime_error(::std::string("B200Renderer: ") + mrt_last_error(nullptr));
        }
    }

    B200Renderer(const B200Renderer &) = delete;
    B200Renderer &operator=(const B200Renderer &) = delete;
    ~B200Renderer() { mrt_destroy(handle_); }

    /* Renderer::renderFrame (Renderer.cpp:53-88). numThreads has no meaning on the GPU. */
    void renderFrame(::std::int32_t *bitmap, ::std::int32_t /*numThreads*/) {
        const int rc = mrt_render_frame(handle_, bitmap);
        if (rc != MRT_OK && rc != MRT_ERR_STOPPED) { throw ::std::runtime_error(::std::string("B200Renderer: ") + mrt_last_error(handle_)); }
    }
    /* Renderer::stopRender (Renderer.cpp:93-99) */
    void stopRender() { mrt_stop(handle_); }
    /* Renderer::getSample (Renderer.cpp:176-178) */
    ::std::int32_t getSample() const { ::std::int32_t s = 0; mrt_get_sample(handle_, &s); return s; }
    /* Renderer::getTotalCastedRays (Renderer.cpp:204-207) */
    ::std::uint64_t getTotalCastedRays() const { ::std::uint64_t r = 0; mrt_get_total_rays(handle_, &r); return r; }
};

/*
 * The render thread and state machine the Android front-end wraps around the blocking renderFrame
 * (System_dependent/Android_JNI/JNI_layer.cpp: State IDLE / BUSY / FINISHED / STOPPED :44 and JNI_layer.hpp:12-14,
 * rtStartRender :406-426, rtStopRender :428-462, the worker lambda of rtRenderIntoBitmap :744-901, rtGetState :904-913),
 * without the JNI calls, over a B200Renderer: startRender() arms the state, renderIntoBitmap() runs renderFrame on a worker
 * thread while the UI thread polls getSample() / getState() and reads the bitmap (create the renderer with
 * MRT_FLAG_PROGRESSIVE to see it build up), stopRender(wait) is rtStopRender.
 */
class B200RenderThread final {
public:
    enum class State { IDLE = 0, BUSY = 1, FINISHED = 2, STOPPED = 3 };   // JNI_layer.hpp:12-14

private:
    B200Renderer &renderer_;
    ::std::atomic<State> state_ {State::IDLE};
    ::std::atomic<bool> finishedRendering_ {true};
    ::std::mutex mutex_ {};
    ::std::condition_variable rendered_ {};
    ::std::unique_ptr<::std::thread> thread_ {};
    ::std::string error_ {};
    double renderingSeconds_ {0.0};

public:
    explicit B200RenderThread(B200Renderer &renderer) : renderer_ {renderer} {}
    B200RenderThread(const B200RenderThread &) = delete;
    B200RenderThread &operator=(const B200RenderThread &) = delete;
    ~B200RenderThread() { stopRender(true); finishRender(); }

    /* rtStartRender (JNI_layer.cpp:406-426) */
    void startRender(bool wait) {
        if (wait) {
            ::std::unique_lock<::std::mutex> lock {mutex_};
            rendered_.wait(lock, [&] { return finishedRendering_.load(); });
        }
        finishedRendering_ = false;
        state_ = State::BUSY;
    }

    /* rtRenderIntoBitmap (JNI_layer.cpp:744-901): `repeats` frames into the caller's bitmap on a worker thread (async) or here */
    void renderIntoBitmap(::std::int32_t *bitmap, ::std::int32_t nThreads, bool async, ::std::int32_t repeats = 1) {
        finishRender();
        auto work = [this, bitmap, nThreads, repeats]() {
            ::std::int32_t rep {repeats};
            const auto t0 = ::std::chrono::steady_clock::now();
            try {
                while (state_ == State::BUSY && rep > 0) { renderer_.renderFrame(bitmap, nThreads); --rep; }
            } catch (const ::std::exception &e) {
                error_ = e.what();
            }
            renderingSeconds_ = ::std::chrono::duration<double>(::std::chrono::steady_clock::now() - t0).count();
            finishedRendering_ = true;
            rendered_.notify_all();
            {
                const ::std::lock_guard<::std::mutex> lock {mutex_};
                if (state_ != State::STOPPED) { state_ = State::FINISHED; }
            }
            state_ = State::IDLE;
        };
        if (async) { thread_.reset(new ::std::thread(work)); } else { work(); }
    }

    /* rtStopRender (JNI_layer.cpp:428-462) */
    void stopRender(bool wait) {
        state_ = State::STOPPED;
        ::std::unique_lock<::std::mutex> lock {mutex_};
        renderer_.stopRender();
        if (wait) {
            while (!finishedRendering_) {
                renderer_.stopRender();
                rendered_.wait_for(lock, ::std::chrono::milliseconds(50), [&] { return finishedRendering_.load(); });
            }
        }
    }

    /* rtFinishRender: join the worker */
    void finishRender() { if (thread_ && thread_->joinable()) { thread_->join(); } thread_.reset(); }
    /* rtGetState (JNI_layer.cpp:904-913) */
    State getState() const { return state_.load(); }
    ::std::int32_t getSample() const { return renderer_.getSample(); }
    double getRenderingSeconds() const { return renderingSeconds_; }
    const ::std::string &getError() const { return error_; }
};

}  // namespace MobileRT

#endif  // MOBILERT_B200_RENDERER_HPP
