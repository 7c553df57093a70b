// TEST INFRASTRUCTURE (oracle) — NOT product code, never linked into libmobilert_b200.so.
//
// Flat extern "C" driver around the UNMODIFIED reference engine (compiled from the sources where
// they lie under /root/reference/app into oracle/_ref/libmobilert_ref.so, see oracle/Makefile).
// It assembles Scene -> Shader -> Renderer exactly in the order the reference's own Linux entry
// does (app/System_dependent/Native/C_wrapper.cpp:76-203) but keeps the Renderer alive so that
// the parity tests can read the post-build primitive arrays, the camera, the shuffled sampler
// tables and per-ray ground truth.  Compiled with -fno-access-control so the
// private members of Shader/Perspective/AreaLight can be read without touching the sources.
#include "Components/Cameras/Orthographic.hpp"
#include "Components/Cameras/Perspective.hpp"
#include "Components/Lights/AreaLight.hpp"
#include "Components/Lights/PointLight.hpp"
#include "Components/Loaders/CameraFactory.hpp"
#include "Components/Loaders/OBJLoader.hpp"
#include "Components/Samplers/Constant.hpp"
#include "Components/Samplers/StaticHaltonSeq.hpp"
#include "Components/Samplers/StaticPCG.hpp"
#include "Components/Samplers/StaticMersenneTwister.hpp"
#include "Components/Samplers/HaltonSeq.hpp"
#include "Components/Samplers/Stratified.hpp"
#include "Components/Samplers/MersenneTwister.hpp"
#include "Components/Samplers/PCG.hpp"
#include "Components/Shaders/DepthMap.hpp"
#include "Components/Shaders/DiffuseMaterial.hpp"
#include "Components/Shaders/NoShadows.hpp"
#include "Components/Shaders/PathTracer.hpp"
#include "Components/Shaders/Whitted.hpp"
#include "MobileRT/Config.hpp"
#include "MobileRT/Renderer.hpp"
#include "MobileRT/Scene.hpp"
#include "Scenes/Scenes.hpp"
#include "C_wrapper.h"
#include <algorithm>
#include <chrono>
#include <cstring>
#include <fstream>
#include <iostream>
#include <sstream>

using namespace MobileRT;

struct RefConfig {
    int width, height, threads, shader, scene, spp, spl, accelerator;
    const char *obj, *mtl, *cam;
    int samplers;   // 0: the Linux front-end's choice (StaticHaltonSeq, C_wrapper.cpp:144-148,176); 1: Android's (StaticPCG, JNI_layer.cpp:601,616-618,636)
    int copies_x, copies_z;   // OBJ scenes: the loaded triangles are copied on a copies_x x copies_z grid (BASELINE config 5's synthetic scene,
                              // same offsets and order as mobileraytracer_b200.loaders.replicate_scene); 0 / 1 = as loaded. Lights are not copied.
};

struct RefHandle {
    ::std::unique_ptr<Renderer> renderer;
    double secsLoad {0}, secsFill {0}, secsCreate {0};
    int cameraKind {0}; // 0 perspective, 1 orthographic
};

static_assert(sizeof(Triangle) == 100, "reference Triangle layout changed");
static_assert(sizeof(Sphere) == 20, "reference Sphere layout changed");
static_assert(sizeof(Plane) == 28, "reference Plane layout changed");

namespace {
// Discards the reference's log lines while a call runs. The sink has no state: the reference's render threads log
// concurrently (LOG_ERROR in checkSystemError when a worker finds errno set), and a shared stringbuf under cout / cerr
// is not thread-safe (it corrupted the heap in about one of four multi-threaded renders).
struct NullBuf final : ::std::streambuf {
    int_type overflow(int_type c) override { return traits_type::not_eof(c); }
    ::std::streamsize xsputn(const char *, ::std::streamsize n) override { return n; }
};
struct Quiet {
    static NullBuf sink;
    ::std::streambuf *o, *e;
    Quiet() : o(::std::cout.rdbuf(&sink)), e(::std::cerr.rdbuf(&sink)) {}
    ~Quiet() { ::std::cout.rdbuf(o); ::std::cerr.rdbuf(e); }
};
NullBuf Quiet::sink;
double now() {
    return ::std::chrono::duration<double>(::std::chrono::steady_clock::now().time_since_epoch()).count();
}
// see ref_grid_cells
template <typename T>
long long gridCells(RegularGrid<T> &g, float *params, unsigned *start, int *items, long long cap) {
    const ::glm::vec3 mn {g.worldBoundaries_.getPointMin()}, mx {g.worldBoundaries_.getPointMax()};
    for (int k = 0; k < 3; ++k) { params[k] = mn[k]; params[3 + k] = mx[k]; params[6 + k] = g.cellSizeInverted_[k]; params[9 + k] = g.cellSize_[k]; }
    long long total {0};
    const T *base {g.primitives_.data()};
    for (size_t c = 0; c < g.grid_.size(); ++c) {
        start[c] = static_cast<unsigned>(total);
        ::std::vector<int> ids {};
        for (const T *prim : g.grid_[c]) { ids.push_back(static_cast<int>(prim - base)); }
        ::std::sort(ids.begin(), ids.end());
        for (const int id : ids) { if (items != nullptr && total < cap) { items[total] = id; } ++total; }
    }
    start[g.grid_.size()] = static_cast<unsigned>(total);
    return total;
}
}

extern "C" {

// Whole-frame path through the reference's own public entry (C_wrapper.h:15). Returns wall seconds.
double ref_raytrace(const RefConfig *c, int32_t *bitmap, int repeats, int verbose) {
    Config config {};
    config.bitmap.assign(static_cast<size_t>(c->width) * c->height, 0);
    config.objFilePath = c->obj ? c->obj : "";
    config.mtlFilePath = c->mtl ? c->mtl : "";
    config.camFilePath = c->cam ? c->cam : "";
    config.width = c->width; config.height = c->height; config.threads = c->threads;
    config.shader = c->shader; config.sceneIndex = c->scene; config.samplesPixel = c->spp;
    config.samplesLight = c->spl; config.repeats = repeats; config.accelerator = c->accelerator;
    config.printStdOut = verbose != 0;
    const double t0 = now();
    RayTrace(config, false);
    const double t1 = now();
    ::std::memcpy(bitmap, config.bitmap.data(), config.bitmap.size() * sizeof(int32_t));
    errno = 0;
    return t1 - t0;
}

void *ref_create(const RefConfig *c) {
    Quiet q;
    errno = 0;
    auto *h = new RefHandle {};
    try {
        const float ratio {static_cast<float>(c->width) / c->height};
        Scene scene {};
        ::std::unique_ptr<Sampler> samplerPixel {};
        ::std::unique_ptr<Shader> shader {};
        ::std::unique_ptr<Camera> camera {};
        ::glm::vec3 maxDist {};
        switch (c->scene) {
            case 0: scene = cornellBox_Scene(::std::move(scene)); camera = cornellBox_Cam(ratio); maxDist = ::glm::vec3 {1, 1, 1}; break;
            case 1: scene = spheres_Scene(::std::move(scene)); camera = spheres_Cam(ratio); maxDist = ::glm::vec3 {8, 8, 8}; h->cameraKind = 1; break;
            case 2: scene = cornellBox2_Scene(::std::move(scene)); camera = cornellBox_Cam(ratio); maxDist = ::glm::vec3 {1, 1, 1}; break;
            case 3: scene = spheres2_Scene(::std::move(scene)); camera = spheres2_Cam(ratio); maxDist = ::glm::vec3 {8, 8, 8}; break;
            default: {
                const double t0 = now();
                ::std::ifstream ifObj {c->obj};
                ::std::ifstream ifMtl {c->mtl};
                ::Components::OBJLoader objLoader {ifObj, ifMtl};
                if (!objLoader.isProcessed()) { delete h; return nullptr; }
                const double t1 = now();
                ::std::unordered_map<::std::string, Texture> texturesCache {};
                const bool pcg {c->samplers == 1};
                if (!objLoader.fillScene(&scene,
                        [pcg]() -> ::std::unique_ptr<Sampler> {
                            if (pcg) { return ::MobileRT::std::make_unique<::Components::StaticPCG>(); }
                            return ::MobileRT::std::make_unique<::Components::StaticHaltonSeq>(); },
                        c->obj, texturesCache)) { delete h; return nullptr; }
                if (c->copies_x > 1 || c->copies_z > 1) {
                    const int nx {c->copies_x > 1 ? c->copies_x : 1}, nz {c->copies_z > 1 ? c->copies_z : 1};
                    const ::std::vector<Triangle> base {scene.triangles_};
                    scene.triangles_.clear();
                    scene.triangles_.reserve(base.size() * static_cast<size_t>(nx * nz));
                    for (int ix {0}; ix < nx; ++ix) {
                        for (int iz {0}; iz < nz; ++iz) {
                            const ::glm::vec3 off {static_cast<float>(ix - nx / 3) * 4200.0F, 0.0F, static_cast<float>(iz - nz / 2) * 5200.0F};
                            for (Triangle t : base) { t.pointA_ += off; scene.triangles_.push_back(t); }
                        }
                    }
                }
                const double t2 = now();
                h->secsLoad = t1 - t0; h->secsFill = t2 - t1;
                ::Components::CameraFactory cameraFactory {};
                ::std::ifstream ifCamera {c->cam};
                ::std::istream iCam {ifCamera.rdbuf()};
                camera = cameraFactory.loadFromFile(iCam, ratio);
                maxDist = ::glm::vec3 {1, 1, 1};
                errno = 0;
            } break;
        }
        if (dynamic_cast<::Components::Orthographic *>(camera.get()) != nullptr) { h->cameraKind = 1; }
        if (c->spp > 1 && c->samplers == 1) {
            samplerPixel = ::MobileRT::std::make_unique<::Components::StaticPCG>();
        } else if (c->spp > 1 && c->samplers == 2) {   // pixel jitter from StaticMersenneTwister (StaticMersenneTwister.cpp:6-16); lights / RR as the Linux front-end
            samplerPixel = ::MobileRT::std::make_unique<::Components::StaticMersenneTwister>();
        } else if (c->spp > 1 && c->samplers == 3) {   // HaltonSeq (HaltonSeq.cpp:10-19): returns 1.0 once its domain is exhausted
            samplerPixel = ::MobileRT::std::make_unique<::Components::HaltonSeq>(static_cast<::std::uint32_t>(c->width), static_cast<::std::uint32_t>(c->height), static_cast<::std::uint32_t>(c->spp));
        } else if (c->spp > 1 && c->samplers == 4) {   // Stratified (Stratified.cpp:10-19)
            samplerPixel = ::MobileRT::std::make_unique<::Components::Stratified>(static_cast<::std::uint32_t>(c->width), static_cast<::std::uint32_t>(c->height), static_cast<::std::uint32_t>(c->spp));
        } else if (c->spp > 1 && c->samplers == 5) {   // MersenneTwister (MersenneTwister.cpp:5-11): thread_local generator, never reset
            samplerPixel = ::MobileRT::std::make_unique<::Components::MersenneTwister>();
        } else if (c->spp > 1 && c->samplers == 6) {   // PCG (PCG.cpp:6-13)
            samplerPixel = ::MobileRT::std::make_unique<::Components::PCG>();
        } else if (c->spp > 1) {
            samplerPixel = ::MobileRT::std::make_unique<::Components::StaticHaltonSeq>();
        } else {
            samplerPixel = ::MobileRT::std::make_unique<::Components::Constant>(0.5F);
        }
        const double t3 = now();
        const auto acc = Shader::Accelerator(c->accelerator);
        switch (c->shader) {
            case 1: shader = ::MobileRT::std::make_unique<::Components::Whitted>(::std::move(scene), c->spl, acc); break;
            case 2: {
                ::std::unique_ptr<Sampler> rr {};
                if (c->samplers == 1) { rr = ::MobileRT::std::make_unique<::Components::StaticPCG>(); }
                else { rr = ::MobileRT::std::make_unique<::Components::StaticHaltonSeq>(); }
                shader = ::MobileRT::std::make_unique<::Components::PathTracer>(::std::move(scene), ::std::move(rr), c->spl, acc);
            } break;
            case 3: shader = ::MobileRT::std::make_unique<::Components::DepthMap>(::std::move(scene), maxDist, acc); break;
            case 4: shader = ::MobileRT::std::make_unique<::Components::DiffuseMaterial>(::std::move(scene), acc); break;
            default: shader = ::MobileRT::std::make_unique<::Components::NoShadows>(::std::move(scene), c->spl, acc); break;
        }
        h->secsCreate = now() - t3;
        h->renderer = ::MobileRT::std::make_unique<Renderer>(::std::move(shader), ::std::move(camera),
                                                              ::std::move(samplerPixel), c->width, c->height, c->spp);
        errno = 0;
    } catch (...) {
        delete h;
        errno = 0;
        return nullptr;
    }
    return h;
}

void ref_destroy(void *p) { delete static_cast<RefHandle *>(p); }

// The next n draws of the renderer's pixel sampler (Renderer.cpp:137-138 calls getSample() twice per pixel-sample), for
// the stateful samplers whose sequence only the sampler object itself can produce (MersenneTwister, PCG).
void ref_pixel_sampler_draws(void *p, int n, float *out) {
    auto *h = static_cast<RefHandle *>(p);
    for (int i = 0; i < n; ++i) { out[i] = h->renderer->samplerPixel_->getSample(); }
    errno = 0;
}
// The seeded std::random_device stand-in counts its draws; setting the counter back makes the next generators and table
// shuffles repeat (used to give two threads' thread_local MersenneTwister / PCG generators the same seed).
unsigned ref_random_device_counter(int set, unsigned value) {
    unsigned &c = ::std::mrt_seeded_random_device::counter();
    const unsigned old = c;
    if (set) { c = value; }
    return old;
}
void ref_pixel_sampler_reset(void *p) { static_cast<RefHandle *>(p)->renderer->samplerPixel_->resetSampling(); errno = 0; }

// renderFrame (app/MobileRT/Renderer.cpp:53) x repeats; returns wall seconds of the loop.
double ref_render(void *p, int32_t *bitmap, int threads, int repeats) {
    auto *h = static_cast<RefHandle *>(p);
    Quiet q;
    errno = 0;
    const double t0 = now();
    try {
        do { h->renderer->renderFrame(bitmap, threads); } while (--repeats > 0);
    } catch (...) {
        errno = 0;
        return -1.0;
    }
    const double t1 = now();
    errno = 0;
    return t1 - t0;
}

uint64_t ref_casted_rays(void *p) { return static_cast<RefHandle *>(p)->renderer->getTotalCastedRays(); }

void ref_times(void *p, double out[3]) {
    auto *h = static_cast<RefHandle *>(p);
    out[0] = h->secsLoad; out[1] = h->secsFill; out[2] = h->secsCreate;
}

// counts: triangles, spheres, planes, materials, lights
void ref_counts(void *p, int out[5]) {
    Shader &s = *static_cast<RefHandle *>(p)->renderer->shader_;
    out[0] = static_cast<int>(s.getTriangles().size());
    out[1] = static_cast<int>(s.getSpheres().size());
    out[2] = static_cast<int>(s.getPlanes().size());
    out[3] = static_cast<int>(s.getMaterials().size());
    out[4] = static_cast<int>(s.getLights().size());
}

// Post-build (BVH-ordered) primitive arrays in the reference's own memory layout
// (Triangle.hpp:18-27 100 B, Sphere 20 B, Plane 28 B).
void ref_triangles(void *p, void *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getTriangles();
    if (!v.empty()) ::std::memcpy(out, static_cast<const void *>(v.data()), v.size() * sizeof(Triangle));
}
void ref_spheres(void *p, void *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getSpheres();
    if (!v.empty()) ::std::memcpy(out, static_cast<const void *>(v.data()), v.size() * sizeof(Sphere));
}
void ref_planes(void *p, void *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getPlanes();
    if (!v.empty()) ::std::memcpy(out, static_cast<const void *>(v.data()), v.size() * sizeof(Plane));
}
// 13 floats per material: Le, Kd, Ks, Kt, ior (Material.hpp:18-38)
void ref_materials(void *p, float *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getMaterials();
    for (size_t i = 0; i < v.size(); ++i) {
        float *o = out + 13 * i;
        for (int k = 0; k < 3; ++k) { o[k] = v[i].Le_[k]; o[3 + k] = v[i].Kd_[k]; o[6 + k] = v[i].Ks_[k]; o[9 + k] = v[i].Kt_[k]; }
        o[12] = v[i].refractiveIndice_;
    }
}
// 16 floats per light: kind(0 point,1 area), A|pos(3), AB(3), AC(3), Le(3), pad(3)
void ref_lights(void *p, float *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getLights();
    for (size_t i = 0; i < v.size(); ++i) {
        float *o = out + 16 * i;
        ::std::memset(o, 0, 16 * sizeof(float));
        if (auto *al = dynamic_cast<::Components::AreaLight *>(v[i].get())) {
            o[0] = 1.0F;
            const ::glm::vec3 A = al->triangle_.getA(), AB = al->triangle_.getAB(), AC = al->triangle_.getAC();
            for (int k = 0; k < 3; ++k) { o[1 + k] = A[k]; o[4 + k] = AB[k]; o[7 + k] = AC[k]; }
        } else if (auto *pl = dynamic_cast<::Components::PointLight *>(v[i].get())) {
            for (int k = 0; k < 3; ++k) { o[1 + k] = pl->position_[k]; }
        }
        for (int k = 0; k < 3; ++k) { o[10 + k] = v[i]->radiance_.Le_[k]; }
    }
}
// Full 100-byte triangle of an area light (normals included), for C-ABI input.
void ref_light_triangle(void *p, int i, void *out) {
    const auto &v = static_cast<RefHandle *>(p)->renderer->shader_->getLights();
    if (auto *al = dynamic_cast<::Components::AreaLight *>(v[static_cast<size_t>(i)].get())) {
        ::std::memcpy(out, static_cast<const void *>(&al->triangle_), sizeof(Triangle));
    }
}
// camera: kind, position(3), direction(3), right(3), up(3), a, b  (a,b = hFov,vFov in radians | sizeH/2,sizeV/2)
void ref_camera(void *p, float *out) {
    auto *h = static_cast<RefHandle *>(p);
    Camera &c = *h->renderer->camera_;
    out[0] = static_cast<float>(h->cameraKind);
    for (int k = 0; k < 3; ++k) { out[1 + k] = c.position_[k]; out[4 + k] = c.direction_[k]; out[7 + k] = c.right_[k]; out[10 + k] = c.up_[k]; }
    if (auto *pc = dynamic_cast<::Components::Perspective *>(&c)) { out[13] = pc->hFov_; out[14] = pc->vFov_; }
    if (auto *oc = dynamic_cast<::Components::Orthographic *>(&c)) { out[13] = oc->sizeH_; out[14] = oc->sizeV_; }
}
// DepthMap's maxPoint (C_wrapper.cpp:80-138)
void ref_depth_max_point(void *p, float *out) {
    if (auto *d = dynamic_cast<::Components::DepthMap *>(static_cast<RefHandle *>(p)->renderer->shader_.get())) {
        for (int k = 0; k < 3; ++k) out[k] = d->maxPoint_[k];
    }
}

// Camera::generateRay for n (u,v,du,dv) tuples.
void ref_generate_rays(void *p, int n, const float *uvdd, float *o, float *d) {
    Camera &c = *static_cast<RefHandle *>(p)->renderer->camera_;
    for (int i = 0; i < n; ++i) {
        Ray r {c.generateRay(uvdd[4 * i], uvdd[4 * i + 1], uvdd[4 * i + 2], uvdd[4 * i + 3])};
        for (int k = 0; k < 3; ++k) { o[3 * i + k] = r.origin_[k]; d[3 * i + k] = r.direction_[k]; }
    }
}

// Per-ray ground truth, restating Shader::rayTrace's geometric part (Shader.cpp:86-111) /
// Shader::shadowTrace (Shader.cpp:132-158) on the reference's own BVH<T> objects.
// origin prim: kind 0 none, 1 plane, 3 triangle (spheres never self-skip, Sphere.cpp:77).
// out kind: 0 miss, 1 plane, 2 sphere, 3 triangle, 4 light.  shadowDist < 0 => closest hit.
void ref_trace_rays(void *p, int n, const float *o, const float *d, const int *originKind, const int *originIdx,
                    const float *shadowDist, float *outT, int *outKind, int *outIdx, int *outMat, float *outNormal) {
    Shader &s = *static_cast<RefHandle *>(p)->renderer->shader_;
    const auto &tris = s.getTriangles();
    const auto &sphs = s.getSpheres();
    const auto &plas = s.getPlanes();
    for (int i = 0; i < n; ++i) {
        const void *prim {nullptr};
        if (originKind != nullptr) {
            if (originKind[i] == 1) prim = &plas[static_cast<size_t>(originIdx[i])];
            if (originKind[i] == 3) prim = &tris[static_cast<size_t>(originIdx[i])];
        }
        const bool shadow {shadowDist != nullptr && shadowDist[i] >= 0.0F};
        Ray ray {::glm::vec3 {d[3 * i], d[3 * i + 1], d[3 * i + 2]}, ::glm::vec3 {o[3 * i], o[3 * i + 1], o[3 * i + 2]}, 1, shadow, prim};
        Intersection isect {::std::move(ray), shadow ? shadowDist[i] : RayLengthMax};
        const float start {isect.length_};
        int kind {0}, idx {-1};
        // the accelerator the Shader was built with (Shader.cpp:89-110,134-155); shadowTrace == trace for the grid (RegularGrid.hpp:308-311)
        const auto tracePlanes = [&](Intersection in) {
            switch (s.accelerator_) {
                case Shader::ACC_NAIVE: return shadow ? s.naivePlanes_.shadowTrace(in) : s.naivePlanes_.trace(in);
                case Shader::ACC_REGULAR_GRID: return shadow ? s.gridPlanes_.shadowTrace(in) : s.gridPlanes_.trace(in);
                default: return shadow ? s.bvhPlanes_.shadowTrace(in) : s.bvhPlanes_.trace(in);
            }
        };
        const auto traceSpheres = [&](Intersection in) {
            switch (s.accelerator_) {
                case Shader::ACC_NAIVE: return shadow ? s.naiveSpheres_.shadowTrace(in) : s.naiveSpheres_.trace(in);
                case Shader::ACC_REGULAR_GRID: return shadow ? s.gridSpheres_.shadowTrace(in) : s.gridSpheres_.trace(in);
                default: return shadow ? s.bvhSpheres_.shadowTrace(in) : s.bvhSpheres_.trace(in);
            }
        };
        const auto traceTriangles = [&](Intersection in) {
            switch (s.accelerator_) {
                case Shader::ACC_NAIVE: return shadow ? s.naiveTriangles_.shadowTrace(in) : s.naiveTriangles_.trace(in);
                case Shader::ACC_REGULAR_GRID: return shadow ? s.gridTriangles_.shadowTrace(in) : s.gridTriangles_.trace(in);
                default: return shadow ? s.bvhTriangles_.shadowTrace(in) : s.bvhTriangles_.trace(in);
            }
        };
        isect = tracePlanes(isect);
        if (isect.length_ < start) { kind = 1; idx = static_cast<int>(static_cast<const Plane *>(isect.primitive_) - plas.data()); }
        float before {isect.length_};
        isect = traceSpheres(isect);
        if (isect.length_ < before) {
            kind = 2; idx = -1;
            for (size_t k = 0; k < sphs.size(); ++k) {
                Intersection probe {Ray {isect.ray_.direction_, isect.ray_.origin_, 1, false, nullptr}};
                probe = sphs[k].intersect(probe);
                if (probe.length_ == isect.length_ && sphs[k].materialIndex_ == isect.materialIndex_) { idx = static_cast<int>(k); break; }
            }
        }
        before = isect.length_;
        isect = traceTriangles(isect);
        if (isect.length_ < before) { kind = 3; idx = static_cast<int>(static_cast<const Triangle *>(isect.primitive_) - tris.data()); }
        if (!shadow) {
            before = isect.length_;
            isect = s.traceLights(isect);
            if (isect.length_ < before) {
                kind = 4; idx = -1;
                const auto &lights = s.getLights();
                for (size_t k = 0; k < lights.size(); ++k) {
                    if (isect.material_ == &lights[k]->radiance_) { idx = static_cast<int>(k); break; }
                }
            }
        }
        outT[i] = isect.length_;
        outKind[i] = kind; outIdx[i] = idx; outMat[i] = isect.materialIndex_;
        if (outNormal != nullptr) { for (int k = 0; k < 3; ++k) outNormal[3 * i + k] = isect.normal_[k]; }
    }
    errno = 0;
}

// The cell lists of the Shader's RegularGrid<T> (kind 0 planes, 1 spheres, 2 triangles; only for a handle created with
// accelerator 2). params: bounds min, max, cellSizeInverted_, cellSize_ (12 floats). start: 32^3 + 1 offsets; items: the
// primitive indices cell after cell, sorted ascending inside a cell (the reference's own order depends on thread timing,
// RegularGrid.hpp:191-203). Returns the number of items, or -1 (items may be null to ask for the count only).
long long ref_grid_cells(void *p, int kind, float *params, unsigned *start, int *items, long long cap) {
    Shader &s = *static_cast<RefHandle *>(p)->renderer->shader_;
    errno = 0;
    if (s.accelerator_ != Shader::ACC_REGULAR_GRID) { return -1; }
    if (kind == 0) { return gridCells(s.gridPlanes_, params, start, items, cap); }
    if (kind == 1) { return gridCells(s.gridSpheres_, params, start, items, cap); }
    return gridCells(s.gridTriangles_, params, start, items, cap);
}

// Shader::rayTrace (Shader.cpp:86) on arbitrary rays of the given depth -> rgb.
void ref_shade_rays(void *p, int n, const float *o, const float *d, int depth, float *rgb) {
    Shader &s = *static_cast<RefHandle *>(p)->renderer->shader_;
    for (int i = 0; i < n; ++i) {
        ::glm::vec3 c {};
        s.rayTrace(&c, Ray {::glm::vec3 {d[3 * i], d[3 * i + 1], d[3 * i + 2]}, ::glm::vec3 {o[3 * i], o[3 * i + 1], o[3 * i + 2]}, depth, false, nullptr});
        for (int k = 0; k < 3; ++k) rgb[3 * i + k] = c[k];
    }
    errno = 0;
}

// ---- primitive-level known-answer entry points (TestTriangle/TestAABB/TestPlane mirrors) ----
// tri = 100-byte reference record; returns t or 1e30; out8 = normal(3), uv(2), point(3)
float ref_triangle_intersect(const void *tri, const float *o, const float *d, float tMax, int selfSkip, float *out8) {
    Triangle t {Triangle::Builder(::glm::vec3 {0, 0, 0}, ::glm::vec3 {1, 0, 0}, ::glm::vec3 {0, 1, 0}).build()};
    ::std::memcpy(static_cast<void *>(&t), tri, sizeof(Triangle));
    Intersection isect {Ray {::glm::vec3 {d[0], d[1], d[2]}, ::glm::vec3 {o[0], o[1], o[2]}, 1, false, selfSkip ? &t : nullptr}, tMax};
    isect = t.intersect(isect);
    if (out8 != nullptr) {
        for (int k = 0; k < 3; ++k) { out8[k] = isect.normal_[k]; out8[5 + k] = isect.point_[k]; }
        out8[3] = isect.texCoords_[0]; out8[4] = isect.texCoords_[1];
    }
    return isect.length_;
}
void ref_triangle_build(const float *A, const float *B, const float *C, int mat, void *out) {
    Triangle t {Triangle::Builder(::glm::vec3 {A[0], A[1], A[2]}, ::glm::vec3 {B[0], B[1], B[2]}, ::glm::vec3 {C[0], C[1], C[2]}).withMaterialIndex(mat).build()};
    ::std::memcpy(out, static_cast<const void *>(&t), sizeof(Triangle));
}
void ref_triangle_aabb(const void *tri, float *out6) {
    Triangle t {Triangle::Builder(::glm::vec3 {0, 0, 0}, ::glm::vec3 {1, 0, 0}, ::glm::vec3 {0, 1, 0}).build()};
    ::std::memcpy(static_cast<void *>(&t), tri, sizeof(Triangle));
    const AABB b {t.getAABB()};
    for (int k = 0; k < 3; ++k) { out6[k] = b.getPointMin()[k]; out6[3 + k] = b.getPointMax()[k]; }
}
float ref_sphere_intersect(const float *center, float radius, const float *o, const float *d, float tMax, float *outN) {
    const Sphere s {::glm::vec3 {center[0], center[1], center[2]}, radius, 0};
    Intersection isect {Ray {::glm::vec3 {d[0], d[1], d[2]}, ::glm::vec3 {o[0], o[1], o[2]}, 1, false, nullptr}, tMax};
    isect = s.intersect(isect);
    if (outN != nullptr) for (int k = 0; k < 3; ++k) outN[k] = isect.normal_[k];
    return isect.length_;
}
float ref_plane_intersect(const float *point, const float *normal, const float *o, const float *d, float tMax) {
    const Plane pl {::glm::vec3 {point[0], point[1], point[2]}, ::glm::vec3 {normal[0], normal[1], normal[2]}, 0};
    Intersection isect {Ray {::glm::vec3 {d[0], d[1], d[2]}, ::glm::vec3 {o[0], o[1], o[2]}, 1, false, nullptr}, tMax};
    isect = pl.intersect(isect);
    return isect.length_;
}
void ref_plane_aabb(const float *point, const float *normal, float *out6) {
    const Plane pl {::glm::vec3 {point[0], point[1], point[2]}, ::glm::vec3 {normal[0], normal[1], normal[2]}, 0};
    const AABB b {pl.getAABB()};
    for (int k = 0; k < 3; ++k) { out6[k] = b.getPointMin()[k]; out6[3 + k] = b.getPointMax()[k]; }
}
int ref_aabb_intersect(const float *mn, const float *mx, const float *o, const float *d) {
    const AABB b {::glm::vec3 {mn[0], mn[1], mn[2]}, ::glm::vec3 {mx[0], mx[1], mx[2]}};
    const Ray r {::glm::vec3 {d[0], d[1], d[2]}, ::glm::vec3 {o[0], o[1], o[2]}, 1, false, nullptr};
    return b.intersect(r) ? 1 : 0;
}
float ref_aabb_area(const float *mn, const float *mx) {
    return AABB {::glm::vec3 {mn[0], mn[1], mn[2]}, ::glm::vec3 {mx[0], mx[1], mx[2]}}.getSurfaceArea();
}
float ref_halton(unsigned index, unsigned base) { return haltonSequence(index, base); }
int32_t ref_incremental_avg(const float *rgb, int32_t avg, int n) { return incrementalAvg(::glm::vec3 {rgb[0], rgb[1], rgb[2]}, avg, n); }
float ref_fresnel(const float *I, const float *N, float ior) { return fresnel(::glm::vec3 {I[0], I[1], I[2]}, ::glm::vec3 {N[0], N[1], N[2]}, ior); }

} // extern "C"
