/*
 * mobilert_b200.h — C-ABI of libmobilert_b200.so: the B200-native (sm_100a CUDA) drop-in for MobileRT's
 * per-pixel ray-tracing hot path (Renderer::renderFrame -> Shader::rayTrace/shadowTrace -> BVH -> intersect -> shade).
 *
 * Every entry point cites the reference interface it replaces (paths relative to the reference's app/ directory).
 * Plain pointers and sizes only; no exceptions cross this boundary; every function returns an mrt_status
 * (0 = ok) unless stated otherwise and leaves errno == 0 on success, because the reference's callers poll errno
 * right after the frame (System_dependent/Native/C_wrapper.cpp:234, Android_JNI/JNI_layer.cpp:821).
 * There is no CPU fallback: without a CUDA device every compute entry point fails with MRT_ERR_CUDA.
 */
#ifndef MOBILERT_B200_H
#define MOBILERT_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum mrt_status {
    MRT_OK = 0,
    MRT_ERR_INVALID = 1,   /* bad argument (null pointer, negative size, unknown enum) */
    MRT_ERR_CUDA = 2,      /* CUDA runtime error or no device; see mrt_last_error() */
    MRT_ERR_OOM = 3,       /* host or device allocation failed (JNI maps this to LowMemoryException, JNI_layer.cpp:706-715) */
    MRT_ERR_STOPPED = 4,   /* mrt_stop() was called while rendering (Renderer::stopRender, MobileRT/Renderer.cpp:93-99) */
    MRT_ERR_OVERFLOW = 5   /* a traversal stack overflowed, or the path tracer's ray queues overflowed even at the largest size the
                              library allocates (an overflowing pass is first repeated with fewer samples, then with doubled queues:
                              mrt_stats::pass_retries); result is invalid */
} mrt_status;

/* ---- scene records: byte-for-byte the reference's own layouts ---------------------------------------------- */
/* MobileRT/Shapes/Triangle.hpp:18-27 (100 bytes) */
typedef struct mrt_triangle {
    float AC[3], AB[3], A[3];
    float nA[3], nB[3], nC[3];
    float tA[2], tB[2], tC[2];
    int32_t material;
} mrt_triangle;
/* MobileRT/Shapes/Sphere.hpp (20 bytes): centre, squared radius, material index */
typedef struct mrt_sphere { float center[3]; float sq_radius; int32_t material; } mrt_sphere;
/* MobileRT/Shapes/Plane.hpp (28 bytes): unit normal, point, material index */
typedef struct mrt_plane { float normal[3]; float point[3]; int32_t material; } mrt_plane;
/* MobileRT/Material.hpp:18-38 without the texture */
typedef struct mrt_material { float Le[3], Kd[3], Ks[3], Kt[3]; float ior; } mrt_material;

enum { MRT_LIGHT_POINT = 0, MRT_LIGHT_AREA = 1 };
/* Components/Lights/PointLight.hpp:12-13, AreaLight.hpp:19-21 (+ Light::radiance_.Le_, MobileRT/Light.hpp:16) */
typedef struct mrt_light {
    int32_t kind;
    float position[3];       /* point light */
    mrt_triangle triangle;   /* area light */
    float Le[3];
} mrt_light;

/* MobileRT/Texture.hpp: a decoded image exactly as stb_image returned it to Texture::createTexture (Texture.cpp:58-118):
 * `channels` (1..4) bytes per texel, rows top to bottom. Texture::loadColor (Texture.cpp:37-48) reads bytes index..index+2. */
typedef struct mrt_texture { const uint8_t *pixels; int32_t width, height, channels; } mrt_texture;

/* MobileRT/Scene.hpp:20-27 flattened. The library copies everything at mrt_create; the caller may free at once. */
typedef struct mrt_scene_desc {
    const mrt_triangle *triangles; int64_t n_triangles;
    const mrt_sphere *spheres;     int64_t n_spheres;
    const mrt_plane *planes;       int64_t n_planes;
    const mrt_material *materials; int64_t n_materials;
    const mrt_light *lights;       int64_t n_lights;
    /* optional (may be NULL / 0): Material::texture_ (Material.hpp:43). material_texture[m] = index into textures or -1.
     * A triangle hit whose interpolated texture coordinates are both >= 0 takes Kd from its material's texture
     * (Shader::rayTrace, Shader.cpp:113-121). */
    const mrt_texture *textures;   int64_t n_textures;
    const int32_t *material_texture;
} mrt_scene_desc;

enum { MRT_CAMERA_PERSPECTIVE = 0, MRT_CAMERA_ORTHOGRAPHIC = 1,
       /* the state of an already constructed camera object instead of its constructor arguments (what a front-end holds:
        * C_wrapper.cpp:76-141 / JNI_layer.cpp:164 keep a Camera, not its lookAt): position = Camera::position_, look_at =
        * Camera::direction_, up = Camera::up_, right = Camera::right_ (MobileRT/Camera.hpp:26-41), a / b = Perspective::hFov_ /
        * vFov_ in radians (Perspective.hpp:16-17) or Orthographic::sizeH_ / sizeV_ (already halved, Orthographic.hpp:19-20).
        * Nothing is recomputed, so the rays are bit-identical to that object's generateRay. */
       MRT_CAMERA_PERSPECTIVE_STATE = 2, MRT_CAMERA_ORTHOGRAPHIC_STATE = 3 };
/* Arguments of Perspective(position, lookAt, up, hFovDeg, vFovDeg) (Components/Cameras/Perspective.cpp:8-14)
 * or Orthographic(position, lookAt, up, sizeH, sizeV) (Components/Cameras/Orthographic.cpp:7-13). */
typedef struct mrt_camera_desc {
    int32_t kind;
    float position[3], look_at[3], up[3];
    float a, b;
    float right[3];   /* only read for the *_STATE kinds */
} mrt_camera_desc;

/* Shader ids of MobileRT::Config::shader (MobileRT/Config.hpp, C_wrapper.cpp:154-194) */
enum { MRT_SHADER_NOSHADOWS = 0, MRT_SHADER_WHITTED = 1, MRT_SHADER_PATHTRACER = 2, MRT_SHADER_DEPTHMAP = 3, MRT_SHADER_DIFFUSE = 4 };
/* MobileRT/Shader.hpp:20-24 plus the new value a front-end selects to get this library */
enum { MRT_ACC_NAIVE = 1, MRT_ACC_REGULAR_GRID = 2, MRT_ACC_BVH = 3, MRT_ACC_B200 = 4 };
/* pixel-jitter sampler (C_wrapper.cpp:144-148 picks Constant(0.5) for spp==1, StaticHaltonSeq otherwise) */
enum { MRT_SAMPLER_AUTO = 0, MRT_SAMPLER_CONSTANT = 1, MRT_SAMPLER_STATIC_HALTON = 2, MRT_SAMPLER_STATIC_MT = 3,
       MRT_SAMPLER_STATIC_PCG = 4 /* the Android front-end's choice (JNI_layer.cpp:601,616-618,636; Components/Samplers/StaticPCG.cpp) */,
       MRT_SAMPLER_HALTON = 5     /* HaltonSeq(width, height, samples) (Components/Samplers/HaltonSeq.cpp:10-19): radical inverse of the
                                     draw counter, 1.0 once Sampler::domainSize_ draws have been made */,
       MRT_SAMPLER_STRATIFIED = 6 /* Stratified(width, height, samples) (Stratified.cpp:10-19): counter / domainSize_, then 1.0 */,
       MRT_SAMPLER_STREAM = 7     /* the draws come from the caller (mrt_set_pixel_stream): MersenneTwister (MersenneTwister.cpp:5-11)
                                     and PCG (PCG.cpp:6-13) keep hidden thread-local generator state, so the host takes the frame's
                                     draws from the sampler object itself and hands them over, index for index */ };

enum {
    MRT_FLAG_FP32_ACCUMULATE = 1,  /* accumulate radiance in fp32 and quantise once (needed for multi-GPU partitions);
                                      default is the reference's 8-bit running average (MobileRT/Utils/Utils.cpp:66-90) */
    MRT_FLAG_SORT_BY_MATERIAL = 2, /* PathTracer: the hits of every depth level are binned by material (counting sort of hit indices)
                                      before they are shaded. Same image. Off by default because it is slower on this design: every
                                      material runs the same shading routine (materials are data, Material.hpp:18-43), so there is no
                                      divergence to sort away, and the permuted order turns the streamed hit / ray records into
                                      gathers (measured A/B: profiles/r2_ab_material_sort.txt) */
    MRT_FLAG_PROGRESSIVE = 8,      /* mrt_render_frame refreshes the caller's HOST bitmap after every sample (wavefront: every pass)
                                      from a second stream, so that a UI thread polling the buffer sees the frame build up as it
                                      does with the reference, whose workers write finished pixels in place (Renderer.cpp:149-157;
                                      JNI_layer.cpp:818-821, Qt mainwindow.cpp:45-66). A refresh is skipped while the previous one
                                      is in flight. The final content is identical with and without the flag. */
    MRT_FLAG_REPLAY_SAMPLERS = 4   /* parity mode for NoShadows / Whitted: every Shader::getLightIndex (Shader.cpp:223-233) and
                                      AreaLight::getPosition (AreaLight.cpp:17-26) draw reads the table entry the reference's
                                      single-thread run reads (two counting passes per sample + host prefix sums, about 3x the
                                      frame time). Up to 8 lights; rejected for the PathTracer, whose ray tree depends on the
                                      draws; needs the reference's tables (mrt_sampler_tables) and mrt_set_replay_counter. */
};

typedef struct mrt_render_desc {
    int32_t width, height;          /* used as given; front-ends round to x16 first (Qt main.cpp:36-44, F8) */
    int32_t samples_pixel;          /* Config::samplesPixel */
    int32_t samples_light;          /* Config::samplesLight */
    int32_t shader;                 /* MRT_SHADER_* */
    int32_t accelerator;            /* MRT_ACC_B200 or MRT_ACC_BVH: the reference's BVH visibility rules (finite plane boxes,
                                       root-box NaN rule); MRT_ACC_NAIVE: Naive<T>'s rules (every plane / sphere tested, caller
                                       order); triangles go through the BVH8 in both. MRT_ACC_REGULAR_GRID: RegularGrid<T>'s rules
                                       (Accelerators/RegularGrid.hpp:112-513: three 32^3 grids built on the host with
                                       T::intersect(AABB), rays walked from the clamped start cell); no BVH is built then and
                                       every ray runs the reference's 3-D DDA, one thread per ray (a compatibility mode: slow). */
    int32_t pixel_sampler;          /* MRT_SAMPLER_* */
    uint32_t seed;                  /* seeds the library's own table shuffles when mrt_sampler_tables is NULL */
    uint32_t flags;                 /* MRT_FLAG_* */
    float depth_max_point[3];       /* DepthMap's maxPoint (Components/Shaders/DepthMap.cpp:7-10) */
    int32_t device;                 /* CUDA device ordinal; -1 = current */
    int32_t light_sampler;          /* table behind the samplers the front-end creates itself, i.e. the area lights of an OBJ scene
                                       (OBJLoader::fillScene's lambda) and the PathTracer's russian roulette: MRT_SAMPLER_AUTO or
                                       _STATIC_HALTON (C_wrapper.cpp:109,176), or MRT_SAMPLER_STATIC_PCG (JNI_layer.cpp:601,636) */
    int32_t tile_begin, tile_stride; /* multi-GPU tile partition: this handle renders tiles tile_begin,
                                       tile_begin + tile_stride, ... of the 256 tiles of Renderer::renderScene's shuffled tile walk
                                       (Renderer.cpp:119-135,189-197) and leaves every other pixel untouched; its queues are sized for
                                       that share. 0 / 0 (or stride 1) = all tiles. The jitter draws of a pixel-sample do not depend
                                       on the partition. Not available together with MRT_FLAG_REPLAY_SAMPLERS. */
} mrt_render_desc;

/* The reference's shuffled tables, when the caller wants the very same sequences (parity runs):
 * tile_order = Renderer.cpp:12 (256 floats); shader_table = Shader.cpp:20 (1<<20 floats, getLightIndex and
 * getCosineSampleHemisphere); static_halton = Components/Samplers/StaticHaltonSeq.cpp:7 (1<<20, pixel jitter, area
 * lights, russian roulette); static_mt = StaticMersenneTwister.cpp:7; static_pcg = StaticPCG.cpp:8. Any pointer may be
 * NULL: the library then fills its own table the way the reference does (Utils.hpp:209-252) from mrt_render_desc::seed. */
typedef struct mrt_sampler_tables {
    const float *tile_order;
    const float *shader_table;
    const float *static_halton;
    const float *static_mt;
    const float *static_pcg;
} mrt_sampler_tables;

typedef struct mrt_stats {
    double build_ms;            /* "Creating Time": acceleration-structure build on the GPU (C_wrapper.cpp:152-198) */
    double upload_ms;           /* host->device copy of the scene */
    double frame_ms;            /* last frame, device time */
    double extend_ms, shadow_ms, shade_ms, other_ms; /* last frame, device time per kernel class */
    uint64_t rays_primary, rays_shadow, rays_secondary; /* last frame; their sum is the reference's "Casted rays" (MobileRT/Ray.cpp:11,26) */
    uint64_t launches;          /* kernels launched by the last frame */
    uint64_t extend_launches, shadow_launches, shade_launches; /* wavefront path tracer: launches per kernel class, last frame */
    uint64_t rays_extend_traced, rays_shadow_traced; /* wavefront: rays the extend / shadow kernels actually traversed
                                                        (a depth-7 ray is constructed and counted but never traced) */
    uint64_t bvh_nodes, bvh_bytes, scene_bytes;
    int64_t  n_primitives;      /* JNI rtInitialize's return value (JNI_layer.cpp:676-680) */
    uint64_t pass_retries;      /* wavefront: passes of the last frame that were repeated with fewer samples (or larger queues)
                                   because a ray queue overflowed; the frame is still complete and correct */
    uint64_t queue_bytes;       /* wavefront: device memory of the ray / shadow queues and path accumulators */
} mrt_stats;

typedef struct mrt_handle mrt_handle;

/* Replaces: Scene -> Shader subclass ctor (-> Shader::initializeAccelerators, MobileRT/Shader.cpp:33-77, BVH build
 * MobileRT/Accelerators/BVH.hpp:125-283) -> Renderer ctor (MobileRT/Renderer.cpp:25-45), i.e. what work_thread
 * (C_wrapper.cpp:76-203) and rtInitialize (JNI_layer.cpp:465-716) do before the first frame.
 * Device memory: the scene, and for MRT_SHADER_PATHTRACER the wavefront's ray / shadow queues and path accumulators, sized so
 * that a frame runs in as few equally large passes as fit 40 GB (about 0.5 KB per pixel and sample of a pass; 31.7 GB for 64 spp
 * at 1920x1072; mrt_stats::queue_bytes). When that much is not free the handle comes up with smaller passes instead of failing. */
int mrt_create(const mrt_scene_desc *scene, const mrt_camera_desc *camera, const mrt_render_desc *render,
               const mrt_sampler_tables *tables, mrt_handle **out);
/* Replaces: ~Renderer / renderer_.reset (C_wrapper.cpp:266, JNI_layer.cpp:406-431). */
void mrt_destroy(mrt_handle *h);

/* Replaces: Renderer::renderFrame(int32* bitmap, int nThreads) (MobileRT/Renderer.cpp:53-88). Blocking. Writes packed
 * ABGR8888 (0xFF000000|B<<16|G<<8|R, MobileRT/Utils/Utils.cpp:66-90) into the caller-owned HOST bitmap of
 * width*height int32. The thread count of the reference signature has no meaning here. */
int mrt_render_frame(mrt_handle *h, int32_t *host_bitmap);
/* Same, into a caller-owned DEVICE buffer (headless replacement of the GL/Qt display path). */
int mrt_render_frame_device(mrt_handle *h, void *device_bitmap);
/* Multi-GPU building block (sample partition): renders samples sample_begin, sample_begin+sample_stride, ...
 * (< samples_pixel) and ADDS their radiance into a caller-owned DEVICE fp32 RGBA buffer (width*height*4 floats;
 * .w counts samples). The buffers of all ranks are summed by one NCCL reduce, then resolved on the root. */
int mrt_render_accumulate_device(mrt_handle *h, void *device_accum_rgba32f, int32_t sample_begin, int32_t sample_stride);
/* Quantises an fp32 RGBA accumulator (sum of per-sample trunc(255*c), .w = sample count) to ABGR8888 on the device. */
int mrt_resolve_device(mrt_handle *h, const void *device_accum_rgba32f, void *device_bitmap);

/* MRT_SAMPLER_STREAM: the pixel sampler's draws of the NEXT frame(s) in the order a single-threaded Renderer::renderScene makes
 * them (Renderer.cpp:137-138: two per pixel-sample; tile walk, samples outermost). n >= 2 * samples_pixel * 256 * (width/16) *
 * (height/16) serves a whole frame; a shorter stream wraps around. HOST array, copied. */
int mrt_set_pixel_stream(mrt_handle *h, const float *draws, int64_t n);

/* Replaces: the camera_ member the front-ends rebuild between frames (JNI_layer.cpp:164; C_wrapper.cpp:76-141). */
int mrt_set_camera(mrt_handle *h, const mrt_camera_desc *camera);

/* Makes every later call on this handle enqueue its kernels and copies on the caller's CUDA stream (a cudaStream_t passed
 * as void*; NULL restores the handle's own stream, so the legacy default stream, whose handle is 0, has to be named by
 * cudaStreamLegacy = (cudaStream_t)0x1), so that a host runtime that owns streams and events (the JNI layer's
 * render thread has none; a PyTorch / NCCL host does) can order and time the frame with its own events. Calls stay
 * blocking: they synchronise that stream before returning. */
int mrt_set_stream(mrt_handle *h, void *cuda_stream);

/* Replaces: Renderer::stopRender (Renderer.cpp:93-99), Renderer::getSample (:176-178),
 * Renderer::getTotalCastedRays (:204-207). Thread-safe; may be called while mrt_render_frame runs. */
/* Shader::getLightIndex's counter is a function-static (Shader.cpp:224): it is never reset and is shared by every
 * renderer of the process. With MRT_FLAG_REPLAY_SAMPLERS the handle mirrors it: starts at 0, advances with every frame
 * it renders; a host that also runs the reference's CPU path in the same process sets the reference's value here first. */
int mrt_set_replay_counter(mrt_handle *h, uint32_t light_index_counter);
int mrt_get_replay_counter(mrt_handle *h, uint32_t *light_index_counter);
int mrt_stop(mrt_handle *h);
int mrt_get_sample(mrt_handle *h, int32_t *sample);
int mrt_get_total_rays(mrt_handle *h, uint64_t *rays);
int mrt_get_stats(mrt_handle *h, mrt_stats *stats);
/* Human-readable description of the last failure on this handle (or of the last failed mrt_create when h == NULL). */
const char *mrt_last_error(mrt_handle *h);

/* Per-ray ground-truth access for parity tests (what Shader::rayTrace's geometric part, Shader.cpp:86-111, and
 * Shader::shadowTrace, :132-158, compute). HOST arrays. origin_kind/origin_idx give the primitive a ray leaves
 * (self-hit skip, Triangle.cpp:64, Plane.cpp:39); shadow_dist[i] >= 0 selects any-hit semantics.
 * out_kind: 0 miss, 1 plane, 2 sphere, 3 triangle, 4 light; out_idx indexes the arrays given to mrt_create.
 * For shadow rays out_kind is 1 when occluded, 0 when not. A PathTracer handle answers through the same persistent
 * extend / shadow kernels its frames use (n up to the queue capacity), any other handle through the per-thread traversal. */
int mrt_trace_rays(mrt_handle *h, int64_t n, const float *origins, const float *directions,
                   const int32_t *origin_kind, const int32_t *origin_idx, const float *shadow_dist,
                   float *out_t, int32_t *out_kind, int32_t *out_idx, int32_t *out_material, float *out_normal);
/* mrt_trace_rays without any allocation. No counterpart in the reference: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * Same results as mrt_trace_rays, through the persistent extend / shadow kernels (the RegularGrid kernels on an
 * MRT_ACC_REGULAR_GRID handle), on any handle. Differences:
 *  - every array is a DEVICE pointer the caller owns (inputs and outputs; origin_kind / origin_idx / shadow_dist / out_normal
 *    may be NULL, origin_kind and origin_idx only together); the call returns after the handle's stream has finished;
 *  - the kernels' ray / hit / shadow queues live in a static device array of the library, MRT_STATIC_TRACE_MAX_RAYS = 65536
 *    rays (7.1 MiB that every GPU which loads the library holds until the process exits); n beyond that is rejected with
 *    MRT_ERR_INVALID and a message naming the limit. The call performs no cudaMalloc / cudaFree / allocating library call;
 *  - every call on a GPU shares that array: calls that overlap on one GPU (two handles, two threads, two streams) race on it.
 *    Serialise them. The array keeps the last call's records; the call clears the counters it needs to start at zero;
 *  - origin_kind[i] may be 0 (no primitive to skip) or 3 (triangle origin_idx[i]); plane origins (1) are not remapped here. */
#define MRT_STATIC_TRACE_MAX_RAYS 65536
int mrt_trace_rays_static(mrt_handle *h, int64_t n, const float *d_origins, const float *d_directions,
                          const int32_t *d_origin_kind, const int32_t *d_origin_idx, const float *d_shadow_dist,
                          float *d_out_t, int32_t *d_out_kind, int32_t *d_out_idx, int32_t *d_out_material, float *d_out_normal);
/* Camera::generateRay (Perspective.cpp:16-28, Orthographic.cpp:15-23) for n (u,v,du,dv) tuples. HOST arrays. */
int mrt_generate_rays(mrt_handle *h, int64_t n, const float *uvdd, float *out_origins, float *out_directions);
/* Shader::rayTrace (Shader.cpp:86) on arbitrary rays of the given depth -> rgb (deterministic shaders). HOST arrays. */
int mrt_shade_rays(mrt_handle *h, int64_t n, const float *origins, const float *directions, int32_t depth, float *out_rgb);

/* Library / device info: returns the number of CUDA devices visible, 0 when none (never falls back to the CPU). */
int mrt_device_count(void);
const char *mrt_version(void);

#ifdef __cplusplus
}
#endif
#endif /* MOBILERT_B200_H */
