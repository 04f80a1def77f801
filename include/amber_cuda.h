/*
 * amber_cuda.h -- C ABI of libamber_cuda.so, the B200 (sm_100a) implementation of
 * amber's path-tracing hot path.
 *
 * This is the drop-in boundary.  Everything above it (amber's C++ scene / sensor /
 * context / algorithm API) stays on the host; everything below it is CUDA.  Plain
 * pointers and sizes only -- no C++ types, no torch types.
 *
 * Each entry point names the reference interface it stands in for (paths relative
 * to the reference tree, file:line).  The C++ adapter that binds these calls behind
 * `amber::rendering::Algorithm<RGB>` lives in include/amber_cuda/adapter.hpp; the
 * ctypes binding used by the tests and bench.py is amber_b200/capi.py.
 *
 * Conventions
 *  - every function returns 0 on success, a negative AMBER_CUDA_E* code on failure;
 *    amber_cuda_last_error() returns the message (the C++ adapter rethrows it as
 *    std::runtime_error, matching the reference's exceptions-only convention,
 *    include/amber/prelude/image.h:73-75).
 *  - the caller owns every host buffer; the context owns all device memory.
 *  - a context is bound to ONE device and is used by one host thread at a time
 *    (the reference's Render() is likewise called from one thread,
 *    src/amber/cli/application.cc:149-152).  Different contexts may be driven from different
 *    threads, also on the same device: the scene and camera of a context live in the device's
 *    __constant__ banks while it renders, so contexts that share a device take turns (a per-device
 *    lock held for the length of one render job or parity call).
 *  - amber_cuda_render_async queues a render on the context's own worker thread and returns;
 *    every other call on that context first waits for the queue to drain.  G contexts on G devices
 *    therefore render concurrently, as the reference's worker threads do
 *    (src/amber/prelude/parallel.cc:29-40, include/amber/rendering/parallel.h:55-68).
 *  - there is no CPU fallback: without a CUDA device amber_cuda_create() fails.
 */
#ifndef AMBER_CUDA_H_
#define AMBER_CUDA_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AMBER_CUDA_ABI_VERSION 2

/* error codes */
#define AMBER_CUDA_OK 0
#define AMBER_CUDA_EINVAL (-1)  /* bad argument / bad scene description */
#define AMBER_CUDA_ECUDA (-2)   /* a CUDA runtime call failed */
#define AMBER_CUDA_ESTATE (-3)  /* call order violated (e.g. render before upload) */
#define AMBER_CUDA_ENODEV (-4)  /* no usable CUDA device: there is no CPU fallback */

/* Primitive kinds: include/amber/scene/primitive_triangle.h:30-35,
 * include/amber/scene/primitive_sphere.h:30-34. */
#define AMBER_CUDA_PRIM_TRIANGLE 0
#define AMBER_CUDA_PRIM_SPHERE 1

/* Material kinds, one per reference factory:
 *   MakeLambertian   include/amber/scene/material_lambertian.h:65-74
 *   MakePhong        include/amber/scene/material_phong.h
 *   MakeSpecular     include/amber/scene/material_specular.h
 *   MakeRefraction   include/amber/scene/material_refraction.h
 *   MakeDiffuseLight include/amber/scene/material_diffuse_light.h:27-29
 *   MakeEye          include/amber/scene/material_eye.h (aperture pass-through) */
#define AMBER_CUDA_MAT_LAMBERTIAN 0
#define AMBER_CUDA_MAT_PHONG 1
#define AMBER_CUDA_MAT_SPECULAR 2
#define AMBER_CUDA_MAT_REFRACTION 3
#define AMBER_CUDA_MAT_DIFFUSE_LIGHT 4
#define AMBER_CUDA_MAT_EYE 5

/* Estimators.
 * REFERENCE: sample-for-sample the estimator of
 *   src/amber/rendering/algorithm_pt.cc:125-160 (emission at every hit, no light
 *   sampling).
 * NEE: next-event estimation; equal in expectation.  Light point sampled as
 *   include/amber/scene/light_set.h:83-104, connection term as the s=1 branch of
 *   include/amber/rendering/bidirectional_path_sampling.h:956-1016; emission on a
 *   BSDF-sampled hit is counted only when the previous vertex could not be
 *   connected (camera, Specular, Refraction, Eye). */
#define AMBER_CUDA_ESTIMATOR_REFERENCE 0
#define AMBER_CUDA_ESTIMATOR_NEE 1
/* LIGHT_TRACING (row N2): sample-for-sample the algorithm of
 *   src/amber/rendering/algorithm_lt.cc:112-163 on the same wavefront kernels: one sample = W*H
 *   paths started on the lights (LightSet::GenerateRay, include/amber/scene/light_set.h:83-104),
 *   scattered with Material::SampleImportance; a path that lands on the aperture is mapped to a
 *   pixel by the lens (BasicThin::Response, src/amber/scene/lens_thin.cc:109-130) and its weight
 *   / (W*H) is added there.  Emitters are triangles or spheres (a sphere is sampled as the reference's
 *   SphereSA does, src/amber/scene/primitive_sphere.cc:116-122). */
#define AMBER_CUDA_ESTIMATOR_LIGHT_TRACING 2
/* BDPT (row N3): sample for sample src/amber/rendering/algorithm_bdpt.cc:116-148 over
 *   include/amber/rendering/bidirectional_path_sampling.h (GenerateLightSubpath,
 *   GenerateEyeSubpath, Connect for every (s, t), balance-heuristic MIS weights, light image) --
 *   one sample = one light subpath + one eye subpath per pixel.  max_depth is the number of
 *   events per subpath (0 = 16, the maximum; the reference has no cap).  Emitters as for light
 *   tracing.  Two implementations with the same results up to float summation order: a wavefront over a
 *   pool of pixel samples (the default; it allocates 5.7 GB of device memory on first use and falls back to
 *   the other one if that fails) and one persistent kernel (environment variable AMBER_CUDA_BDPT=megakernel). */
#define AMBER_CUDA_ESTIMATOR_BDPT 3

typedef struct amber_cuda_ctx amber_cuda_ctx;

/* One primitive.  The index in the array is the primitive id reported by
 * amber_cuda_cast (creation order, as objects are emplaced in
 * src/amber/etude/cornel_box.cc:62-196). */
typedef struct amber_cuda_primitive {
  int32_t kind;     /* AMBER_CUDA_PRIM_* */
  int32_t material; /* index into materials[] */
  float data[9];    /* triangle: v0 v1 v2 (xyz each); sphere: centre xyz, radius */
} amber_cuda_primitive;

typedef struct amber_cuda_material {
  int32_t kind; /* AMBER_CUDA_MAT_* */
  float rgb[3]; /* kd / ks / emitted radiance; ignored for REFRACTION and EYE */
  float param;  /* PHONG: exponent; REFRACTION: index of refraction */
} amber_cuda_material;

/* Lens kinds: MakeThinLens (include/amber/scene/lens_thin.h:30-38; the lens of the built-in
 * scene and of the importer) and MakePinholeLens (include/amber/scene/lens_pinhole.h:32-37). */
#define AMBER_CUDA_LENS_THIN 0
#define AMBER_CUDA_LENS_PINHOLE 1

/* The arguments of MakeThinLens, or (kind = PINHOLE) of MakePinholeLens: transform and
 * sensor_distance; the thin-lens fields are then ignored.  A zero-initialised tail is a thin lens. */
typedef struct amber_cuda_lens {
  float transform[16]; /* row-major 4x4, lens-local -> world (Matrix4 ctor order) */
  float focal_length;
  float focus_distance;
  float radius;
  int32_t n_blades;
  int32_t kind;          /* AMBER_CUDA_LENS_* */
  float sensor_distance; /* PINHOLE only */
} amber_cuda_lens;

typedef struct amber_cuda_scene_desc {
  uint32_t n_primitives;
  const amber_cuda_primitive* primitives;
  uint32_t n_materials;
  const amber_cuda_material* materials;
  amber_cuda_lens lens;
} amber_cuda_scene_desc;

typedef struct amber_cuda_render_params {
  uint64_t sample_begin; /* first global sample index of this batch */
  uint32_t sample_count; /* samples per pixel to add in this call */
  uint32_t max_depth;    /* bounce cap (0 = 2^16); the reference has none, algorithm_pt.cc:137 */
  uint64_t seed;
  int32_t estimator;     /* AMBER_CUDA_ESTIMATOR_* */
  uint32_t pool_paths;   /* resident path slots (0 = default) */
  /* Pixel window: only pixels [pixel_begin, pixel_begin + pixel_count) of the frame (index x + y*W,
   * i.e. whole rows when both are multiples of W) are sampled; pixel_count = 0 is the whole frame.
   * The random stream is keyed by the pixel's index in the FULL frame, so a window holds exactly the
   * values the full render has there.  Path tracing estimators only. */
  uint32_t pixel_begin;
  uint32_t pixel_count;
  /* The k-th sample of this call has global index sample_begin + k * sample_stride (0 = 1): device d
   * of G renders {d, d + G, d + 2G, ...}, so that consecutive calls on one device continue one
   * sequence (see amber_cuda_render_async) and the union over the devices is {0, 1, 2, ...}. */
  uint32_t sample_stride;
  uint32_t reserved;
} amber_cuda_render_params;

typedef struct amber_cuda_stats {
  uint64_t paths;           /* camera paths started since the last film clear */
  uint64_t closest_rays;    /* rays through `extend` */
  uint64_t shadow_rays;     /* rays through `connect` */
  uint64_t truncated_paths; /* paths cut by max_depth */
  uint64_t phong_capped;    /* Phong rejection loops that hit the retry cap */
  uint64_t kernel_launches; /* our kernels launched since context creation */
  double last_render_ms;    /* device time of the last amber_cuda_render */
  double extend_ms;         /* per-kernel device time inside the last render (profiling mode) */
  double shade_ms;
  double connect_ms;
  double raygen_ms;
  uint64_t extend_launches; /* launches of `extend` inside the last render */
  uint32_t accel;           /* 0 = constant-memory list, 1 = wide BVH */
  uint32_t bvh_nodes;
  uint64_t light_splats;    /* light tracing: film contributions through the lens */
  /* wide-BVH traversal work of the last render (counted in profiling mode only): node steps (one
   * 128-byte node read each), triangles tested (one 48-byte record each), by `extend` and `connect` */
  uint64_t bvh_node_visits, bvh_triangle_tests;
  uint64_t bvh_node_visits_connect, bvh_triangle_tests_connect;
  uint32_t bvh_node_bytes, bvh_triangle_bytes; /* size of one node / one leaf triangle record in HBM */
  uint64_t chained_renders; /* render calls absorbed into a running wavefront (no pool drain between them) */
  double scene_build_ms;    /* host time of the last upload_scene: flattening + acceleration build */
  double scene_upload_ms;   /* ... and its host->device copies */
} amber_cuda_stats;

/* ABI version of the loaded library. */
int amber_cuda_abi_version(void);

/* Message of the last failure on this context (ctx may be NULL for a failed create). */
const char* amber_cuda_last_error(const amber_cuda_ctx* ctx);

/* Create a context on CUDA device `device`.  Replaces nothing in the reference (it
 * has no device); plays the role of constructing the algorithm object,
 * MakeRGBPathTracing(), src/amber/rendering/algorithm_pt.cc:76-80. */
int amber_cuda_create(amber_cuda_ctx** out, int device);
int amber_cuda_destroy(amber_cuda_ctx* ctx);

/* Upload a scene and build the acceleration structure.  Replaces
 * scene::Scene<RGB>::Create<BVH> (include/amber/scene/scene.h:169-193) and
 * BVH::BVH/BuildBVH (include/amber/raytracer/acceleration_bvh.h:135-180).
 * Scenes too large for constant memory get a 4-wide BVH, built on the host (binned SAH; default) or, with the
 * environment variable AMBER_CUDA_BVH_BUILDER=lbvh, on the device (20 ms instead of 0.4 s for 1.3 M triangles, for a
 * tree that traces ~9 % slower). */
int amber_cuda_upload_scene(amber_cuda_ctx* ctx, const amber_cuda_scene_desc* scene);

/* Film geometry.  Replaces rendering::Sensor::Sensor
 * (src/amber/rendering/sensor.cc:28-38); clears the film. */
int amber_cuda_set_sensor(amber_cuda_ctx* ctx, uint32_t width, uint32_t height,
                          float film_width, float film_height);

/* Zero the film and the sample count. */
int amber_cuda_film_clear(amber_cuda_ctx* ctx);

/* Trace sample_count samples for every pixel and add them to the device film.
 * Replaces the body of PathTracing<RGB>::Render / Thread::operator() / Thread::Render
 * (src/amber/rendering/algorithm_pt.cc:82-160): ray generation, extend, shade,
 * connect.  Asynchronous work is complete when the call returns. */
int amber_cuda_render(amber_cuda_ctx* ctx, const amber_cuda_render_params* params);

/* The same, asynchronously: the render is queued on the context's worker thread and the call returns.
 * A queued call whose parameters continue the running one (same estimator, seed, depth cap, pool,
 * window and stride; sample_begin = where the previous call ends) is ABSORBED into the running
 * wavefront: the path pool is not drained and refilled between the two, which is what makes small
 * batches of Context::Iterate() (include/amber/rendering/context.h:26-32) as cheap as one big call.
 * amber_cuda_wait blocks until everything queued is done and returns the first error of the queue
 * (then forgets it); amber_cuda_pending returns the number of queued + running calls;
 * amber_cuda_wait_pending blocks until at most max_pending are left.  Every other entry point
 * implies amber_cuda_wait. */
int amber_cuda_render_async(amber_cuda_ctx* ctx, const amber_cuda_render_params* params);
int amber_cuda_wait(amber_cuda_ctx* ctx);
int amber_cuda_pending(amber_cuda_ctx* ctx);
int amber_cuda_wait_pending(amber_cuda_ctx* ctx, int max_pending);

/* Sum the films (and sample counts) of contexts 1 .. n-1 into context 0: the reduction of the
 * per-worker images, Accumulator::operator+= (include/amber/prelude/accumulator.h:136-166), for
 * contexts of ONE process on different devices.  One kernel on context 0's device reads the other
 * films through peer memory (NVLink); without peer access they are staged through a copy.  All
 * contexts must have the same sensor; the other films are left as they are. */
int amber_cuda_film_reduce(amber_cuda_ctx* const* ctxs, int n);

/* Mean image to the host: host_rgb[(x + y*W)*3 + c] = film / samples.  Replaces
 * Accumulator::Mean (include/amber/prelude/accumulator.h:88-95) and the Image
 * returned by Render (layout include/amber/prelude/image.h:117-124).
 * samples_done may be NULL. */
int amber_cuda_resolve(amber_cuda_ctx* ctx, float* host_rgb, uint64_t* samples_done);

/* Mean image through the reference's post-process, to the host as 8-bit RGB:
 * host_rgb8[(x + y*W)*3 + c] = Gamma(2.2)(Filmic(film / samples)), the image the reference
 * writes as <output>.png.  Replaces postprocess::Filmic::operator()
 * (src/amber/postprocess/filmic.cc:28-67) and postprocess::Gamma::operator()
 * (src/amber/postprocess/gamma.cc:36-52) as applied at src/amber/cli/application.cc:98-106.
 * flip_x != 0 mirrors the columns as ExportPNG does (src/amber/cli/image.cc:59-71). */
int amber_cuda_resolve_ldr(amber_cuda_ctx* ctx, uint8_t* host_rgb8, int flip_x, uint64_t* samples_done);

/* The raw (un-normalised) device film for the multi-process film reduction done by the host (ONE
 * NCCL all-reduce over this buffer): W*H float4 (r,g,b,unused) sums followed by one float4 that
 * carries the sample count as (count & 0xffff, (count >> 16) & 0xffff, count >> 32, 0) -- exact in
 * float under summation over up to 256 ranks.  n_floats = 4 * (W*H + 1).  After the all-reduce,
 * amber_cuda_film_set_samples(ctx, AMBER_CUDA_SAMPLES_IN_FILM) makes resolve take the count from
 * that element on the device (no host round trip); any other value sets the count directly. */
#define AMBER_CUDA_SAMPLES_IN_FILM UINT64_MAX
int amber_cuda_film_device(amber_cuda_ctx* ctx, void** device_ptr, uint64_t* n_floats,
                           uint64_t* samples_done);
int amber_cuda_film_set_samples(amber_cuda_ctx* ctx, uint64_t samples_done);

/* Closest hit for n rays in HOST memory: origins/directions are n*3 floats
 * (directions must be unit, as prelude::Ray stores them, include/amber/prelude/ray.h:52-62).
 * prim[i] = primitive id or -1, t[i] = hit distance (NaN on a miss, as
 * include/amber/prelude/hit.h:62-66).  Replaces scene::Scene::Cast ->
 * Acceleration::Cast (include/amber/scene/scene.h:236-244,
 * include/amber/raytracer/acceleration.h:46-51). */
int amber_cuda_cast(amber_cuda_ctx* ctx, const float* origins, const float* directions,
                    uint64_t n, int32_t* prim, float* t);

/* Same, device-resident SoA float4 rays (xyz + pad) and float4 hit records
 * (t, prim id bits, u, v): the `extend` kernel exactly as the wavefront launches it.
 * elapsed_ms (may be NULL) receives the CUDA-event time of the launch. */
int amber_cuda_cast_device(amber_cuda_ctx* ctx, const void* ray_origin4,
                           const void* ray_direction4, uint64_t n, void* hit4,
                           float* elapsed_ms);

/* Closest hit for n rays WITHOUT ANY ALLOCATION.  This entry point has no counterpart in the reference; it was
 * added for callers that cannot call cudaMalloc (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid it).  Same result as amber_cuda_cast, but inputs and outputs
 * are caller-owned DEVICE buffers on the context's device (origins / directions: n*3 floats, unit directions;
 * prim: n int32, -1 on a miss; t: n floats, NaN on a miss) and the packed rays and hit records in between live in a
 * static device array of the library sized for AMBER_CUDA_CAST_STATIC_MAX rays (3 MiB, held by every GPU that loads
 * the library until the process exits).  n beyond that is refused with AMBER_CUDA_EINVAL.  The call makes no
 * cudaMalloc / cudaFree or other allocating call; it runs on the context's stream and returns when the results
 * are in device_prim / device_t (the caller synchronises its own producer of the inputs first).  Every call on
 * a device shares the one array, which keeps the last call's contents (each call overwrites what it reads);
 * calls that overlap on the same GPU -- from different contexts or processes' threads -- race on it, and
 * contexts that share a device already take turns (see the conventions above). */
#define AMBER_CUDA_CAST_STATIC_MAX 65536
int amber_cuda_cast_static(amber_cuda_ctx* ctx, const float* device_origins3, const float* device_directions3,
                           uint64_t n, int32_t* device_prim, float* device_t);

/* Per-vertex unit checks (parity entry points; each runs one tiny kernel).
 * sample_light: n BSDF samples for material `material` at normal/dir_out with the
 * scripted uniforms xi[n][8]; outputs direction, rgb weight and draws consumed.
 * Replaces Material<RGB>::SampleLight (include/amber/scene/material.h:31-88). */
int amber_cuda_sample_light(amber_cuda_ctx* ctx, int32_t material, uint64_t n,
                            const float* normal3, const float* dir_out3, const float* xi8,
                            float* dir_in3, float* weight3, int32_t* draws);

/* sample_importance: the same for Material<RGB>::SampleImportance, the light-tracing
 * direction (include/amber/scene/material.h:31-88; src/amber/scene/material_refraction.cc:222-263
 * is the one material that differs from SampleLight). */
int amber_cuda_sample_importance(amber_cuda_ctx* ctx, int32_t material, uint64_t n,
                                 const float* normal3, const float* dir_out3, const float* xi8,
                                 float* dir_in3, float* weight3, int32_t* draws);

/* light_rays: n rays leaving the lights with scripted uniforms xi[n][5] (light choice, point u,
 * v, direction r0, r1); weight3 = Irradiance / PDFArea.  Replaces scene::Scene::GenerateLightRay
 * (include/amber/scene/scene.h:225-234) -> LightSet::GenerateRay
 * (include/amber/scene/light_set.h:83-104). */
int amber_cuda_light_rays(amber_cuda_ctx* ctx, uint64_t n, const float* xi5, float* origin3,
                          float* direction3, float* weight3);

/* lens_response: for n rays arriving at aperture point position3[i] from direction dir_out3[i]
 * (unit, pointing away from the lens), the pixel x + y*W they land on (-1: none) and the
 * response value.  Replaces scene::Scene::Response (include/amber/scene/scene.h:298-307) ->
 * BasicThin::Response (src/amber/scene/lens_thin.cc:109-130) + Sensor::ResponsePixel
 * (src/amber/rendering/sensor.cc:52-65). */
int amber_cuda_lens_response(amber_cuda_ctx* ctx, uint64_t n, const float* position3,
                             const float* dir_out3, int32_t* pixel, float* value);

/* normalize: unit3[i] = vector3[i] / |vector3[i]| with the device's exact normalisation (IEEE sqrt and
 * three IEEE divisions, bit for bit the host's).  Replaces prelude::Normalize
 * (include/amber/prelude/vector3.h:309-316); a unit check for the parity tests.  With denominator != NULL
 * the result is vector3[i] / denominator[i] instead (the same division sequence, e.g. the barycentrics of
 * src/amber/scene/primitive_triangle.cc:108-118 over det). */
int amber_cuda_normalize(amber_cuda_ctx* ctx, uint64_t n, const float* vector3, const float* denominator,
                         float* unit3);

/* eye_rays: n camera rays for pixel (px[i], py[i]) with scripted uniforms xi[n][5].
 * Replaces scene::Scene::GenerateEyeRay (include/amber/scene/scene.h:210-223,
 * src/amber/scene/lens_thin.cc:70-107). */
int amber_cuda_eye_rays(amber_cuda_ctx* ctx, uint64_t n, const uint32_t* px,
                        const uint32_t* py, const float* xi5, float* origin3,
                        float* direction3, float* weight);

/* Counters; see amber_cuda_stats.  profile != 0 makes the next renders time each
 * kernel class with CUDA events (slower: one event pair per launch). */
int amber_cuda_get_stats(amber_cuda_ctx* ctx, amber_cuda_stats* out);
int amber_cuda_set_profiling(amber_cuda_ctx* ctx, int profile);

#ifdef __cplusplus
}
#endif

#endif /* AMBER_CUDA_H_ */
