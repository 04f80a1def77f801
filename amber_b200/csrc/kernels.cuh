// kernels.cuh -- the wavefront path tracer: raygen, extend, shade, connect, resolve.
//
// One pool of P resident path slots.  Per wavefront step:
//   raygen   : slots freed by the previous shade take the next (pixel, sample) and a
//              thin-lens camera ray             (algorithm_pt.cc:112-135, lens_thin.cc:70-107)
//   bookkeep : one thread advances the path counter and clears the queue counters
//   extend   : closest hit for every live slot  (scene.h:236-244 -> acceleration Cast)
//   shade    : emission, [light sampling -> shadow queue], BSDF sample, Russian
//              roulette, next ray in place      (algorithm_pt.cc:137-157)
//   connect  : any-hit for the shadow queue + film add
// `shade` sorts each 512-slot tile by hit class in shared memory so that a warp shades one
// material; freed slots are compacted into `q_new` and slots with a pending shadow ray into
// `q_shadow`: per-thread bit masks, warp shuffles, one smem scan and ONE global atomic per
// queue per tile.
// The same wavefront runs amber's light tracer (algorithm_lt.cc:112-163): raygen<true> starts paths
// on the lights, shade<true> scatters importance and splats paths that land on the aperture.
// Bidirectional path tracing is one kernel of its own (bdpt.cuh, included at the end).
//
// Layout in HBM (all float4 / 16-byte accesses):
//   ray[2i], ray[2i+1]       origin xyz + live flag, direction xyz             32 B per slot
//   state[2i], state[2i+1]   throughput rgb ; (pixel, sample, depth|flags, -)  32 B per slot
//   hit[i]                   (t, primitive | -1 miss | -2 empty slot)           8 B per slot
//   shadow[2i], shadow[2i+1] direction + t_max ; contribution rgb + pixel      32 B per slot
//   q_new[j], q_shadow[j]    slot ids                                           4 B per entry
// A slot's ray, state and shadow record each fill exactly one 32-byte sector, so the
// permuted accesses of raygen / shade / connect never touch a partial sector.
#pragma once

#include "bvh.cuh"
#include "intersect.cuh"
#include "materials.cuh"

namespace amber_cuda {

constexpr int kBlock = 256;
constexpr uint32_t kFlagCountEmission = 1u << 31;  // emission at the next hit counts
constexpr uint32_t kDepthMask = 0xffffffu;
constexpr int32_t kEstimatorNee = 1, kEstimatorLightTracing = 2;  // AMBER_CUDA_ESTIMATOR_*

struct Control {
  unsigned long long next_path;    // next path index of this render call
  unsigned long long total_paths;  // W * H * sample_count
  unsigned int n_new;              // entries in q_new
  unsigned int n_shadow;           // entries in the shadow queue
  unsigned int n_alive;            // survivors of the last shade
  unsigned int done;
  unsigned int cursor_extend, cursor_connect, cursor_cast;  // persistent BVH kernels: next item to fetch
  unsigned int base_pixel;         // next_path % n_pixels, kept by bookkeep so raygen divides 32-bit numbers
  unsigned long long paths, closest_rays, shadow_rays, truncated, phong_capped;
  unsigned long long splats;  // light tracing: film contributions through the lens
  unsigned long long base_sample;  // next_path / n_pixels
  unsigned int n_pixels;
  unsigned int n_marked;  // free-list entries [0, n_marked) are known to hold no ray (ray.w == 0)
  unsigned int raygen_blocks_done;  // blocks of the running raygen launch that have finished (the last one does the bookkeeping)
  unsigned long long cursor_bdpt;  // bdpt_kernel: next (pixel, sample) pair index
  // wavefront BDPT (bdpt.cuh): items waiting for their connections, by iteration parity; items that found no pair left
  unsigned int bdpt_n_connect[2], bdpt_done, bdpt_connect_cursor, bdpt_n_evaluate, bdpt_pad;
  // profiling mode: traversal work of the BVH kernels: extend node steps, extend triangle tests, connect
  // node steps, connect triangle tests
  unsigned long long bvh_work[4];
};

struct PathPool {
  float4* ray;      // [2 * n_slots]
  float4* state;    // [2 * n_slots]
  float2* hit;      // [n_slots]
  uint32_t* q_new;  // [n_slots] slots freed by the last shade
  float4* shadow;   // [2 * n_slots] per-slot shadow ray: direction + t_max ; contribution + pixel
  uint32_t* q_shadow;  // [n_slots] slots whose shadow ray is pending
  uint32_t n_slots;
};

struct RenderArgs {
  PathPool pool;
  SceneView scene;
  Control* ctl;
  float4* film;
  uint64_t seed;
  uint64_t sample_begin;
  uint32_t n_pixels;      // pixels sampled by this render: the window, or the whole frame
  uint32_t max_depth;
  int32_t estimator;
  float inv_n_pixels;  // light tracing: each splat is divided by image.Size(), algorithm_lt.cc:146
  uint32_t pixel_begin;   // first pixel of the window (index in the full frame)
  uint32_t sample_stride; // global sample index = sample_begin + ordinal * sample_stride
};

__constant__ CameraDev c_camera;  // single translation unit: defined here

// Scripted uniforms for the unit-check entry points (amber_cuda_sample_light / _eye_rays).
struct ScriptedRng {
  const float* xi;
  int n;
  int draws;
  __device__ __forceinline__ float next() {
    // past the scripted list: a Weyl sequence, identical in the CPU checkers
    const float v = draws < n ? xi[draws]
                              : static_cast<float>((static_cast<uint32_t>(draws) * 2654435769u) >> 8) * (1.0f / 16777216.0f);
    draws++;
    return v;
  }
};

// ---------------------------------------------------------------------------------
// Camera ray: PixelBound::Uniform (src/amber/rendering/sensor.cc:111-120) +
// BasicThin::GenerateRay (src/amber/scene/lens_thin.cc:70-107) incl. PDFArea/PDFDirection
// (:132-148).  Five uniforms in the compiled reference's order: blade, u, v, film y, film x.
template <class Rng>
__device__ __forceinline__ void generate_eye_ray(uint32_t px, uint32_t py, Rng& rng, float3& origin,
                                                 float3& direction, float& weight, float3* lens_normal = nullptr) {
  // Origin and direction are computed with the host's operations one by one (no FMA
  // contraction, IEEE sqrt / division): the camera sits 3-5 units from the scene, where
  // one ulp of the direction moves the first hit by more than kEPS, so only bit-identical
  // camera rays give the reference's first-hit behaviour on curved surfaces.
  const LensDev& lens = c_camera.lens;
  const SensorDev& sensor = c_camera.sensor;
  if (lens.kind == 1) {
    // BasicPinhole::GenerateRay / PDFArea / PDFDirection, src/amber/scene/lens_pinhole.cc:48-69,
    // 87-108: the ray from the lens origin through the film point mirrored about it; weight =
    // 1 / kDiracDelta / (p_area / geometry factor).  Two uniforms (film y, film x).
    const float sv = __fdiv_rn(__fadd_rn(static_cast<float>(py), rng.next()), static_cast<float>(sensor.height));
    const float su = __fdiv_rn(__fadd_rn(static_cast<float>(px), rng.next()), static_cast<float>(sensor.width));
    const float3 film_point = v3(-__fmul_rn(su - 0.5f, sensor.film_width), -__fmul_rn(sv - 0.5f, sensor.film_height),
                                 -lens.sensor_distance);
    origin = v3(lens.origin[0], lens.origin[1], lens.origin[2]);
    direction = normalize_exact(mat3_apply_exact(lens.global, film_point));
    const float3 back = mat3_apply_exact(lens.local, direction);
    const float3 point = (lens.sensor_distance / back.z) * back;
    const float geometry_factor = back.z * back.z / dot(point, point);
    weight = 1.0f / (sensor.inv_area / geometry_factor);
    // the leading event's normal (BDPT): Normalize(global(0, 0, -1)), lens_pinhole.cc:62
    if (lens_normal) *lens_normal = normalize_exact(mat3_apply_exact(lens.global, v3(0.0f, 0.0f, -1.0f)));
    return;
  }
  const float nb = static_cast<float>(lens.n_blades);
  int blade = static_cast<int>(floorf(__fmul_rn(rng.next(), nb)));
  blade = min(lens.n_blades - 1, blade);
  const float* b = lens.blades[blade];
  if (lens_normal) *lens_normal = v3(b[9], b[10], b[11]);  // the leading event's normal (BDPT), lens_thin.cc:99
  // Triangle::SampleSurfacePoint, src/amber/scene/primitive_triangle.cc:136-150
  float u = rng.next();
  float v = rng.next();
  if (__fadd_rn(u, v) >= 1.0f) {
    u = 1.0f - u;
    v = 1.0f - v;
  }
  const float w0 = __fsub_rn(__fsub_rn(1.0f, u), v);
  origin = v3(__fadd_rn(__fadd_rn(__fmul_rn(w0, b[0]), __fmul_rn(u, b[3])), __fmul_rn(v, b[6])),
              __fadd_rn(__fadd_rn(__fmul_rn(w0, b[1]), __fmul_rn(u, b[4])), __fmul_rn(v, b[7])),
              __fadd_rn(__fadd_rn(__fmul_rn(w0, b[2]), __fmul_rn(u, b[5])), __fmul_rn(v, b[8])));
  const float3 rel = v3(origin.x - lens.origin[0], origin.y - lens.origin[1], origin.z - lens.origin[2]);
  const float3 aperture = mat3_apply_exact(lens.local, rel);
  // The reference builds Vector2((x + xi)/W, (y + xi)/H) (sensor.cc:114-117); argument
  // evaluation order is unspecified and the g++ build used as the oracle draws Y first.
  const float sv = __fdiv_rn(__fadd_rn(static_cast<float>(py), rng.next()), static_cast<float>(sensor.height));
  const float su = __fdiv_rn(__fadd_rn(static_cast<float>(px), rng.next()), static_cast<float>(sensor.width));
  const float3 sensor_point = v3(__fmul_rn(su - 0.5f, sensor.film_width), __fmul_rn(sv - 0.5f, sensor.film_height),
                                 lens.sensor_distance);
  const float k = lens.focus_over_sensor;  // -focus_distance / sensor_distance
  const float3 dir_local = normalize_exact(v3(__fsub_rn(__fmul_rn(k, sensor_point.x), aperture.x),
                                              __fsub_rn(__fmul_rn(k, sensor_point.y), aperture.y),
                                              __fsub_rn(__fmul_rn(k, sensor_point.z), aperture.z)));
  direction = normalize_exact(mat3_apply_exact(lens.global, dir_local));
  // weight = factor / PDFArea / PDFDirection (lens_thin.cc:96-105,132-148); the host
  // evaluates the pow() calls in double, float is within 1e-6 relative of it
  const float3 to_sensor = normalize(sensor_point - aperture);
  const float ratio = to_sensor.z / dir_local.z;
  const float r2 = ratio * ratio;
  const float factor = r2 * r2;
  const float* L = lens.local;
  const float dz = L[6] * direction.x + L[7] * direction.y + L[8] * direction.z;
  const float dz2 = dz * dz;
  const float pdf_direction =
      sensor.size_over_area * (lens.sensor_distance * lens.sensor_distance) / (dz2 * dz2);
  weight = factor / lens.p_area / pdf_direction;
}

// ---------------------------------------------------------------------------------
// Light ray (light tracing, row N2): LightSet::GenerateRay, include/amber/scene/light_set.h:83-104.
// Light by lower_bound over the cumulative powers of Uniform(total power); point by
// Triangle::SampleSurfacePoint (primitive_triangle.cc:136-150); direction by HemispherePSA about
// the triangle's stored normal (sampling.h:234-265); weight = Irradiance / PDFArea (host-side,
// in LightDev).  Five uniforms, like the camera vertex.
// An emitting sphere: Sphere::SampleSurfacePoint (src/amber/scene/primitive_sphere.cc:115-122), normal =
// SphereSA (include/amber/prelude/sampling.h:185-199) exactly as written there -- cos_theta = Uniform(-1, 1),
// sin_theta = sqrt(1 - cos_theta^2) >= 0, (cos_theta cos phi, cos_theta sin phi, sin_theta): it only reaches
// the z >= 0 half of the sphere; light tracing and BDPT reproduce the reference.
template <class Rng>
__device__ __forceinline__ void generate_light_ray(const SceneView& scene, Rng& rng, float3& origin,
                                                   float3& direction, float3& weight, int* light_index = nullptr,
                                                   float3* surface_normal = nullptr) {
  const float target = __fmul_rn(rng.next(), scene.total_power);
  int lo = 0, hi = scene.n_lights - 1;
  while (lo < hi) {  // first light whose cumulative power >= target
    const int mid = (lo + hi) >> 1;
    if (__ldg(&scene.lights[mid].v0.w) < target) lo = mid + 1; else hi = mid;
  }
  if (light_index) *light_index = lo;
  const float4* L = reinterpret_cast<const float4*>(scene.lights + lo);
  const float4 v0 = __ldg(L), v1 = __ldg(L + 1), v2 = __ldg(L + 2), nl = __ldg(L + 3);
  const float radius = __ldg(L + 5).x;
  float u = rng.next();
  float v = rng.next();
  float3 axis;
  if (radius > 0.0f) {
    const float r0 = __fadd_rn(__fmul_rn(u, 2.0f), -1.0f);  // Uniform(min, max): sampler() * (max - min) + min
    const float sin_theta = __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(r0, r0)));
    float sin_phi, cos_phi;
    sincosf(__fmul_rn(2.0f * kPiF, v), &sin_phi, &cos_phi);
    axis = v3(__fmul_rn(r0, cos_phi), __fmul_rn(r0, sin_phi), sin_theta);
    origin = v3(__fadd_rn(v0.x, __fmul_rn(radius, axis.x)), __fadd_rn(v0.y, __fmul_rn(radius, axis.y)),
                __fadd_rn(v0.z, __fmul_rn(radius, axis.z)));
  } else {
    if (__fadd_rn(u, v) >= 1.0f) {
      u = 1.0f - u;
      v = 1.0f - v;
    }
    const float w0 = __fsub_rn(__fsub_rn(1.0f, u), v);
    origin = v3(__fadd_rn(__fadd_rn(__fmul_rn(w0, v0.x), __fmul_rn(u, v1.x)), __fmul_rn(v, v2.x)),
                __fadd_rn(__fadd_rn(__fmul_rn(w0, v0.y), __fmul_rn(u, v1.y)), __fmul_rn(v, v2.y)),
                __fadd_rn(__fadd_rn(__fmul_rn(w0, v0.z), __fmul_rn(u, v1.z)), __fmul_rn(v, v2.z)));
    axis = v3(nl);
  }
  if (surface_normal) *surface_normal = axis;
  float3 bu, bv;
  orthonormal_basis(axis, bu, bv);
  const float r0 = rng.next();
  const float r1 = rng.next();
  direction = local_to_world(bu, bv, axis, __fsqrt_rn(1.0f - r0), __fsqrt_rn(r0), __fmul_rn(2.0f * kPiF, r1));
  weight = v3(v1.w, v2.w, nl.w);
}

// Lens response (light tracing): BasicThin::Response, src/amber/scene/lens_thin.cc:109-130, with
// Sensor::ResponsePixel / UVToPixel, src/amber/rendering/sensor.cc:52-65,122-129.  `position` is a
// point on the aperture, `dir_out` points from it back along the arriving ray.  The pixel is a
// discrete decision, so the projection onto the film uses the host's operations one by one.
__device__ __forceinline__ bool lens_response(float3 position, float3 dir_out, uint32_t& pixel, float& value) {
  const LensDev& lens = c_camera.lens;
  const SensorDev& sensor = c_camera.sensor;
  const float3 direction = mat3_apply_exact(lens.local, dir_out);
  const float k_direction = __fdiv_rn(lens.sensor_distance, direction.z);
  float3 aperture = v3(0.0f, 0.0f, 0.0f);
  float3 sensor_point;
  if (lens.kind == 1) {  // BasicPinhole::Response, lens_pinhole.cc:71-85: no facing test, value 1
    sensor_point = scale_exact(k_direction, direction);
  } else {
    if (direction.z >= 0.0f) return false;
    aperture = mat3_apply_exact(
        lens.local, v3(position.x - lens.origin[0], position.y - lens.origin[1], position.z - lens.origin[2]));
    const float k_aperture = __fdiv_rn(-lens.sensor_distance, lens.focus_distance);
    sensor_point = add_exact(scale_exact(k_aperture, aperture), scale_exact(k_direction, direction));
  }
  const float u = __fadd_rn(__fdiv_rn(sensor_point.x, sensor.film_width), 0.5f);
  const float v = __fadd_rn(__fdiv_rn(sensor_point.y, sensor.film_height), 0.5f);
  if (fminf(u, v) < 0.0f || fmaxf(u, v) >= 1.0f) return false;
  const uint32_t ix = min(sensor.width - 1u, static_cast<uint32_t>(__fmul_rn(u, static_cast<float>(sensor.width))));
  const uint32_t iy = min(sensor.height - 1u, static_cast<uint32_t>(__fmul_rn(v, static_cast<float>(sensor.height))));
  pixel = ix + iy * sensor.width;
  if (lens.kind == 1) {
    value = 1.0f;
    return true;
  }
  const float ratio = normalize(sensor_point - aperture).z / direction.z;
  const float r2 = ratio * ratio;
  value = r2 * r2;
  return true;
}

// Grid-stride over the q_new entries: the host does not know n_new, so the grid is sized for the
// device (resident blocks), not for the pool -- a launch over every slot spent 40 % of its stall samples
// in blocks that only read n_new and left.
#ifndef AMBER_QUEUE_BLOCKS
#define AMBER_QUEUE_BLOCKS 6
#endif
constexpr int kQueueBlocks = AMBER_QUEUE_BLOCKS;  // resident blocks per SM of the grid-stride queue kernels
#ifndef AMBER_RAYGEN_BLOCKS
#define AMBER_RAYGEN_BLOCKS 4
#endif
#ifndef AMBER_RAYGEN_LOOP
#define AMBER_RAYGEN_LOOP 1
#endif
#ifndef AMBER_FUSE_BOOKKEEP
#define AMBER_FUSE_BOOKKEEP 1
#endif
// Between raygen and extend: advances the render's counters by what raygen has just started and resets the per-iteration
// queues.  Run by the LAST block of raygen to finish (AMBER_FUSE_BOOKKEEP, the default) instead of a launch of its own.
__device__ __forceinline__ void bookkeep(Control* ctl) {
  unsigned long long generated = 0;
  if (ctl->next_path < ctl->total_paths) {
    const unsigned long long left = ctl->total_paths - ctl->next_path;
    generated = ctl->n_new < left ? ctl->n_new : left;
  }
  ctl->next_path += generated;
  ctl->base_pixel = static_cast<unsigned int>(ctl->next_path % ctl->n_pixels);
  ctl->base_sample = ctl->next_path / ctl->n_pixels;
  ctl->paths += generated;
  ctl->shadow_rays += ctl->n_shadow;
  // finished: every path of the render has been started and none is alive (n_alive: survivors of the last shade)
  ctl->done = (generated == 0 && ctl->next_path >= ctl->total_paths && ctl->n_alive == 0) ? 1u : 0u;
  ctl->n_new -= static_cast<unsigned int>(generated);  // the free list keeps what raygen did not use ...
  ctl->n_marked = ctl->n_new;                           // ... and raygen has marked all of that empty
  ctl->n_shadow = 0;
  ctl->n_alive = 0;
  ctl->cursor_extend = 0;
  ctl->cursor_connect = 0;
}

// Experiment: with the constant-memory list the closest hit of a path's NEXT ray is found by the
// kernel that makes the ray (raygen, shade) and `extend` is not launched.  Off: see DESIGN.md section 4.4.
#ifndef AMBER_FUSE_EXTEND
#define AMBER_FUSE_EXTEND 0
#endif
template <bool kLightTracing>
__global__ void __launch_bounds__(kBlock, AMBER_RAYGEN_BLOCKS) raygen_kernel(const RenderArgs args) {
  // q_new is the pool's FREE LIST: entries [0, n_new) are slots without a path.  This launch starts
  // `generated` = min(n_new, paths left) paths in the LAST `generated` entries; what is left at the front
  // stays on the list for a later launch (bookkeep shortens n_new by `generated`, shade appends the slots it
  // frees) -- so a render call absorbed into a wavefront that has partly drained finds the free slots.
  // A slot whose path left the scene still holds that path's ray (w = 1): when such a slot stays on the list
  // unused it is marked empty here (entries [n_marked, first_entry): only the ones appended since the last
  // launch; nothing to do while every free slot is reused at once, the steady state).
  const Control* ctl = args.ctl;
  if (ctl->done) return;  // the render has finished: the iterations queued behind the last useful one cost a launch each
  const unsigned long long next_path = ctl->next_path, total_paths = ctl->total_paths;
  const uint32_t n_free = ctl->n_new;
  const unsigned long long left = next_path < total_paths ? total_paths - next_path : 0ull;
  const uint32_t generated = left < n_free ? static_cast<uint32_t>(left) : n_free;
  const uint32_t first_entry = n_free - generated;
  const uint32_t n_new = AMBER_RAYGEN_LOOP ? generated : min(generated, (blockIdx.x + 1) * kBlock);
  const uint32_t n_marked = ctl->n_marked;
  const unsigned long long base_pixel = ctl->base_pixel, base_ordinal = ctl->base_sample;
  const uint32_t width = c_camera.sensor.width;
  const uint32_t* __restrict__ entries = args.pool.q_new + first_entry;
  for (uint32_t j = blockIdx.x * kBlock + threadIdx.x; j < n_new; j += gridDim.x * kBlock) {
    const uint32_t slot = entries[j];
    // Never taken; it makes the slot id a dependency at the TOP of the iteration.  Without it the compiler
    // keeps the load in flight until the stores at the bottom and reuses its address registers for them, and
    // the next iteration's load then waits on stores still queued in the LSU (ncu: lg_throttle 0.19 -> 1.04
    // stalls per issue): 52.0 us per launch against 45.9 us with this line (1.68 M new paths).
    if (slot >= args.pool.n_slots) continue;
    float4* ray = args.pool.ray + 2 * static_cast<size_t>(slot);
    // path = next_path + j = (base_sample * n_pixels + base_pixel) + j: the 64-bit division is done once
    // per launch by `bookkeep`, a thread divides the 32-bit (base_pixel + j)
    const unsigned long long offset = base_pixel + j;
    uint32_t pixel, wrapped;
    if (offset >> 32) {
      pixel = static_cast<uint32_t>(offset % args.n_pixels);
      wrapped = static_cast<uint32_t>(offset / args.n_pixels);
    } else {
      wrapped = static_cast<uint32_t>(offset) / args.n_pixels;
      pixel = static_cast<uint32_t>(offset) - wrapped * args.n_pixels;
    }
    const uint32_t sample = static_cast<uint32_t>(args.sample_begin + (base_ordinal + wrapped) * args.sample_stride);
    pixel += args.pixel_begin;  // the stream is keyed by the pixel's index in the full frame
    CameraRng rng;
    rng.init(args.seed, pixel, sample);
    float3 o, d, w3;
    if constexpr (kLightTracing) {
      // one sample = W*H light paths (algorithm_lt.cc:112-122); `pixel` is only the path's stream key
      generate_light_ray(args.scene, rng, o, d, w3);
    } else {
      float w;
      generate_eye_ray(pixel % width, pixel / width, rng, o, d, w);
      w3 = v3(w, w, w);
    }
    ray[0] = make_float4(o.x, o.y, o.z, 1.0f);
    ray[1] = make_float4(d.x, d.y, d.z, 0.0f);
    float4* state = args.pool.state + 2 * static_cast<size_t>(slot);
    state[0] = make_float4(w3.x, w3.y, w3.z, 0.0f);
    state[1] = make_float4(__uint_as_float(pixel), __uint_as_float(sample), __uint_as_float(kFlagCountEmission), 0.0f);
#if AMBER_FUSE_EXTEND
    if (!args.scene.bvh_nodes) {
      Ray first;
      first.o = o;
      first.d = d;
      const HitRecord h = ConstList::closest(args.scene, first, true);
      args.pool.hit[slot] = make_float2(h.t, __int_as_float(h.prim));
    }
#endif
  }
  // (after the main loop: in front of it the main loop's loads of the control block are ordered behind these stores)
  for (uint32_t j = n_marked + blockIdx.x * kBlock + threadIdx.x; j < first_entry; j += gridDim.x * kBlock)
    args.pool.ray[2 * static_cast<size_t>(args.pool.q_new[j])] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#if AMBER_FUSE_BOOKKEEP
  // The last block to arrive here advances the counters: by then every block has used the values it read above.
  // (One launch per iteration less: a one-thread kernel costs its ~2.6 us plus the gaps either side of it.)
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&args.ctl->raygen_blocks_done, 1u) == gridDim.x - 1) {
      args.ctl->raygen_blocks_done = 0;
      bookkeep(args.ctl);
    }
  }
#endif
}

// Fills q_new with every slot (start of a render call: all slots are free).
__global__ void __launch_bounds__(kBlock) pool_reset_kernel(const RenderArgs args) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= args.pool.n_slots) return;
  args.pool.q_new[i] = i;
  args.pool.ray[2 * static_cast<size_t>(i)] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
#if AMBER_FUSE_EXTEND
  args.pool.hit[i] = make_float2(CUDART_NAN_F, __int_as_float(-2));
#endif
}

#if !AMBER_FUSE_BOOKKEEP
__global__ void bookkeep_kernel(Control* ctl) { bookkeep(ctl); }
#endif

// ---------------------------------------------------------------------------------
// Every thread reaches Accel::closest (the BVH walk votes across the warp); threads without a
// ray pass active = false.
template <class Accel>
__global__ void __launch_bounds__(kBlock) extend_kernel(const RenderArgs args) {
  if (args.ctl->done) return;
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  const bool in_range = i < args.pool.n_slots;
  float4 o = make_float4(0.0f, 0.0f, 0.0f, 0.0f), d = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
  if (in_range) {
    const float4* r = args.pool.ray + 2 * static_cast<size_t>(i);
    o = r[0];
    if (o.w != 0.0f) d = r[1];
  }
  const bool active = in_range && o.w != 0.0f;
  Ray ray;
  ray.o = v3(o);
  ray.d = v3(d);
  const HitRecord hit = Accel::closest(args.scene, ray, active);
  if (active) {
    args.pool.hit[i] = make_float2(hit.t, __int_as_float(hit.prim));
  } else if (in_range) {  // no path in this slot: tell `shade` so (it sorts the tile by hit class)
    args.pool.hit[i] = make_float2(CUDART_NAN_F, __int_as_float(-2));
  }
}

// `extend` on caller-supplied rays (amber_cuda_cast / amber_cuda_cast_device).
template <class Accel>
__global__ void __launch_bounds__(kBlock) cast_kernel(const SceneView scene, const float4* __restrict__ ray_o,
                                                      const float4* __restrict__ ray_d, uint64_t n,
                                                      float4* __restrict__ hits) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kBlock + threadIdx.x;
  const bool active = i < n;
  Ray ray;
  ray.o = active ? v3(ray_o[i]) : v3(0.0f, 0.0f, 0.0f);
  ray.d = active ? v3(ray_d[i]) : v3(0.0f, 0.0f, 1.0f);
  const HitRecord hit = Accel::closest(scene, ray, active);
  if (active) hits[i] = make_float4(hit.t, __int_as_float(hit.prim), 0.0f, 0.0f);
}

// ---------------------------------------------------------------------------------
// Large scenes: `extend` / cast as persistent kernels over the 4-wide BVH (bvh.cuh,
// trace_persistent).  A source hands rays to idle lanes and takes finished hits back.
// Spheres (constant memory) are tested when a ray is loaded and bound the walk.
struct SphereBound {
  int sphere;  // index of the closest sphere or -1
  __device__ __forceinline__ float bound(const SceneView& scene, const Ray& ray) {
    const HitRecord h = ConstList::closest(scene, ray, true);
    sphere = h.prim;  // c_scene.n_tris == 0 on this path: prim = sphere index
    return sphere >= 0 ? h.t : CUDART_MAX_NORMAL_F;
  }
  __device__ __forceinline__ HitRecord resolve(const SceneView& scene, const HitRecord& hit, bool found) const {
    HitRecord out;
    out.prim = found ? hit.prim : (sphere >= 0 ? scene.n_tris + sphere : -1);
    out.t = out.prim < 0 ? CUDART_NAN_F : hit.t;
    return out;
  }
};

struct ExtendSource {
  const RenderArgs& args;
  SphereBound spheres;
  __device__ __forceinline__ uint32_t count() const { return args.pool.n_slots; }
  __device__ __forceinline__ bool load(uint32_t i, Ray& ray, float& t_best) {
    const float4* r = args.pool.ray + 2 * static_cast<size_t>(i);
    const float4 o = r[0];
    if (o.w == 0.0f) {  // no path in this slot
      args.pool.hit[i] = make_float2(CUDART_NAN_F, __int_as_float(-2));
      return false;
    }
    ray.o = v3(o);
    ray.d = v3(r[1]);
    t_best = spheres.bound(args.scene, ray);
    return true;
  }
  __device__ __forceinline__ void store(uint32_t i, const HitRecord& hit, bool found) const {
    const HitRecord out = spheres.resolve(args.scene, hit, found);
    args.pool.hit[i] = make_float2(out.t, __int_as_float(out.prim));
  }
};

template <>
__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) extend_kernel<WideBvh>(const RenderArgs args) {
  if (args.ctl->done) return;
  ExtendSource source{args, {-1}};
  WideBvh::trace_persistent<false>(args.scene, source, &args.ctl->cursor_extend);
}

// profiling mode: the same kernel with the traversal work counted (amber_cuda_stats.bvh_*)
__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) extend_bvh_counting_kernel(const RenderArgs args) {
  if (args.ctl->done) return;
  ExtendSource source{args, {-1}};
  WideBvh::trace_persistent<false, true>(args.scene, source, &args.ctl->cursor_extend, &args.ctl->bvh_work[0]);
}

struct CastSource {
  const SceneView& scene;
  const float4* ray_o;
  const float4* ray_d;
  float4* hits;
  uint32_t n;
  SphereBound spheres;
  __device__ __forceinline__ uint32_t count() const { return n; }
  __device__ __forceinline__ bool load(uint32_t i, Ray& ray, float& t_best) {
    ray.o = v3(ray_o[i]);
    ray.d = v3(ray_d[i]);
    t_best = spheres.bound(scene, ray);
    return true;
  }
  __device__ __forceinline__ void store(uint32_t i, const HitRecord& hit, bool found) const {
    const HitRecord out = spheres.resolve(scene, hit, found);
    hits[i] = make_float4(out.t, __int_as_float(out.prim), 0.0f, 0.0f);
  }
};

__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) cast_bvh_kernel(const SceneView scene, const float4* __restrict__ ray_o,
                                                          const float4* __restrict__ ray_d, uint32_t n,
                                                          float4* __restrict__ hits, unsigned int* cursor) {
  CastSource source{scene, ray_o, ray_d, hits, n, {-1}};
  WideBvh::trace_persistent<false>(scene, source, cursor);
}

// ---------------------------------------------------------------------------------
// Surface point and normal of a hit, as the reference builds its Hit:
//   triangle: v0 + u E1 + v E2, stored normal   (primitive_triangle.cc:127)
//   sphere  : o + t d, normalize(p - centre)    (primitive_sphere.cc:90-96, hit.h:68-77)
__device__ __forceinline__ void surface_point(const SceneView& scene, const Ray& ray, const HitRecord& hit,
                                              float3& position, float3& normal, int& material) {
  if (hit.prim < scene.n_tris) {
    const float4* tri = reinterpret_cast<const float4*>(scene.tri_shade + hit.prim);
    const float4 a = __ldg(tri), b = __ldg(tri + 1), c = __ldg(tri + 2);
    // The traversal found WHICH triangle; the barycentrics that place the hit point are
    // recomputed here once per hit with the reference's own Moeller-Trumbore sequence
    // (primitive_triangle.cc:100-127), separately rounded and with IEEE division, so the
    // vertex position -- the origin of the next ray -- is the host's bit for bit.
    const float3 v0 = v3(a), E1 = v3(b), E2 = v3(c);
    const float3 P = cross_exact(ray.d, E2);
    const float det = dot_exact(P, E1);
    const float3 T = v3(ray.o.x - v0.x, ray.o.y - v0.y, ray.o.z - v0.z);
    const float3 Q = cross_exact(T, E1);
    // u = (P.T) / det, v = (Q.d) / det: two IEEE divisions over one denominator (vec.cuh, div3_exact;
    // the third lane of it is a constant and folds away)
    const float3 uv = div3_exact(v3(dot_exact(P, T), dot_exact(Q, ray.d), 1.0f), det);
    const float u = uv.x, v = uv.y;
    position = add_exact(add_exact(v0, scale_exact(u, E1)), scale_exact(v, E2));
    normal = v3(a.w, b.w, c.w);
    material = __ldg(scene.tri_material + hit.prim);
  } else {
    const int s = hit.prim - scene.n_tris;
    const float4 sphere = __ldg(scene.spheres + s);
    position = madd_exact(ray.o, hit.t, ray.d);
    normal = normalize_exact(v3(position.x - sphere.x, position.y - sphere.y, position.z - sphere.z));
    material = __ldg(scene.sphere_material + s);
  }
}

__device__ __forceinline__ MaterialDev load_material(const SceneView& scene, int index) {
  const float4* p = reinterpret_cast<const float4*>(scene.materials + index);
  const float4 a = __ldg(p), b = __ldg(p + 1);
  MaterialDev m;
  m.kind = __float_as_int(a.x);
  m.r = a.y; m.g = a.z; m.b = a.w;
  m.param = b.x; m.r0 = b.y;
  m.nee = __float_as_int(b.z);
  m.pad = 0;
  return m;
}

struct ShadowRay {
  float3 d;
  float t_max;
  float3 contribution;
};

// Light sampling for next-event estimation.  Light choice and surface point as
// LightSet::GenerateRay (include/amber/scene/light_set.h:83-104: power CDF
// lower_bound, Triangle::SampleSurfacePoint); contribution as the s = 1 connection
// of include/amber/rendering/bidirectional_path_sampling.h:956-1016:
//   Le(y -> x) / pdf_area * throughput * BSDF(x) * |cos_x cos_y| / r^2.
// Three uniforms: light choice, u, v (fixed slots 3, 4, 5 of the vertex).
__device__ __forceinline__ bool sample_direct(const SceneView& scene, const MaterialDev& m, float3 x,
                                              float3 n, float3 wo, float3 throughput, float xi_light,
                                              float u, float v, ShadowRay& shadow) {
  const float target = xi_light * scene.total_power;
  int lo = 0, hi = scene.n_lights - 1;
  while (lo < hi) {  // first light whose cumulative power >= target
    const int mid = (lo + hi) >> 1;
    if (__ldg(&scene.lights[mid].v0.w) < target) lo = mid + 1; else hi = mid;
  }
  const float4* L = reinterpret_cast<const float4*>(scene.lights + lo);
  // an emitting sphere is not light-sampled: its record carries a zero normal, so cos_y below is 0 and no shadow
  // ray is queued when the power CDF picks one; its emission is counted on every BSDF-sampled hit instead
  // (shade_kernel, sampled_by_nee) -- same expectation
  const float4 v0 = __ldg(L), v1 = __ldg(L + 1), v2 = __ldg(L + 2), nl = __ldg(L + 3), le = __ldg(L + 4);
  if (u + v >= 1.0f) {
    u = 1.0f - u;
    v = 1.0f - v;
  }
  const float w0 = 1.0f - u - v;
  const float3 y = v3(w0 * v0.x + u * v1.x + v * v2.x, w0 * v0.y + u * v1.y + v * v2.y,
                      w0 * v0.z + u * v1.z + v * v2.z);
  const float3 xy = y - x;
  const float dist2 = dot(xy, xy);
  const float inv_dist = rsqrtf(dist2);
  const float3 wi = inv_dist * xy;
  const float cos_y = -dot(wi, v3(nl));  // light emits on its normal side only
  if (!(cos_y > 0.0f)) return false;
  const float3 f = bsdf_eval(m, n, wo, wi);
  const float g = fabsf(cos_y * dot(wi, n)) / dist2;
  shadow.contribution = throughput * v3(le) * f * g;
  if (!(max3(shadow.contribution) > 0.0f)) return false;
  shadow.d = wi;
  shadow.t_max = dist2 * inv_dist * (1.0f - 1e-4f);
  return true;
}

__device__ __forceinline__ void film_add(float4* film, uint32_t pixel, float3 value) {
  atomicAdd(&film[pixel], make_float4(value.x, value.y, value.z, 0.0f));
}

// ---------------------------------------------------------------------------------
// shade.  One block owns a TILE of kShadeTile consecutive path slots and works in three
// phases:
//   1. classify: every slot's hit record -> class = (material kind, triangle|sphere),
//      or MISS / DEAD; a counting sort of the tile's slot indices by class in shared
//      memory (warp match + one smem atomic per distinct class per warp).
//   2. shade in sorted order: warps take 32-entry chunks of the sorted tile from a shared
//      counter, so a warp runs ONE material's code on 32 lanes instead of six materials'
//      code on a few lanes each, and no warp idles while another works through an
//      expensive class.  Ray / state records are 32-byte sectors, so the permuted accesses
//      inside the tile's window still move whole sectors.
//   3. queues: each thread has one bit per chunk for "slot freed" and "shadow ray
//      emitted"; one block scan and ONE global atomic per queue per tile hand out the
//      q_new / q_shadow ranges.
// The shadow ray itself stays in the slot (pool.shadow[2 slot..]); its origin is the
// vertex position, which phase 2 has just written to ray[2 slot].
#ifndef AMBER_SHADE_MIN_BLOCKS
#define AMBER_SHADE_MIN_BLOCKS 8
#endif
#ifndef AMBER_SHADE_ROUNDS
#define AMBER_SHADE_ROUNDS 4
#endif
constexpr int kShadeRounds = AMBER_SHADE_ROUNDS;

__device__ __forceinline__ void cp_async_16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(gmem));
}
__device__ __forceinline__ void cp_async_8(void* smem, const void* gmem) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(smem))), "l"(gmem));
}
// atomicAdd on a __shared__ counter as ONE ATOMS.ADD: the C++ atomicAdd is expanded by the compiler into a
// warp-aggregated sequence (vote, leader, popc, shuffle: ~15 instructions), which the callers here
// have already done by hand (one elected lane per warp).
__device__ __forceinline__ uint32_t atom_shared_add(uint32_t* counter, uint32_t value) {
  uint32_t old;
  asm volatile("atom.shared.add.u32 %0, [%1], %2;"
               : "=r"(old)
               : "r"(static_cast<uint32_t>(__cvta_generic_to_shared(counter))), "r"(value)
               : "memory");
  return old;
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
__device__ __forceinline__ void cp_async_wait_all_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }
#ifndef AMBER_SHADE_BLOCK
#define AMBER_SHADE_BLOCK 128
#endif
constexpr int kShadeBlock = AMBER_SHADE_BLOCK;  // threads per block of `shade`
constexpr int kShadeTile = kShadeRounds * kShadeBlock;
constexpr int kClassMiss = 14;  // hit.prim == -1: the path left the scene, algorithm_pt.cc:140-142
constexpr int kClassDead = 15;  // hit.prim == -2: slot holds no path
constexpr int32_t kPrimMiss = -1, kPrimDead = -2;

// kLightTracing selects the light-tracing variant (row N2) at compile time, so the path tracer's
// instantiation keeps its register allocation.
template <bool kLightTracing>
__global__ void __launch_bounds__(kShadeBlock, AMBER_SHADE_MIN_BLOCKS) shade_kernel(const RenderArgs args) {
  __shared__ uint16_t sorted[kShadeTile];
  __shared__ float4 stage[2][4][kShadeBlock];  // double-buffered ray.o, ray.d, throughput, path ids
  __shared__ float2 stage_hit[2][kShadeBlock];
  __shared__ uint32_t class_count[16], class_offset[16];
  __shared__ uint32_t warp_scan[kShadeBlock / 32];
  __shared__ uint32_t queue_base[2];
  __shared__ uint32_t next_chunk;
  if (args.ctl->done) return;  // (block-uniform: before any barrier)
  const PathPool& pool = args.pool;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const uint32_t tile_base = blockIdx.x * kShadeTile;
  const uint32_t lane_lt = (1u << lane) - 1u;
  if (tid < 16) class_count[tid] = 0;
  if (tid == 0) next_chunk = 0;
  __syncthreads();

  // ---- phase 1: classify + counting sort ------------------------------------------
  uint32_t keys[kShadeRounds / 2];  // two (class:4 | rank:12) halves per register
  int prims[kShadeRounds];
#pragma unroll
  for (int r = 0; r < kShadeRounds; r++) {  // all of the tile's hit records in flight at once
    const uint32_t i = tile_base + r * kShadeBlock + tid;
    prims[r] = i < pool.n_slots ? __float_as_int(pool.hit[i].y) : kPrimDead;
  }
#pragma unroll
  for (int r = 0; r < kShadeRounds; r++) {
    const int prim = prims[r];
    const int cls = prim >= 0 ? __ldg(args.scene.prim_class + prim) : (prim == kPrimMiss ? kClassMiss : kClassDead);
    const uint32_t peers = __match_any_sync(0xffffffffu, cls);
    const int leader = __ffs(peers) - 1;
    uint32_t base = 0;
    if (lane == leader) base = atom_shared_add(&class_count[cls], __popc(peers));
    base = __shfl_sync(0xffffffffu, base, leader);
    const uint32_t key = (static_cast<uint32_t>(cls) << 12) | (base + __popc(peers & lane_lt));
    if (r & 1) keys[r >> 1] |= key << 16; else keys[r >> 1] = key;
  }
  __syncthreads();
  if (warp == 0) {
    const uint32_t count = lane < 16 ? class_count[lane] : 0u;
    uint32_t scan = count;
#pragma unroll
    for (int delta = 1; delta < 16; delta <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, scan, delta);
      if (lane >= delta) scan += up;
    }
    if (lane < 16) class_offset[lane] = scan - count;
  }
  __syncthreads();
#pragma unroll
  for (int r = 0; r < kShadeRounds; r++) {
    const uint32_t key = (keys[r >> 1] >> ((r & 1) * 16)) & 0xffffu;
    sorted[class_offset[key >> 12] + (key & 0xfffu)] = static_cast<uint16_t>(r * kShadeBlock + tid);
  }
  __syncthreads();
  const uint32_t n_hit = class_offset[kClassMiss];   // entries [0, n_hit): surface vertices
  const uint32_t n_live = class_offset[kClassDead];  // entries [n_hit, n_live): misses

  // ---- phase 2: shade in class order ------------------------------------------------
  // The sorted tile is cut into chunks of 32 entries; a warp takes the next chunk from a shared
  // counter until none is left, so warps that drew cheap classes (misses, emitters, mirrors) keep
  // taking work instead of waiting at the barrier for the warps that drew refraction and Phong.
  // Each thread streams the records of ITS entry of the warp's NEXT chunk (ray, state, hit: 72 B)
  // into shared memory with cp.async while it shades the current one: the DRAM latency of the
  // permuted gathers is paid once per tile.  Staging is thread-private: no barrier involved.
  static_assert(kShadeTile <= 1024, "one bit per chunk in the 32-bit freed / shadow masks");
  const uint32_t n_chunks = (n_live + 31u) >> 5;
  auto take_chunk = [&]() {
    uint32_t c = 0;
    if (lane == 0) c = atom_shared_add(&next_chunk, 1u);
    return __shfl_sync(0xffffffffu, c, 0);
  };
  auto stage_records = [&](uint32_t chunk, int b) {
    const uint32_t jn = chunk * 32u + lane;
    if (chunk < n_chunks && jn < n_hit) {
      const size_t next = tile_base + sorted[jn];
      cp_async_16(&stage[b][0][tid], pool.ray + 2 * next);
      cp_async_16(&stage[b][1][tid], pool.ray + 2 * next + 1);
      cp_async_16(&stage[b][2][tid], pool.state + 2 * next);
      cp_async_16(&stage[b][3][tid], pool.state + 2 * next + 1);
      cp_async_8(&stage_hit[b][tid], pool.hit + next);
    }
    cp_async_commit();
  };
  uint32_t freed_mask = 0, shadow_mask = 0;  // bit c: my entry of chunk c (sorted[32 c + lane])
  uint32_t chunk = take_chunk();
  int buf = 0;
  stage_records(chunk, buf);
#pragma unroll 1
  for (; chunk < n_chunks; buf ^= 1) {
    const uint32_t r = chunk;  // bit index of this chunk in the masks
    const uint32_t j = chunk * 32u + lane;
    // this chunk's records were copied during the previous one; start the next chunk's copies now
    chunk = take_chunk();
    stage_records(chunk, buf ^ 1);
    cp_async_wait_all_but_one();
    if (j >= n_live) continue;
    if (j >= n_hit) {  // miss: the slot is free again
      freed_mask |= 1u << r;
#if AMBER_FUSE_EXTEND
      if (!args.scene.bvh_nodes) pool.hit[tile_base + sorted[j]] = make_float2(CUDART_NAN_F, __int_as_float(kPrimDead));
#endif
      continue;
    }
    const uint32_t slot = tile_base + sorted[j];
    float4* ray_slot = pool.ray + 2 * static_cast<size_t>(slot);
    float4* state_slot = pool.state + 2 * static_cast<size_t>(slot);
    const float4 o4 = stage[buf][0][tid];
    const float4 d4 = stage[buf][1][tid];
    const float4 t4 = stage[buf][2][tid];
    const float4 m4 = stage[buf][3][tid];
    const float2 h2 = stage_hit[buf][tid];
    HitRecord hit;
    hit.t = h2.x;
    hit.prim = __float_as_int(h2.y);
    Ray ray;
    ray.o = v3(o4);
    ray.d = v3(d4);
    const uint32_t pixel = __float_as_uint(m4.x);
    const uint32_t sample = __float_as_uint(m4.y);
    const uint32_t flags = __float_as_uint(m4.z);
    const uint32_t depth = (flags & kDepthMask) + 1u;  // this is the depth-th surface vertex
    float3 throughput = v3(t4);
    float3 x, n;
    int material_index;
    surface_point(args.scene, ray, hit, x, n, material_index);
    const MaterialDev m = load_material(args.scene, material_index);
    const float3 wo = -ray.d;
    VertexRng rng;
    rng.init(args.seed, pixel, sample, depth);  // one Philox block, same place for every lane

    constexpr bool light_tracing = kLightTracing;
    if constexpr (kLightTracing) {
      // a light path that lands on the aperture is measured there (algorithm_lt.cc:141-148) and
      // then goes on through it like any other surface; nothing else contributes
      if (m.kind == 5) {
        uint32_t film_pixel;
        float response;
        if (lens_response(x, wo, film_pixel, response)) {
          film_add(args.film, film_pixel, (response * args.inv_n_pixels) * throughput);
          atomicAdd(&args.ctl->splats, 1ull);
        }
      }
    } else if (m.kind == 4) {  // emission, algorithm_pt.cc:144
      const bool sampled_by_nee = args.estimator == kEstimatorNee && hit.prim < args.scene.n_tris;
      if (!sampled_by_nee || (flags & kFlagCountEmission)) {
        const float3 le = emitted_radiance(m, n, wo);
        if (le.x > 0.0f || le.y > 0.0f || le.z > 0.0f) film_add(args.film, pixel, throughput * le);
      }
    }
    const bool connect = !kLightTracing && args.estimator == kEstimatorNee && m.nee && args.scene.n_lights > 0;
    if (connect) {
      float lu, lv;
      rng.light_point(lu, lv);
      ShadowRay shadow;
      if (sample_direct(args.scene, m, x, n, wo, throughput, rng.light_choice(), lu, lv, shadow)) {
        shadow_mask |= 1u << r;
        float4* out = pool.shadow + 2 * static_cast<size_t>(slot);
        out[0] = make_float4(shadow.d.x, shadow.d.y, shadow.d.z, shadow.t_max);
        out[1] = make_float4(shadow.contribution.x, shadow.contribution.y, shadow.contribution.z,
                             __uint_as_float(pixel));
      }
    }

    // BSDF sample + Russian roulette, algorithm_pt.cc:145-156
    const Scatter scatter = sample_light(m, n, wo, rng, light_tracing);
    const float p = fminf(kRussianRoulette, max3(scatter.weight));
    if (scatter.capped) atomicAdd(&args.ctl->phong_capped, 1ull);
    bool terminated = false;
    if (!(rng.rr() < p)) {
      terminated = true;
    } else if (depth >= args.max_depth) {
      terminated = true;
      atomicAdd(&args.ctl->truncated, 1ull);
    }
    if (terminated) {
      freed_mask |= 1u << r;
      // the vertex stays readable for `connect` (shadow origin); w = 0 marks the slot free
      ray_slot[0] = make_float4(x.x, x.y, x.z, 0.0f);
#if AMBER_FUSE_EXTEND
      if (!args.scene.bvh_nodes) pool.hit[slot] = make_float2(CUDART_NAN_F, __int_as_float(kPrimDead));
#endif
    } else {
      const float inv_p = __fdividef(1.0f, p);  // scales the throughput only
      throughput = throughput * (inv_p * scatter.weight);
      ray_slot[0] = make_float4(x.x, x.y, x.z, 1.0f);
      ray_slot[1] = make_float4(scatter.dir.x, scatter.dir.y, scatter.dir.z, 0.0f);
      state_slot[0] = make_float4(throughput.x, throughput.y, throughput.z, 0.0f);
      state_slot[1] = make_float4(m4.x, m4.y, __uint_as_float(depth | (connect ? 0u : kFlagCountEmission)), 0.0f);
#if AMBER_FUSE_EXTEND
      if (!args.scene.bvh_nodes) {
        Ray next;
        next.o = x;
        next.d = scatter.dir;
        const HitRecord h = ConstList::closest(args.scene, next, true);
        pool.hit[slot] = make_float2(h.t, __int_as_float(h.prim));
      }
#endif
    }
  }

  // ---- phase 3: queue ranges, one atomic per queue per tile --------------------------
  const uint32_t mine = static_cast<uint32_t>(__popc(freed_mask)) | (static_cast<uint32_t>(__popc(shadow_mask)) << 16);
  uint32_t scan = mine;  // two 16-bit inclusive scans side by side (tile size < 65536)
#pragma unroll
  for (int delta = 1; delta < 32; delta <<= 1) {
    const uint32_t up = __shfl_up_sync(0xffffffffu, scan, delta);
    if (lane >= delta) scan += up;
  }
  if (lane == 31) warp_scan[warp] = scan;
  __syncthreads();
  if (warp == 0) {
    const uint32_t total = lane < kShadeBlock / 32 ? warp_scan[lane] : 0u;
    uint32_t wscan = total;
#pragma unroll
    for (int delta = 1; delta < kShadeBlock / 32; delta <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, wscan, delta);
      if (lane >= delta) wscan += up;
    }
    if (lane < kShadeBlock / 32) warp_scan[lane] = wscan - total;
    if (lane == kShadeBlock / 32 - 1) {
      const uint32_t n_freed = wscan & 0xffffu, n_shadow = wscan >> 16;
      queue_base[0] = n_freed ? atomicAdd(&args.ctl->n_new, n_freed) : 0u;
      queue_base[1] = n_shadow ? atomicAdd(&args.ctl->n_shadow, n_shadow) : 0u;
      const uint32_t live = n_live;
      if (live) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(live));
      if (live > n_freed) atomicAdd(&args.ctl->n_alive, live - n_freed);
    }
  }
  __syncthreads();
  const uint32_t before = warp_scan[warp] + scan - mine;
  uint32_t freed_at = queue_base[0] + (before & 0xffffu);
  uint32_t shadow_at = queue_base[1] + (before >> 16);
  while (freed_mask) {
    const int r = __ffs(freed_mask) - 1;
    freed_mask &= freed_mask - 1u;
    pool.q_new[freed_at++] = tile_base + sorted[r * 32 + lane];
  }
  while (shadow_mask) {
    const int r = __ffs(shadow_mask) - 1;
    shadow_mask &= shadow_mask - 1u;
    pool.q_shadow[shadow_at++] = tile_base + sorted[r * 32 + lane];
  }
}

// ---------------------------------------------------------------------------------
// connect: any-hit along the shadow rays queued by `shade`; unoccluded ones add their
// contribution to the film.  Queue entry = slot id; the slot holds the ray.
template <class Accel>
__global__ void __launch_bounds__(kBlock, kQueueBlocks) connect_kernel(const RenderArgs args) {
  // grid-stride over the shadow queue (the grid is sized for the device, see raygen_kernel); the loop
  // bound is warp-uniform so that Accel::occluded may vote
  const uint32_t n_shadow = args.ctl->n_shadow;
  for (uint32_t base = blockIdx.x * kBlock; base < n_shadow; base += gridDim.x * kBlock) {
    const uint32_t j = base + threadIdx.x;
    const bool active = j < n_shadow;
    float4 o = make_float4(0.0f, 0.0f, 0.0f, 0.0f), d = make_float4(0.0f, 0.0f, 1.0f, 0.0f);
    float4 c = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (active) {
      const uint32_t slot = args.pool.q_shadow[j];
      o = args.pool.ray[2 * static_cast<size_t>(slot)];
      const float4* in = args.pool.shadow + 2 * static_cast<size_t>(slot);
      d = in[0];
      c = in[1];  // same 32-byte sector as in[0]: requested now, not after the visibility test
    }
    Ray ray;
    ray.o = v3(o);
    ray.d = v3(d);
    if (Accel::occluded(args.scene, ray, d.w, active) || !active) continue;
    film_add(args.film, __float_as_uint(c.w), v3(c));
  }
}

struct ConnectSource {
  const RenderArgs& args;
  uint32_t slot;
  __device__ __forceinline__ uint32_t count() const { return args.ctl->n_shadow; }
  __device__ __forceinline__ bool load(uint32_t j, Ray& ray, float& t_best) {
    slot = args.pool.q_shadow[j];
    const float4 o = args.pool.ray[2 * static_cast<size_t>(slot)];
    const float4 d = args.pool.shadow[2 * static_cast<size_t>(slot)];
    ray.o = v3(o);
    ray.d = v3(d);
    t_best = d.w;
    return !ConstList::occluded(args.scene, ray, d.w, true);  // a sphere in the way: nothing to add
  }
  __device__ __forceinline__ void store(uint32_t, const HitRecord&, bool found) const {
    if (found) return;
    const float4 c = args.pool.shadow[2 * static_cast<size_t>(slot) + 1];
    film_add(args.film, __float_as_uint(c.w), v3(c));
  }
};

template <>
__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) connect_kernel<WideBvh>(const RenderArgs args) {
  ConnectSource source{args, 0u};
  WideBvh::trace_persistent<true>(args.scene, source, &args.ctl->cursor_connect);
}

__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) connect_bvh_counting_kernel(const RenderArgs args) {
  ConnectSource source{args, 0u};
  WideBvh::trace_persistent<true, true>(args.scene, source, &args.ctl->cursor_connect, &args.ctl->bvh_work[2]);
}

// Accumulator::Mean (include/amber/prelude/accumulator.h:88-95): film / samples,
// packed to the reference's Image<RGB> layout (x + y*W, 3 floats).
// The sample count either comes from the host (inv_samples) or, after an external all-reduce of the
// film, from the film's trailing element (count_in_film != 0: three 16-bit digits as floats).
__device__ __forceinline__ float film_inv_samples(const float4* film, uint32_t n_pixels, float inv_samples,
                                                  int count_in_film) {
  if (!count_in_film) return inv_samples;
  const float4 tail = film[n_pixels];
  const double count = static_cast<double>(tail.x) + 65536.0 * static_cast<double>(tail.y) +
                       4294967296.0 * static_cast<double>(tail.z);
  return count > 0.0 ? 1.0f / static_cast<float>(static_cast<unsigned long long>(count)) : 0.0f;
}

__global__ void __launch_bounds__(kBlock) resolve_kernel(const float4* __restrict__ film, uint32_t n_pixels,
                                                         float inv_samples, int count_in_film,
                                                         float* __restrict__ rgb) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n_pixels) return;
  inv_samples = film_inv_samples(film, n_pixels, inv_samples, count_in_film);
  const float4 v = film[i];
  rgb[3 * i + 0] = v.x * inv_samples;
  rgb[3 * i + 1] = v.y * inv_samples;
  rgb[3 * i + 2] = v.z * inv_samples;
}

// Film reduction across the contexts of one process (amber_cuda_film_reduce): dst += sum of the source
// films, which live on OTHER devices and are read through peer memory over NVLink -- the in-process
// counterpart of the NCCL all-reduce, Accumulator::operator+= (include/amber/prelude/accumulator.h:136-166).
constexpr int kMaxReduceSources = 15;
struct FilmSources {
  const float4* film[kMaxReduceSources];
  int n;
};
__global__ void __launch_bounds__(kBlock) film_sum_kernel(float4* __restrict__ dst, const FilmSources sources,
                                                          uint32_t n_pixels) {
  for (uint32_t i = blockIdx.x * kBlock + threadIdx.x; i < n_pixels; i += gridDim.x * kBlock) {
    float4 v = dst[i];
    for (int k = 0; k < sources.n; k++) {
      const float4 a = sources.film[k][i];
      v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
    }
    dst[i] = v;
  }
}

// Resolve fused with the reference's post-process (src/amber/cli/application.cc:98-106):
// film / samples -> Filmic (src/amber/postprocess/filmic.cc:28-67: Hable's curve, exposure 16,
// white point 0.70) -> Gamma 2.2 (src/amber/postprocess/gamma.cc:36-52) -> 8 bit.  The curve is
// evaluated with the host's float operations one by one; only powf can differ from glibc's by
// an ulp, i.e. by one 8-bit level on the rare value that sits on a level boundary.  flip_x
// mirrors columns as ExportPNG does (src/amber/cli/image.cc:59-71, `Width - i - 1`).
__device__ __forceinline__ float filmic_map(float h) {
  const float kA = 0.22f, kB = 0.30f, kC = 0.10f, kD = 0.20f, kE = 0.01f, kF = 0.30f;
  const float num = __fadd_rn(__fmul_rn(h, __fadd_rn(__fmul_rn(h, kA), __fmul_rn(kB, kC))), __fmul_rn(kD, kE));
  const float den = __fadd_rn(__fmul_rn(h, __fadd_rn(__fmul_rn(h, kA), kB)), __fmul_rn(kD, kF));
  return __fsub_rn(__fdiv_rn(num, den), __fdiv_rn(kE, kF));
}

__device__ __forceinline__ uint8_t tonemap_channel(float value, float white) {
  const float mapped = __fdiv_rn(filmic_map(__fmul_rn(value, 16.0f)), white);
  const float p = powf(mapped, __fdiv_rn(1.0f, 2.2f));
  const float clamped = p < 1.0f ? p : 1.0f;  // std::min<float>(1, p): a NaN becomes 1
  return static_cast<uint8_t>(__fmul_rn(255.0f, clamped));
}

__global__ void __launch_bounds__(kBlock) resolve_tonemap_kernel(const float4* __restrict__ film, uint32_t width,
                                                                 uint32_t height, float inv_samples, int count_in_film,
                                                                 int flip_x, uint8_t* __restrict__ rgb8) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= width * height) return;
  inv_samples = film_inv_samples(film, width * height, inv_samples, count_in_film);
  const uint32_t x = i % width, y = i / width;
  const float4 v = film[i];
  const float white = filmic_map(0.70f);
  const uint32_t out = (flip_x ? width - 1 - x : x) + y * width;
  rgb8[3 * out + 0] = tonemap_channel(v.x * inv_samples, white);
  rgb8[3 * out + 1] = tonemap_channel(v.y * inv_samples, white);
  rgb8[3 * out + 2] = tonemap_channel(v.z * inv_samples, white);
}

// ---------------------------------------------------------------------------------
// helpers for the host-buffer entry points
__global__ void pack_rays_kernel(const float* __restrict__ o3, const float* __restrict__ d3, uint64_t n,
                                 float4* __restrict__ o4, float4* __restrict__ d4) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  o4[i] = make_float4(o3[3 * i], o3[3 * i + 1], o3[3 * i + 2], 1.0f);
  d4[i] = make_float4(d3[3 * i], d3[3 * i + 1], d3[3 * i + 2], 0.0f);
}

__global__ void unpack_hits_kernel(const float4* __restrict__ hits, uint64_t n,
                                   const int32_t* __restrict__ internal_to_prim, int32_t* __restrict__ prim,
                                   float* __restrict__ t) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4 h = hits[i];
  const int internal = __float_as_int(h.y);
  prim[i] = internal < 0 ? -1 : internal_to_prim[internal];
  t[i] = h.x;
}

__global__ void sample_light_kernel(const SceneView scene, int material, bool importance, uint64_t n,
                                    const float* __restrict__ normal, const float* __restrict__ dir_out,
                                    const float* __restrict__ xi8, float* __restrict__ dir_in,
                                    float* __restrict__ weight, int32_t* __restrict__ draws) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const MaterialDev m = load_material(scene, material);
  ScriptedRng rng{xi8 + 8 * i, 8, 0};
  const Scatter s = sample_light(m, v3(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]),
                                 v3(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), rng, importance);
  dir_in[3 * i] = s.dir.x; dir_in[3 * i + 1] = s.dir.y; dir_in[3 * i + 2] = s.dir.z;
  weight[3 * i] = s.weight.x; weight[3 * i + 1] = s.weight.y; weight[3 * i + 2] = s.weight.z;
  draws[i] = rng.draws;
}

__global__ void eye_rays_kernel(uint64_t n, const uint32_t* __restrict__ px, const uint32_t* __restrict__ py,
                                const float* __restrict__ xi5, float* __restrict__ origin,
                                float* __restrict__ direction, float* __restrict__ weight) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ScriptedRng rng{xi5 + 5 * i, 5, 0};
  float3 o, d;
  float w;
  generate_eye_ray(px[i], py[i], rng, o, d, w);
  origin[3 * i] = o.x; origin[3 * i + 1] = o.y; origin[3 * i + 2] = o.z;
  direction[3 * i] = d.x; direction[3 * i + 1] = d.y; direction[3 * i + 2] = d.z;
  weight[i] = w;
}

__global__ void light_rays_kernel(const SceneView scene, uint64_t n, const float* __restrict__ xi5,
                                  float* __restrict__ origin, float* __restrict__ direction,
                                  float* __restrict__ weight) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  ScriptedRng rng{xi5 + 5 * i, 5, 0};
  float3 o, d, w;
  generate_light_ray(scene, rng, o, d, w);
  origin[3 * i] = o.x; origin[3 * i + 1] = o.y; origin[3 * i + 2] = o.z;
  direction[3 * i] = d.x; direction[3 * i + 1] = d.y; direction[3 * i + 2] = d.z;
  weight[3 * i] = w.x; weight[3 * i + 1] = w.y; weight[3 * i + 2] = w.z;
}

// Normalize (include/amber/prelude/vector3.h:309-316) on caller-supplied vectors: the exact
// normalisation every sampled direction and basis vector goes through (vec.cuh, normalize_exact).
// With a denominator array: the three IEEE quotients in3[i] / denominator[i] (div3_exact) instead.
__global__ void normalize_kernel(uint64_t n, const float* __restrict__ in3, const float* __restrict__ denominator,
                                 float* __restrict__ out3) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float3 a = v3(in3[3 * i], in3[3 * i + 1], in3[3 * i + 2]);
  const float3 u = denominator ? div3_exact(a, denominator[i]) : normalize_exact(a);
  out3[3 * i] = u.x; out3[3 * i + 1] = u.y; out3[3 * i + 2] = u.z;
}

__global__ void lens_response_kernel(uint64_t n, const float* __restrict__ position,
                                     const float* __restrict__ dir_out, int32_t* __restrict__ pixel,
                                     float* __restrict__ value) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint32_t p = 0;
  float f = 0.0f;
  const bool ok = lens_response(v3(position[3 * i], position[3 * i + 1], position[3 * i + 2]),
                                v3(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), p, f);
  pixel[i] = ok ? static_cast<int32_t>(p) : -1;
  value[i] = ok ? f : 0.0f;
}

}  // namespace amber_cuda

#include "bdpt.cuh"
