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
// Terminated slots are compacted into `q_new`, shadow rays into the shadow queue:
// __ballot_sync/__popc inside the warp, one smem scan and ONE global atomic per block.
//
// Layout in HBM (all float4 / 16-byte accesses):
//   ray[2i], ray[2i+1]     origin xyz + live flag, direction xyz             32 B per slot
//   state[2i], state[2i+1] throughput rgb ; (pixel, sample, depth|flags, -)  32 B per slot
//   hit[i]                 (t, primitive)                                     8 B per slot
//   shadow[3j..3j+2]       origin + t_max, direction + pixel, contribution   48 B per entry
// A slot's ray and state each fill exactly one 32-byte sector, so the scattered writes of
// raygen (slots arrive in queue order) never touch a partial sector.
#pragma once

#include "intersect.cuh"
#include "materials.cuh"

namespace amber_cuda {

constexpr int kBlock = 256;
constexpr uint32_t kFlagCountEmission = 1u << 31;  // emission at the next hit counts
constexpr uint32_t kDepthMask = 0xffffffu;

struct Control {
  unsigned long long next_path;    // next path index of this render call
  unsigned long long total_paths;  // W * H * sample_count
  unsigned int n_new;              // entries in q_new
  unsigned int n_shadow;           // entries in the shadow queue
  unsigned int n_alive;            // survivors of the last shade
  unsigned int done;
  unsigned long long paths, closest_rays, shadow_rays, truncated, phong_capped;
};

struct PathPool {
  float4* ray;      // [2 * n_slots]
  float4* state;    // [2 * n_slots]
  float2* hit;      // [n_slots]
  uint32_t* q_new;  // [n_slots]
  float4* shadow;   // [3 * n_slots]
  uint32_t n_slots;
};

struct RenderArgs {
  PathPool pool;
  SceneView scene;
  Control* ctl;
  float4* film;
  uint64_t seed;
  uint64_t sample_begin;
  uint32_t n_pixels;
  uint32_t max_depth;
  int32_t estimator;
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
// Block-aggregated queue slot allocation: returns this thread's index in the queue
// (valid only if `want`).  Warp ballot + popc, smem scan over the warps, one atomicAdd
// per block.
__device__ __forceinline__ uint32_t queue_alloc(bool want, unsigned int* counter, uint32_t* smem /*[kBlock/32]*/) {
  const unsigned int ballot = __ballot_sync(0xffffffffu, want);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = __popc(ballot);
  __syncthreads();
  if (warp == 0) {
    const uint32_t count = lane < kBlock / 32 ? smem[lane] : 0u;
    uint32_t scan = count;
#pragma unroll
    for (int delta = 1; delta < kBlock / 32; delta <<= 1) {
      const uint32_t up = __shfl_up_sync(0xffffffffu, scan, delta);
      if (lane >= delta) scan += up;
    }
    const uint32_t total = __shfl_sync(0xffffffffu, scan, kBlock / 32 - 1);
    uint32_t base = 0;
    if (lane == 0 && total) base = atomicAdd(counter, total);
    base = __shfl_sync(0xffffffffu, base, 0);
    if (lane < kBlock / 32) smem[lane] = base + scan - count;
  }
  __syncthreads();
  const uint32_t index = smem[warp] + __popc(ballot & ((1u << lane) - 1u));
  __syncthreads();  // smem is reused by the next call
  return index;
}

// ---------------------------------------------------------------------------------
// Camera ray: PixelBound::Uniform (src/amber/rendering/sensor.cc:111-120) +
// BasicThin::GenerateRay (src/amber/scene/lens_thin.cc:70-107) incl. PDFArea/PDFDirection
// (:132-148).  Five uniforms in the compiled reference's order: blade, u, v, film y, film x.
template <class Rng>
__device__ __forceinline__ void generate_eye_ray(uint32_t px, uint32_t py, Rng& rng, float3& origin,
                                                 float3& direction, float& weight) {
  // Origin and direction are computed with the host's operations one by one (no FMA
  // contraction, IEEE sqrt / division): the camera sits 3-5 units from the scene, where
  // one ulp of the direction moves the first hit by more than kEPS, so only bit-identical
  // camera rays give the reference's first-hit behaviour on curved surfaces.
  const LensDev& lens = c_camera.lens;
  const SensorDev& sensor = c_camera.sensor;
  const float nb = static_cast<float>(lens.n_blades);
  int blade = static_cast<int>(floorf(__fmul_rn(rng.next(), nb)));
  blade = min(lens.n_blades - 1, blade);
  const float* b = lens.blades[blade];
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

__global__ void __launch_bounds__(kBlock) raygen_kernel(const RenderArgs args) {
  const uint32_t j = blockIdx.x * kBlock + threadIdx.x;
  const Control* ctl = args.ctl;
  if (j >= ctl->n_new) return;
  const uint32_t slot = args.pool.q_new[j];
  const unsigned long long path = ctl->next_path + j;
  float4* ray = args.pool.ray + 2 * static_cast<size_t>(slot);
  if (path >= ctl->total_paths) {
    ray[0] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    return;
  }
  const uint32_t pixel = static_cast<uint32_t>(path % args.n_pixels);
  const uint32_t sample = static_cast<uint32_t>(args.sample_begin + path / args.n_pixels);
  CameraRng rng;
  rng.init(args.seed, pixel, sample);
  const uint32_t width = c_camera.sensor.width;
  float3 o, d;
  float w;
  generate_eye_ray(pixel % width, pixel / width, rng, o, d, w);
  ray[0] = make_float4(o.x, o.y, o.z, 1.0f);
  ray[1] = make_float4(d.x, d.y, d.z, 0.0f);
  float4* state = args.pool.state + 2 * static_cast<size_t>(slot);
  state[0] = make_float4(w, w, w, 0.0f);
  state[1] = make_float4(__uint_as_float(pixel), __uint_as_float(sample), __uint_as_float(kFlagCountEmission), 0.0f);
}

// Fills q_new with every slot (start of a render call: all slots are free).
__global__ void __launch_bounds__(kBlock) pool_reset_kernel(const RenderArgs args) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= args.pool.n_slots) return;
  args.pool.q_new[i] = i;
  args.pool.ray[2 * static_cast<size_t>(i)] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
}

__global__ void bookkeep_kernel(Control* ctl) {
  unsigned long long generated = 0;
  if (ctl->next_path < ctl->total_paths) {
    const unsigned long long left = ctl->total_paths - ctl->next_path;
    generated = ctl->n_new < left ? ctl->n_new : left;
  }
  ctl->next_path += generated;
  ctl->paths += generated;
  ctl->shadow_rays += ctl->n_shadow;
  ctl->done = (generated == 0 && ctl->n_alive == 0) ? 1u : 0u;
  ctl->n_new = 0;
  ctl->n_shadow = 0;
  ctl->n_alive = 0;
}

// ---------------------------------------------------------------------------------
template <class Accel>
__global__ void __launch_bounds__(kBlock) extend_kernel(const RenderArgs args) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= args.pool.n_slots) return;
  const float4* r = args.pool.ray + 2 * static_cast<size_t>(i);
  const float4 o = r[0];
  if (o.w == 0.0f) return;
  const float4 d = r[1];
  Ray ray;
  ray.o = v3(o);
  ray.d = v3(d);
  const HitRecord hit = Accel::closest(args.scene, ray);
  args.pool.hit[i] = make_float2(hit.t, __int_as_float(hit.prim));
}

// `extend` on caller-supplied rays (amber_cuda_cast / amber_cuda_cast_device).
template <class Accel>
__global__ void __launch_bounds__(kBlock) cast_kernel(const SceneView scene, const float4* __restrict__ ray_o,
                                                      const float4* __restrict__ ray_d, uint64_t n,
                                                      float4* __restrict__ hits) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * kBlock + threadIdx.x;
  if (i >= n) return;
  Ray ray;
  ray.o = v3(ray_o[i]);
  ray.d = v3(ray_d[i]);
  const HitRecord hit = Accel::closest(scene, ray);
  hits[i] = make_float4(hit.t, __int_as_float(hit.prim), 0.0f, 0.0f);
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
    const float u = __fdiv_rn(dot_exact(P, T), det);
    const float3 Q = cross_exact(T, E1);
    const float v = __fdiv_rn(dot_exact(Q, ray.d), det);
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
  float3 o, d;
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
  shadow.o = x;
  shadow.d = wi;
  shadow.t_max = dist2 * inv_dist * (1.0f - 1e-4f);
  return true;
}

__device__ __forceinline__ void film_add(float4* film, uint32_t pixel, float3 value) {
  atomicAdd(&film[pixel], make_float4(value.x, value.y, value.z, 0.0f));
}

__global__ void __launch_bounds__(kBlock) shade_kernel(const RenderArgs args) {
  __shared__ uint32_t smem[kBlock / 32];
  __shared__ unsigned int block_live, block_alive;
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  const PathPool& pool = args.pool;
  if (threadIdx.x == 0) {
    block_live = 0;
    block_alive = 0;
  }
  bool live = false, terminated = false, has_shadow = false;
  ShadowRay shadow;
  uint32_t pixel = 0;
  if (i < pool.n_slots) {
    float4* ray_slot = pool.ray + 2 * static_cast<size_t>(i);
    const float4 o4 = ray_slot[0];
    live = o4.w != 0.0f;
    if (live) {
      const float2 h2 = pool.hit[i];
      HitRecord hit;
      hit.t = h2.x;
      hit.prim = __float_as_int(h2.y);
      if (hit.prim < 0) {
        terminated = true;  // algorithm_pt.cc:140-142
      } else {
        const float4 d4 = ray_slot[1];
        float4* state_slot = pool.state + 2 * static_cast<size_t>(i);
        const float4 t4 = state_slot[0];
        const float4 m4 = state_slot[1];
        Ray ray;
        ray.o = v3(o4);
        ray.d = v3(d4);
        pixel = __float_as_uint(m4.x);
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

        // emission, algorithm_pt.cc:144
        if (m.kind == 4) {
          const bool sampled_by_nee = args.estimator == 1 && hit.prim < args.scene.n_tris;
          if (!sampled_by_nee || (flags & kFlagCountEmission)) {
            const float3 le = emitted_radiance(m, n, wo);
            if (le.x > 0.0f || le.y > 0.0f || le.z > 0.0f) film_add(args.film, pixel, throughput * le);
          }
        }
        const bool connect = args.estimator == 1 && m.nee && args.scene.n_lights > 0;
        if (connect) {
          float lu, lv;
          rng.light_point(lu, lv);
          has_shadow = sample_direct(args.scene, m, x, n, wo, throughput, rng.light_choice(), lu, lv, shadow);
        }

        // BSDF sample + Russian roulette, algorithm_pt.cc:145-156
        const Scatter scatter = sample_light(m, n, wo, rng);
        const float p = fminf(kRussianRoulette, max3(scatter.weight));
        if (scatter.capped) atomicAdd(&args.ctl->phong_capped, 1ull);
        if (!(rng.rr() < p)) {
          terminated = true;
        } else if (depth >= args.max_depth) {
          terminated = true;
          atomicAdd(&args.ctl->truncated, 1ull);
        } else {
          const float inv_p = 1.0f / p;
          throughput = throughput * (inv_p * scatter.weight);
          ray_slot[0] = make_float4(x.x, x.y, x.z, 1.0f);
          ray_slot[1] = make_float4(scatter.dir.x, scatter.dir.y, scatter.dir.z, 0.0f);
          state_slot[0] = make_float4(throughput.x, throughput.y, throughput.z, 0.0f);
          state_slot[1] = make_float4(m4.x, m4.y, __uint_as_float(depth | (connect ? 0u : kFlagCountEmission)), 0.0f);
        }
      }
    }
  }
  // queues: freed slots -> q_new, shadow rays -> shadow queue; counters per block
  const uint32_t new_index = queue_alloc(terminated, &args.ctl->n_new, smem);
  if (terminated) pool.q_new[new_index] = i;
  const uint32_t shadow_index = queue_alloc(has_shadow, &args.ctl->n_shadow, smem);
  if (has_shadow) {
    float4* out = pool.shadow + 3 * static_cast<size_t>(shadow_index);
    out[0] = make_float4(shadow.o.x, shadow.o.y, shadow.o.z, shadow.t_max);
    out[1] = make_float4(shadow.d.x, shadow.d.y, shadow.d.z, __uint_as_float(pixel));
    out[2] = make_float4(shadow.contribution.x, shadow.contribution.y, shadow.contribution.z, 0.0f);
  }
  // statistics: rays extended this step, survivors (queue_alloc's barriers order the init above)
  const unsigned int live_ballot = __ballot_sync(0xffffffffu, live);
  const unsigned int alive_ballot = __ballot_sync(0xffffffffu, live && !terminated);
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&block_live, __popc(live_ballot));
    atomicAdd(&block_alive, __popc(alive_ballot));
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    if (block_live) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(block_live));
    if (block_alive) atomicAdd(&args.ctl->n_alive, block_alive);
  }
}

// ---------------------------------------------------------------------------------
template <class Accel>
__global__ void __launch_bounds__(kBlock) connect_kernel(const RenderArgs args) {
  const uint32_t j = blockIdx.x * kBlock + threadIdx.x;
  if (j >= args.ctl->n_shadow) return;
  const float4* in = args.pool.shadow + 3 * static_cast<size_t>(j);
  const float4 o = in[0];
  const float4 d = in[1];
  Ray ray;
  ray.o = v3(o);
  ray.d = v3(d);
  if (Accel::occluded(args.scene, ray, o.w)) return;
  const float4 c = in[2];
  film_add(args.film, __float_as_uint(d.w), v3(c));
}

// Accumulator::Mean (include/amber/prelude/accumulator.h:88-95): film / samples,
// packed to the reference's Image<RGB> layout (x + y*W, 3 floats).
__global__ void __launch_bounds__(kBlock) resolve_kernel(const float4* __restrict__ film, uint32_t n_pixels,
                                                         float inv_samples, float* __restrict__ rgb) {
  const uint32_t i = blockIdx.x * kBlock + threadIdx.x;
  if (i >= n_pixels) return;
  const float4 v = film[i];
  rgb[3 * i + 0] = v.x * inv_samples;
  rgb[3 * i + 1] = v.y * inv_samples;
  rgb[3 * i + 2] = v.z * inv_samples;
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

__global__ void sample_light_kernel(const SceneView scene, int material, uint64_t n,
                                    const float* __restrict__ normal, const float* __restrict__ dir_out,
                                    const float* __restrict__ xi8, float* __restrict__ dir_in,
                                    float* __restrict__ weight, int32_t* __restrict__ draws) {
  const uint64_t i = static_cast<uint64_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const MaterialDev m = load_material(scene, material);
  ScriptedRng rng{xi8 + 8 * i, 8, 0};
  const Scatter s = sample_light(m, v3(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]),
                                 v3(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), rng);
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

}  // namespace amber_cuda
