// bdpt.cuh -- bidirectional path tracing (row N3): include/amber/rendering/bidirectional_path_sampling.h
// and src/amber/rendering/algorithm_bdpt.cc:116-148 as ONE kernel, one thread per (pixel, sample).
//
// Why not the wavefront: a BDPT sample is two short subpaths (2.8 events each on the Cornell box)
// followed by (s + 1)(t + 1) connections whose MIS weights need every event of both subpaths; the
// state is two event arrays, not one ray.  A thread therefore keeps both subpaths in local memory
// (L1-resident, <= 2 x 16 events x 64 B), traces with the same intersection code as `extend`
// (constant-memory list, or a per-thread walk of the 4-wide BVH), and adds its contributions to the
// film with float4 atomics: the eye image at its own pixel, the light image where the lens maps it.
//
// Random numbers: each subpath draws from one sequential Philox stream, counter
// (pixel, sample, kSubpath | {0 light, 1 eye}, block) -- the reference's generators interleave
// BSDF and Russian-roulette draws in one call order, which a thread simply replays
// (oracle/philox_stream.h has the same convention).
//
// Discrete decisions (hit points, sampled directions, Russian roulette, the refraction branch, the
// response pixel) use the host's arithmetic operation by operation, as everywhere on the path;
// connection terms and MIS weights only scale a contribution and use fast math.
#pragma once

namespace amber_cuda {

constexpr int kBdptMaxEvents = 16;  // events per subpath (the reference has no cap: truncations are rare)
constexpr int32_t kObjectLens = -2;  // the lens's own aperture object: never equal to an object found by a cast
constexpr int32_t kEstimatorBdpt = 3;

struct SequentialRng {
  uint32_t key0, key1, pixel, sample, id, i;
  uint4 block;
  __device__ __forceinline__ void init(uint64_t seed, uint32_t pixel_index, uint32_t sample_index, uint32_t stream_id) {
    key0 = static_cast<uint32_t>(seed);
    key1 = static_cast<uint32_t>(seed >> 32);
    pixel = pixel_index;
    sample = sample_index;
    id = stream_id;
    i = 0;
  }
  __device__ __forceinline__ float next() {
    const uint32_t k = i++;
    if ((k & 3u) == 0) block = philox4x32_10(key0, key1, pixel, sample, id, k >> 2);
    const uint32_t w = k & 3u;
    return philox_uniform(w == 0 ? block.x : w == 1 ? block.y : w == 2 ? block.z : block.w);
  }
};

// SubpathEvent, bidirectional_path_sampling.h:46-90,281-315
struct BdptEvent {
  int32_t object;    // internal primitive index; kObjectLens on the eye's leading event
  int32_t material;  // material index; -1 = Eye (the eye's leading event)
  float3 position, normal, direction_out, weight;
  float geometry_factor, p_backward, p_forward;  // NaN on a leading event
  __device__ __forceinline__ bool leading() const { return isnan(geometry_factor); }
};

enum : int { kSurfaceDiffuse = 0, kSurfaceSpecular = 1, kSurfaceLight = 2, kSurfaceEye = 3 };
__device__ __forceinline__ int surface_of_kind(int kind) {
  return kind <= 1 ? kSurfaceDiffuse : kind == 4 ? kSurfaceLight : kind == 5 ? kSurfaceEye : kSurfaceSpecular;
}
__device__ __forceinline__ int surface_of(const SceneView& scene, int material) {
  if (material < 0) return kSurfaceEye;
  return surface_of_kind(__float_as_int(__ldg(reinterpret_cast<const float4*>(scene.materials + material)).x));
}

// Closest hit of one lane's ray, callable from divergent code.
template <class Accel>
__device__ __forceinline__ HitRecord closest_lane(const SceneView& scene, const Ray& ray);
template <>
__device__ __forceinline__ HitRecord closest_lane<ConstList>(const SceneView& scene, const Ray& ray) {
  return ConstList::closest(scene, ray, true);
}
template <>
__device__ __forceinline__ HitRecord closest_lane<WideBvh>(const SceneView& scene, const Ray& ray) {
  SphereBound spheres;
  HitRecord hit;
  hit.t = spheres.bound(scene, ray);
  hit.prim = -1;
  bool found = false;
  int32_t stack[kBvhStackMax];
  int sp = 0;
  const float3 idir = v3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
  const float3 oidir = v3(ray.o.x * idir.x, ray.o.y * idir.y, ray.o.z * idir.z);
  int32_t current = scene.bvh_root;
  while (current != kEmptyChild) {
    if (current >= 0) {
      current = WideBvh::node_step(scene, current, idir, oidir, hit.t, stack, 1, sp);
      continue;
    }
    const int first = bvh_leaf_first(current), count = bvh_leaf_count(current);
    for (int k = 0; k < count; k++) {
      const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + first + k);
      if (intersect_triangle(__ldg(tri), __ldg(tri + 1), __ldg(tri + 2), ray, hit.t, hit.t)) {
        hit.prim = first + k;
        found = true;
      }
    }
    current = sp > 0 ? stack[--sp] : kEmptyChild;
  }
  return spheres.resolve(scene, hit, found);
}

// Refraction, src/amber/scene/material_refraction.cc:41-175: the Fresnel term of the branch
// (reflection iff wo and wi lie on one side) and its sampling probability, for radiance
// (BSDF / PDFLight: ior^2 scaling) or importance (AdjointBSDF / PDFImportance).
__device__ __forceinline__ void refraction_terms(const MaterialDev& m, float3 n, float3 wo, float3 wi, bool importance,
                                                 float& term, float& pdf) {
  const float cos_o = dot(wo, n);
  const float ior = cos_o > 0.0f ? 1.0f / m.param : m.param;
  const float ior2 = ior * ior;
  if (1.0f - (1.0f - cos_o * cos_o) * ior2 < 0.0f) {  // total internal reflection: kDiracDelta
    term = 1.0f;
    pdf = 1.0f;
    return;
  }
  const float rho_r = schlick(m.r0, fabsf(cos_o));
  const float rho_t = importance ? 1.0f - rho_r : (1.0f - rho_r) * ior2;
  const float total = importance ? 1.0f : rho_r + rho_t;
  const bool reflect = cos_o * dot(wi, n) > 0.0f;
  term = reflect ? rho_r : rho_t;
  pdf = ((reflect ? rho_r : rho_t) / total + 0.5f) * 0.5f;
}

// Material::BSDF (adjoint = false) / AdjointBSDF (true) for every material kind:
// material_{lambertian,phong,specular}.cc (symmetric), material_refraction.cc,
// material_diffuse_light.h:141-161 (0), material_eye.h:97-117 (kDiracDelta).
__device__ __forceinline__ float3 event_bsdf(const SceneView& scene, int material, float3 n, float3 wo, float3 wi,
                                             bool adjoint) {
  if (material < 0) return v3(1.0f, 1.0f, 1.0f);
  const MaterialDev m = load_material(scene, material);
  const float3 rho = v3(m.r, m.g, m.b);
  switch (m.kind) {
    case 0:
    case 1: return bsdf_eval(m, n, wo, wi);
    case 2: return dot(wo, n) * dot(wi, n) <= 0.0f ? v3(0.0f, 0.0f, 0.0f) : rho;
    case 3: {
      float term, pdf;
      refraction_terms(m, n, wo, wi, adjoint, term, pdf);
      return term * rho;
    }
    case 5: return v3(1.0f, 1.0f, 1.0f);
    default: return v3(0.0f, 0.0f, 0.0f);
  }
}

// Material::PDFLight (importance = false) / PDFImportance (true)
__device__ __forceinline__ float event_pdf(const SceneView& scene, int material, float3 n, float3 wo, float3 wi,
                                           bool importance) {
  if (material < 0) return 1.0f;
  const MaterialDev m = load_material(scene, material);
  switch (m.kind) {
    case 0: return 1.0f / kPiF;  // material_lambertian.cc:51-59
    case 1: {                    // material_phong.cc:64-79
      const float3 refl = perfect_reflection(wo, n, dot(wo, n));
      const float cos_alpha = fmaxf(0.0f, dot(wi, refl));
      return (m.param + 1.0f) / (2.0f * kPiF) * powf(cos_alpha, m.param) / fabsf(dot(wi, n));
    }
    case 3: {
      float term, pdf;
      refraction_terms(m, n, wo, wi, importance, term, pdf);
      return pdf;
    }
    default: return 1.0f;  // Specular, DiffuseLight, Eye: kDiracDelta
  }
}

__device__ __forceinline__ float min_rr(float x) { return x < kRussianRoulette ? x : kRussianRoulette; }  // std::min(kRR, x)

// LightSet::PDFArea, include/amber/scene/light_set.h:106-111: Sum(Irradiance) / total power
__device__ __forceinline__ float light_pdf_area(const SceneView& scene, int material) {
  const MaterialDev m = load_material(scene, material);
  return (m.r * kPiF + m.g * kPiF + m.b * kPiF) / scene.total_power;
}

// Lens::PDFDirection: BasicThin src/amber/scene/lens_thin.cc:138-148, BasicPinhole lens_pinhole.cc:93-108
__device__ __forceinline__ float eye_pdf_direction(float3 direction) {
  const LensDev& lens = c_camera.lens;
  const SensorDev& sensor = c_camera.sensor;
  const float3 d = mat3_apply_exact(lens.local, direction);
  if (lens.kind == 1) {
    const float3 point = (lens.sensor_distance / d.z) * d;
    return sensor.inv_area / (d.z * d.z / dot(point, point));
  }
  const float dz2 = d.z * d.z;
  return sensor.size_over_area * (lens.sensor_distance * lens.sensor_distance) / (dz2 * dz2);
}

// prelude::GeometryFactor, include/amber/prelude/geometry.h:49-75
__device__ __forceinline__ float geometry_factor(float3 px, float3 py, float3 nx, float3 ny) {
  const float3 xy = py - px;
  const float d2 = dot(xy, xy);
  const float3 dir = rsqrtf(d2) * xy;
  return fabsf(dot(dir, nx) * dot(dir, ny) / d2);
}

struct BdptCounters {
  uint32_t casts, splats;
};

// GenerateLightSubpath (kImportance = true, :726-801) / GenerateEyeSubpath (false, :821-899)
template <class Accel, bool kImportance>
__device__ __forceinline__ int generate_subpath(const RenderArgs& args, uint32_t pixel, SequentialRng& rng, int cap,
                                                BdptEvent* out, BdptCounters& count) {
  const SceneView& scene = args.scene;
  Ray ray;
  float3 weight, normal;
  if (kImportance) {
    int light_index;
    generate_light_ray(scene, rng, ray.o, ray.d, weight, &light_index);
    const float4* L = reinterpret_cast<const float4*>(scene.lights + light_index);
    normal = v3(__ldg(L + 3));
    out[0].object = -3 - light_index;
    out[0].material = __float_as_int(__ldg(L + 4).w);
  } else {
    const uint32_t width = c_camera.sensor.width;
    float w;
    generate_eye_ray(pixel % width, pixel / width, rng, ray.o, ray.d, w, &normal);
    weight = v3(w, w, w);
    out[0].object = kObjectLens;
    out[0].material = -1;
  }
  out[0].position = ray.o;
  out[0].normal = normal;
  out[0].direction_out = normal;  // SubpathEvent(object, leading), :281-294
  out[0].weight = weight;
  out[0].geometry_factor = out[0].p_backward = out[0].p_forward = CUDART_NAN_F;
  int n = 1;
  for (int i = 1; i < cap; i++) {
    const HitRecord hit = closest_lane<Accel>(scene, ray);
    count.casts++;
    if (hit.prim < 0) break;
    float3 x, hn;
    int material_index;
    surface_point(scene, ray, hit, x, hn, material_index);
    const MaterialDev m = load_material(scene, material_index);
    const float3 wo = -ray.d;
    const Scatter scatter = sample_light(m, hn, wo, rng, kImportance);
    const float3 wi = scatter.dir;
    // forward = the transport being traced, backward = the other one
    const float p_forward = event_pdf(scene, material_index, hn, wo, wi, kImportance);
    const float p_rr_forward = min_rr(max3(scatter.weight));
    const float p_backward = event_pdf(scene, material_index, hn, wi, wo, !kImportance);
    const float3 bsdf_backward = event_bsdf(scene, material_index, hn, wi, wo, !kImportance);
    const float p_rr_backward = min_rr(max3((1.0f / p_backward) * bsdf_backward));
    BdptEvent& e = out[n++];
    e.object = hit.prim;
    e.material = material_index;
    e.position = x;
    e.normal = hn;
    e.direction_out = wo;
    e.weight = weight;
    e.geometry_factor = fabsf(dot(wo, normal) * dot(wo, hn) / (hit.t * hit.t));
    e.p_backward = p_backward * p_rr_backward;
    e.p_forward = p_forward * p_rr_forward;
    if (!(rng.next() < p_rr_forward)) break;
    ray.o = x;
    ray.d = wi;
    normal = hn;
    weight = weight * ((1.0f / p_rr_forward) * scatter.weight);
  }
  return n;
}

struct BdptContribution {
  float3 value;
  bool has_pixel;
  uint32_t pixel;
};

// Connect, :901-1017
template <class Accel>
__device__ __forceinline__ BdptContribution bdpt_connect(const SceneView& scene, const BdptEvent* le, const BdptEvent* ee,
                                                         BdptCounters& count) {
  BdptContribution c;
  c.value = v3(0.0f, 0.0f, 0.0f);
  c.has_pixel = false;
  c.pixel = 0;
  if (!le && !ee) return c;
  if (!le) {  // s = 0: the eye subpath found a light
    if (surface_of(scene, ee->material) != kSurfaceLight) return c;
    c.value = ee->weight * emitted_radiance(load_material(scene, ee->material), ee->normal, ee->direction_out);
    return c;
  }
  if (!ee) {  // t = 0: the light subpath found the aperture
    float response;
    if (surface_of(scene, le->material) != kSurfaceEye) return c;
    if (!lens_response(le->position, le->direction_out, c.pixel, response)) return c;
    c.has_pixel = true;
    c.value = response * le->weight;
    return c;
  }
  if (!le->leading() && surface_of(scene, le->material) != kSurfaceDiffuse) return c;
  if (!ee->leading() && surface_of(scene, ee->material) != kSurfaceDiffuse) return c;
  Ray shadow;
  shadow.o = le->position;
  shadow.d = normalize_exact(v3(ee->position.x - le->position.x, ee->position.y - le->position.y,
                                ee->position.z - le->position.z));
  const float3 to_light = -shadow.d;
  float3 value;
  if (le->leading()) {  // s = 1: Radiance / LightPDFArea
    const float4 le_over_pdf = __ldg(reinterpret_cast<const float4*>(scene.lights + (-3 - le->object)) + 4);
    value = dot(shadow.d, le->normal) > 0.0f ? v3(le_over_pdf) : v3(0.0f, 0.0f, 0.0f);
  } else {
    value = le->weight * event_bsdf(scene, le->material, le->normal, le->direction_out, shadow.d, true);
  }
  if (ee->leading()) {  // t = 1
    float response;
    if (!lens_response(ee->position, to_light, c.pixel, response)) return c;
    c.has_pixel = true;
    value = (response / c_camera.lens.p_area) * value;
  } else {
    value = value * (ee->weight * event_bsdf(scene, ee->material, ee->normal, ee->direction_out, to_light, false));
  }
  if (!(max3(value) > 0.0f)) return c;
  // "XXX rough check" (:1006-1010): the first object along the shadow ray must be the eye
  // endpoint's object.  The eye's leading event carries the lens's own aperture object, which
  // no cast ever returns, so t = 1 connections contribute nothing in the reference -- nor here.
  const HitRecord hit = closest_lane<Accel>(scene, shadow);
  count.casts++;
  if (hit.prim != ee->object) return c;
  c.value = geometry_factor(le->position, ee->position, le->normal, ee->normal) * value;
  return c;
}

// BidirectionalPathSamplingBuffer::Buffer + MISWeight(BalanceHeuristic), :423-697 and
// include/amber/rendering/multiple_importance_sampling.h:58-69
__device__ __forceinline__ float bdpt_mis_weight(const SceneView& scene, const BdptEvent* L, int s, const BdptEvent* E,
                                                 int t) {
  float gf[2 * kBdptMaxEvents], p_light[2 * kBdptMaxEvents], p_importance[2 * kBdptMaxEvents];
  float p_technique[2 * kBdptMaxEvents + 1];
  int n_gf = 0, n_pl = 0, n_pi = 0;
  const int n = s + t + 1;
  const bool joined = s > 0 && t > 0;
  for (int i = 1; i < s; i++) gf[n_gf++] = L[i].geometry_factor;
  if (joined) gf[n_gf++] = geometry_factor(L[s - 1].position, E[t - 1].position, L[s - 1].normal, E[t - 1].normal);
  for (int i = t - 1; i >= 1; i--) gf[n_gf++] = E[i].geometry_factor;

  for (int i = 1; i < s; i++) p_light[n_pl++] = L[i].p_backward;
  if (joined) {
    const BdptEvent& le = L[s - 1];
    const BdptEvent& ee = E[t - 1];
    const float3 din = normalize(le.position - ee.position);
    const float3 bsdf = event_bsdf(scene, ee.material, ee.normal, ee.direction_out, din, false);
    const float pdf = event_pdf(scene, ee.material, ee.normal, ee.direction_out, din, false);
    p_light[n_pl++] = pdf * min_rr(max3((1.0f / pdf) * bsdf));
  }
  for (int i = t - 2; i >= 0; i--) p_light[n_pl++] = E[i].p_forward;
  {
    const BdptEvent* z0 = t == 0 ? &L[s - 1] : &E[0];
    const BdptEvent* z1 = t == 0 ? &L[s - 2] : t == 1 ? &L[s - 1] : &E[1];
    p_light[n_pl - 1] = eye_pdf_direction(normalize(z1->position - z0->position));
  }

  for (int i = 0; i + 1 < s; i++) p_importance[n_pi++] = L[i].p_forward;
  if (joined) {
    const BdptEvent& le = L[s - 1];
    const BdptEvent& ee = E[t - 1];
    const float3 din = normalize(ee.position - le.position);
    const float3 bsdf = event_bsdf(scene, le.material, le.normal, le.direction_out, din, true);
    const float pdf = event_pdf(scene, le.material, le.normal, le.direction_out, din, true);
    p_importance[n_pi++] = pdf * min_rr(max3((1.0f / pdf) * bsdf));
  }
  for (int i = t - 1; i >= 1; i--) p_importance[n_pi++] = E[i].p_backward;
  p_importance[0] = 1.0f / kPiF;  // LightSet::PDFDirection, light_set.h:113-122

  const float eye_pdf_area = c_camera.lens.p_area;
  for (int i = 0; i < n; i++) p_technique[i] = 1.0f;
  for (int i = s - 1; i >= 0; i--) {
    float p = p_technique[i + 1];
    if (t == 0 && i == n - 2) p *= eye_pdf_area; else p *= p_light[i] * gf[i];
    if (s > 0 && i == 0) p /= light_pdf_area(scene, L[0].material); else p /= p_importance[i - 1] * gf[i - 1];
    p_technique[i] = p;
  }
  for (int i = s + 1; i < n; i++) {
    float p = p_technique[i - 1];
    if (s == 0 && i == 1) p *= light_pdf_area(scene, E[t - 1].material); else p *= p_importance[i - 2] * gf[i - 2];
    if (t > 0 && i == n - 1) p /= eye_pdf_area; else p /= p_light[i - 1] * gf[i - 1];
    p_technique[i] = p;
  }
  // DisableSpecularConnection, :661-697
  for (int i = 1; i + 1 < s; i++) {
    if (surface_of(scene, L[i].material) != kSurfaceDiffuse) p_technique[i] = p_technique[i + 1] = 0.0f;
  }
  for (int i = 1; i + 1 < t; i++) {
    if (surface_of(scene, E[i].material) != kSurfaceDiffuse) p_technique[n - i - 1] = p_technique[n - i - 2] = 0.0f;
  }
  float sum = 0.0f;
  for (int i = 0; i < n; i++) sum += p_technique[i];
  return 1.0f / sum;
}

// Thread::operator() / Thread::Render / Combine: one (pixel, sample) per thread, grid-stride.
template <class Accel>
__global__ void __launch_bounds__(128) bdpt_kernel(const RenderArgs args, unsigned long long total_pairs, int cap) {
  BdptEvent L[kBdptMaxEvents], E[kBdptMaxEvents];
  BdptCounters count;
  count.casts = 0;
  count.splats = 0;
  uint32_t pairs = 0;
  const unsigned long long stride = static_cast<unsigned long long>(gridDim.x) * blockDim.x;
  for (unsigned long long pair = static_cast<unsigned long long>(blockIdx.x) * blockDim.x + threadIdx.x; pair < total_pairs;
       pair += stride) {
    const uint32_t pixel = static_cast<uint32_t>(pair % args.n_pixels);
    const uint32_t sample = static_cast<uint32_t>(args.sample_begin + pair / args.n_pixels);
    SequentialRng rng;
    rng.init(args.seed, pixel, sample, 0x40000000u);
    const int s_max = generate_subpath<Accel, true>(args, pixel, rng, cap, L, count);
    rng.init(args.seed, pixel, sample, 0x40000001u);
    const int t_max = generate_subpath<Accel, false>(args, pixel, rng, cap, E, count);
    float3 measurement = v3(0.0f, 0.0f, 0.0f);
    for (int s = 0; s <= s_max; s++) {
      for (int t = 0; t <= t_max; t++) {
        BdptContribution c = bdpt_connect<Accel>(args.scene, s ? &L[s - 1] : nullptr, t ? &E[t - 1] : nullptr, count);
        if (!(max3(c.value) > 0.0f)) continue;
        const float weight = bdpt_mis_weight(args.scene, L, s, E, t);
        if (c.has_pixel) {  // light image: / sensor.Size(), :1054-1056
          film_add(args.film, c.pixel, (weight * args.inv_n_pixels) * c.value);
          count.splats++;
        } else {
          measurement = measurement + weight * c.value;
        }
      }
    }
    if (max3(measurement) > 0.0f) film_add(args.film, pixel, measurement);
    pairs++;
  }
  // counters: one atomic per warp
  const uint32_t casts = __reduce_add_sync(0xffffffffu, count.casts);
  const uint32_t splats = __reduce_add_sync(0xffffffffu, count.splats);
  const uint32_t n_pairs = __reduce_add_sync(0xffffffffu, pairs);
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
    atomicAdd(&args.ctl->splats, static_cast<unsigned long long>(splats));
    atomicAdd(&args.ctl->paths, static_cast<unsigned long long>(n_pairs));
  }
}

}  // namespace amber_cuda
