// bdpt.cuh -- bidirectional path tracing (row N3): include/amber/rendering/bidirectional_path_sampling.h
// and src/amber/rendering/algorithm_bdpt.cc:116-148 as ONE kernel of persistent warps whose lanes each
// carry one (pixel, sample) through a small state machine (see bdpt_kernel).
//
// Why not the wavefront: a BDPT sample is two short subpaths (2.8 events each on the Cornell box)
// followed by (s + 1)(t + 1) connections whose MIS weights need every event of both subpaths; the
// state is two event arrays, not one ray.  A thread therefore keeps both subpaths in local memory
// (<= 2 x 16 events x 80 B), traces with the same intersection code as `extend`
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
  // out of line: next() has some 16 call sites in the subpath generator, a Philox block is ~90 instructions
  // by value in and out: a member function called out of line would pin the whole generator in local memory
  static __device__ __noinline__ uint4 philox_block(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2,
                                                    uint32_t c3) {
    return philox4x32_10(k0, k1, c0, c1, c2, c3);
  }
  __device__ __forceinline__ void refill(uint32_t index) { block = philox_block(key0, key1, pixel, sample, id, index); }
  __device__ __forceinline__ float next() {
    const uint32_t k = i++;
    if ((k & 3u) == 0) refill(k >> 2);
    const uint32_t w = k & 3u;
    return philox_uniform(w == 0 ? block.x : w == 1 ? block.y : w == 2 ? block.z : block.w);
  }
};

// SubpathEvent, bidirectional_path_sampling.h:46-90,281-315, as five float4 so that local-memory
// accesses are 128-bit and the MIS weight (which walks every event of both subpaths for every
// connection) reads ONE of them per event: `mis` = geometry factor, backward pdf, forward pdf
// (NaN on a leading event) and material | surface type << 28.
struct BdptEvent {
  float4 mis;      // geometry_factor, p_backward, p_forward, bits
  float4 pos_obj;  // position, object (internal primitive index; kObjectLens / -3 - light on a leading event)
  float4 normal4;
  float4 dout4;    // direction_out
  float4 weight4;
  __device__ __forceinline__ bool leading() const { return isnan(mis.x); }
  __device__ __forceinline__ int32_t object() const { return __float_as_int(pos_obj.w); }
  __device__ __forceinline__ int32_t material() const {  // -1 = Eye (the eye's leading event)
    const int32_t bits = __float_as_int(mis.w);
    return bits < 0 ? -1 : (bits & 0x0fffffff);
  }
  __device__ __forceinline__ int surface() const { return (__float_as_int(mis.w) >> 28) & 3; }
  __device__ __forceinline__ float3 position() const { return v3(pos_obj); }
  __device__ __forceinline__ float3 normal() const { return v3(normal4); }
  __device__ __forceinline__ float3 direction_out() const { return v3(dout4); }
  __device__ __forceinline__ float3 weight() const { return v3(weight4); }
};
// bits of a leading eye event: sign bit set (material -1), surface Eye
__device__ __forceinline__ float event_bits(int material, int surface) {
  return __int_as_float(material < 0 ? static_cast<int32_t>(0x80000000u | (3u << 28)) : (material | (surface << 28)));
}

enum : int { kSurfaceDiffuse = 0, kSurfaceSpecular = 1, kSurfaceLight = 2, kSurfaceEye = 3 };
__device__ __forceinline__ int surface_of_kind(int kind) {
  return kind <= 1 ? kSurfaceDiffuse : kind == 4 ? kSurfaceLight : kind == 5 ? kSurfaceEye : kSurfaceSpecular;
}
__device__ __forceinline__ int surface_of(const SceneView& scene, int material) {
  if (material < 0) return kSurfaceEye;
  return surface_of_kind(__float_as_int(__ldg(reinterpret_cast<const float4*>(scene.materials + material)).x));
}

// Closest hit of one lane's ray, callable from divergent code.
//
// Code size: the first version of this kernel inlined everything and came to 176 KB of SASS -- far beyond
// the instruction caches, and ncu showed 86 % of its stall samples as "no instruction" with 9.5 % of the
// issue slots used.  Every large piece that has more than one call site (the cast, subpath generation, the
// BSDF / pdf evaluators with their powf, the connection, the MIS weight) is therefore ONE out-of-line
// function (__noinline__): a call costs a few moves, an instruction-cache miss costs hundreds of cycles.
template <class Accel>
__device__ __forceinline__ HitRecord closest_lane_inline(const SceneView& scene, const Ray& ray);
template <>
__device__ __forceinline__ HitRecord closest_lane_inline<ConstList>(const SceneView& scene, const Ray& ray) {
  return ConstList::closest(scene, ray, true);
}
#if AMBER_BVH_WIDTH == 8
constexpr int kLaneStackMax = 64;  // node groups: one entry per level (configure_bvh_launch refuses deeper trees)
template <>
__device__ __forceinline__ HitRecord closest_lane_inline<WideBvh>(const SceneView& scene, const Ray& ray) {
  SphereBound spheres;
  HitRecord hit;
  hit.t = spheres.bound(scene, ray);
  hit.prim = -1;
  uint2 stack[kLaneStackMax];
  const bool found = WideBvh::walk_lane<false>(scene, ray, hit, stack);
  return spheres.resolve(scene, hit, found);
}
#else
template <>
__device__ __forceinline__ HitRecord closest_lane_inline<WideBvh>(const SceneView& scene, const Ray& ray) {
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

#endif

// kCopy: each kernel that casts gets its own copy of the out-of-line walk (ptxas 12.9 crashes on a module in which two
// of the wavefront kernels share the BVH instantiation).
template <class Accel, int kCopy = 0>
__device__ __noinline__ HitRecord closest_lane(const SceneView& scene, const Ray& ray) {
  return closest_lane_inline<Accel>(scene, ray);
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
__device__ __noinline__ float3 event_bsdf(const MaterialDev m, float3 n, float3 wo, float3 wi, bool adjoint) {
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
__device__ __noinline__ float event_pdf(const MaterialDev m, float3 n, float3 wo, float3 wi, bool importance) {
  switch (m.kind) {
    case 0: return 1.0f / kPiF;  // material_lambertian.cc:51-59
    case 1: {                    // material_phong.cc:64-79
      const float3 refl = perfect_reflection(wo, n, dot(wo, n));
      const float cos_alpha = fmaxf(0.0f, dot(wi, refl));
      return (m.param + 1.0f) / (2.0f * kPiF) * pow_scale_only(cos_alpha, m.param) / fabsf(dot(wi, n));
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

// The leading event of a light subpath (LightSet::GenerateRay) / an eye subpath (Lens::GenerateRay): run
// once per subpath, kept out of the generator's loop body (code size, see closest_lane).
__device__ __forceinline__ void bdpt_leading_light(const SceneView& scene, SequentialRng& rng, Ray& ray, float3& weight,
                                                float3& normal, int32_t& object, float& bits) {
  int light_index;
  generate_light_ray(scene, rng, ray.o, ray.d, weight, &light_index, &normal);
  const float4* L = reinterpret_cast<const float4*>(scene.lights + light_index);
  object = -3 - light_index;
  bits = event_bits(__float_as_int(__ldg(L + 4).w), kSurfaceLight);
}
__device__ __forceinline__ void bdpt_leading_eye(uint32_t pixel, SequentialRng& rng, Ray& ray, float3& weight,
                                              float3& normal, int32_t& object, float& bits) {
  const uint32_t width = c_camera.sensor.width;
  float w;
  generate_eye_ray(pixel % width, pixel / width, rng, ray.o, ray.d, w, &normal);
  weight = v3(w, w, w);
  object = kObjectLens;
  bits = event_bits(-1, kSurfaceEye);
}

// The state one lane carries while it grows a subpath (GenerateLightSubpath, :726-801, kImportance = true;
// GenerateEyeSubpath, :821-899, false): the ray about to be cast, the weight and normal of the event it
// leaves, and what Combine needs to know about the events so far.
struct BdptSubpath {
  Ray ray;
  float3 weight, normal;
  int n;                 // events so far
  uint32_t specular;     // bit k: event k's surface is not Diffuse (bit 0 set: a leading event never is)
  uint32_t eye_surface;  // bit k: (non-leading) event k lies on an aperture triangle
  bool ends_on_light;    // the last event is an emitter (only the last one can be: a light absorbs)
};

// The leading event: SubpathEvent(object, leading), :281-294 -- direction_out = normal, NaN geometry
// factor and pdfs.
__device__ __forceinline__ void bdpt_begin_subpath(const RenderArgs& args, bool importance, uint32_t pixel,
                                                   SequentialRng& rng, BdptEvent* out, BdptSubpath& sp) {
  int32_t object;
  float bits;
  if (importance) {
    bdpt_leading_light(args.scene, rng, sp.ray, sp.weight, sp.normal, object, bits);
  } else {
    bdpt_leading_eye(pixel, rng, sp.ray, sp.weight, sp.normal, object, bits);
  }
  out[0].mis = make_float4(CUDART_NAN_F, CUDART_NAN_F, CUDART_NAN_F, bits);
  out[0].pos_obj = make_float4(sp.ray.o.x, sp.ray.o.y, sp.ray.o.z, __int_as_float(object));
  out[0].normal4 = make_float4(sp.normal.x, sp.normal.y, sp.normal.z, 0.0f);
  out[0].dout4 = out[0].normal4;
  out[0].weight4 = make_float4(sp.weight.x, sp.weight.y, sp.weight.z, 0.0f);
  sp.specular = 1u;
  sp.eye_surface = 0u;
  sp.ends_on_light = false;
  sp.n = 1;
}

// One iteration of the generators' loop: cast (bdpt_grow_subpath), then append the event and play Russian roulette
// (bdpt_grow_event).  Both return false when the subpath is complete (miss, roulette, or `cap` events).
__device__ __forceinline__ bool bdpt_grow_event(const SceneView& scene, const bool kImportance, SequentialRng& rng, int cap,
                                                BdptEvent* out, BdptSubpath& sp, const HitRecord& hit) {
  if (hit.prim < 0) return false;
  const Ray ray = sp.ray;
  float3 x, hn;
  int material_index;
  surface_point(scene, ray, hit, x, hn, material_index);
  const MaterialDev m = load_material(scene, material_index);
  const float3 wo = -ray.d;
  const Scatter scatter = sample_light(m, hn, wo, rng, kImportance);
  const float3 wi = scatter.dir;
  // forward = the transport being traced, backward = the other one
  const float p_forward = event_pdf(m, hn, wo, wi, kImportance);
  const float p_rr_forward = min_rr(max3(scatter.weight));
  const float p_backward = event_pdf(m, hn, wi, wo, !kImportance);
  const float3 bsdf_backward = event_bsdf(m, hn, wi, wo, !kImportance);
  const float p_rr_backward = min_rr(max3((1.0f / p_backward) * bsdf_backward));
  const int surface = surface_of_kind(m.kind);
  if (surface != kSurfaceDiffuse) sp.specular |= 1u << sp.n;
  if (surface == kSurfaceEye) sp.eye_surface |= 1u << sp.n;
  sp.ends_on_light = surface == kSurfaceLight;
  BdptEvent& e = out[sp.n++];
  e.mis = make_float4(fabsf(dot(wo, sp.normal) * dot(wo, hn) / (hit.t * hit.t)), p_backward * p_rr_backward,
                      p_forward * p_rr_forward, event_bits(material_index, surface));
  e.pos_obj = make_float4(x.x, x.y, x.z, __int_as_float(hit.prim));
  e.normal4 = make_float4(hn.x, hn.y, hn.z, 0.0f);
  e.dout4 = make_float4(wo.x, wo.y, wo.z, 0.0f);
  e.weight4 = make_float4(sp.weight.x, sp.weight.y, sp.weight.z, 0.0f);
  if (!(rng.next() < p_rr_forward)) return false;
  sp.ray.o = x;
  sp.ray.d = wi;
  sp.normal = hn;
  sp.weight = sp.weight * ((1.0f / p_rr_forward) * scatter.weight);
  return sp.n < cap;
}
template <class Accel, int kCopy = 0>
__device__ __forceinline__ bool bdpt_grow_subpath(const SceneView& scene, const bool kImportance, SequentialRng& rng,
                                               int cap, BdptEvent* out, BdptSubpath& sp, BdptCounters& count) {
  if (sp.n >= cap) return false;
  const HitRecord hit = closest_lane<Accel, kCopy>(scene, sp.ray);
  count.casts++;
  return bdpt_grow_event(scene, kImportance, rng, cap, out, sp, hit);
}

struct BdptContribution {
  float3 value;
  bool has_pixel;
  uint32_t pixel;
  // the connection edge, for the MIS weight: geometry factor, pdf (x rr) of sampling the light
  // endpoint from the eye endpoint (radiance transport) and the eye endpoint from the light endpoint
  float edge_gf, edge_p_light, edge_p_importance;
};

__device__ __forceinline__ BdptContribution empty_contribution() {
  BdptContribution c;
  c.value = v3(0.0f, 0.0f, 0.0f);
  c.has_pixel = false;
  c.pixel = 0;
  c.edge_gf = c.edge_p_light = c.edge_p_importance = 0.0f;
  return c;
}

// Connect, :901-1017, s = 0 (:910-921): the eye subpath found a light.
__device__ __forceinline__ BdptContribution bdpt_connect_s0(const SceneView& scene, const BdptEvent* ee) {
  BdptContribution c = empty_contribution();
  if (ee->surface() != kSurfaceLight) return c;
  c.value = ee->weight() * emitted_radiance(load_material(scene, ee->material()), ee->normal(), ee->direction_out());
  return c;
}

// Connect, t = 0 (:923-936): the light subpath found the aperture.
__device__ __forceinline__ BdptContribution bdpt_connect_t0(const BdptEvent* le) {
  BdptContribution c = empty_contribution();
  float response;
  if (le->surface() != kSurfaceEye || le->leading()) return c;
  if (!lens_response(le->position(), le->direction_out(), c.pixel, response)) return c;
  c.has_pixel = true;
  c.value = response * le->weight();
  return c;
}

// Connect, s >= 1 and t >= 1 (:938-1016).  le / ee are register copies of the two endpoints.
// kProbe: stop in front of the visibility cast and return the unoccluded value -- "is this candidate worth a cast?"
// given: the hit of the visibility ray when a traversal kernel has already found it (wavefront on the BVH)
template <class Accel, int kCopy = 0, bool kProbe = false>
__device__ __forceinline__ BdptContribution bdpt_connect(const SceneView& scene, const BdptEvent* le, const BdptEvent* ee,
                                                      BdptCounters& count, const HitRecord* given = nullptr) {
  BdptContribution c = empty_contribution();
  if (!le->leading() && le->surface() != kSurfaceDiffuse) return c;
  if (!ee->leading() && ee->surface() != kSurfaceDiffuse) return c;
  const float3 lp = le->position(), ep = ee->position();
  Ray shadow;
  shadow.o = lp;
  shadow.d = normalize_exact(v3(ep.x - lp.x, ep.y - lp.y, ep.z - lp.z));
  const float3 to_light = -shadow.d;
  float3 value;
  float3 adjoint_bsdf = v3(0.0f, 0.0f, 0.0f);
  MaterialDev lm;
  if (le->leading()) {  // s = 1: Radiance / LightPDFArea
    const float4 le_over_pdf = __ldg(reinterpret_cast<const float4*>(scene.lights + (-3 - le->object())) + 4);
    value = dot(shadow.d, le->normal()) > 0.0f ? v3(le_over_pdf) : v3(0.0f, 0.0f, 0.0f);
  } else {
    lm = load_material(scene, le->material());
    adjoint_bsdf = event_bsdf(lm, le->normal(), le->direction_out(), shadow.d, true);
    value = le->weight() * adjoint_bsdf;
  }
  if (ee->leading()) {  // t = 1
    float response;
    if (!lens_response(ep, to_light, c.pixel, response)) return c;
    c.has_pixel = true;
    value = (response / c_camera.lens.p_area) * value;
  } else {
    // (the eye endpoint's BSDF is evaluated again below for the MIS edge; kept separate for clarity)
    value = value * (ee->weight() * event_bsdf(load_material(scene, ee->material()), ee->normal(), ee->direction_out(),
                                               to_light, false));
  }
  if (!(max3(value) > 0.0f)) return c;
  // "XXX rough check" (:1006-1010): the first object along the shadow ray must be the eye
  // endpoint's object.  The eye's leading event carries the lens's own aperture object, which
  // no cast ever returns, so t = 1 connections contribute nothing in the reference -- nor here.
  if (ee->leading()) return c;
  if (kProbe) {
    c.value = value;
    return c;
  }
  HitRecord hit;
  if (given) {
    hit = *given;
  } else {
    hit = closest_lane<Accel, kCopy>(scene, shadow);
    count.casts++;
  }
  if (hit.prim != ee->object()) return c;
  c.edge_gf = geometry_factor(lp, ep, le->normal(), ee->normal());
  c.value = c.edge_gf * value;
  {  // BufferLightProbability / BufferImportanceProbability at the connection, :519-535, :574-590
    const MaterialDev em = load_material(scene, ee->material());
    const float3 bsdf = event_bsdf(em, ee->normal(), ee->direction_out(), to_light, false);
    const float pdf = event_pdf(em, ee->normal(), ee->direction_out(), to_light, false);
    c.edge_p_light = pdf * min_rr(max3((1.0f / pdf) * bsdf));
    if (le->leading()) {  // DiffuseLight: AdjointBSDF 0, PDFImportance kDiracDelta
      c.edge_p_importance = 0.0f;
    } else {
      const float pdf_i = event_pdf(lm, le->normal(), le->direction_out(), shadow.d, true);
      c.edge_p_importance = pdf_i * min_rr(max3((1.0f / pdf_i) * adjoint_bsdf));
    }
  }
  return c;
}

// BidirectionalPathSamplingBuffer::Buffer + MISWeight(BalanceHeuristic), :423-697 and
// include/amber/rendering/multiple_importance_sampling.h:58-69, without the buffers: with the full
// path x_0 .. x_{s+t-1} (x_k = L[k] for k < s, E[s+t-1-k] otherwise) the reference's arrays are views
// of the events' own three numbers,
//   gf[j]           edge (x_j, x_j+1):  L[j+1].gf | the connection | E[s+t-1-j].gf
//   p_light[j]      x_j sampled from x_j+1 by radiance transport:   L[j+1].p_backward | connection | E[s+t-2-j].p_forward
//   p_importance[j] x_j+1 sampled from x_j by importance transport: L[j].p_forward | connection | E[s+t-1-j].p_backward
// with p_light[last] = the lens's directional pdf and p_importance[0] = the light's (1/pi).  The
// technique ratios are accumulated outward from technique s as running products; a technique is
// disabled when it would connect at a specular interior event (:661-697).
__device__ __forceinline__ float bdpt_mis_weight(const SceneView& scene, const BdptEvent* L, int s, uint32_t spec_l,
                                                 const BdptEvent* E, int t, uint32_t spec_e, const BdptContribution& c) {
  const int n = s + t + 1, last = s + t - 2;  // last edge index
  const float eye_pdf_area = c_camera.lens.p_area;
  // technique i is disabled if a light interior event a in [1, s-2] is specular and i in {a, a+1},
  // or an eye interior event b in [1, t-2] is specular and i in {n-b-1, n-b-2}
  uint64_t disabled = 0;
  {
    const uint32_t li = s >= 3 ? spec_l & ((1u << (s - 1)) - 2u) : 0u;  // bits 1 .. s-2
    disabled |= static_cast<uint64_t>(li) | (static_cast<uint64_t>(li) << 1);
    uint32_t ei = t >= 3 ? spec_e & ((1u << (t - 1)) - 2u) : 0u;
    while (ei) {
      const int b = __ffs(ei) - 1;
      ei &= ei - 1u;
      disabled |= (1ull << (n - b - 1)) | (1ull << (n - b - 2));
    }
  }
  // the lens's directional pdf replaces p_light[last] (:541-547)
  float eye_dir_pdf;
  {
    const float3 z0 = t == 0 ? L[s - 1].position() : E[0].position();
    const float3 z1 = t == 0 ? L[s - 2].position() : t == 1 ? L[s - 1].position() : E[1].position();
    eye_dir_pdf = eye_pdf_direction(normalize(z1 - z0));
  }
  auto edge = [&](int j, float& gf, float& p_light, float& p_importance) {
    if (j < s - 1) {
      const float4 a = L[j + 1].mis;
      gf = a.x;
      p_light = a.y;
      p_importance = L[j].mis.z;
    } else if (j == s - 1 && s > 0 && t > 0) {
      gf = c.edge_gf;
      p_light = c.edge_p_light;
      p_importance = c.edge_p_importance;
    } else {
      const float4 a = E[s + t - 1 - j].mis;
      gf = a.x;
      p_importance = a.y;
      p_light = E[s + t - 2 - j].mis.z;
    }
    if (j == last) p_light = eye_dir_pdf;
    if (j == 0) p_importance = 1.0f / kPiF;  // LightSet::PDFDirection, light_set.h:113-122
  };
  float sum = 1.0f;  // technique s itself
  {
    float p = 1.0f;
    float gf_i = 0.0f, pl_i = 0.0f, pi_i = 0.0f;
    if (s - 1 <= last && s >= 1) edge(s - 1, gf_i, pl_i, pi_i);
    for (int i = s - 1; i >= 0; i--) {  // techniques with a shorter light subpath
      float gf_m = 0.0f, pl_m = 0.0f, pi_m = 0.0f;
      if (i >= 1) edge(i - 1, gf_m, pl_m, pi_m);
      p *= (t == 0 && i == n - 2) ? eye_pdf_area : pl_i * gf_i;
      p /= (i == 0) ? light_pdf_area(scene, L[0].material()) : pi_m * gf_m;
      if (!((disabled >> i) & 1ull)) sum += p;
      gf_i = gf_m; pl_i = pl_m; pi_i = pi_m;
    }
  }
  {
    float p = 1.0f;
    for (int i = s + 1; i < n; i++) {  // techniques with a longer light subpath
      float gf_a = 0.0f, pl_a = 0.0f, pi_a = 0.0f, gf_b = 0.0f, pl_b = 0.0f, pi_b = 0.0f;
      if (i >= 2) edge(i - 2, gf_a, pl_a, pi_a);
      if (i - 1 <= last) edge(i - 1, gf_b, pl_b, pi_b);
      p *= (s == 0 && i == 1) ? light_pdf_area(scene, E[t - 1].material()) : pi_a * gf_a;
      p /= (t > 0 && i == n - 1) ? eye_pdf_area : pl_b * gf_b;
      if (!((disabled >> i) & 1ull)) sum += p;
    }
  }
  // technique s is disabled too when it connects at a specular interior event
  if ((disabled >> s) & 1ull) sum -= 1.0f;
  return 1.0f / sum;
}

// index of the k-th set bit of mask (k < popc(mask)); __fns is a 160-instruction library loop
__device__ __forceinline__ int nth_set_bit(uint32_t mask, int k) {
  for (; k > 0; k--) mask &= mask - 1u;
  return __ffs(mask) - 1;
}

// Thread::operator() / Thread::Render / Combine (algorithm_bdpt.cc:116-148, :1019-1068): one
// (pixel, sample) per LANE at a time, persistent warps.
//
// A pixel sample goes through four kinds of step -- start a light subpath, grow a subpath by one event
// (either side: the same code), start the eye subpath, evaluate one connection candidate -- and the
// number of steps differs from sample to sample (2.8 events per subpath on average, up to 16; 0 to
// 200 candidates).  Run sample by sample in lockstep, a warp spends most of its instructions with a few
// lanes active (ncu: 7 of 32).  Here every lane is a small state machine; per iteration the warp votes
// for the step most lanes are waiting for and runs that step for those lanes; a lane whose sample is
// finished takes the next pair index from a global counter.  No lane ever waits for another lane's longer
// subpath or longer candidate list, only for its kind of step to come up.
//
// Which (s, t) can contribute follows from the surface types alone (Connect, :910-950): s = 0 needs the
// eye subpath to END on a light; t = 0 needs a light-subpath event on the aperture; t = 1 never passes
// the visibility test (see bdpt_connect); every other pair needs a Diffuse (or leading light) endpoint
// on each side -- the PRODUCT of two event sets.  A lane enumerates its candidates by index in
// Combine's order: s = 0, the t = 0 ones by increasing s, then the product.
enum : int { kBdptStartLight = 0, kBdptGrow = 1, kBdptStartEye = 2, kBdptConnect = 3, kBdptDone = 4 };

// Who votes.  On the constant-memory list every step takes about the same time, and a BLOCK-wide vote (all
// warps of a 512-thread block run the same step, two blocks per SM) lets the warps share the instruction
// cache lines of that step: 231 -> 314 M pixel samples/s (128 / 256 / 512 / 1024 threads: 246 / 285 / 314 /
// 310).  On the 4-wide BVH a step's duration varies with the walk, a block-wide barrier per step makes every
// warp wait for the slowest (117 -> 92 M/s at 512 threads), so there each WARP votes for itself.
template <class Accel>
struct BdptLaunch {
  static constexpr bool kBlockVote = !Accel::kPersistent;
  // list: one 512-thread block per SM at 123 registers (fewer spills and a smaller local-memory working
  // set than two blocks at 64 registers: 436 -> 460 M/s); BVH: 8 blocks of 128 threads at 64 registers
  static constexpr int kThreads = kBlockVote ? 512 : 128;
  static constexpr int kBlocks = kBlockVote ? 1 : 8;  // BVH at 4 / 5 / 6 blocks: 110 / 113 / 116 M/s against 122
};
template <class Accel>
__global__ void __launch_bounds__(BdptLaunch<Accel>::kThreads, BdptLaunch<Accel>::kBlocks)
    bdpt_kernel(const RenderArgs args, unsigned long long total_pairs, int cap) {
  BdptEvent events[2 * kBdptMaxEvents];
  BdptEvent* const L = events;
  BdptEvent* const E = events + kBdptMaxEvents;
  BdptCounters count;
  count.casts = 0;
  count.splats = 0;
  uint32_t pairs = 0;
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t lane_lt = (1u << lane) - 1u;
  constexpr uint32_t kFull = 0xffffffffu;

  int phase = kBdptStartLight;
  bool eye_side = false;  // kBdptGrow: which subpath is growing
  uint32_t pixel = 0, sample = 0;
  SequentialRng rng;
  BdptSubpath sp;
  int s_max = 0, t_max = 0;
  uint32_t spec_l = 0, eye_l = 0, s_set = 0, t_set = 0, spec_e = 0;
  int n_s0 = 0, n_t0 = 0, n_s = 0, candidates = 0, next_candidate = 0;
  float3 measurement = v3(0.0f, 0.0f, 0.0f);

  constexpr bool kBlockVote = BdptLaunch<Accel>::kBlockVote;
  __shared__ uint32_t votes[3][4];
  uint32_t iteration = 0;
  if (kBlockVote) {
    if (threadIdx.x < 12) (&votes[0][0])[threadIdx.x] = 0;
    __syncthreads();
  }
  for (;;) {
    const uint32_t want_light = __ballot_sync(kFull, phase == kBdptStartLight);
    const uint32_t want_grow = __ballot_sync(kFull, phase == kBdptGrow);
    const uint32_t want_eye = __ballot_sync(kFull, phase == kBdptStartEye);
    const uint32_t want_connect = __ballot_sync(kFull, phase == kBdptConnect);
    int n_light, n_grow, n_eye, n_connect;
    if (kBlockVote) {
      // every warp of the block runs the SAME step.  Three vote buffers, one barrier per iteration:
      // buffer (i+1)%3 was last read in iteration i-2, which every warp finished before it passed
      // barrier i-1.
      if (threadIdx.x < 4) votes[(iteration + 1) % 3][threadIdx.x] = 0;
      if (lane == 0) {
        if (want_light) atomicAdd(&votes[iteration % 3][0], __popc(want_light));
        if (want_grow) atomicAdd(&votes[iteration % 3][1], __popc(want_grow));
        if (want_eye) atomicAdd(&votes[iteration % 3][2], __popc(want_eye));
        if (want_connect) atomicAdd(&votes[iteration % 3][3], __popc(want_connect));
      }
      __syncthreads();
      n_light = votes[iteration % 3][0];
      n_grow = votes[iteration % 3][1];
      n_eye = votes[iteration % 3][2];
      n_connect = votes[iteration % 3][3];
      iteration++;
    } else {
      n_light = __popc(want_light);
      n_grow = __popc(want_grow);
      n_eye = __popc(want_eye);
      n_connect = __popc(want_connect);
    }
    if (!(n_light | n_grow | n_eye | n_connect)) break;
    // the step most lanes wait for (ties: grow, connect, start eye, start light -- work in flight first)
    int step = kBdptGrow, best = n_grow;
    if (n_connect > best) { step = kBdptConnect; best = n_connect; }
    if (n_eye > best) { step = kBdptStartEye; best = n_eye; }
    if (n_light > best) { step = kBdptStartLight; best = n_light; }

    if (step == kBdptStartLight) {
      if (phase == kBdptStartLight) {
        // next pair index: one atomic per warp
        const int leader = __ffs(want_light) - 1;
        unsigned long long base = 0;
        if (static_cast<int>(lane) == leader)
          base = atomicAdd(&args.ctl->cursor_bdpt, static_cast<unsigned long long>(__popc(want_light)));
        base = __shfl_sync(want_light, base, leader);
        const unsigned long long pair = base + __popc(want_light & lane_lt);
        if (pair >= total_pairs) {
          phase = kBdptDone;
        } else {
          pixel = static_cast<uint32_t>(pair % args.n_pixels);
          sample = static_cast<uint32_t>(args.sample_begin + (pair / args.n_pixels) * args.sample_stride);
          pairs++;
          measurement = v3(0.0f, 0.0f, 0.0f);
          rng.init(args.seed, pixel, sample, 0x40000000u);
          bdpt_begin_subpath(args, true, pixel, rng, L, sp);
          eye_side = false;
          phase = kBdptGrow;
        }
      }
    } else if (step == kBdptStartEye) {
      if (phase == kBdptStartEye) {
        rng.init(args.seed, pixel, sample, 0x40000001u);
        bdpt_begin_subpath(args, false, pixel, rng, E, sp);
        eye_side = true;
        phase = kBdptGrow;
      }
    } else if (step == kBdptGrow) {
      if (phase == kBdptGrow) {
        const bool more = bdpt_grow_subpath<Accel>(args.scene, !eye_side, rng, cap, eye_side ? E : L, sp, count);
        if (!more) {
          if (!eye_side) {
            s_max = sp.n;
            spec_l = sp.specular;
            eye_l = sp.eye_surface;
            phase = kBdptStartEye;
          } else {
            t_max = sp.n;
            spec_e = sp.specular;
            // the candidate list of this sample
            n_s0 = (t_max >= 2 && sp.ends_on_light) ? 1 : 0;
            n_t0 = __popc(eye_l);
            s_set = (~spec_l | 1u) & ((s_max >= 32 ? 0u : (1u << s_max)) - 1u);   // bit k: s = k + 1
            t_set = (~spec_e & ~1u) & ((t_max >= 32 ? 0u : (1u << t_max)) - 1u);  // bit k: t = k + 1 >= 2
            n_s = __popc(s_set);
            candidates = n_s0 + n_t0 + n_s * __popc(t_set);
            next_candidate = 0;
            phase = candidates > 0 ? kBdptConnect : kBdptStartLight;
          }
        }
      }
    } else {
      if (phase == kBdptConnect) {
        const int j = next_candidate++;
        BdptContribution c;
        int s, t;
        if (j < n_s0) {
          s = 0;
          t = t_max;
          c = bdpt_connect_s0(args.scene, &E[t_max - 1]);
        } else if (j < n_s0 + n_t0) {
          s = nth_set_bit(eye_l, j - n_s0) + 1;  // event index + 1
          t = 0;
          c = bdpt_connect_t0(&L[s - 1]);
        } else {
          const int k = j - n_s0 - n_t0;
          s = nth_set_bit(s_set, k % n_s) + 1;
          t = nth_set_bit(t_set, k / n_s) + 1;
          // the endpoints are read in place: a copy would be made in local memory (the callee is out of
          // line), 160 B loaded and stored again per candidate
          c = bdpt_connect<Accel>(args.scene, &L[s - 1], &E[t - 1], count);
        }
        if (max3(c.value) > 0.0f) {
          const float weight = bdpt_mis_weight(args.scene, L, s, spec_l, E, t, spec_e, c);
          if (c.has_pixel) {  // light image: / sensor.Size(), :1054-1056
            film_add(args.film, c.pixel, (weight * args.inv_n_pixels) * c.value);
            count.splats++;
          } else {
            measurement = measurement + weight * c.value;
          }
        }
        if (next_candidate >= candidates) {
          if (max3(measurement) > 0.0f) film_add(args.film, pixel, measurement);
          phase = kBdptStartLight;
        }
      }
    }
  }
  // counters: one atomic per warp
  const uint32_t casts = __reduce_add_sync(kFull, count.casts);
  const uint32_t splats = __reduce_add_sync(kFull, count.splats);
  const uint32_t n_pairs = __reduce_add_sync(kFull, pairs);
  if (lane == 0) {
    atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
    atomicAdd(&args.ctl->splats, static_cast<unsigned long long>(splats));
    atomicAdd(&args.ctl->paths, static_cast<unsigned long long>(n_pairs));
  }
}

// =====================================================================================================
// The same algorithm as a WAVEFRONT: a pool of items, each one (pixel, sample) pair in flight,
// state in global memory, two kernels per iteration.
//   bdpt_grow_kernel     one thread per item, every item of the pool: take a new pair / start the eye subpath where
//                        due, then ONE step of the subpath generator (cast, event, roulette) -- the state machine's
//                        "start light", "start eye" and "grow" steps, which every item is in at any time except while it
//                        waits for its connections, so no queue is needed and every lane of the launch has work.
//                        (Splitting it into a cast kernel and an event kernel whose blocks deal their items out in order
//                        of the material hit was measured: 134 + 143 us against 190 us for the one kernel -- the split
//                        pays the state traffic twice and the start steps still diverge.)
//   bdpt_connect_kernel  the items whose eye subpath ended in this iteration (a queue): the candidates of the 32 items
//                        of a warp are POOLED and dealt out lane by lane (prefix sums in shared memory), so a sample
//                        with 200 candidates does not hold up one with 2; measurements go to the film one by one.
// The subpath state is a structure of arrays (coalesced: thread i <-> item i); the events are an array of structures per
// item (80-byte events read through pointers by the same connection / MIS code as the one-kernel version).
struct BdptPool {
  float4* state;        // [kBdptStateWords][n_items]
  BdptEvent* events;    // [n_items][2 * kBdptMaxEvents]: light subpath, eye subpath
  uint32_t* q_connect;  // [n_items]
  uint2* q_evaluate;    // [evaluate_capacity] candidates that passed the probe: item, s | t << 8
  float4* evaluate_ray; // [2 * evaluate_capacity] BVH only: the candidate's visibility ray (direction.w = 1) or no ray (0)
  float2* evaluate_hit; // [evaluate_capacity] BVH only: what bdpt_trace_evaluate_kernel found along it
  uint32_t evaluate_capacity;
  uint32_t n_items;     // a multiple of kBdptGrowBlock
};
// state words: 0 ray.o + weight.x | 1 ray.d + weight.y | 2 normal + weight.z | 3 pixel, sample, rng index, packed |
//              4 specular, eye_surface (growing subpath), spec_l, eye_l (finished light subpath) | 5 the rng's Philox block |
//              6 (BVH only) the hit of this iteration's cast: t, primitive
// packed: events so far [0,5) | eye side 5 | ends on light 6 | phase [7,9) | s_max [9,14)
constexpr int kBdptStateWords = 7;
constexpr int kBdptGrowBlock = 128;
enum : uint32_t { kWaveStartLight = 0, kWaveGrow = 1, kWaveStartEye = 2, kWaveDone = 3 };

#ifndef AMBER_BDPT_GROW_BLOCKS
#define AMBER_BDPT_GROW_BLOCKS 6
#endif
#ifndef AMBER_BDPT_CONNECT_BLOCKS
#define AMBER_BDPT_CONNECT_BLOCKS 6
#endif
#ifndef AMBER_BDPT_EVENT_BLOCKS
#define AMBER_BDPT_EVENT_BLOCKS 6  // 80 registers, no spills (8 / 10 blocks at 64 / 48 registers: no faster)
#endif
#ifndef AMBER_BDPT_COPY_ENDPOINTS
#define AMBER_BDPT_COPY_ENDPOINTS 1
#endif
constexpr uint32_t kPackedCount = 31u, kPackedEyeSide = 32u, kPackedEndsOnLight = 64u;

// (Constant-memory list only: on the BVH the step is split into begin / trace / event kernels, below.  Dealing a block's
// items to its threads in order of the step they take -- start light, start eye, grow light, grow eye -- cost the list
// 3 %: 841 -> 818 M pixel samples/s.)
template <class Accel>
__global__ void __launch_bounds__(kBdptGrowBlock, AMBER_BDPT_GROW_BLOCKS) bdpt_grow_kernel(const RenderArgs args, const BdptPool pool,
                                                                                            unsigned long long total_pairs, int cap,
                                                                                            uint32_t parity) {
  constexpr uint32_t kFull = 0xffffffffu;
  const uint32_t lane = threadIdx.x & 31u, lane_lt = (1u << lane) - 1u;
  const size_t n = pool.n_items;
  const uint32_t i = blockIdx.x * kBdptGrowBlock + threadIdx.x;  // < n_items: the grid is exact
  // the whole state is requested at once, before the phase (in `ids`) says which words this step needs: one round trip
  // to memory instead of two
  uint4 ids = reinterpret_cast<const uint4*>(pool.state + 3 * n)[i];
  const float4 state_a = pool.state[i], state_b = pool.state[n + i], state_c = pool.state[2 * n + i];
  const uint4 state_masks = reinterpret_cast<const uint4*>(pool.state + 4 * n)[i];
  const uint4 state_block = reinterpret_cast<const uint4*>(pool.state + 5 * n)[i];
  uint32_t pixel = ids.x, sample = ids.y;
  uint32_t phase = (ids.w >> 7) & 3u;
  BdptCounters count;
  count.casts = 0;
  count.splats = 0;
  uint32_t pairs = 0, newly_done = 0;
  // new pairs: one atomic per BLOCK (60 k atomics on one address per launch of a 2 M-item pool when every warp asks)
  constexpr int kWarps = kBdptGrowBlock / 32;
  __shared__ uint32_t warp_count[kWarps];
  __shared__ unsigned long long block_pair_base;
  const uint32_t warp = threadIdx.x >> 5;
  const uint32_t want = __ballot_sync(kFull, phase == kWaveStartLight);
  if (lane == 0) warp_count[warp] = __popc(want);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) total += warp_count[w];
    block_pair_base = total ? atomicAdd(&args.ctl->cursor_bdpt, static_cast<unsigned long long>(total)) : 0ull;
  }
  __syncthreads();
  if (phase == kWaveStartLight) {
    uint32_t before = __popc(want & lane_lt);
#pragma unroll
    for (int w = 0; w < kWarps; w++)
      if (static_cast<uint32_t>(w) < warp) before += warp_count[w];
    const unsigned long long pair = block_pair_base + before;
    if (pair >= total_pairs) {
      phase = kWaveDone;
      newly_done = 1;
    } else {
      pixel = static_cast<uint32_t>(pair % args.n_pixels);
      sample = static_cast<uint32_t>(args.sample_begin + (pair / args.n_pixels) * args.sample_stride);
      pairs = 1;
    }
  }
  if (i == 0) {  // for this iteration's connect kernels
    args.ctl->bdpt_connect_cursor = 0;
    args.ctl->bdpt_n_evaluate = 0;
  }
  const bool active = phase != kWaveDone;
  BdptEvent* const L = pool.events + static_cast<size_t>(i) * (2 * kBdptMaxEvents);
  BdptEvent* const E = L + kBdptMaxEvents;
  SequentialRng rng;
  BdptSubpath sp;
  bool eye_side = false, more = false;
  uint32_t s_max = (ids.w >> 9) & 31u, spec_l = 0, eye_l = 0;
  if (active) {
    if (phase == kWaveStartLight) {
      rng.init(args.seed, pixel, sample, 0x40000000u);
      bdpt_begin_subpath(args, true, pixel, rng, L, sp);
    } else if (phase == kWaveStartEye) {
      const uint4 masks = state_masks;
      spec_l = masks.z;
      eye_l = masks.w;
      rng.init(args.seed, pixel, sample, 0x40000001u);
      bdpt_begin_subpath(args, false, pixel, rng, E, sp);
      eye_side = true;
    } else {
      const float4 a = state_a, b = state_b, c = state_c;
      const uint4 masks = state_masks;
      sp.ray.o = v3(a);
      sp.ray.d = v3(b);
      sp.normal = v3(c);
      sp.weight = v3(a.w, b.w, c.w);
      sp.n = static_cast<int>(ids.w & kPackedCount);
      sp.specular = masks.x;
      sp.eye_surface = masks.y;
      spec_l = masks.z;
      eye_l = masks.w;
      sp.ends_on_light = false;  // (a subpath that goes on did not end on a light)
      eye_side = (ids.w & kPackedEyeSide) != 0;
      rng.init(args.seed, pixel, sample, eye_side ? 0x40000001u : 0x40000000u);
      rng.i = ids.z;
      rng.block = state_block;
    }
    more = bdpt_grow_subpath<Accel, 1>(args.scene, !eye_side, rng, cap, eye_side ? E : L, sp, count);
  }
  bool to_connect = false;
  if (active) {
    if (more) {
      phase = kWaveGrow;
    } else if (!eye_side) {
      s_max = static_cast<uint32_t>(sp.n);
      spec_l = sp.specular;
      eye_l = sp.eye_surface;
      phase = kWaveStartEye;
    } else {
      // the candidate list of this sample (see bdpt_kernel): anything to connect?
      const int t_max = sp.n;
      const uint32_t s_set = (~spec_l | 1u) & ((s_max >= 32u ? 0u : (1u << s_max)) - 1u);
      const uint32_t t_set = (~sp.specular & ~1u) & ((t_max >= 32 ? 0u : (1u << t_max)) - 1u);
      const int candidates = ((t_max >= 2 && sp.ends_on_light) ? 1 : 0) + __popc(eye_l) + __popc(s_set) * __popc(t_set);
      to_connect = candidates > 0;
      phase = kWaveStartLight;
    }
  }
  const uint32_t queued = __ballot_sync(kFull, to_connect);
  if (queued) {  // (per warp: a block-wide count would make every warp wait here for the block's longest cast)
    const int leader = __ffs(queued) - 1;
    uint32_t base = 0;
    if (static_cast<int>(lane) == leader) base = atomicAdd(&args.ctl->bdpt_n_connect[parity], __popc(queued));
    base = __shfl_sync(kFull, base, leader);
    if (to_connect) pool.q_connect[base + __popc(queued & lane_lt)] = i;
  }
  if (active) {
    pool.state[i] = make_float4(sp.ray.o.x, sp.ray.o.y, sp.ray.o.z, sp.weight.x);
    pool.state[n + i] = make_float4(sp.ray.d.x, sp.ray.d.y, sp.ray.d.z, sp.weight.y);
    pool.state[2 * n + i] = make_float4(sp.normal.x, sp.normal.y, sp.normal.z, sp.weight.z);
    reinterpret_cast<uint4*>(pool.state + 4 * n)[i] = make_uint4(sp.specular, sp.eye_surface, spec_l, eye_l);
    reinterpret_cast<uint4*>(pool.state + 5 * n)[i] = rng.block;
    ids = make_uint4(pixel, sample, rng.i,
                     static_cast<uint32_t>(sp.n) | (eye_side ? kPackedEyeSide : 0u) | (sp.ends_on_light ? kPackedEndsOnLight : 0u) |
                         (phase << 7) | (s_max << 9));
  } else {
    ids.w = (ids.w & ~(3u << 7)) | (kWaveDone << 7);
  }
  reinterpret_cast<uint4*>(pool.state + 3 * n)[i] = ids;
  const uint32_t casts = __reduce_add_sync(kFull, count.casts);
  const uint32_t n_pairs = __reduce_add_sync(kFull, pairs);
  const uint32_t n_done = __reduce_add_sync(kFull, newly_done);
  if (lane == 0) {
    if (casts) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
    if (n_pairs) atomicAdd(&args.ctl->paths, static_cast<unsigned long long>(n_pairs));
    if (n_done) atomicAdd(&args.ctl->bdpt_done, n_done);
  }
}

// ---- the grow step in three launches, for the BVH -------------------------------------------------------------
// On the BVH the cast is most of a grow step, and a per-thread walk of the tree inside a lock-step launch runs at a quarter
// of the rate of the path tracer's warp-cooperative traversal (bvh.cuh, trace_persistent: lanes of a warp kept together
// by votes, finished lanes refilled).  So there the step is: begin (new pairs, leading events) -> trace (the persistent
// traversal over the items' rays, hits to state word 6) -> event (the rest of the generator's loop body).
constexpr int32_t kPrimCapped = -3;  // state word 6 of an item whose subpath already holds `cap` events: no cast was made

__global__ void __launch_bounds__(kBdptGrowBlock) bdpt_begin_kernel(const RenderArgs args, const BdptPool pool, unsigned long long total_pairs) {
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr int kWarps = kBdptGrowBlock / 32;
  __shared__ uint32_t warp_count[kWarps];
  __shared__ unsigned long long block_pair_base;
  const uint32_t lane = threadIdx.x & 31u, lane_lt = (1u << lane) - 1u, warp = threadIdx.x >> 5;
  const size_t n = pool.n_items;
  // (dealing the block's items to its threads by what they start -- light, eye, nothing -- was measured: 196 us per
  // launch against 171)
  const uint32_t i = blockIdx.x * kBdptGrowBlock + threadIdx.x;
  uint4 ids = reinterpret_cast<const uint4*>(pool.state + 3 * n)[i];
  uint32_t phase = (ids.w >> 7) & 3u;
  uint32_t pairs = 0, newly_done = 0;
  const uint32_t want = __ballot_sync(kFull, phase == kWaveStartLight);
  if (lane == 0) warp_count[warp] = __popc(want);
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t total = 0;
#pragma unroll
    for (int w = 0; w < kWarps; w++) total += warp_count[w];
    block_pair_base = total ? atomicAdd(&args.ctl->cursor_bdpt, static_cast<unsigned long long>(total)) : 0ull;
  }
  __syncthreads();
  if (i == 0) {  // for this iteration's trace and connect kernels
    args.ctl->cursor_cast = 0;
    args.ctl->cursor_connect = 0;
    args.ctl->bdpt_connect_cursor = 0;
    args.ctl->bdpt_n_evaluate = 0;
  }
  bool begin = phase == kWaveStartEye;
  if (phase == kWaveStartLight) {
    uint32_t before = __popc(want & lane_lt);
#pragma unroll
    for (int w = 0; w < kWarps; w++)
      if (static_cast<uint32_t>(w) < warp) before += warp_count[w];
    const unsigned long long pair = block_pair_base + before;
    if (pair >= total_pairs) {
      newly_done = 1;
      ids.w = (ids.w & ~(3u << 7)) | (kWaveDone << 7);
      reinterpret_cast<uint4*>(pool.state + 3 * n)[i] = ids;
    } else {
      ids.x = static_cast<uint32_t>(pair % args.n_pixels);
      ids.y = static_cast<uint32_t>(args.sample_begin + (pair / args.n_pixels) * args.sample_stride);
      pairs = 1;
      begin = true;
    }
  }
  if (begin) {
    const bool eye_side = phase == kWaveStartEye;
    BdptEvent* const L = pool.events + static_cast<size_t>(i) * (2 * kBdptMaxEvents);
    SequentialRng rng;
    BdptSubpath sp;
    rng.init(args.seed, ids.x, ids.y, eye_side ? 0x40000001u : 0x40000000u);
    bdpt_begin_subpath(args, !eye_side, ids.x, rng, eye_side ? L + kBdptMaxEvents : L, sp);
    uint4 masks = make_uint4(sp.specular, sp.eye_surface, 0u, 0u);
    if (eye_side) {  // the finished light subpath's masks stay
      const uint4 old = reinterpret_cast<const uint4*>(pool.state + 4 * n)[i];
      masks.z = old.z;
      masks.w = old.w;
    }
    pool.state[i] = make_float4(sp.ray.o.x, sp.ray.o.y, sp.ray.o.z, sp.weight.x);
    pool.state[n + i] = make_float4(sp.ray.d.x, sp.ray.d.y, sp.ray.d.z, sp.weight.y);
    pool.state[2 * n + i] = make_float4(sp.normal.x, sp.normal.y, sp.normal.z, sp.weight.z);
    reinterpret_cast<uint4*>(pool.state + 4 * n)[i] = masks;
    reinterpret_cast<uint4*>(pool.state + 5 * n)[i] = rng.block;
    ids.z = rng.i;
    ids.w = static_cast<uint32_t>(sp.n) | (eye_side ? kPackedEyeSide : 0u) | (kWaveGrow << 7) | (ids.w & (31u << 9));
    reinterpret_cast<uint4*>(pool.state + 3 * n)[i] = ids;
  }
  const uint32_t n_pairs = __reduce_add_sync(kFull, pairs);
  const uint32_t n_done = __reduce_add_sync(kFull, newly_done);
  if (lane == 0) {
    if (n_pairs) atomicAdd(&args.ctl->paths, static_cast<unsigned long long>(n_pairs));
    if (n_done) atomicAdd(&args.ctl->bdpt_done, n_done);
  }
}

struct BdptCastSource {
  const RenderArgs& args;
  const BdptPool& pool;
  int cap;
  SphereBound spheres;
  uint32_t casts;
  __device__ __forceinline__ uint32_t count() const { return pool.n_items; }
  __device__ __forceinline__ bool load(uint32_t item, Ray& ray, float& t_best) {
    const size_t n = pool.n_items;
    const uint32_t packed = reinterpret_cast<const uint4*>(pool.state + 3 * n)[item].w;
    if (((packed >> 7) & 3u) != kWaveGrow) return false;
    if (static_cast<int>(packed & kPackedCount) >= cap) {
      pool.state[6 * n + item] = make_float4(0.0f, __int_as_float(kPrimCapped), 0.0f, 0.0f);
      return false;
    }
    ray.o = v3(pool.state[item]);
    ray.d = v3(pool.state[n + item]);
    t_best = spheres.bound(args.scene, ray);
    casts++;
    return true;
  }
  __device__ __forceinline__ void store(uint32_t item, const HitRecord& hit, bool found) const {
    const HitRecord out = spheres.resolve(args.scene, hit, found);
    pool.state[6 * static_cast<size_t>(pool.n_items) + item] = make_float4(out.t, __int_as_float(out.prim), 0.0f, 0.0f);
  }
};

__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) bdpt_trace_kernel(const RenderArgs args, const BdptPool pool, int cap) {
  BdptCastSource source{args, pool, cap, {-1}, 0u};
  WideBvh::trace_persistent<false>(args.scene, source, &args.ctl->cursor_cast);
  const uint32_t casts = __reduce_add_sync(0xffffffffu, source.casts);
  if ((threadIdx.x & 31) == 0 && casts) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
}

__global__ void __launch_bounds__(kBdptGrowBlock, AMBER_BDPT_EVENT_BLOCKS) bdpt_event_kernel(const RenderArgs args, const BdptPool pool, int cap, uint32_t parity) {
  constexpr uint32_t kFull = 0xffffffffu;
  const uint32_t i = blockIdx.x * kBdptGrowBlock + threadIdx.x;
  const uint32_t lane = threadIdx.x & 31u, lane_lt = (1u << lane) - 1u;
  const size_t n = pool.n_items;
  uint4 ids = reinterpret_cast<const uint4*>(pool.state + 3 * n)[i];
  const float4 a = pool.state[i], b = pool.state[n + i], c = pool.state[2 * n + i], h = pool.state[6 * n + i];
  uint4 masks = reinterpret_cast<const uint4*>(pool.state + 4 * n)[i];
  const uint4 block = reinterpret_cast<const uint4*>(pool.state + 5 * n)[i];
  bool to_connect = false;
  if (((ids.w >> 7) & 3u) == kWaveGrow) {
    BdptEvent* const L = pool.events + static_cast<size_t>(i) * (2 * kBdptMaxEvents);
    const bool eye_side = (ids.w & kPackedEyeSide) != 0;
    uint32_t s_max = (ids.w >> 9) & 31u;
    BdptSubpath sp;
    sp.ray.o = v3(a);
    sp.ray.d = v3(b);
    sp.normal = v3(c);
    sp.weight = v3(a.w, b.w, c.w);
    sp.n = static_cast<int>(ids.w & kPackedCount);
    sp.specular = masks.x;
    sp.eye_surface = masks.y;
    sp.ends_on_light = (ids.w & kPackedEndsOnLight) != 0;
    SequentialRng rng;
    rng.init(args.seed, ids.x, ids.y, eye_side ? 0x40000001u : 0x40000000u);
    rng.i = ids.z;
    rng.block = block;
    HitRecord hit;
    hit.t = h.x;
    hit.prim = __float_as_int(h.y);
    const bool more = hit.prim != kPrimCapped &&
                      bdpt_grow_event(args.scene, !eye_side, rng, cap, eye_side ? L + kBdptMaxEvents : L, sp, hit);
    uint32_t phase = kWaveGrow;
    if (!more) {
      if (!eye_side) {
        s_max = static_cast<uint32_t>(sp.n);
        masks.z = sp.specular;
        masks.w = sp.eye_surface;
        phase = kWaveStartEye;
      } else {
        const int t_max = sp.n;
        const uint32_t s_set = (~masks.z | 1u) & ((s_max >= 32u ? 0u : (1u << s_max)) - 1u);
        const uint32_t t_set = (~sp.specular & ~1u) & ((t_max >= 32 ? 0u : (1u << t_max)) - 1u);
        const int candidates = ((t_max >= 2 && sp.ends_on_light) ? 1 : 0) + __popc(masks.w) + __popc(s_set) * __popc(t_set);
        to_connect = candidates > 0;
        phase = kWaveStartLight;
      }
    }
    pool.state[i] = make_float4(sp.ray.o.x, sp.ray.o.y, sp.ray.o.z, sp.weight.x);
    pool.state[n + i] = make_float4(sp.ray.d.x, sp.ray.d.y, sp.ray.d.z, sp.weight.y);
    pool.state[2 * n + i] = make_float4(sp.normal.x, sp.normal.y, sp.normal.z, sp.weight.z);
    reinterpret_cast<uint4*>(pool.state + 4 * n)[i] = make_uint4(sp.specular, sp.eye_surface, masks.z, masks.w);
    reinterpret_cast<uint4*>(pool.state + 5 * n)[i] = rng.block;
    ids.z = rng.i;
    ids.w = static_cast<uint32_t>(sp.n) | (eye_side ? kPackedEyeSide : 0u) | (sp.ends_on_light ? kPackedEndsOnLight : 0u) |
            (phase << 7) | (s_max << 9);
    reinterpret_cast<uint4*>(pool.state + 3 * n)[i] = ids;
  }
  const uint32_t queued = __ballot_sync(kFull, to_connect);
  if (queued) {
    const int leader = __ffs(queued) - 1;
    uint32_t base = 0;
    if (static_cast<int>(lane) == leader) base = atomicAdd(&args.ctl->bdpt_n_connect[parity], __popc(queued));
    base = __shfl_sync(kFull, base, leader);
    if (to_connect) pool.q_connect[base + __popc(queued & lane_lt)] = i;
  }
}

constexpr int kBdptConnectBlock = 128;

// One connection candidate of an item, start to finish: Connect for (s, t), the MIS weight, the film.
template <class Accel, int kCopy>
__device__ __forceinline__ void bdpt_evaluate_candidate(const RenderArgs& args, const BdptPool& pool, uint32_t item, int s, int t,
                                                        BdptCounters& count, const HitRecord* given = nullptr) {
  const size_t n = pool.n_items;
  const BdptEvent* const L = pool.events + static_cast<size_t>(item) * (2 * kBdptMaxEvents);
  const BdptEvent* const E = L + kBdptMaxEvents;
  BdptContribution contribution;
  if (s == 0) contribution = bdpt_connect_s0(args.scene, &E[t - 1]);
  else if (t == 0) contribution = bdpt_connect_t0(&L[s - 1]);
  else {
#if AMBER_BDPT_COPY_ENDPOINTS
    const BdptEvent le = L[s - 1], ee = E[t - 1];  // ten 16-byte loads in flight at once instead of one per early-out test
    contribution = bdpt_connect<Accel, kCopy>(args.scene, &le, &ee, count, given);
#else
    contribution = bdpt_connect<Accel, kCopy>(args.scene, &L[s - 1], &E[t - 1], count, given);
#endif
  }
  if (!(max3(contribution.value) > 0.0f)) return;
  const uint4 masks = reinterpret_cast<const uint4*>(pool.state + 4 * n)[item];  // spec_e, -, spec_l, eye_l
  const float weight = bdpt_mis_weight(args.scene, L, s, masks.z, E, t, masks.x, contribution);
  if (contribution.has_pixel) {  // light image: / sensor.Size(), :1054-1056
    film_add(args.film, contribution.pixel, (weight * args.inv_n_pixels) * contribution.value);
    count.splats++;
  } else if (weight == weight) {
    // (a NaN weight -- a 0 * inf among the pdf ratios of a degenerate path -- never reaches the film in the
    // one-kernel version either: there it makes the sample's whole measurement NaN, which fails `> 0`)
    const uint32_t pixel = reinterpret_cast<const uint4*>(pool.state + 3 * n)[item].x;
    film_add(args.film, pixel, weight * contribution.value);
  }
}

// Connections, step 1: the candidates of the items whose eye subpath ended in this iteration.  A warp takes 32 queued
// items, pools their candidates behind a prefix sum and deals them out lane by lane; a lane PROBES its candidate --
// everything Connect does in front of the visibility cast -- and the one in five that is worth a cast goes onto the
// evaluation queue.  (Evaluating in place ran the cast and the MIS weight with a fifth of the lanes: 152 us per
// iteration of a 1 M-item pool against 50 + 50 for probe + evaluate.)
template <class Accel>
__global__ void __launch_bounds__(kBdptConnectBlock, AMBER_BDPT_CONNECT_BLOCKS) bdpt_connect_kernel(const RenderArgs args, const BdptPool pool,
                                                                                                    uint32_t parity) {
  constexpr uint32_t kFull = 0xffffffffu;
  constexpr int kWarps = kBdptConnectBlock / 32;
  __shared__ uint32_t s_first[kWarps][32];   // index of an item's first candidate in the warp's pooled list
  __shared__ uint32_t s_item[kWarps][32], s_counts[kWarps][32];
  __shared__ uint32_t s_spec_l[kWarps][32], s_spec_e[kWarps][32], s_eye_l[kWarps][32];
  const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5, lane_lt = (1u << lane) - 1u;
  const uint32_t n_queued = args.ctl->bdpt_n_connect[parity];
  // the other counter is the one the next iteration's grow kernel appends to; its last reader was the previous connect
  if (blockIdx.x == 0 && threadIdx.x == 0) args.ctl->bdpt_n_connect[parity ^ 1u] = 0;
  const size_t n = pool.n_items;
  BdptCounters count;
  count.casts = 0;
  count.splats = 0;
  // chunks of 32 queue entries are taken from a counter (the grow kernel of this iteration reset it)
  for (;;) {
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(&args.ctl->bdpt_connect_cursor, 32u);
    base = __shfl_sync(kFull, base, 0);
    if (base >= n_queued) break;
    const bool have = base + lane < n_queued;
    const uint32_t item = have ? pool.q_connect[base + lane] : 0u;
    uint32_t candidates = 0;
    if (have) {
      const uint4 ids = reinterpret_cast<const uint4*>(pool.state + 3 * n)[item];
      const uint4 masks = reinterpret_cast<const uint4*>(pool.state + 4 * n)[item];
      const uint32_t t_max = ids.w & 31u, s_max = (ids.w >> 9) & 31u;
      const uint32_t spec_e = masks.x, spec_l = masks.z, eye_l = masks.w;
      const uint32_t n_s0 = (t_max >= 2u && ((ids.w >> 6) & 1u)) ? 1u : 0u;
      const uint32_t s_set = (~spec_l | 1u) & ((s_max >= 32u ? 0u : (1u << s_max)) - 1u);
      const uint32_t t_set = (~spec_e & ~1u) & ((t_max >= 32u ? 0u : (1u << t_max)) - 1u);
      candidates = n_s0 + __popc(eye_l) + __popc(s_set) * __popc(t_set);
      s_item[warp][lane] = item;
      s_counts[warp][lane] = s_max | (t_max << 8) | (n_s0 << 16);
      s_spec_l[warp][lane] = spec_l;
      s_spec_e[warp][lane] = spec_e;
      s_eye_l[warp][lane] = eye_l;
    }
    uint32_t scan = candidates;
#pragma unroll
    for (int delta = 1; delta < 32; delta <<= 1) {
      const uint32_t up = __shfl_up_sync(kFull, scan, delta);
      if (lane >= static_cast<uint32_t>(delta)) scan += up;
    }
    s_first[warp][lane] = scan - candidates;
    const uint32_t total = __shfl_sync(kFull, scan, 31);
    __syncwarp();
    for (uint32_t c0 = 0; c0 < total; c0 += 32u) {
      const uint32_t c = c0 + lane;
      bool worth = false;
      uint32_t owner_item = 0;
      int s = 0, t = 0;
      float3 ray_o = v3(0.0f, 0.0f, 0.0f), ray_d = v3(0.0f, 0.0f, 0.0f);  // BVH: the visibility ray of a product candidate
      float ray_reach = 0.0f;                                              // ... and the distance to its eye endpoint
      if (c < total) {
        // owner: the last item whose first candidate is <= c (items without candidates share their index with the next)
        int owner = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1)
          if (owner + step < 32 && s_first[warp][owner + step] <= c) owner += step;
        const int j = static_cast<int>(c - s_first[warp][owner]);
        owner_item = s_item[warp][owner];
        const BdptEvent* const L = pool.events + static_cast<size_t>(owner_item) * (2 * kBdptMaxEvents);
        const BdptEvent* const E = L + kBdptMaxEvents;
        const uint32_t counts = s_counts[warp][owner];
        const int s_max = counts & 0xff, t_max = (counts >> 8) & 0xff, n_s0 = (counts >> 16) & 1;
        const uint32_t spec_l = s_spec_l[warp][owner], spec_e = s_spec_e[warp][owner], eye_l = s_eye_l[warp][owner];
        const int n_t0 = __popc(eye_l);
        if (j < n_s0) {  // the eye subpath ended on a light (that is what n_s0 says)
          s = 0;
          t = t_max;
          worth = true;
        } else if (j < n_s0 + n_t0) {  // a light-subpath event on the aperture
          s = nth_set_bit(eye_l, j - n_s0) + 1;
          t = 0;
          worth = true;
        } else {
          const uint32_t s_set = (~spec_l | 1u) & ((s_max >= 32 ? 0u : (1u << s_max)) - 1u);
          const uint32_t t_set = (~spec_e & ~1u) & ((t_max >= 32 ? 0u : (1u << t_max)) - 1u);
          const int n_s = __popc(s_set), k = j - n_s0 - n_t0;
          s = nth_set_bit(s_set, k % n_s) + 1;
          t = nth_set_bit(t_set, k / n_s) + 1;
#if AMBER_BDPT_COPY_ENDPOINTS
          const BdptEvent le = L[s - 1], ee = E[t - 1];
          worth = max3(bdpt_connect<Accel, 3, true>(args.scene, &le, &ee, count).value) > 0.0f;
          if (Accel::kPersistent && worth) {  // as bdpt_connect builds it
            const float3 lp = le.position(), ep = ee.position();
            ray_o = lp;
            ray_d = normalize_exact(v3(ep.x - lp.x, ep.y - lp.y, ep.z - lp.z));
            ray_reach = sqrtf(dot(ep - lp, ep - lp));
          }
#else
          worth = max3(bdpt_connect<Accel, 3, true>(args.scene, &L[s - 1], &E[t - 1], count).value) > 0.0f;
          if (Accel::kPersistent && worth) {
            const float3 lp = L[s - 1].position(), ep = E[t - 1].position();
            ray_o = lp;
            ray_d = normalize_exact(v3(ep.x - lp.x, ep.y - lp.y, ep.z - lp.z));
            ray_reach = sqrtf(dot(ep - lp, ep - lp));
          }
#endif
        }
      }
      const uint32_t keep = __ballot_sync(kFull, worth);
      if (keep) {
        uint32_t at = 0;
        if (lane == 0) at = atomicAdd(&args.ctl->bdpt_n_evaluate, __popc(keep));
        at = __shfl_sync(kFull, at, 0) + __popc(keep & lane_lt);
        if (worth) {
          if (at < pool.evaluate_capacity) {
            pool.q_evaluate[at] = make_uint2(owner_item, static_cast<uint32_t>(s) | (static_cast<uint32_t>(t) << 8));
            if (Accel::kPersistent) {
              pool.evaluate_ray[2 * static_cast<size_t>(at)] = make_float4(ray_o.x, ray_o.y, ray_o.z, ray_reach);
              pool.evaluate_ray[2 * static_cast<size_t>(at) + 1] = make_float4(ray_d.x, ray_d.y, ray_d.z, (s > 0 && t > 0) ? 1.0f : 0.0f);
            }
          } else {
            bdpt_evaluate_candidate<Accel, 3>(args, pool, owner_item, s, t, count);  // queue full: in place
          }
        }
      }
    }
    __syncwarp();
  }
  const uint32_t casts = __reduce_add_sync(kFull, count.casts);
  const uint32_t splats = __reduce_add_sync(kFull, count.splats);
  if (lane == 0) {
    if (casts) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
    if (splats) atomicAdd(&args.ctl->splats, static_cast<unsigned long long>(splats));
  }
}

// BVH: the visibility rays of the candidates that passed the probe, through the warp-cooperative traversal.
struct BdptShadowSource {
  const RenderArgs& args;
  const BdptPool& pool;
  uint32_t n_queued;
  SphereBound spheres;
  uint32_t casts;
  __device__ __forceinline__ uint32_t count() const { return n_queued; }
  __device__ __forceinline__ bool load(uint32_t j, Ray& ray, float& t_best) {
    const float4 d = pool.evaluate_ray[2 * static_cast<size_t>(j) + 1];
    if (d.w == 0.0f) return false;  // s = 0 or t = 0: nothing to cast
    const float4 o = pool.evaluate_ray[2 * static_cast<size_t>(j)];
    ray.o = v3(o);
    ray.d = v3(d);
    // Connect only asks whether the FIRST object along the ray is the eye endpoint's: nothing beyond the endpoint
    // (o.w away, plus slack for its own rounding) can be that first object, so the walk need not go there
    t_best = fminf(spheres.bound(args.scene, ray), o.w * 1.001f + 1e-4f);
    casts++;
    return true;
  }
  __device__ __forceinline__ void store(uint32_t j, const HitRecord& hit, bool found) const {
    const HitRecord out = spheres.resolve(args.scene, hit, found);
    pool.evaluate_hit[j] = make_float2(out.t, __int_as_float(out.prim));
  }
};
__global__ void __launch_bounds__(kBlock, AMBER_BVH_MIN_BLOCKS) bdpt_trace_evaluate_kernel(const RenderArgs args, const BdptPool pool) {
  BdptShadowSource source{args, pool, min(args.ctl->bdpt_n_evaluate, pool.evaluate_capacity), {-1}, 0u};
  WideBvh::trace_persistent<false>(args.scene, source, &args.ctl->cursor_connect);
  const uint32_t casts = __reduce_add_sync(0xffffffffu, source.casts);
  if ((threadIdx.x & 31) == 0 && casts) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
}

// Connections, step 2: one thread per candidate that passed the probe -- the visibility cast, the MIS weight, the film.
template <class Accel>
__global__ void __launch_bounds__(kBdptConnectBlock, AMBER_BDPT_CONNECT_BLOCKS) bdpt_evaluate_kernel(const RenderArgs args, const BdptPool pool) {
  constexpr uint32_t kFull = 0xffffffffu;
  const uint32_t n_queued = min(args.ctl->bdpt_n_evaluate, pool.evaluate_capacity);
  BdptCounters count;
  count.casts = 0;
  count.splats = 0;
  for (uint32_t base = blockIdx.x * kBdptConnectBlock; base < n_queued; base += gridDim.x * kBdptConnectBlock) {
    const uint32_t j = base + threadIdx.x;
    if (j < n_queued) {
      const uint2 candidate = pool.q_evaluate[j];
      const int s = static_cast<int>(candidate.y & 0xffu), t = static_cast<int>(candidate.y >> 8);
      if (Accel::kPersistent && s > 0 && t > 0) {  // bdpt_trace_evaluate_kernel has cast the visibility ray
        const float2 found = pool.evaluate_hit[j];
        HitRecord hit;
        hit.t = found.x;
        hit.prim = __float_as_int(found.y);
        bdpt_evaluate_candidate<Accel, 2>(args, pool, candidate.x, s, t, count, &hit);
      } else {
        bdpt_evaluate_candidate<Accel, 2>(args, pool, candidate.x, s, t, count);
      }
    }
  }
  const uint32_t casts = __reduce_add_sync(kFull, count.casts);
  const uint32_t splats = __reduce_add_sync(kFull, count.splats);
  if ((threadIdx.x & 31u) == 0) {
    if (casts) atomicAdd(&args.ctl->closest_rays, static_cast<unsigned long long>(casts));
    if (splats) atomicAdd(&args.ctl->splats, static_cast<unsigned long long>(splats));
  }
}

}  // namespace amber_cuda
