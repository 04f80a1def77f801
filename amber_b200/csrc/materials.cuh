// materials.cuh -- BSDF sampling / evaluation and emission, one switch on the
// material tag instead of the reference's virtual Material<RGB> interface
// (include/amber/scene/material.h:31-88).  Each function cites what it restates.
#pragma once

#include "philox.cuh"
#include "scene_device.cuh"

namespace amber_cuda {

constexpr int kPhongMaxTries = 64;

struct Scatter {
  float3 dir;     // direction_in, used as the next ray direction as-is (ray.h:52-62)
  float3 weight;  // f cos / pdf (include/amber/rendering/scatter.h:45-48)
  bool capped;    // Phong rejection loop hit kPhongMaxTries
};

// PerfectReflection, include/amber/prelude/geometry.h:38-47: 2 (d.n) n - d
__device__ __forceinline__ float3 perfect_reflection(float3 incident, float3 normal, float signed_cos) {
  const float k = 2.0f * signed_cos;
  return v3(__fsub_rn(__fmul_rn(k, normal.x), incident.x), __fsub_rn(__fmul_rn(k, normal.y), incident.y),
            __fsub_rn(__fmul_rn(k, normal.z), incident.z));
}

// dir = u sin(theta) cos(phi) + v sin(theta) sin(phi) + w cos(theta), evaluated in the
// host's order ((u*sin)*cos + (v*sin)*sin) + w*cos without contraction
// (include/amber/prelude/sampling.h:248-252); only sincosf itself can differ from the
// host's libm, by an ulp.
__device__ __forceinline__ float3 local_to_world(float3 u, float3 v, float3 w, float sin_theta,
                                                 float cos_theta, float phi) {
  float s, c;
  sincosf(phi, &s, &c);
  const float3 a = scale_exact(c, scale_exact(sin_theta, u));
  const float3 b = scale_exact(s, scale_exact(sin_theta, v));
  return add_exact(add_exact(a, b), scale_exact(cos_theta, w));
}

// Schlick, src/amber/scene/material_refraction.cc:271-275.  The host evaluates
// r0 + (1 - r0) * pow(1 - cos, 5) in double (std::pow(float, int)) and rounds once.
__device__ __forceinline__ float schlick(float r0, float cos_theta) {
  const double x = static_cast<double>(1.0f - cos_theta);
  const double x2 = x * x;
  return static_cast<float>(static_cast<double>(r0) + static_cast<double>(1.0f - r0) * (x2 * x2 * x));
}

// Material::SampleLight for radiance transport.  `wo` = direction_out = -ray.d.
// `rng.next()` yields the BSDF-sampling uniforms in the reference's call order.
//
// Lambertian (material_lambertian.cc:61-70: HemispherePSA, sampling.h:234-265, cos = sqrt(r0),
// sin = sqrt(1-r0), phi = 2 pi r1, weight kd) and Phong (material_phong.cc:81-106: CosinePower,
// sampling.h:267-300, cos = r0^(1/(e+1)), rejected until on the side of wo, weight
// ks (e+2)/(e+1)|cos_i|) share one lobe-sampling block -- basis, compose, test -- so that a
// warp holding both kinds runs it once.  The reference rebuilds the basis on every Phong
// retry; it only depends on the mirror direction, so it is hoisted.  The retry loop is
// unbounded there; here it is capped (the cap is counted and never hit in practice).
//
// importance = true gives Material::SampleImportance, the light-tracing direction (row N2):
// Lambertian, Phong and Specular sit behind SymmetricBasicMaterialForwarder, whose
// SampleImportance IS Sample (include/amber/scene/material_basic.h:340-351), Eye passes through
// and DiffuseLight returns weight 0 either way; only Refraction changes
// (material_refraction.cc:222-263: no ior^2 scaling of the transmitted weight, probabilities
// (rho + 1/2) / 2 without the normalisation).
template <class Rng>
__device__ __forceinline__ Scatter sample_light(const MaterialDev& m, float3 n, float3 wo, Rng& rng,
                                                bool importance = false) {
  Scatter s;
  s.capped = false;
  s.dir = v3(0.0f, 0.0f, 0.0f);
  s.weight = v3(0.0f, 0.0f, 0.0f);
  const float3 rho = v3(m.r, m.g, m.b);
  const float cos_o = dot_exact(wo, n);
  float3 axis = n;
  bool lobe = false;
  const bool phong = m.kind == 1;
  switch (m.kind) {
    case 0:
      axis = cos_o > 0.0f ? n : -n;
      lobe = true;
      break;
    case 1:
      axis = perfect_reflection(wo, n, cos_o);
      lobe = true;
      break;
    case 2:  // Specular: src/amber/scene/material_specular.cc:62-70
      s.dir = perfect_reflection(wo, n, cos_o);
      s.weight = rho;
      break;
    case 3: {  // Refraction: src/amber/scene/material_refraction.cc:177-220
      const float ior = cos_o > 0.0f ? __fdiv_rn(1.0f, m.param) : m.param;
      const float ior2 = __fmul_rn(ior, ior);
      const float cos2_beta = __fsub_rn(1.0f, __fmul_rn(__fsub_rn(1.0f, __fmul_rn(cos_o, cos_o)), ior2));
      const float3 dir_r = perfect_reflection(wo, n, cos_o);
      if (cos2_beta < 0.0f) {  // total internal reflection
        s.dir = dir_r;
        s.weight = rho;
        break;
      }
      const float cos_alpha = fabsf(cos_o);
      const float cos_beta = __fsqrt_rn(cos2_beta);
      const float k = __fadd_rn(cos_o < 0.0f ? cos_beta : -cos_beta, __fmul_rn(ior, cos_o));
      const float3 dir_t = add_exact(scale_exact(-ior, wo), scale_exact(k, n));
      const float rho_r = schlick(m.r0, cos_alpha);
      const float rho_t = importance ? 1.0f - rho_r : __fmul_rn(1.0f - rho_r, ior2);
      const float total = rho_r + rho_t;
      const float p_r = __fdiv_rn((importance ? rho_r : __fdiv_rn(rho_r, total)) + 0.5f, 2.0f);
      const float p_t = __fdiv_rn((importance ? rho_t : __fdiv_rn(rho_t, total)) + 0.5f, 2.0f);
      if (rng.next() < p_r) {
        s.dir = dir_r;
        s.weight = __fdiv_rn(rho_r, p_r) * rho;
      } else {
        s.dir = dir_t;
        s.weight = __fdiv_rn(rho_t, p_t) * rho;
      }
      break;
    }
    case 5:  // Eye: include/amber/scene/material_eye.h:146-155, pass-through
      s.dir = -wo;
      s.weight = v3(1.0f, 1.0f, 1.0f);
      break;
    default:  // DiffuseLight: include/amber/scene/material_diffuse_light.h:185-194, weight 0
      break;
  }
  if (lobe) {
    float3 u, v;
    orthonormal_basis(axis, u, v);
    const float inv_e1 = __fdiv_rn(1.0f, m.param + 1.0f);
    float r0 = rng.next();
    float r1 = rng.next();
    for (int attempt = 1;; attempt++) {
      float cos_theta, sin_theta;
      if (phong) {
        cos_theta = powf(r0, inv_e1);
        sin_theta = __fsqrt_rn(__fsub_rn(1.0f, __fmul_rn(cos_theta, cos_theta)));
      } else {
        cos_theta = __fsqrt_rn(r0);
        sin_theta = __fsqrt_rn(1.0f - r0);
      }
      s.dir = local_to_world(u, v, axis, sin_theta, cos_theta, __fmul_rn(2.0f * kPiF, r1));
      if (!phong) {
        s.weight = rho;
        break;
      }
      const float cos_i = dot_exact(s.dir, n);
      if (cos_o * cos_i > 0.0f) {
        s.weight = ((m.param + 2.0f) / (m.param + 1.0f) * fabsf(cos_i)) * rho;
        break;
      }
      if (attempt >= kPhongMaxTries) {
        s.capped = true;
        break;
      }
      r0 = rng.next();
      r1 = rng.next();
    }
  }
  return s;
}

// x^y for a value that only scales a contribution (the Phong lobe in BSDF / pdf evaluation): exp2(y log2 x)
// on the MUFU unit, ~6 instructions against ~150 for powf; relative error ~4e-5 at y = 256.  Sampled
// DIRECTIONS keep powf (sample_light): they decide where the path goes.
__device__ __forceinline__ float pow_scale_only(float x, float y) {
  return x > 0.0f ? exp2f(y * __log2f(x)) : (y == 0.0f ? 1.0f : 0.0f);
}

// Material::BSDF for the connectable (SurfaceType::Diffuse) materials, w.r.t.
// projected solid angle.  Lambertian: material_lambertian.cc:34-49;
// Phong: material_phong.cc:39-61.  Others return 0 (never connected,
// include/amber/rendering/bidirectional_path_sampling.h:944-952).
__device__ __forceinline__ float3 bsdf_eval(const MaterialDev& m, float3 n, float3 wo, float3 wi) {
  const float cos_o = dot(wo, n);
  const float cos_i = dot(wi, n);
  const float3 rho = v3(m.r, m.g, m.b);
  if (cos_o * cos_i <= 0.0f) return v3(0.0f, 0.0f, 0.0f);
  if (m.kind == 0) return (1.0f / kPiF) * rho;
  if (m.kind == 1) {
    const float3 refl = perfect_reflection(wi, n, cos_i);
    const float cos_alpha = fmaxf(0.0f, dot(refl, wo));
    return ((m.param + 2.0f) / (2.0f * kPiF) * pow_scale_only(cos_alpha, m.param)) * rho;
  }
  return v3(0.0f, 0.0f, 0.0f);
}

// DiffuseLight::Radiance, material_diffuse_light.h:127-139: one-sided.
__device__ __forceinline__ float3 emitted_radiance(const MaterialDev& m, float3 n, float3 wo) {
  if (m.kind != 4 || !(dot(wo, n) > 0.0f)) return v3(0.0f, 0.0f, 0.0f);
  return v3(m.r, m.g, m.b);
}

}  // namespace amber_cuda
