/*
 * amber_oracle.c -- TEST INFRASTRUCTURE.  Not part of the product.
 *
 * CPU restatement of the reference's path-tracing hot path, in plain C, float for
 * float in the reference's operation order (compile with -ffp-contract=off: the
 * reference is built without FMA, CMakeLists.txt:5).  Citations are to files under
 * /root/reference.  See amber_oracle.h for how it is pinned.
 */
#include "amber_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#include "philox_stream.h"

/* include/amber/constants.h:25-28 */
static const long double kPI = 3.141592653589793238462643383279503L;
static const long double kEPS = 1e-6L;
static const long double kRussianRoulette = 0.9375L;
#define PI_F ((float)kPI)

typedef struct { float x, y, z; } vec3;

/* ---- include/amber/prelude/vector3.h:170-342 ------------------------------------- */
static vec3 V(float x, float y, float z) { vec3 v = {x, y, z}; return v; }
static vec3 add(vec3 a, vec3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static vec3 sub(vec3 a, vec3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static vec3 mul(vec3 a, vec3 b) { return V(a.x * b.x, a.y * b.y, a.z * b.z); }
static vec3 scale(float s, vec3 a) { return V(s * a.x, s * a.y, s * a.z); } /* Vector3(s) * a */
static vec3 neg(vec3 a) { return V(-a.x, -a.y, -a.z); }
static float dot(vec3 u, vec3 v) { return u.x * v.x + u.y * v.y + u.z * v.z; }
static float length(vec3 v) { return sqrtf(dot(v, v)); }
static vec3 normalize(vec3 v) { const float l = length(v); return V(v.x / l, v.y / l, v.z / l); }
static vec3 cross(vec3 u, vec3 v) {
  return V(u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x);
}
static float max3(vec3 v) { return fmaxf(v.x, fmaxf(v.y, v.z)); }
static void orthonormal_basis(vec3 w, vec3* u, vec3* v) { /* vector3.h:330-342 */
  *u = normalize(cross(w, fabsf(w.x) < fabsf(w.y) ? V(1, 0, 0) : V(0, 1, 0)));
  *v = normalize(cross(w, *u));
}

/* ---- uniform source: the Philox stream, or a scripted list ------------------------ */
typedef struct {
  philox_stream* stream; /* if non-NULL */
  const float* xi;
  int n, used;
  int camera; /* stream only: 1 while the camera vertex is being sampled */
} sampler;
static float scripted(sampler* s) {
  /* past the scripted list: a Weyl sequence, identical in the harness and the kernels */
  const float v = s->used < s->n ? s->xi[s->used]
                                 : (float)(((uint32_t)s->used * 2654435769u) >> 8) * (1.0f / 16777216.0f);
  s->used++;
  return v;
}
/* Uniform<float>: static_cast<float>(sampler()), sampling.h:156-161.  On the stream the call
 * is a camera uniform or the next BSDF-sampling uniform of the current vertex. */
static float uniform(sampler* s) {
  if (!s->stream) return scripted(s);
  return s->camera ? philox_stream_next_camera(s->stream) : philox_stream_next_bsdf(s->stream);
}
static float uniform_rr(sampler* s) { return s->stream ? philox_stream_rr(s->stream) : scripted(s); }
static float uniform_light(sampler* s, uint32_t i) {
  return s->stream ? philox_stream_light(s->stream, i) : scripted(s);
}

/* ---- scene ----------------------------------------------------------------------- */
typedef struct {
  int kind; /* AMBER_CUDA_PRIM_* */
  int material;
  vec3 v0, v1, v2, normal; /* triangle; normal = Normalize(Cross(v1-v0, v2-v0)), primitive_triangle.cc:59-68 */
  vec3 center;             /* sphere */
  float radius;
} prim;

typedef struct {
  int prim_index;
  float cumulative_power;
  float pdf_area;
} light;

struct oracle_scene {
  uint32_t n_prims, n_materials, n_lights;
  prim* prims;
  amber_cuda_material* materials;
  float* r0; /* Refraction: Fresnel(ior) */
  light* lights;
  float total_power;
  /* BasicThin, src/amber/scene/lens_thin.cc:32-57 */
  vec3 lens_origin;
  float global[9], local[9];
  float focus_distance, sensor_distance, p_area;
  int lens_kind; /* AMBER_CUDA_LENS_*; PINHOLE: BasicPinhole, src/amber/scene/lens_pinhole.cc:31-42 */
  int n_blades;
  vec3 blades[16][4]; /* v0 v1 v2 normal */
};

static vec3 mat3_apply(const float* m, vec3 v) { /* include/amber/prelude/matrix3.h:83-90 */
  return V(m[0] * v.x + m[1] * v.y + m[2] * v.z, m[3] * v.x + m[4] * v.y + m[5] * v.z,
           m[6] * v.x + m[7] * v.y + m[8] * v.z);
}

static float triangle_area(vec3 v0, vec3 v1, vec3 v2) { /* primitive_triangle.cc:130-134 */
  return length(cross(sub(v1, v0), sub(v2, v0))) / 2;
}

oracle_scene* oracle_scene_create(const amber_cuda_scene_desc* desc) {
  oracle_scene* s = (oracle_scene*)calloc(1, sizeof *s);
  uint32_t i;
  s->n_prims = desc->n_primitives;
  s->n_materials = desc->n_materials;
  s->prims = (prim*)calloc(desc->n_primitives ? desc->n_primitives : 1, sizeof(prim));
  s->materials = (amber_cuda_material*)calloc(desc->n_materials, sizeof(amber_cuda_material));
  s->r0 = (float*)calloc(desc->n_materials, sizeof(float));
  memcpy(s->materials, desc->materials, desc->n_materials * sizeof(amber_cuda_material));
  for (i = 0; i < desc->n_materials; i++) {
    amber_cuda_material* m = &s->materials[i];
    if (m->kind == AMBER_CUDA_MAT_REFRACTION || m->kind == AMBER_CUDA_MAT_EYE) {
      m->rgb[0] = m->rgb[1] = m->rgb[2] = 1.0f; /* Radiant(1), material_refraction.h:94-101 */
    }
    if (m->kind == AMBER_CUDA_MAT_REFRACTION) {
      /* BasicRefraction::Fresnel, material_refraction.cc:265-269: std::pow(float, int) -> double */
      s->r0[i] = (float)pow((double)((m->param - 1) / (m->param + 1)), 2.0);
    }
  }
  for (i = 0; i < desc->n_primitives; i++) {
    const amber_cuda_primitive* in = &desc->primitives[i];
    prim* p = &s->prims[i];
    p->kind = in->kind;
    p->material = in->material;
    if (in->kind == AMBER_CUDA_PRIM_TRIANGLE) {
      p->v0 = V(in->data[0], in->data[1], in->data[2]);
      p->v1 = V(in->data[3], in->data[4], in->data[5]);
      p->v2 = V(in->data[6], in->data[7], in->data[8]);
      p->normal = normalize(cross(sub(p->v1, p->v0), sub(p->v2, p->v0)));
    } else {
      p->center = V(in->data[0], in->data[1], in->data[2]);
      p->radius = in->data[3];
    }
  }
  /* emitting objects sorted by power, LightSet::LightSet (include/amber/scene/light_set.h:61-81);
   * Power = Sum(SurfaceArea * Irradiance) (include/amber/scene/object.h:98-103),
   * Irradiance = radiance * pi (material_diffuse_light.h:119-124); SurfaceArea: triangle
   * primitive_triangle.cc:130-134, sphere 4 pi r^2 (primitive_sphere.cc:109-113) */
  {
    float (*tmp)[3] = (float (*)[3])calloc(desc->n_primitives ? desc->n_primitives : 1, sizeof *tmp);
    uint32_t n = 0;
    float cumulative = 0;
    for (i = 0; i < desc->n_primitives; i++) {
      const prim* p = &s->prims[i];
      const amber_cuda_material* m = &s->materials[p->material];
      if (m->kind != AMBER_CUDA_MAT_DIFFUSE_LIGHT) continue;
      {
        const float area = p->kind == AMBER_CUDA_PRIM_TRIANGLE ? triangle_area(p->v0, p->v1, p->v2)
                                                               : 4 * PI_F * p->radius * p->radius;
        const float ir = m->rgb[0] * PI_F, ig = m->rgb[1] * PI_F, ib = m->rgb[2] * PI_F;
        const float power = area * ir + area * ig + area * ib;
        if (!(power > 0)) continue;
        tmp[n][0] = power;
        tmp[n][1] = (float)i;
        tmp[n][2] = ir + ig + ib;
        n++;
      }
    }
    /* ascending power; equal powers keep creation order */
    {
      uint32_t a, b;
      for (a = 1; a < n; a++) { /* insertion sort is stable */
        float key[3];
        memcpy(key, tmp[a], sizeof key);
        b = a;
        while (b > 0 && tmp[b - 1][0] > key[0]) {
          memcpy(tmp[b], tmp[b - 1], sizeof key);
          b--;
        }
        memcpy(tmp[b], key, sizeof key);
      }
    }
    s->n_lights = n;
    s->lights = (light*)calloc(n ? n : 1, sizeof(light));
    s->total_power = 0;
    for (i = 0; i < n; i++) s->total_power += tmp[i][0];
    for (i = 0; i < n; i++) {
      cumulative += tmp[i][0];
      s->lights[i].prim_index = (int)tmp[i][1];
      s->lights[i].cumulative_power = cumulative;
      s->lights[i].pdf_area = tmp[i][2] / s->total_power; /* LightSet::PDFArea, light_set.h:106-111 */
    }
    free(tmp);
  }
  /* lens */
  {
    const amber_cuda_lens* in = &desc->lens;
    const float* t = in->transform;
    const float e11 = t[0], e12 = t[1], e13 = t[2], e21 = t[4], e22 = t[5], e23 = t[6], e31 = t[8],
                e32 = t[9], e33 = t[10];
    const float g[9] = {e11, e12, e13, e21, e22, e23, e31, e32, e33};
    /* Matrix3::Determinant / Inverse, include/amber/prelude/matrix3.h:110-145 */
    const float det = e11 * (e22 * e33 - e23 * e32) - e21 * (e12 * e33 - e13 * e32) + e31 * (e12 * e23 - e13 * e22);
    const float dinv = 1 / det;
    const float inv[9] = {+dinv * (e22 * e33 - e23 * e32), -dinv * (e12 * e33 - e13 * e32), +dinv * (e12 * e23 - e13 * e22),
                          -dinv * (e21 * e33 - e23 * e31), +dinv * (e11 * e33 - e13 * e31), -dinv * (e11 * e23 - e13 * e21),
                          +dinv * (e21 * e32 - e22 * e31), -dinv * (e11 * e32 - e12 * e31), +dinv * (e11 * e22 - e12 * e21)};
    int b;
    memcpy(s->global, g, sizeof g);
    memcpy(s->local, inv, sizeof inv);
    s->lens_origin = V(t[3], t[7], t[11]); /* transform(Vector3()), matrix4.h:100-106 */
    s->lens_kind = in->kind;
    if (in->kind == AMBER_CUDA_LENS_PINHOLE) {
      s->sensor_distance = in->sensor_distance;
      s->n_blades = 0;
      s->p_area = 1; /* kDiracDelta */
      return s;
    }
    s->focus_distance = in->focus_distance;
    s->sensor_distance = 1 / (1 / in->focal_length - 1 / in->focus_distance);
    s->n_blades = in->n_blades;
    for (b = 0; b < in->n_blades && b < 16; b++) {
      const float nb = (float)in->n_blades; /* size_t -> float in `2*pi / n_blades * i` */
      const float alpha = 2 * PI_F / nb * (float)b;
      const float beta = 2 * PI_F / nb * (float)(b + 1);
      s->blades[b][0] = add(s->lens_origin, mat3_apply(g, scale(in->radius, V(cosf(alpha), sinf(alpha), 0))));
      s->blades[b][1] = add(s->lens_origin, mat3_apply(g, scale(in->radius, V(cosf(beta), sinf(beta), 0))));
      s->blades[b][2] = s->lens_origin;
      s->blades[b][3] = normalize(cross(sub(s->blades[b][1], s->blades[b][0]), sub(s->blades[b][2], s->blades[b][0])));
    }
    s->p_area = 1;
    s->p_area /= triangle_area(s->blades[0][0], s->blades[0][1], s->blades[0][2]) * (float)in->n_blades;
  }
  return s;
}

void oracle_scene_destroy(oracle_scene* s) {
  if (!s) return;
  free(s->prims);
  free(s->materials);
  free(s->r0);
  free(s->lights);
  free(s);
}

/* ---- intersection ------------------------------------------------------------------ */
typedef struct {
  float t; /* NaN on a miss, include/amber/prelude/hit.h:62-66 */
  vec3 position, normal;
} hit_t;

/* Triangle::Intersect, src/amber/scene/primitive_triangle.cc:97-128 */
static int triangle_intersect(const prim* p, vec3 o, vec3 d, hit_t* hit) {
  const vec3 E1 = sub(p->v1, p->v0);
  const vec3 E2 = sub(p->v2, p->v0);
  const vec3 P = cross(d, E2);
  const float det = dot(P, E1);
  const vec3 T = sub(o, p->v0);
  const float u = dot(P, T) / det;
  vec3 Q;
  float v, t;
  if (u > 1 || u < 0) return 0;
  Q = cross(T, E1);
  v = dot(Q, d) / det;
  if (v > 1 || v < 0) return 0;
  if (u + v > 1) return 0;
  t = dot(Q, E2) / det;
  if ((long double)t < kEPS) return 0;
  if (!isfinite(t)) return 0; /* Hit::operator bool, hit.h:90-94 */
  hit->t = t;
  hit->position = add(add(p->v0, scale(u, E1)), scale(v, E2));
  hit->normal = p->normal;
  return 1;
}

/* Sphere::Intersect, src/amber/scene/primitive_sphere.cc:75-107 +
 * SolveQuadratic, include/amber/prelude/algebra.h:31-52 */
static int sphere_intersect(const prim* p, vec3 o, vec3 d, hit_t* hit) {
  const float a = 1;
  const float b = -2 * dot(sub(p->center, o), d);
  const vec3 oc = sub(p->center, o);
  const float c = dot(oc, oc) - p->radius * p->radius;
  const float disc = b * b - 4 * a * c;
  float sqrt_d, alpha, beta, t;
  if (disc < 0) return 0;
  sqrt_d = sqrtf(disc);
  alpha = -b - sqrt_d;
  beta = -b + sqrt_d;
  if (fabsf(alpha) < fabsf(beta)) {
    alpha = c / beta * 2;
    beta /= 2 * a;
  } else {
    beta = c / alpha * 2;
    alpha /= 2 * a;
  }
  if ((long double)alpha > kEPS) t = alpha;
  else if ((long double)beta > kEPS) t = beta;
  else return 0;
  if (!isfinite(t)) return 0;
  hit->t = t;
  hit->position = add(o, scale(t, d));
  hit->normal = normalize(sub(hit->position, p->center)); /* Hit ctor normalises, hit.h:68-77 */
  return 1;
}

/* raytracer::List::Cast / the leaf loop of BVH::Node::Cast
 * (include/amber/raytracer/acceleration_list.h:51-68, acceleration_bvh.h:343-356):
 * closest hit, strict `<`, starting from numeric_limits<float>::max (acceleration.h:46-51). */
static int scene_cast(const oracle_scene* s, vec3 o, vec3 d, hit_t* closest) {
  float distance = 3.402823466e+38f;
  int object = -1;
  uint32_t i;
  for (i = 0; i < s->n_prims; i++) {
    hit_t h;
    const prim* p = &s->prims[i];
    const int ok = p->kind == AMBER_CUDA_PRIM_TRIANGLE ? triangle_intersect(p, o, d, &h)
                                                       : sphere_intersect(p, o, d, &h);
    if (ok && h.t < distance) {
      distance = h.t;
      *closest = h;
      object = (int)i;
    }
  }
  if (object < 0) closest->t = NAN;
  return object;
}

/* ---- sampling routines, include/amber/prelude/sampling.h ---------------------------- */
static vec3 local_dir(vec3 u, vec3 v, vec3 w, float sin_theta, float cos_theta, float phi) {
  /* u * sin_theta * cos(phi) + v * sin_theta * sin(phi) + w * cos_theta, sampling.h:248-252 */
  const vec3 a = scale(cosf(phi), scale(sin_theta, u));
  const vec3 b = scale(sinf(phi), scale(sin_theta, v));
  const vec3 c = scale(cos_theta, w);
  return add(add(a, b), c);
}

static vec3 hemisphere_psa(vec3 w, sampler* s) { /* sampling.h:234-265 */
  vec3 u, v;
  float r0, r1;
  orthonormal_basis(w, &u, &v);
  r0 = uniform(s);
  r1 = uniform(s);
  return local_dir(u, v, w, sqrtf(1 - r0), sqrtf(r0), 2 * PI_F * r1);
}

static vec3 cosine_power(vec3 w, float exponent, sampler* s) { /* sampling.h:267-300 */
  vec3 u, v;
  float r0, r1, cos_theta, sin_theta;
  orthonormal_basis(w, &u, &v);
  r0 = uniform(s);
  r1 = uniform(s);
  cos_theta = powf(r0, 1 / (exponent + 1));
  sin_theta = sqrtf(1 - cos_theta * cos_theta);
  return local_dir(u, v, w, sin_theta, cos_theta, 2 * PI_F * r1);
}

/* PerfectReflection, include/amber/prelude/geometry.h:38-47 */
static vec3 perfect_reflection(vec3 incident, vec3 normal, float signed_cos) {
  return sub(scale(2 * signed_cos, normal), incident);
}

/* BasicRefraction::Schlick, material_refraction.cc:271-275: std::pow(float, int) is double */
static float schlick(float r0, float cos_theta) {
  return (float)((double)r0 + (double)(1 - r0) * pow((double)(1 - cos_theta), 5.0));
}

typedef struct {
  vec3 dir, weight;
} scatter_t;

/* Material<RGB>::SampleLight through the Basic*Forwarder (material_basic.h:233-245,327-338):
 * weight = scalar weight * rho. */
static scatter_t material_sample_light(const amber_cuda_material* m, float r0, vec3 n, vec3 wo, sampler* s) {
  scatter_t out;
  const vec3 rho = V(m->rgb[0], m->rgb[1], m->rgb[2]);
  float scalar = 0;
  out.dir = V(0, 0, 0);
  switch (m->kind) {
    case AMBER_CUDA_MAT_LAMBERTIAN: { /* material_lambertian.cc:61-70 */
      const vec3 w = dot(wo, n) > 0 ? n : neg(n);
      out.dir = hemisphere_psa(w, s);
      scalar = 1;
      break;
    }
    case AMBER_CUDA_MAT_PHONG: { /* material_phong.cc:81-106 */
      const vec3 refl = perfect_reflection(wo, n, dot(wo, n));
      const float cos_o = dot(wo, n);
      for (;;) {
        const vec3 dir = cosine_power(refl, m->param, s);
        const float cos_i = dot(dir, n);
        if (cos_o * cos_i <= 0) continue;
        out.dir = dir;
        scalar = (m->param + 2) / (m->param + 1) * fabsf(cos_i);
        break;
      }
      break;
    }
    case AMBER_CUDA_MAT_SPECULAR: /* material_specular.cc:62-70 */
      out.dir = perfect_reflection(wo, n, dot(wo, n));
      scalar = 1;
      break;
    case AMBER_CUDA_MAT_REFRACTION: { /* material_refraction.cc:177-220 */
      const float signed_cos_alpha = dot(wo, n);
      const float ior = signed_cos_alpha > 0 ? 1 / m->param : m->param;
      const float squared_cos_beta = 1 - (1 - signed_cos_alpha * signed_cos_alpha) * (ior * ior);
      const vec3 dir_r = perfect_reflection(wo, n, signed_cos_alpha);
      if (squared_cos_beta < 0) {
        out.dir = dir_r;
        scalar = 1;
        break;
      }
      {
        const float cos_alpha = fabsf(signed_cos_alpha);
        const float cos_beta = sqrtf(squared_cos_beta);
        const vec3 dir_t = add(scale(-ior, wo),
                               scale((signed_cos_alpha < 0 ? 1 : -1) * cos_beta + ior * signed_cos_alpha, n));
        const float rho_r = schlick(r0, cos_alpha);
        const float rho_t = (1 - rho_r) * (ior * ior);
        const float rho_sum = rho_r + rho_t;
        const float p_r = (rho_r / rho_sum + 0.5f) / 2;
        const float p_t = (rho_t / rho_sum + 0.5f) / 2;
        if (uniform(s) < p_r) {
          out.dir = dir_r;
          scalar = rho_r / p_r;
        } else {
          out.dir = dir_t;
          scalar = rho_t / p_t;
        }
      }
      break;
    }
    case AMBER_CUDA_MAT_EYE: /* material_eye.h:146-155 */
      out.dir = neg(wo);
      out.weight = V(1, 1, 1);
      return out;
    default: /* DiffuseLight::SampleLight returns Scatter(): weight 0, material_diffuse_light.h:185-194 */
      out.weight = V(0, 0, 0);
      return out;
  }
  out.weight = scale(scalar, rho); /* scatter.weight * rho_ */
  return out;
}

/* Material<RGB>::SampleImportance (light tracing, row N2).  Lambertian / Phong / Specular sit
 * behind SymmetricBasicMaterialForwarder, whose SampleImportance IS Sample
 * (material_basic.h:340-351); Eye passes through (material_eye.h:157-166); DiffuseLight returns
 * Scatter() (material_diffuse_light.h:196-205).  Only Refraction differs
 * (material_refraction.cc:222-263): no ior^2 radiance scaling of the transmitted weight and
 * sampling probabilities (rho + 1/2) / 2 without the normalisation by rho_r + rho_t. */
static scatter_t material_sample_importance(const amber_cuda_material* m, float r0, vec3 n, vec3 wo, sampler* s) {
  scatter_t out;
  if (m->kind != AMBER_CUDA_MAT_REFRACTION) return material_sample_light(m, r0, n, wo, s);
  {
    const vec3 rho = V(m->rgb[0], m->rgb[1], m->rgb[2]);
    const float signed_cos_alpha = dot(wo, n);
    const float ior = signed_cos_alpha > 0 ? 1 / m->param : m->param;
    const float squared_cos_beta = 1 - (1 - signed_cos_alpha * signed_cos_alpha) * (ior * ior);
    const vec3 dir_r = perfect_reflection(wo, n, signed_cos_alpha);
    float scalar;
    if (squared_cos_beta < 0) {
      out.dir = dir_r;
      scalar = 1;
    } else {
      const float cos_alpha = fabsf(signed_cos_alpha);
      const float cos_beta = sqrtf(squared_cos_beta);
      const vec3 dir_t = add(scale(-ior, wo),
                             scale((signed_cos_alpha < 0 ? 1 : -1) * cos_beta + ior * signed_cos_alpha, n));
      const float rho_r = schlick(r0, cos_alpha);
      const float rho_t = 1 - rho_r;
      const float p_r = (rho_r + 0.5f) / 2;
      const float p_t = (rho_t + 0.5f) / 2;
      if (uniform(s) < p_r) {
        out.dir = dir_r;
        scalar = rho_r / p_r;
      } else {
        out.dir = dir_t;
        scalar = rho_t / p_t;
      }
    }
    out.weight = scale(scalar, rho);
  }
  return out;
}

/* Material<RGB>::BSDF: Lambertian material_lambertian.cc:34-49, Phong material_phong.cc:39-61,
 * Specular material_specular.cc:34-49 (kDiracDelta = 1), others 0 / not connectable. */
static vec3 material_bsdf(const amber_cuda_material* m, vec3 n, vec3 wo, vec3 wi) {
  const vec3 rho = V(m->rgb[0], m->rgb[1], m->rgb[2]);
  const float cos_o = dot(wo, n), cos_i = dot(wi, n);
  if (m->kind == AMBER_CUDA_MAT_LAMBERTIAN) {
    if (cos_o * cos_i <= 0) return V(0, 0, 0);
    return scale(1 / PI_F, rho);
  }
  if (m->kind == AMBER_CUDA_MAT_PHONG) {
    vec3 refl;
    float cos_alpha;
    if (cos_o * cos_i <= 0) return V(0, 0, 0);
    refl = perfect_reflection(wi, n, cos_i);
    cos_alpha = fmaxf(0, dot(refl, wo));
    return scale((m->param + 2) / (2 * PI_F) * powf(cos_alpha, m->param), rho);
  }
  return V(0, 0, 0);
}

/* DiffuseLight::Radiance, material_diffuse_light.h:127-139 (0 for every other material,
 * include/amber/scene/material.h:101-109) */
static vec3 material_radiance(const amber_cuda_material* m, vec3 n, vec3 wo) {
  if (m->kind != AMBER_CUDA_MAT_DIFFUSE_LIGHT) return V(0, 0, 0);
  if (dot(wo, n) <= 0) return V(0, 0, 0);
  return V(m->rgb[0], m->rgb[1], m->rgb[2]);
}

/* Triangle::SampleSurfacePoint, primitive_triangle.cc:136-150 */
static vec3 triangle_point(vec3 v0, vec3 v1, vec3 v2, float u, float v) {
  if (u + v >= 1) {
    u = 1 - u;
    v = 1 - v;
  }
  return add(add(scale(1 - u - v, v0), scale(u, v1)), scale(v, v2));
}
static vec3 triangle_sample(vec3 v0, vec3 v1, vec3 v2, sampler* s) {
  const float u = uniform(s);
  const float v = uniform(s);
  return triangle_point(v0, v1, v2, u, v);
}

/* PixelBound::Uniform (src/amber/rendering/sensor.cc:94-101,111-120) +
 * BasicThin::GenerateRay / PDFArea / PDFDirection (src/amber/scene/lens_thin.cc:70-107,132-148) */
static void generate_eye_ray(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                             uint32_t px, uint32_t py, sampler* s, vec3* origin, vec3* direction,
                             float* weight) {
  const float nb = (float)sc->n_blades;
  float f;
  size_t pos;
  vec3 aperture_origin, aperture_point, sensor_point, dir_local, dir_world, dir_back;
  float su, sv;
  double factor, pdf_direction_d;
  float pdf_direction;
  if (sc->lens_kind == AMBER_CUDA_LENS_PINHOLE) {
    /* BasicPinhole::GenerateRay / PDFArea / PDFDirection, src/amber/scene/lens_pinhole.cc:48-69,
     * 87-108: ray through the lens origin, weight = 1 / kDiracDelta / (p_area / geometry factor) */
    vec3 point;
    float geometry_factor;
    sv = ((float)py + uniform(s)) / (float)h; /* Y first, as below */
    su = ((float)px + uniform(s)) / (float)w;
    sensor_point = V((su - 0.5f) * film_w, (sv - 0.5f) * film_h, sc->sensor_distance);
    dir_world = normalize(mat3_apply(sc->global, neg(sensor_point)));
    dir_back = mat3_apply(sc->local, dir_world);
    point = scale(sc->sensor_distance / dir_back.z, dir_back);
    geometry_factor = dir_back.z * dir_back.z / dot(point, point);
    *origin = sc->lens_origin;
    *direction = dir_world;
    *weight = 1 / 1.0f / ((1 / (film_w * film_h)) / geometry_factor);
    return;
  }
  f = floorf(uniform(s) * nb);
  pos = (size_t)f;
  if (pos > (size_t)sc->n_blades - 1) pos = (size_t)sc->n_blades - 1;
  aperture_origin = triangle_sample(sc->blades[pos][0], sc->blades[pos][1], sc->blades[pos][2], s);
  aperture_point = mat3_apply(sc->local, sub(aperture_origin, sc->lens_origin));
  /* Vector2((x + xi) / W, (y + xi) / H), sensor.cc:114-117: the order in which the two
   * constructor arguments are evaluated is unspecified in C++; g++ (the compiler that
   * builds oracle/_ref) evaluates them right to left, so the Y uniform is drawn first. */
  sv = ((float)py + uniform(s)) / (float)h;
  su = ((float)px + uniform(s)) / (float)w;
  sensor_point = V((su - 0.5f) * film_w, (sv - 0.5f) * film_h, sc->sensor_distance);
  dir_local = normalize(sub(scale(-sc->focus_distance / sc->sensor_distance, sensor_point), aperture_point));
  factor = pow((double)(normalize(sub(sensor_point, aperture_point)).z / dir_local.z), 4.0);
  dir_world = normalize(mat3_apply(sc->global, dir_local)); /* Ray(origin, Vector3) normalises, ray.h:52-56 */
  dir_back = mat3_apply(sc->local, dir_world);
  /* sensor.Size() / sensor.SceneArea() * pow(sd, 2) / pow(dz, 4): float, then double */
  pdf_direction_d = (double)((float)((unsigned long)w * (unsigned long)h) / (film_w * film_h)) *
                    pow((double)sc->sensor_distance, 2.0) / pow((double)dir_back.z, 4.0);
  pdf_direction = (float)pdf_direction_d;
  *origin = aperture_origin;
  *direction = dir_world;
  *weight = (float)(factor / sc->p_area / pdf_direction);
}

/* BasicThin::Response (src/amber/scene/lens_thin.cc:109-130) + Sensor::ResponsePixel / UVToPixel
 * (src/amber/rendering/sensor.cc:52-65,122-129): the pixel a ray ARRIVING at the aperture point
 * `position` from direction `dir_out` (pointing away from the lens) lands on, and the factor
 * (cos at the sensor / cos at the lens)^4.  Returns 0 when the ray comes from behind the lens or
 * misses the film. */
static int lens_response(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                         vec3 position, vec3 dir_out, uint32_t* px, uint32_t* py, float* value) {
  const vec3 direction = mat3_apply(sc->local, dir_out);
  vec3 aperture_point, sensor_point;
  float u, v;
  unsigned long ix, iy;
  if (sc->lens_kind == AMBER_CUDA_LENS_PINHOLE) {
    /* BasicPinhole::Response, lens_pinhole.cc:71-85: no facing test, value 1 */
    aperture_point = V(0, 0, 0);
    sensor_point = scale(sc->sensor_distance / direction.z, direction);
  } else {
    if (direction.z >= 0) return 0;
    aperture_point = mat3_apply(sc->local, sub(position, sc->lens_origin));
    sensor_point = add(scale(-sc->sensor_distance / sc->focus_distance, aperture_point),
                       scale(sc->sensor_distance / direction.z, direction));
  }
  u = sensor_point.x / film_w + 0.5f;
  v = sensor_point.y / film_h + 0.5f;
  if (fminf(u, v) < 0 || fmaxf(u, v) >= 1) return 0;
  /* std::min<pixel_size_type>(W - 1, uv.X() * W): float product, truncated to unsigned */
  ix = (unsigned long)(u * (float)(unsigned long)w);
  iy = (unsigned long)(v * (float)(unsigned long)h);
  *px = ix < (unsigned long)w - 1 ? (uint32_t)ix : w - 1;
  *py = iy < (unsigned long)h - 1 ? (uint32_t)iy : h - 1;
  if (sc->lens_kind == AMBER_CUDA_LENS_PINHOLE) {
    *value = 1;
    return 1;
  }
  /* std::pow(float, int) is double; PixelValue<real_type> stores float */
  *value = (float)pow((double)(normalize(sub(sensor_point, aperture_point)).z / direction.z), 4.0);
  return 1;
}

/* Object::SampleSurfacePoint of a light: Triangle::SampleSurfacePoint (primitive_triangle.cc:136-150,
 * the stored normal) or Sphere::SampleSurfacePoint (primitive_sphere.cc:115-122): normal = SphereSA
 * (include/amber/prelude/sampling.h:185-199: cos_theta = Uniform(-1, 1), sin_theta = sqrt(1 - cos^2),
 * (cos_theta cos phi, cos_theta sin phi, sin_theta) -- as written it only reaches the z >= 0 half of the
 * sphere, and not uniformly; restated as it is), origin = center + radius * normal.  Two uniforms. */
static vec3 light_surface_point(const prim* lp, sampler* s, vec3* normal) {
  if (lp->kind == AMBER_CUDA_PRIM_TRIANGLE) {
    *normal = lp->normal;
    return triangle_sample(lp->v0, lp->v1, lp->v2, s);
  } else {
    const float r0 = uniform(s) * (1.0f - -1.0f) + -1.0f; /* Uniform(min, max): sampler() * (max - min) + min */
    const float r1 = uniform(s);
    const float cos_theta = r0;
    const float sin_theta = sqrtf(1 - r0 * r0);
    const float phi = 2 * PI_F * r1;
    *normal = V(cos_theta * cosf(phi), cos_theta * sinf(phi), sin_theta);
    return add(lp->center, scale(lp->radius, *normal));
  }
}

/* LightSet::GenerateRay (include/amber/scene/light_set.h:83-104): light by lower_bound over the
 * cumulative powers of Uniform(total), point by Object::SampleSurfacePoint, direction by
 * HemispherePSA about the surface normal there, weight = Irradiance / PDFArea. */
static void generate_light_ray(const oracle_scene* sc, sampler* s, vec3* origin, vec3* direction, vec3* weight) {
  const float target = uniform(s) * sc->total_power;
  uint32_t li = 0;
  const light* l;
  const prim* lp;
  const amber_cuda_material* lm;
  while (li + 1 < sc->n_lights && sc->lights[li].cumulative_power < target) li++;
  l = &sc->lights[li];
  lp = &sc->prims[l->prim_index];
  lm = &sc->materials[lp->material];
  {
    vec3 normal;
    *origin = light_surface_point(lp, s, &normal);
    *direction = hemisphere_psa(normal, s);
  }
  *weight = V(lm->rgb[0] * PI_F / l->pdf_area, lm->rgb[1] * PI_F / l->pdf_area, lm->rgb[2] * PI_F / l->pdf_area);
}

/* ---- exported unit entry points --------------------------------------------------- */
typedef struct {
  const oracle_scene* scene;
  uint64_t n;
  const float *origins, *directions;
  int32_t* prim;
  float* t;
  int worker, workers;
} cast_job;

static void* cast_worker(void* arg) {
  cast_job* job = (cast_job*)arg;
  uint64_t i;
  for (i = (uint64_t)job->worker; i < job->n; i += (uint64_t)job->workers) {
    hit_t h;
    const vec3 o = V(job->origins[3 * i], job->origins[3 * i + 1], job->origins[3 * i + 2]);
    const vec3 d = V(job->directions[3 * i], job->directions[3 * i + 1], job->directions[3 * i + 2]);
    job->prim[i] = scene_cast(job->scene, o, d, &h);
    job->t[i] = h.t;
  }
  return NULL;
}

void oracle_cast(const oracle_scene* scene, uint64_t n, const float* origins, const float* directions,
                 int32_t* prim, float* t, int threads) {
  pthread_t tid[64];
  cast_job jobs[64];
  int k;
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  for (k = 0; k < threads; k++) {
    cast_job j = {scene, n, origins, directions, prim, t, k, threads};
    jobs[k] = j;
    pthread_create(&tid[k], NULL, cast_worker, &jobs[k]);
  }
  for (k = 0; k < threads; k++) pthread_join(tid[k], NULL);
}

void oracle_sample_light(const amber_cuda_material* material, uint64_t n, const float* normal,
                         const float* dir_out, const float* xi8, float* dir_in, float* weight,
                         int32_t* draws) {
  amber_cuda_material m = *material;
  float r0 = 0;
  uint64_t i;
  if (m.kind == AMBER_CUDA_MAT_REFRACTION || m.kind == AMBER_CUDA_MAT_EYE) m.rgb[0] = m.rgb[1] = m.rgb[2] = 1.0f;
  if (m.kind == AMBER_CUDA_MAT_REFRACTION) r0 = (float)pow((double)((m.param - 1) / (m.param + 1)), 2.0);
  for (i = 0; i < n; i++) {
    sampler s = {NULL, xi8 + 8 * i, 8, 0, 0};
    const scatter_t sc = material_sample_light(&m, r0, V(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]),
                                               V(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), &s);
    dir_in[3 * i] = sc.dir.x; dir_in[3 * i + 1] = sc.dir.y; dir_in[3 * i + 2] = sc.dir.z;
    weight[3 * i] = sc.weight.x; weight[3 * i + 1] = sc.weight.y; weight[3 * i + 2] = sc.weight.z;
    draws[i] = s.used;
  }
}

void oracle_bsdf(const amber_cuda_material* material, uint64_t n, const float* normal,
                 const float* dir_out, const float* dir_in, float* bsdf, float* radiance) {
  uint64_t i;
  for (i = 0; i < n; i++) {
    const vec3 nn = V(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
    const vec3 wo = V(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]);
    const vec3 wi = V(dir_in[3 * i], dir_in[3 * i + 1], dir_in[3 * i + 2]);
    const vec3 f = material_bsdf(material, nn, wo, wi);
    const vec3 le = material_radiance(material, nn, wo);
    bsdf[3 * i] = f.x; bsdf[3 * i + 1] = f.y; bsdf[3 * i + 2] = f.z;
    radiance[3 * i] = le.x; radiance[3 * i + 1] = le.y; radiance[3 * i + 2] = le.z;
  }
}

void oracle_eye_rays(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                     uint64_t n, const uint32_t* px, const uint32_t* py, const float* xi5,
                     float* origin, float* direction, float* weight) {
  uint64_t i;
  for (i = 0; i < n; i++) {
    sampler s = {NULL, xi5 + 5 * i, 5, 0, 0};
    vec3 o, d;
    generate_eye_ray(scene, w, h, film_w, film_h, px[i], py[i], &s, &o, &d, &weight[i]);
    origin[3 * i] = o.x; origin[3 * i + 1] = o.y; origin[3 * i + 2] = o.z;
    direction[3 * i] = d.x; direction[3 * i + 1] = d.y; direction[3 * i + 2] = d.z;
  }
}

/* ---- the path tracer ------------------------------------------------------------------
 * REFERENCE estimator: PathTracing<RGB>::Thread::Render, src/amber/rendering/algorithm_pt.cc:125-160,
 *   plus a bounce cap (the reference has none, :137).
 * NEE estimator: equal in expectation.  At a connectable vertex (SurfaceType::Diffuse:
 *   Lambertian, Phong) one light point is sampled as LightSet::GenerateRay does
 *   (include/amber/scene/light_set.h:83-104) and connected with the s = 1 term of
 *   include/amber/rendering/bidirectional_path_sampling.h:956-1016; the emission of an
 *   emitting TRIANGLE found by the next BSDF-sampled ray is then not counted (it was just
 *   sampled), while after a camera / Specular / Refraction / Eye vertex it is. */
typedef struct {
  uint64_t closest_rays, shadow_rays, truncated, paths;
} counters;

static void trace_path(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                       uint32_t px, uint32_t py, philox_stream* stream, uint32_t max_depth, int estimator,
                       double rgb[3], counters* count) {
  sampler s = {stream, NULL, 0, 0, 1};
  vec3 origin, direction, weight3, measurement = V(0, 0, 0);
  float weight;
  int count_emission = 1;
  uint32_t depth;
  generate_eye_ray(sc, w, h, film_w, film_h, px, py, &s, &origin, &direction, &weight);
  weight3 = V(weight, weight, weight);
  count->paths++;
  for (depth = 1;; depth++) {
    hit_t hit;
    const int object = scene_cast(sc, origin, direction, &hit);
    const amber_cuda_material* m;
    const prim* p;
    vec3 wo;
    scatter_t scatter;
    float p_rr;
    int connectable;
    philox_stream_begin_vertex(stream, depth);
    s.camera = 0;
    count->closest_rays++;
    if (object < 0) break;
    p = &sc->prims[object];
    m = &sc->materials[p->material];
    wo = neg(direction);
    /* measurement += weight * Radiance(...), algorithm_pt.cc:144 */
    {
      const int sampled_by_nee = estimator == AMBER_CUDA_ESTIMATOR_NEE && p->kind == AMBER_CUDA_PRIM_TRIANGLE;
      if (!sampled_by_nee || count_emission) {
        measurement = add(measurement, mul(weight3, material_radiance(m, hit.normal, wo)));
      }
    }
    connectable = estimator == AMBER_CUDA_ESTIMATOR_NEE && sc->n_lights > 0 &&
                  (m->kind == AMBER_CUDA_MAT_LAMBERTIAN || m->kind == AMBER_CUDA_MAT_PHONG);
    if (connectable) {
      /* light choice: lower_bound over cumulative powers of Uniform(total), light_set.h:87-91 */
      const float target = uniform_light(&s, 0) * sc->total_power;
      uint32_t li = 0;
      const light* l;
      const prim* lp;
      vec3 y, xy, wi, contribution;
      float dist2, dist, cos_y, g;
      while (li + 1 < sc->n_lights && sc->lights[li].cumulative_power < target) li++;
      l = &sc->lights[li];
      lp = &sc->prims[l->prim_index];
      /* an emitting SPHERE is not light-sampled (the reference's own sphere sampling only covers half
       * of it): when the power CDF picks one the vertex has no shadow ray, and a sphere's emission is
       * counted on every BSDF-sampled hit instead (sampled_by_nee above) -- same expectation */
      if (lp->kind == AMBER_CUDA_PRIM_TRIANGLE) {
        const float lu = uniform_light(&s, 1);
        const float lv = uniform_light(&s, 2);
        y = triangle_point(lp->v0, lp->v1, lp->v2, lu, lv);
      } else {
        y = hit.position;
      }
      xy = sub(y, hit.position);
      dist2 = dot(xy, xy);
      dist = sqrtf(dist2);
      wi = V(xy.x / dist, xy.y / dist, xy.z / dist);
      cos_y = lp->kind == AMBER_CUDA_PRIM_TRIANGLE ? -dot(wi, lp->normal) : -1.0f;
      if (cos_y > 0) { /* one-sided emitter, material_diffuse_light.h:133-135 */
        const amber_cuda_material* lm = &sc->materials[lp->material];
        const vec3 le_over_pdf = V(lm->rgb[0] / l->pdf_area, lm->rgb[1] / l->pdf_area, lm->rgb[2] / l->pdf_area);
        /* GeometryFactor, include/amber/prelude/geometry.h:65-75 */
        g = fabsf(cos_y * dot(wi, hit.normal)) / dist2;
        contribution = scale(g, mul(mul(weight3, le_over_pdf), material_bsdf(m, hit.normal, wo, wi)));
        if (max3(contribution) > 0) {
          hit_t blocker;
          count->shadow_rays++;
          /* visible iff nothing lies in (kEPS, dist * (1 - 1e-4)) along x -> y */
          if (scene_cast(sc, hit.position, wi, &blocker) < 0 || !(blocker.t < dist * (1.0f - 1e-4f))) {
            measurement = add(measurement, contribution);
          }
        }
      }
    }
    scatter = material_sample_light(m, sc->r0[p->material], hit.normal, wo, &s);
    /* p = min(kRussianRoulette, Max(weight)); stop if Uniform >= p, algorithm_pt.cc:148-153 */
    p_rr = fminf((float)kRussianRoulette, max3(scatter.weight));
    if (uniform_rr(&s) >= p_rr) break;
    if (depth >= max_depth) {
      count->truncated++;
      break;
    }
    origin = hit.position; /* Ray(hit.Position(), scatter.DirectionIn()): no re-normalisation, ray.h:58-62 */
    direction = scatter.dir;
    weight3 = mul(weight3, V(scatter.weight.x / p_rr, scatter.weight.y / p_rr, scatter.weight.z / p_rr));
    count_emission = !connectable;
  }
  rgb[0] += measurement.x;
  rgb[1] += measurement.y;
  rgb[2] += measurement.z;
}

typedef struct {
  const oracle_scene* scene;
  uint32_t w, h;
  float film_w, film_h;
  uint64_t sample_begin, seed;
  uint32_t sample_count, max_depth;
  int estimator, worker, workers;
  float* rgb;
  counters count;
} render_job;

static void* render_worker(void* arg) {
  render_job* job = (render_job*)arg;
  uint32_t x, y;
  memset(&job->count, 0, sizeof job->count);
  for (y = (uint32_t)job->worker; y < job->h; y += (uint32_t)job->workers) {
    for (x = 0; x < job->w; x++) {
      double sum[3] = {0, 0, 0};
      uint64_t s;
      int c;
      for (s = job->sample_begin; s < job->sample_begin + job->sample_count; s++) {
        philox_stream stream;
        philox_stream_reset(&stream, job->seed, x + y * job->w, s);
        trace_path(job->scene, job->w, job->h, job->film_w, job->film_h, x, y, &stream, job->max_depth,
                   job->estimator, sum, &job->count);
      }
      for (c = 0; c < 3; c++) job->rgb[((size_t)x + (size_t)y * job->w) * 3 + c] = (float)(sum[c] / job->sample_count);
    }
  }
  return NULL;
}

void oracle_render(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                   uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_depth,
                   int estimator, int threads, float* rgb, uint64_t* stats) {
  pthread_t tid[64];
  render_job jobs[64];
  int k;
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  if (max_depth == 0) max_depth = 1u << 16;
  for (k = 0; k < threads; k++) {
    render_job j;
    memset(&j, 0, sizeof j);
    j.scene = scene; j.w = w; j.h = h; j.film_w = film_w; j.film_h = film_h;
    j.sample_begin = sample_begin; j.seed = seed; j.sample_count = sample_count; j.max_depth = max_depth;
    j.estimator = estimator; j.worker = k; j.workers = threads; j.rgb = rgb;
    jobs[k] = j;
    pthread_create(&tid[k], NULL, render_worker, &jobs[k]);
  }
  if (stats) memset(stats, 0, 4 * sizeof(uint64_t));
  for (k = 0; k < threads; k++) {
    pthread_join(tid[k], NULL);
    if (stats) {
      stats[0] += jobs[k].count.closest_rays;
      stats[1] += jobs[k].count.shadow_rays;
      stats[2] += jobs[k].count.truncated;
      stats[3] += jobs[k].count.paths;
    }
  }
}

/* ---- light tracing (row N2) ---------------------------------------------------------
 * LightTracing<RGB>::Thread::Render, src/amber/rendering/algorithm_lt.cc:124-163: a path starts
 * on a light, and whenever it lands on the aperture (an Eye surface) the lens maps it to a pixel
 * and the path weight is splatted there; it then goes on through the aperture like any other
 * surface.  One iteration = w*h such paths, each splat divided by w*h (:112-122,146).
 * film: w*h*3 doubles.  Same stream convention as the path tracer: vertex 0 = the light's five
 * sequential uniforms, vertex k >= 1 = the k-th surface interaction. */
static void trace_light_path(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                             philox_stream* stream, uint32_t max_depth, double* film, counters* count) {
  sampler s = {stream, NULL, 0, 0, 1};
  vec3 origin, direction, weight3;
  const float n_pixels = (float)((unsigned long)w * (unsigned long)h);
  uint32_t depth;
  generate_light_ray(sc, &s, &origin, &direction, &weight3);
  count->paths++;
  for (depth = 1;; depth++) {
    hit_t hit;
    const int object = scene_cast(sc, origin, direction, &hit);
    const amber_cuda_material* m;
    const prim* p;
    vec3 wo;
    scatter_t scatter;
    float p_rr;
    philox_stream_begin_vertex(stream, depth);
    s.camera = 0;
    count->closest_rays++;
    if (object < 0) break;
    p = &sc->prims[object];
    m = &sc->materials[p->material];
    wo = neg(direction);
    if (m->kind == AMBER_CUDA_MAT_EYE) {
      uint32_t px, py;
      float value;
      if (lens_response(sc, w, h, film_w, film_h, hit.position, wo, &px, &py, &value)) {
        /* image[pixel] += weight * response.Value() / image.Size() */
        double* out = film + ((size_t)px + (size_t)py * w) * 3;
        out[0] += (double)(weight3.x * value / n_pixels);
        out[1] += (double)(weight3.y * value / n_pixels);
        out[2] += (double)(weight3.z * value / n_pixels);
        count->shadow_rays++; /* counts splats here */
      }
    }
    scatter = material_sample_importance(m, sc->r0[p->material], hit.normal, wo, &s);
    p_rr = fminf((float)kRussianRoulette, max3(scatter.weight));
    if (uniform_rr(&s) >= p_rr) break;
    if (depth >= max_depth) {
      count->truncated++;
      break;
    }
    origin = hit.position;
    direction = scatter.dir;
    weight3 = mul(weight3, V(scatter.weight.x / p_rr, scatter.weight.y / p_rr, scatter.weight.z / p_rr));
  }
}

typedef struct {
  const oracle_scene* scene;
  uint32_t w, h;
  float film_w, film_h;
  uint64_t sample_begin, seed;
  uint32_t sample_count, max_depth;
  int worker, workers;
  double* film;
  counters count;
} lt_job;

static void* lt_worker(void* arg) {
  lt_job* job = (lt_job*)arg;
  const uint64_t n_pixels = (uint64_t)job->w * job->h;
  uint64_t s, i;
  memset(&job->count, 0, sizeof job->count);
  for (s = job->sample_begin; s < job->sample_begin + job->sample_count; s++) {
    for (i = (uint64_t)job->worker; i < n_pixels; i += (uint64_t)job->workers) {
      philox_stream stream;
      philox_stream_reset(&stream, job->seed, (uint32_t)i, s);
      trace_light_path(job->scene, job->w, job->h, job->film_w, job->film_h, &stream, job->max_depth,
                       job->film, &job->count);
    }
  }
  return NULL;
}

/* stats[0] closest rays, [1] splats, [2] truncated paths, [3] paths */
void oracle_render_lt(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                      uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_depth,
                      int threads, float* rgb, uint64_t* stats) {
  pthread_t tid[64];
  lt_job jobs[64];
  const size_t n_values = (size_t)w * h * 3;
  size_t v;
  int k;
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  if (max_depth == 0) max_depth = 1u << 16;
  if (stats) memset(stats, 0, 4 * sizeof(uint64_t));
  if (scene->n_lights == 0) { /* LightSet::GenerateRay on an empty set is undefined in the reference */
    memset(rgb, 0, n_values * sizeof(float));
    return;
  }
  for (k = 0; k < threads; k++) {
    lt_job j;
    memset(&j, 0, sizeof j);
    j.scene = scene; j.w = w; j.h = h; j.film_w = film_w; j.film_h = film_h;
    j.sample_begin = sample_begin; j.seed = seed; j.sample_count = sample_count; j.max_depth = max_depth;
    j.worker = k; j.workers = threads;
    j.film = (double*)calloc(n_values, sizeof(double));
    jobs[k] = j;
    pthread_create(&tid[k], NULL, lt_worker, &jobs[k]);
  }
  for (k = 0; k < threads; k++) {
    pthread_join(tid[k], NULL);
    if (stats) {
      stats[0] += jobs[k].count.closest_rays;
      stats[1] += jobs[k].count.shadow_rays;
      stats[2] += jobs[k].count.truncated;
      stats[3] += jobs[k].count.paths;
    }
  }
  for (v = 0; v < n_values; v++) {
    double sum = 0;
    for (k = 0; k < threads; k++) sum += jobs[k].film[v];
    rgb[v] = (float)(sum / sample_count);
  }
  for (k = 0; k < threads; k++) free(jobs[k].film);
}

/* Unit checks for the two functions light tracing adds to the path. */
void oracle_sample_importance(const amber_cuda_material* material, uint64_t n, const float* normal,
                              const float* dir_out, const float* xi8, float* dir_in, float* weight,
                              int32_t* draws) {
  uint64_t i;
  amber_cuda_material m = *material;
  float r0 = 0;
  if (m.kind == AMBER_CUDA_MAT_REFRACTION || m.kind == AMBER_CUDA_MAT_EYE) m.rgb[0] = m.rgb[1] = m.rgb[2] = 1.0f;
  if (m.kind == AMBER_CUDA_MAT_REFRACTION) r0 = (float)pow((double)((m.param - 1) / (m.param + 1)), 2.0);
  for (i = 0; i < n; i++) {
    sampler s = {NULL, xi8 + 8 * i, 8, 0, 0};
    const scatter_t sc = material_sample_importance(&m, r0, V(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]),
                                                    V(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), &s);
    dir_in[3 * i] = sc.dir.x; dir_in[3 * i + 1] = sc.dir.y; dir_in[3 * i + 2] = sc.dir.z;
    weight[3 * i] = sc.weight.x; weight[3 * i + 1] = sc.weight.y; weight[3 * i + 2] = sc.weight.z;
    if (draws) draws[i] = s.used;
  }
}

/* pixel[i] = x + y*w, or -1 when the lens does not respond */
void oracle_lens_response(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                          uint64_t n, const float* position, const float* dir_out, int32_t* pixel,
                          float* value) {
  uint64_t i;
  for (i = 0; i < n; i++) {
    uint32_t px = 0, py = 0;
    float f = 0;
    const int ok = lens_response(scene, w, h, film_w, film_h,
                                 V(position[3 * i], position[3 * i + 1], position[3 * i + 2]),
                                 V(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]), &px, &py, &f);
    pixel[i] = ok ? (int32_t)(px + py * w) : -1;
    value[i] = ok ? f : 0.0f;
  }
}

/* n light rays with scripted uniforms xi5[n][5] (choice, u, v, direction 2) */
void oracle_light_rays(const oracle_scene* scene, uint64_t n, const float* xi5, float* origin,
                       float* direction, float* weight) {
  uint64_t i;
  for (i = 0; i < n; i++) {
    sampler s = {NULL, xi5 + 5 * i, 5, 0, 0};
    vec3 o, d, wgt;
    generate_light_ray(scene, &s, &o, &d, &wgt);
    origin[3 * i] = o.x; origin[3 * i + 1] = o.y; origin[3 * i + 2] = o.z;
    direction[3 * i] = d.x; direction[3 * i + 1] = d.y; direction[3 * i + 2] = d.z;
    weight[3 * i] = wgt.x; weight[3 * i + 1] = wgt.y; weight[3 * i + 2] = wgt.z;
  }
}

/* ---- bidirectional path tracing (row N3) ----------------------------------------------
 * include/amber/rendering/bidirectional_path_sampling.h restated: SubpathEvent (:46-90,281-315),
 * GenerateLightSubpath (:726-801), GenerateEyeSubpath (:821-899), Connect (:901-1017),
 * BidirectionalPathSamplingBuffer (:448-697: geometry factors, light / importance
 * probabilities, technique probabilities, specular connections disabled), Combine (:1019-1068)
 * with BalanceHeuristic (multiple_importance_sampling.h:58-69), driven per pixel as
 * src/amber/rendering/algorithm_bdpt.cc:116-148.  Each subpath draws from its own sequential
 * stream (philox_stream.h).  Everything in float, in the reference's operation order. */
#define BDPT_MAX_EVENTS 64
#define OBJECT_LENS (-2) /* the lens's own aperture object: never equal to an object found by Cast */

typedef struct {
  int object;   /* primitive index, or OBJECT_LENS on the eye's leading event */
  int material; /* material index; -1 = Eye (leading eye event) */
  vec3 position, normal, direction_out, weight;
  float geometry_factor, p_backward, p_forward; /* NaN on a leading event */
} event_t;

enum { SURFACE_DIFFUSE, SURFACE_SPECULAR, SURFACE_LIGHT, SURFACE_EYE };
static int surface_of(const oracle_scene* sc, int material) {
  int kind;
  if (material < 0) return SURFACE_EYE;
  kind = sc->materials[material].kind;
  if (kind == AMBER_CUDA_MAT_LAMBERTIAN || kind == AMBER_CUDA_MAT_PHONG) return SURFACE_DIFFUSE;
  if (kind == AMBER_CUDA_MAT_DIFFUSE_LIGHT) return SURFACE_LIGHT;
  if (kind == AMBER_CUDA_MAT_EYE) return SURFACE_EYE;
  return SURFACE_SPECULAR;
}
static int event_is_leading(const event_t* e) { return isnan(e->geometry_factor); }

/* Refraction, material_refraction.cc:41-109: Fresnel term of the branch (reflection if wo and wi
 * are on the same side), with (radiance) or without (importance) the ior^2 scaling. */
static float refraction_term(const amber_cuda_material* m, float r0, vec3 n, vec3 wo, vec3 wi, int adjoint) {
  const float signed_cos_alpha = dot(wo, n);
  const float ior = signed_cos_alpha > 0 ? 1 / m->param : m->param;
  const float squared_cos_beta = 1 - (1 - signed_cos_alpha * signed_cos_alpha) * (ior * ior);
  float rho_r, rho_t, signed_cos_i;
  if (squared_cos_beta < 0) return 1; /* kDiracDelta */
  rho_r = schlick(r0, fabsf(signed_cos_alpha));
  rho_t = adjoint ? 1 - rho_r : (1 - rho_r) * (ior * ior);
  signed_cos_i = dot(wi, n);
  return signed_cos_alpha * signed_cos_i > 0 ? rho_r * 1 : rho_t * 1;
}
/* PDFLight (:111-143) / PDFImportance (:145-175) */
static float refraction_pdf(const amber_cuda_material* m, float r0, vec3 n, vec3 wo, vec3 wi, int importance) {
  const float signed_cos_alpha = dot(wo, n);
  const float ior = signed_cos_alpha > 0 ? 1 / m->param : m->param;
  const float squared_cos_beta = 1 - (1 - signed_cos_alpha * signed_cos_alpha) * (ior * ior);
  float rho_r, rho_t, rho, signed_cos_i;
  if (squared_cos_beta < 0) return 1;
  rho_r = schlick(r0, fabsf(signed_cos_alpha));
  signed_cos_i = dot(wi, n);
  if (importance) {
    rho_t = 1 - rho_r;
    return signed_cos_alpha * signed_cos_i > 0 ? (rho_r + 0.5f) / 2 * 1.0f : (rho_t + 0.5f) / 2 * 1.0f;
  }
  rho_t = (1 - rho_r) * (ior * ior);
  rho = rho_r + rho_t;
  return signed_cos_alpha * signed_cos_i > 0 ? (rho_r / rho + 0.5f) / 2 * 1.0f : (rho_t / rho + 0.5f) / 2 * 1.0f;
}

/* Material::BSDF (adjoint = 0) / Material::AdjointBSDF (adjoint = 1) of every material kind */
static vec3 event_bsdf(const oracle_scene* sc, int material, vec3 n, vec3 wo, vec3 wi, int adjoint) {
  const amber_cuda_material* m;
  vec3 rho;
  if (material < 0) return V(1, 1, 1); /* Eye: Radiant(kDiracDelta), material_eye.h:97-117 */
  m = &sc->materials[material];
  rho = V(m->rgb[0], m->rgb[1], m->rgb[2]);
  switch (m->kind) {
    case AMBER_CUDA_MAT_LAMBERTIAN:
    case AMBER_CUDA_MAT_PHONG: return material_bsdf(m, n, wo, wi);
    case AMBER_CUDA_MAT_SPECULAR: /* material_specular.cc:36-50 */
      if (dot(wo, n) * dot(wi, n) <= 0) return V(0, 0, 0);
      return scale(1.0f, rho);
    case AMBER_CUDA_MAT_REFRACTION: return scale(refraction_term(m, sc->r0[material], n, wo, wi, adjoint), rho);
    case AMBER_CUDA_MAT_EYE: return V(1, 1, 1);
    default: return V(0, 0, 0); /* DiffuseLight, material_diffuse_light.h:141-161 */
  }
}
/* Material::PDFLight (importance = 0) / PDFImportance (importance = 1) */
static float event_pdf(const oracle_scene* sc, int material, vec3 n, vec3 wo, vec3 wi, int importance) {
  const amber_cuda_material* m;
  if (material < 0) return 1;
  m = &sc->materials[material];
  switch (m->kind) {
    case AMBER_CUDA_MAT_LAMBERTIAN: return 1 / PI_F; /* material_lambertian.cc:51-59 */
    case AMBER_CUDA_MAT_PHONG: {                     /* material_phong.cc:64-79 */
      const vec3 refl = perfect_reflection(wo, n, dot(wo, n));
      const float cos_alpha = fmaxf(0, dot(wi, refl));
      const float cos_i = fabsf(dot(wi, n));
      return (m->param + 1) / (2 * PI_F) * powf(cos_alpha, m->param) / cos_i;
    }
    case AMBER_CUDA_MAT_REFRACTION: return refraction_pdf(m, sc->r0[material], n, wo, wi, importance);
    default: return 1; /* Specular, DiffuseLight, Eye: kDiracDelta */
  }
}
static float min_rr(float x) { return x < (float)kRussianRoulette ? x : (float)kRussianRoulette; } /* std::min(kRR, x) */
static vec3 div3(vec3 a, float b) { return V(a.x / b, a.y / b, a.z / b); }

static float light_pdf_area(const oracle_scene* sc, int material) { /* LightSet::PDFArea, light_set.h:106-111 */
  const amber_cuda_material* m = &sc->materials[material];
  return (m->rgb[0] * PI_F + m->rgb[1] * PI_F + m->rgb[2] * PI_F) / sc->total_power;
}
/* Lens::PDFDirection: BasicThin lens_thin.cc:138-148, BasicPinhole lens_pinhole.cc:93-108 */
static float eye_pdf_direction(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h, vec3 dir) {
  const vec3 d = mat3_apply(sc->local, dir);
  if (sc->lens_kind == AMBER_CUDA_LENS_PINHOLE) {
    const vec3 point = scale(sc->sensor_distance / d.z, d);
    const float geometry_factor = d.z * d.z / dot(point, point);
    return (1 / (film_w * film_h)) / geometry_factor;
  }
  return (float)((double)((float)((unsigned long)w * (unsigned long)h) / (film_w * film_h)) *
                 pow((double)sc->sensor_distance, 2.0) / pow((double)d.z, 4.0));
}
/* prelude::GeometryFactor, geometry.h:49-75 */
static float geometry_factor_dir(vec3 direction, float squared_distance, vec3 nx, vec3 ny) {
  return fabsf(dot(direction, nx) * dot(direction, ny) / squared_distance);
}
static float geometry_factor_points(vec3 px, vec3 py, vec3 nx, vec3 ny) {
  const vec3 xy = sub(py, px);
  const float squared_distance = dot(xy, xy);
  const float distance = sqrtf(squared_distance);
  return geometry_factor_dir(V(xy.x / distance, xy.y / distance, xy.z / distance), squared_distance, nx, ny);
}

/* GenerateLightSubpath (importance = 1) / GenerateEyeSubpath (importance = 0); returns the events */
static int generate_subpath(const oracle_scene* sc, int importance, uint32_t w, uint32_t h, float film_w,
                            float film_h, uint32_t px, uint32_t py, sampler* s, int cap, event_t* out) {
  vec3 origin, direction, weight, normal;
  int n = 0, i;
  if (cap == 0) return 0;
  if (importance) {
    /* the light's leading event: generate_light_ray + which light it was */
    const float target = uniform(s) * sc->total_power;
    uint32_t li = 0;
    const light* l;
    const prim* lp;
    const amber_cuda_material* lm;
    while (li + 1 < sc->n_lights && sc->lights[li].cumulative_power < target) li++;
    l = &sc->lights[li];
    lp = &sc->prims[l->prim_index];
    lm = &sc->materials[lp->material];
    origin = light_surface_point(lp, s, &normal);
    direction = hemisphere_psa(normal, s);
    weight = V(lm->rgb[0] * PI_F / l->pdf_area, lm->rgb[1] * PI_F / l->pdf_area, lm->rgb[2] * PI_F / l->pdf_area);
    out[n].object = l->prim_index;
    out[n].material = lp->material;
  } else {
    float wgt;
    size_t blade;
    /* the blade is the first camera uniform: its normal is the leading event's normal */
    {
      sampler probe = *s;
      philox_stream probe_stream;
      if (s->stream) {
        probe_stream = *s->stream;
        probe.stream = &probe_stream;
      }
      blade = sc->lens_kind == AMBER_CUDA_LENS_PINHOLE ? 0 : (size_t)floorf(uniform(&probe) * (float)sc->n_blades);
      if (sc->lens_kind != AMBER_CUDA_LENS_PINHOLE && blade > (size_t)sc->n_blades - 1) blade = (size_t)sc->n_blades - 1;
    }
    generate_eye_ray(sc, w, h, film_w, film_h, px, py, s, &origin, &direction, &wgt);
    weight = V(wgt, wgt, wgt);
    /* BasicThin: the blade's normal (lens_thin.cc:99); BasicPinhole: Normalize(global(0, 0, -1)) (lens_pinhole.cc:62) */
    normal = sc->lens_kind == AMBER_CUDA_LENS_PINHOLE ? normalize(mat3_apply(sc->global, V(0, 0, -1))) : sc->blades[blade][3];
    out[n].object = OBJECT_LENS;
    out[n].material = -1;
  }
  out[n].position = origin;
  out[n].normal = normal;
  out[n].direction_out = normal; /* SubpathEvent(object, leading), :281-294 */
  out[n].weight = weight;
  out[n].geometry_factor = out[n].p_backward = out[n].p_forward = NAN;
  n++;
  for (i = 1; i < cap && n < BDPT_MAX_EVENTS; i++) {
    hit_t hit;
    const int object = scene_cast(sc, origin, direction, &hit);
    const prim* p;
    const amber_cuda_material* m;
    vec3 wo, wi, bsdf_backward;
    scatter_t scatter;
    float p_forward, p_rr_forward, p_backward, p_rr_backward;
    event_t* e;
    if (object < 0) break;
    p = &sc->prims[object];
    m = &sc->materials[p->material];
    wo = neg(direction);
    scatter = importance ? material_sample_importance(m, sc->r0[p->material], hit.normal, wo, s)
                         : material_sample_light(m, sc->r0[p->material], hit.normal, wo, s);
    wi = scatter.dir;
    /* forward = the transport being traced, backward = the other one */
    p_forward = event_pdf(sc, p->material, hit.normal, wo, wi, importance);
    p_rr_forward = min_rr(max3(scatter.weight));
    p_backward = event_pdf(sc, p->material, hit.normal, wi, wo, !importance);
    bsdf_backward = event_bsdf(sc, p->material, hit.normal, wi, wo, !importance);
    p_rr_backward = min_rr(max3(div3(bsdf_backward, p_backward)));
    e = &out[n++];
    e->object = object;
    e->material = p->material;
    e->position = hit.position;
    e->normal = hit.normal;
    e->direction_out = wo;
    e->weight = weight;
    e->geometry_factor = geometry_factor_dir(wo, hit.t * hit.t, normal, hit.normal);
    e->p_backward = p_backward * p_rr_backward;
    e->p_forward = p_forward * p_rr_forward;
    if (uniform(s) >= p_rr_forward) break;
    origin = hit.position;
    direction = wi;
    normal = hit.normal;
    weight = mul(weight, div3(scatter.weight, p_rr_forward));
  }
  return n;
}

typedef struct {
  vec3 value;
  int has_pixel;
  uint32_t px, py;
} contribution_t;

/* Connect, :901-1017 */
static contribution_t bdpt_connect(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                                   const event_t* le, const event_t* ee) {
  contribution_t c;
  vec3 shadow_dir, to_light;
  c.value = V(0, 0, 0);
  c.has_pixel = 0;
  c.px = c.py = 0;
  if (!le && !ee) return c;
  if (!le) { /* s = 0: the eye subpath found a light */
    if (surface_of(sc, ee->material) != SURFACE_LIGHT) return c;
    c.value = mul(ee->weight, material_radiance(&sc->materials[ee->material], ee->normal, ee->direction_out));
    return c;
  }
  if (!ee) { /* t = 0: the light subpath found the aperture */
    float response;
    if (surface_of(sc, le->material) != SURFACE_EYE) return c;
    if (!lens_response(sc, w, h, film_w, film_h, le->position, le->direction_out, &c.px, &c.py, &response)) return c;
    c.has_pixel = 1;
    c.value = mul(le->weight, V(response, response, response));
    return c;
  }
  if (!event_is_leading(le) && surface_of(sc, le->material) != SURFACE_DIFFUSE) return c;
  if (!event_is_leading(ee) && surface_of(sc, ee->material) != SURFACE_DIFFUSE) return c;
  c.value = V(1, 1, 1);
  shadow_dir = normalize(sub(ee->position, le->position)); /* Ray(origin, Vector3) normalises */
  to_light = neg(shadow_dir);
  if (event_is_leading(le)) { /* s = 1 */
    c.value = mul(c.value, div3(material_radiance(&sc->materials[le->material], le->normal, shadow_dir),
                                light_pdf_area(sc, le->material)));
  } else {
    c.value = mul(c.value, mul(le->weight, event_bsdf(sc, le->material, le->normal, le->direction_out, shadow_dir, 1)));
  }
  if (event_is_leading(ee)) { /* t = 1 */
    float response;
    if (!lens_response(sc, w, h, film_w, film_h, ee->position, to_light, &c.px, &c.py, &response)) {
      c.value = V(0, 0, 0);
      return c;
    }
    c.has_pixel = 1;
    c.value = div3(mul(c.value, V(response, response, response)), sc->p_area);
  } else {
    c.value = mul(c.value, mul(ee->weight, event_bsdf(sc, ee->material, ee->normal, ee->direction_out, to_light, 0)));
  }
  if (!(max3(c.value) > 0)) {
    c.value = V(0, 0, 0);
    return c;
  }
  { /* "XXX rough check": the first object along the shadow ray must be the eye endpoint's object */
    hit_t hit;
    const int object = scene_cast(sc, le->position, shadow_dir, &hit);
    if (object != ee->object) {
      c.value = V(0, 0, 0);
      return c;
    }
  }
  c.value = scale(geometry_factor_points(le->position, ee->position, le->normal, ee->normal), c.value);
  return c;
}

/* BidirectionalPathSamplingBuffer::Buffer + MISWeight(BalanceHeuristic), :423-697 */
static float bdpt_mis_weight(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                             const event_t* L, int s, const event_t* E, int t) {
  float gf[2 * BDPT_MAX_EVENTS + 2], p_light[2 * BDPT_MAX_EVENTS + 2], p_importance[2 * BDPT_MAX_EVENTS + 2];
  float p_technique[2 * BDPT_MAX_EVENTS + 2];
  int n_gf = 0, n_pl = 0, n_pi = 0, i;
  const int n = s + t + 1;
  float sum;
  /* geometry factors */
  for (i = 1; i < s; i++) gf[n_gf++] = L[i].geometry_factor;
  if (s > 0 && t > 0) gf[n_gf++] = geometry_factor_points(L[s - 1].position, E[t - 1].position, L[s - 1].normal, E[t - 1].normal);
  for (i = t - 1; i >= 1; i--) gf[n_gf++] = E[i].geometry_factor;
  /* light probabilities */
  for (i = 1; i < s; i++) p_light[n_pl++] = L[i].p_backward;
  if (s > 0 && t > 0) {
    const event_t* le = &L[s - 1];
    const event_t* ee = &E[t - 1];
    const vec3 din = normalize(sub(le->position, ee->position));
    const vec3 bsdf = event_bsdf(sc, ee->material, ee->normal, ee->direction_out, din, 0);
    const float pdf = event_pdf(sc, ee->material, ee->normal, ee->direction_out, din, 0);
    p_light[n_pl++] = pdf * min_rr(max3(div3(bsdf, pdf)));
  }
  for (i = t - 2; i >= 0; i--) p_light[n_pl++] = E[i].p_forward;
  {
    const event_t* z0;
    const event_t* z1;
    if (t == 0) { z0 = &L[s - 1]; z1 = &L[s - 2]; }
    else if (t == 1) { z0 = &E[0]; z1 = &L[s - 1]; }
    else { z0 = &E[0]; z1 = &E[1]; }
    p_light[n_pl - 1] = eye_pdf_direction(sc, w, h, film_w, film_h, normalize(sub(z1->position, z0->position)));
  }
  /* importance probabilities */
  for (i = 0; i + 1 < s; i++) p_importance[n_pi++] = L[i].p_forward;
  if (s > 0 && t > 0) {
    const event_t* le = &L[s - 1];
    const event_t* ee = &E[t - 1];
    const vec3 din = normalize(sub(ee->position, le->position));
    const vec3 bsdf = event_bsdf(sc, le->material, le->normal, le->direction_out, din, 1);
    const float pdf = event_pdf(sc, le->material, le->normal, le->direction_out, din, 1);
    p_importance[n_pi++] = pdf * min_rr(max3(div3(bsdf, pdf)));
  }
  for (i = t - 1; i >= 1; i--) p_importance[n_pi++] = E[i].p_backward;
  p_importance[0] = (float)(1 / kPI); /* LightSet::PDFDirection, light_set.h:113-122 */
  /* technique probabilities relative to the one that was used (index s) */
  for (i = 0; i < n; i++) p_technique[i] = 1;
  for (i = s - 1; i >= 0; i--) {
    float p = p_technique[i + 1];
    if (t == 0 && i == n - 2) p *= sc->p_area; else p *= p_light[i] * gf[i];
    if (s > 0 && i == 0) p /= light_pdf_area(sc, L[0].material); else p /= p_importance[i - 1] * gf[i - 1];
    p_technique[i] = p;
  }
  for (i = s + 1; i < n; i++) {
    float p = p_technique[i - 1];
    if (s == 0 && i == 1) p *= light_pdf_area(sc, E[t - 1].material); else p *= p_importance[i - 2] * gf[i - 2];
    if (t > 0 && i == n - 1) p /= sc->p_area; else p /= p_light[i - 1] * gf[i - 1];
    p_technique[i] = p;
  }
  /* DisableSpecularConnection */
  for (i = 1; i + 1 < s; i++) {
    if (surface_of(sc, L[i].material) != SURFACE_DIFFUSE) p_technique[i] = p_technique[i + 1] = 0;
  }
  for (i = 1; i + 1 < t; i++) {
    if (surface_of(sc, E[i].material) != SURFACE_DIFFUSE) p_technique[n - i - 1] = p_technique[n - i - 2] = 0;
  }
  sum = 0;
  for (i = 0; i < n; i++) sum = sum + p_technique[i];
  (void)n_gf;
  return 1 / sum;
}

typedef struct {
  uint64_t light_events, eye_events, splats;
} bdpt_counters;

/* Thread::Render + Combine for one pixel sample; film: w*h*3 doubles */
static void bdpt_sample(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h, uint32_t px,
                        uint32_t py, uint64_t seed, uint64_t sample, int cap, double* film, bdpt_counters* count) {
  event_t L[BDPT_MAX_EVENTS], E[BDPT_MAX_EVENTS];
  philox_stream stream;
  sampler smp = {&stream, NULL, 0, 0, 1};
  const float n_pixels = (float)((unsigned long)w * (unsigned long)h);
  int s_max, t_max, s, t;
  vec3 measurement = V(0, 0, 0);
  philox_stream_reset(&stream, seed, px + py * w, sample);
  philox_stream_begin_sequence(&stream, PHILOX_SUBPATH_LIGHT);
  s_max = generate_subpath(sc, 1, w, h, film_w, film_h, px, py, &smp, cap, L);
  philox_stream_begin_sequence(&stream, PHILOX_SUBPATH_EYE);
  t_max = generate_subpath(sc, 0, w, h, film_w, film_h, px, py, &smp, cap, E);
  count->light_events += (uint64_t)s_max;
  count->eye_events += (uint64_t)t_max;
  for (s = 0; s <= s_max; s++) {
    for (t = 0; t <= t_max; t++) {
      contribution_t c = bdpt_connect(sc, w, h, film_w, film_h, s ? &L[s - 1] : NULL, t ? &E[t - 1] : NULL);
      float weight;
      if (!(max3(c.value) > 0)) continue;
      weight = bdpt_mis_weight(sc, w, h, film_w, film_h, L, s, E, t);
      c.value = scale(weight, c.value);
      if (c.has_pixel) { /* light image: / sensor.Size() */
        double* to = film + ((size_t)c.px + (size_t)c.py * w) * 3;
        to[0] += (double)(c.value.x / n_pixels);
        to[1] += (double)(c.value.y / n_pixels);
        to[2] += (double)(c.value.z / n_pixels);
        count->splats++;
      } else {
        measurement = add(measurement, c.value);
      }
    }
  }
  {
    double* to = film + ((size_t)px + (size_t)py * w) * 3;
    to[0] += (double)measurement.x;
    to[1] += (double)measurement.y;
    to[2] += (double)measurement.z;
  }
}

typedef struct {
  const oracle_scene* scene;
  uint32_t w, h;
  float film_w, film_h;
  uint64_t sample_begin, seed;
  uint32_t sample_count;
  int cap, worker, workers;
  double* film;
  bdpt_counters count;
} bdpt_job;

static void* bdpt_worker(void* arg) {
  bdpt_job* job = (bdpt_job*)arg;
  const uint64_t n_pixels = (uint64_t)job->w * job->h;
  uint64_t s, i;
  memset(&job->count, 0, sizeof job->count);
  for (s = job->sample_begin; s < job->sample_begin + job->sample_count; s++) {
    for (i = (uint64_t)job->worker; i < n_pixels; i += (uint64_t)job->workers) {
      bdpt_sample(job->scene, job->w, job->h, job->film_w, job->film_h, (uint32_t)(i % job->w), (uint32_t)(i / job->w),
                  job->seed, s, job->cap, job->film, &job->count);
    }
  }
  return NULL;
}

/* stats[0] light-subpath events, [1] eye-subpath events, [2] light-image splats.
 * max_vertices: events per subpath, 1..64 (0 = 64; the reference has no cap). */
void oracle_render_bdpt(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                        uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_vertices,
                        int threads, float* rgb, uint64_t* stats) {
  pthread_t tid[64];
  bdpt_job jobs[64];
  const size_t n_values = (size_t)w * h * 3;
  size_t v;
  int k;
  if (threads < 1) threads = 1;
  if (threads > 64) threads = 64;
  if (max_vertices == 0 || max_vertices > BDPT_MAX_EVENTS) max_vertices = BDPT_MAX_EVENTS;
  if (stats) memset(stats, 0, 3 * sizeof(uint64_t));
  for (k = 0; k < threads; k++) {
    bdpt_job j;
    memset(&j, 0, sizeof j);
    j.scene = scene; j.w = w; j.h = h; j.film_w = film_w; j.film_h = film_h;
    j.sample_begin = sample_begin; j.seed = seed; j.sample_count = sample_count; j.cap = (int)max_vertices;
    j.worker = k; j.workers = threads;
    j.film = (double*)calloc(n_values, sizeof(double));
    jobs[k] = j;
    pthread_create(&tid[k], NULL, bdpt_worker, &jobs[k]);
  }
  for (k = 0; k < threads; k++) {
    pthread_join(tid[k], NULL);
    if (stats) {
      stats[0] += jobs[k].count.light_events;
      stats[1] += jobs[k].count.eye_events;
      stats[2] += jobs[k].count.splats;
    }
  }
  for (v = 0; v < n_values; v++) {
    double sum = 0;
    for (k = 0; k < threads; k++) sum += jobs[k].film[v];
    rgb[v] = (float)(sum / sample_count);
  }
  for (k = 0; k < threads; k++) free(jobs[k].film);
}

/* ---- test hooks -------------------------------------------------------------------- */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  philox4x32_10(ctr, key, out);
}

void oracle_stream(uint64_t seed, uint32_t pixel, uint64_t sample, uint32_t n, float* out) {
  philox_stream s;
  uint32_t i;
  philox_stream_reset(&s, seed, pixel, sample);
  for (i = 0; i < n; i++) out[i] = philox_stream_next_camera(&s); /* vertex 0, slots 0.. */
}

/* The fixed ray set of the intersection-parity test: ray i of the first
 * half is the camera ray of a uniformly random pixel (stream (seed, i, 0)); ray n + i is the
 * secondary ray leaving its hit point in the BSDF-sampled direction -- it starts ON a surface,
 * which exercises the t >= kEPS rule, grazing, mirror and refracted directions.  Where the
 * camera ray misses, the secondary ray repeats it. */
void oracle_ray_set(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                    uint64_t seed, uint64_t n, float* origins, float* directions) {
  uint64_t i;
  for (i = 0; i < n; i++) {
    philox_stream stream;
    sampler s = {&stream, NULL, 0, 0, 1};
    vec3 o, d, o2, d2;
    float weight;
    uint32_t px, py;
    hit_t hit;
    int object;
    philox_stream_reset(&stream, seed, (uint32_t)i, 0);
    px = (uint32_t)(uniform(&s) * (float)w);
    py = (uint32_t)(uniform(&s) * (float)h);
    if (px >= w) px = w - 1;
    if (py >= h) py = h - 1;
    generate_eye_ray(sc, w, h, film_w, film_h, px, py, &s, &o, &d, &weight);
    object = scene_cast(sc, o, d, &hit);
    o2 = o;
    d2 = d;
    if (object >= 0) {
      const prim* p = &sc->prims[object];
      scatter_t scatter;
      philox_stream_begin_vertex(&stream, 1);
      s.camera = 0;
      scatter =
          material_sample_light(&sc->materials[p->material], sc->r0[p->material], hit.normal, neg(d), &s);
      if (scatter.dir.x != 0 || scatter.dir.y != 0 || scatter.dir.z != 0) {
        o2 = hit.position;
        d2 = scatter.dir;
      }
    }
    origins[3 * i] = o.x; origins[3 * i + 1] = o.y; origins[3 * i + 2] = o.z;
    directions[3 * i] = d.x; directions[3 * i + 1] = d.y; directions[3 * i + 2] = d.z;
    origins[3 * (n + i)] = o2.x; origins[3 * (n + i) + 1] = o2.y; origins[3 * (n + i) + 2] = o2.z;
    directions[3 * (n + i)] = d2.x; directions[3 * (n + i) + 1] = d2.y; directions[3 * (n + i) + 2] = d2.z;
  }
}

/* Debug: the vertices of one path, REFERENCE estimator.  out[k] = {prim, t, x, y, z, draws
 * consumed after the vertex, throughput.r, p_rr}; returns the number of vertices. */
int oracle_trace_path(const oracle_scene* sc, uint32_t w, uint32_t h, float film_w, float film_h,
                      uint32_t px, uint32_t py, uint64_t sample, uint64_t seed, uint32_t max_depth,
                      int estimator, float* out, int max_vertices) {
  philox_stream stream;
  sampler s = {&stream, NULL, 0, 0, 1};
  vec3 origin, direction, weight3;
  float weight;
  int n = 0;
  uint32_t depth;
  philox_stream_reset(&stream, seed, px + py * w, sample);
  generate_eye_ray(sc, w, h, film_w, film_h, px, py, &s, &origin, &direction, &weight);
  weight3 = V(weight, weight, weight);
  for (depth = 1; n < max_vertices; depth++) {
    hit_t hit;
    const int object = scene_cast(sc, origin, direction, &hit);
    const amber_cuda_material* m;
    const prim* p;
    scatter_t scatter;
    float p_rr;
    float* v = out + 8 * n;
    philox_stream_begin_vertex(&stream, depth);
    s.camera = 0;
    v[0] = (float)object; v[1] = hit.t; v[2] = v[3] = v[4] = 0; v[5] = (float)stream.draws; v[6] = weight3.x; v[7] = 0;
    n++;
    if (object < 0) break;
    v[2] = hit.position.x; v[3] = hit.position.y; v[4] = hit.position.z;
    p = &sc->prims[object];
    m = &sc->materials[p->material];
    if (estimator == AMBER_CUDA_ESTIMATOR_NEE && sc->n_lights > 0 &&
        (m->kind == AMBER_CUDA_MAT_LAMBERTIAN || m->kind == AMBER_CUDA_MAT_PHONG)) {
      /* the light-sampling slots are fixed: nothing to skip */
    }
    scatter = material_sample_light(m, sc->r0[p->material], hit.normal, neg(direction), &s);
    p_rr = fminf((float)kRussianRoulette, max3(scatter.weight));
    v[7] = p_rr;
    if (uniform_rr(&s) >= p_rr) break;
    v[5] = (float)stream.draws;
    if (depth >= max_depth) break;
    origin = hit.position;
    direction = scatter.dir;
    weight3 = mul(weight3, V(scatter.weight.x / p_rr, scatter.weight.y / p_rr, scatter.weight.z / p_rr));
  }
  return n;
}

/* ---- post-process: Filmic + Gamma -> 8 bit (row N4) ------------------------------------
 * Filmic::operator() / Filmic::Map, src/amber/postprocess/filmic.cc:28-67: Hable's curve with
 * exposure 16 and white point 0.70, every operation in float (hdr_value_type = std::float_t);
 * Gamma::operator(), src/amber/postprocess/gamma.cc:36-52: 255 * min(1, pow(x, 1 / 2.2)),
 * truncated to uint_fast8_t.  A NaN (pow of a tiny negative Map value) becomes 255 through
 * std::min's operand order, as in the reference. */
static float filmic_map(float h) {
  const float kA = 0.22f, kB = 0.30f, kC = 0.10f, kD = 0.20f, kE = 0.01f, kF = 0.30f;
  return (h * (h * kA + kB * kC) + kD * kE) / (h * (h * kA + kB) + kD * kF) - kE / kF;
}

void oracle_tonemap(uint64_t n_values, const float* hdr, uint8_t* ldr) {
  const float kW = 0.70f, kExposure = 16.0f, gamma = 2.2f;
  const float white = filmic_map(kW);
  uint64_t i;
  for (i = 0; i < n_values; i++) {
    const float mapped = filmic_map(hdr[i] * kExposure) / white;
    const float p = powf(mapped, 1 / gamma);
    const float clamped = p < 1.0f ? p : 1.0f; /* std::min<float>(1, p) */
    ldr[i] = (uint8_t)(255 * clamped);
  }
}
