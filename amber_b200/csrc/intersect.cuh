// intersect.cuh -- ray/triangle, ray/sphere and the constant-memory list traversal.
#pragma once

#include <math_constants.h>

#include "scene_device.cuh"

namespace amber_cuda {

__constant__ ConstScene c_scene;  // single translation unit: defined here

struct Ray {
  float3 o, d;
};

// Triangle test.  Accepts iff u >= 0, v >= 0, u + v <= 1 (which implies u,v <= 1),
// t > kEPS and t < t_best -- the inclusive edge rules and the strict `<` closest-hit
// update of the reference (src/amber/scene/primitive_triangle.cc:108-125,
// include/amber/raytracer/acceleration_bvh.h:346-351).  Every comparison is written
// so that a NaN (n.d == 0, the reference's det == 0 case) rejects, as
// include/amber/prelude/hit.h:90-94 does.
__device__ __forceinline__ bool intersect_triangle(const float4 n_v0x, const float4 nu_v0y,
                                                   const float4 nv_v0z, const Ray& ray,
                                                   float t_best, float& t_out) {
  const float3 T = v3(ray.o.x - n_v0x.w, ray.o.y - nu_v0y.w, ray.o.z - nv_v0z.w);
  const float3 n = v3(n_v0x);
  const float nd = dot(n, ray.d);
  const float nT = dot(n, T);
  const float t = __fdividef(-nT, nd);
  const float3 h = v3(fmaf(t, ray.d.x, T.x), fmaf(t, ray.d.y, T.y), fmaf(t, ray.d.z, T.z));
  const float u = dot(v3(nu_v0y), h);
  const float v = dot(v3(nv_v0z), h);
  const bool hit = (u >= 0.0f) & (v >= 0.0f) & (u + v <= 1.0f) & (t > kEpsF) & (t < t_best);
  if (hit) t_out = t;
  return hit;
}

// Sphere test, op for op src/amber/scene/primitive_sphere.cc:75-107 with
// SolveQuadratic (include/amber/prelude/algebra.h:31-52): a = 1, b = -2 (c-o).d,
// c = |c-o|^2 - r^2; cancellation-safe roots; the nearer root if > kEPS, else the
// farther one if > kEPS.  A ray leaving a sphere re-tests that sphere with c equal to
// pure rounding noise, and whether the root clears kEPS depends on that noise: the
// arithmetic here is therefore the reference's, operation by operation, with no FMA
// contraction and IEEE sqrt / division, so t is bit-identical to the host's.
__device__ __forceinline__ bool intersect_sphere(const float4 sphere, const Ray& ray, float t_best,
                                                 float& t_out) {
  const float3 oc = v3(sphere.x - ray.o.x, sphere.y - ray.o.y, sphere.z - ray.o.z);
  const float b = -2.0f * dot_exact(oc, ray.d);
  const float c = __fsub_rn(dot_exact(oc, oc), __fmul_rn(sphere.w, sphere.w));
  const float disc = __fsub_rn(__fmul_rn(b, b), 4.0f * c);
  if (!(disc >= 0.0f)) return false;
  // Origin outside (c > 0) and the centre behind the ray (b > 0): both roots are negative in the
  // reference's arithmetic too (alpha = -b - sqrt_d < 0 keeps the larger magnitude, so its branch
  // gives beta = 2 c / alpha < 0), so the sqrt and the division are skipped -- same result.
  if (c > 0.0f && b > 0.0f) return false;
  const float sqrt_d = __fsqrt_rn(disc);
  float alpha = -b - sqrt_d;
  float beta = -b + sqrt_d;
  // SolveQuadratic's two branches (algebra.h:43-49) as ONE division: a warp whose lanes take both
  // sides would otherwise run the IEEE division sequence twice
  const bool by_beta = fabsf(alpha) < fabsf(beta);
  const float den = by_beta ? beta : alpha;
  const float quotient = __fdiv_rn(c, den) * 2.0f;
  const float half = den * 0.5f;
  alpha = by_beta ? quotient : half;
  beta = by_beta ? half : quotient;
  const float t = alpha > kEpsF ? alpha : beta;
  const bool hit = (t > kEpsF) & (t < t_best);
  if (hit) t_out = t;
  return hit;
}

// ---------------------------------------------------------------------------------
// ConstList: every lane walks the same primitive list in __constant__ memory.
// Closest-hit semantics of raytracer::List / a single-leaf BVH node
// (include/amber/raytracer/acceleration_list.h:51-68, acceleration_bvh.h:343-356).
struct ConstList {
  static constexpr bool kPersistent = false;
  // One plane group: the ray/plane intersection once, then two dot products per entry
  // (scene_device.cuh, PlaneGroup).  Acceptance is the reference's: inclusive edges, t > kEPS,
  // t < t_best (see intersect_triangle); a NaN or infinite t (n.d == 0) rejects the whole
  // group.  Returns the first accepting triangle of the group or -1 -- triangles of one group
  // share t, so after the first acceptance the strict `<` of the reference's closest-hit
  // update rejects the rest.
  static __device__ __forceinline__ int intersect_group(int g, const Ray& ray, float t_best, float& t_out) {
    const float4 n_gx = c_scene.group[g].n_gx;
    const float4 g_range = c_scene.group[g].g_range;
    const float3 T = v3(ray.o.x - n_gx.w, ray.o.y - g_range.x, ray.o.z - g_range.y);
    const float3 n = v3(n_gx);
    float inv_nd;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv_nd) : "f"(dot(n, ray.d)));
    const float t = -dot(n, T) * inv_nd;
    bool open = (t > kEpsF) & (t < t_best);
    const float3 h = v3(fmaf(t, ray.d.x, T.x), fmaf(t, ray.d.y, T.y), fmaf(t, ray.d.z, T.z));
    const int firsts = __float_as_int(g_range.z), counts = __float_as_int(g_range.w);
    int e = firsts & 0xffff, tri = firsts >> 16;
    const int quads_end = e + (counts & 0xffff), tris_end = quads_end + (counts >> 16);
    int found = -1;
    for (; e < quads_end; e++, tri += 2) {
      const float4 a = c_scene.entry[e].nu_cu, b = c_scene.entry[e].nv_cv;
      const float s = fmaf(a.x, h.x, fmaf(a.y, h.y, fmaf(a.z, h.z, a.w)));
      const float r = fmaf(b.x, h.x, fmaf(b.y, h.y, fmaf(b.z, h.z, b.w)));
      if (open & (s >= 0.0f) & (s <= 1.0f) & (r >= 0.0f) & (r <= 1.0f)) {
        found = tri + (s + r > 1.0f ? 1 : 0);
        open = false;
      }
    }
    for (; e < tris_end; e++, tri++) {
      const float4 a = c_scene.entry[e].nu_cu, b = c_scene.entry[e].nv_cv;
      const float u = fmaf(a.x, h.x, fmaf(a.y, h.y, fmaf(a.z, h.z, a.w)));
      const float v = fmaf(b.x, h.x, fmaf(b.y, h.y, fmaf(b.z, h.z, b.w)));
      if (open & (u >= 0.0f) & (v >= 0.0f) & (u + v <= 1.0f)) {
        found = tri;
        open = false;
      }
    }
    if (found >= 0) t_out = t;
    return found;
  }

  // The entries perpendicular to one axis (scene_device.cuh, AxisEntry).  ok / oa / ob (dk / da / db)
  // are the ray origin (direction) along that axis and the two others, inv_dk = 1 / dk.  Same
  // acceptance rules as intersect_group; updates the closest hit in place (kAnyHit: returns as
  // soon as anything is hit).
  template <bool kAnyHit>
  static __device__ __forceinline__ bool scan_axis(int first, int count, float ok, float oa, float ob, float inv_dk,
                                                   float da, float db, float& t_best, int& prim) {
    // the winning entry and its s + q are kept in registers; the triangle id is looked up once
    // after the loop instead of once per entry
    int best = -1;
    float best_sum = 0.0f;
#pragma unroll 1
    for (int e = first; e < first + count; e++) {
      const AxisEntry& E = c_scene.axis_entry[e];
      const float t = (E.r_k - ok) * inv_dk;
      const float ha = fmaf(t, da, oa - E.r_a), hb = fmaf(t, db, ob - E.r_b);
      const float s = fmaf(E.nu_a, ha, E.nu_b * hb);
      const float q = fmaf(E.nv_a, ha, E.nv_b * hb);
      const float sum = s + q;
      if ((t > kEpsF) & (t < t_best) & (s >= 0.0f) & (q >= 0.0f) & (s <= 1.0f) & (q <= 1.0f) & (sum <= E.limit)) {
        t_best = t;
        best = e;
        best_sum = sum;
        if (kAnyHit) return true;
      }
    }
    if (best < 0) return false;
    prim = c_scene.axis_tri[best] + (best_sum > 1.0f ? 1 : 0);
    return true;
  }

  // scan_axis over the PAIRED unclustered entries of one axis (scene_device.cuh, AxisPair): the arithmetic
  // of two entries in packed fp32 instructions, the acceptance tests one entry after the other in list
  // order (so the strict `<` of the closest-hit update sees them as the flat loop does).
  // `first_entry` is the index of the axis's first entry in axis_tri[].
  template <bool kAnyHit>
  static __device__ __forceinline__ bool scan_axis_pairs(int first_pair, int n_pairs, int first_entry, float ok,
                                                         float oa, float ob, float inv_dk, float da, float db,
                                                         float& t_best, int& prim) {
    const float2 nok2 = make_float2(-ok, -ok), oa2 = make_float2(oa, oa), ob2 = make_float2(ob, ob);
    const float2 inv2 = make_float2(inv_dk, inv_dk), da2 = make_float2(da, da), db2 = make_float2(db, db);
    int best = -1;
    float best_sum = 0.0f;
#pragma unroll 1
    for (int p = 0; p < n_pairs; p++) {
      const AxisPair& P = c_scene.axis_pair[first_pair + p];
      const float2 t = __fmul2_rn(__fadd2_rn(P.r_k, nok2), inv2);
      const float2 ha = __ffma2_rn(t, da2, __fadd2_rn(oa2, P.neg_r_a));
      const float2 hb = __ffma2_rn(t, db2, __fadd2_rn(ob2, P.neg_r_b));
      const float2 s = __ffma2_rn(P.nu_a, ha, __fmul2_rn(P.nu_b, hb));
      const float2 q = __ffma2_rn(P.nv_a, ha, __fmul2_rn(P.nv_b, hb));
      const float2 sum = __fadd2_rn(s, q);
      if ((t.x > kEpsF) & (t.x < t_best) & (s.x >= 0.0f) & (q.x >= 0.0f) & (s.x <= 1.0f) & (q.x <= 1.0f) &
          (sum.x <= P.limit.x)) {
        t_best = t.x;
        best = 2 * p;
        best_sum = sum.x;
        if (kAnyHit) return true;
      }
      if ((t.y > kEpsF) & (t.y < t_best) & (s.y >= 0.0f) & (q.y >= 0.0f) & (s.y <= 1.0f) & (q.y <= 1.0f) &
          (sum.y <= P.limit.y)) {
        t_best = t.y;
        best = 2 * p + 1;
        best_sum = sum.y;
        if (kAnyHit) return true;
      }
    }
    if (best < 0) return false;
    prim = c_scene.axis_tri[first_entry + best] + (best_sum > 1.0f ? 1 : 0);
    return true;
  }

  // All constant-memory triangles: axis-perpendicular entries, then the general plane groups.
  template <bool kAnyHit>
  static __device__ __forceinline__ bool scan_triangles(const Ray& ray, float& t_best, int& prim) {
    bool any = false;
    const int nx = c_scene.n_axis_entries[0], ny = c_scene.n_axis_entries[1], nz = c_scene.n_axis_entries[2];
    if (nx + ny + nz + c_scene.n_clusters > 0) {
      float ix, iy, iz;  // 1 / 0 = inf: t = +-inf or NaN rejects the plane, as n.d == 0 does elsewhere
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(ix) : "f"(ray.d.x));
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iy) : "f"(ray.d.y));
      asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iz) : "f"(ray.d.z));
      const int px = (nx + 1) >> 1, py = (ny + 1) >> 1, pz = (nz + 1) >> 1;  // pairs per axis
      any |= scan_axis_pairs<kAnyHit>(0, px, 0, ray.o.x, ray.o.y, ray.o.z, ix, ray.d.y, ray.d.z, t_best, prim);
      if (kAnyHit && any) return true;
      any |= scan_axis_pairs<kAnyHit>(px, py, nx, ray.o.y, ray.o.z, ray.o.x, iy, ray.d.z, ray.d.x, t_best, prim);
      if (kAnyHit && any) return true;
      any |= scan_axis_pairs<kAnyHit>(px + py, pz, nx + ny, ray.o.z, ray.o.x, ray.o.y, iz, ray.d.x, ray.d.y, t_best, prim);
      if (kAnyHit && any) return true;
      // clusters: one bounding-rectangle test per plane, entries only if some lane lands inside.
      // The vote only ever ADDS work (a lane that lands inside always scans), so which lanes take
      // part in it does not matter.
      const int n_clusters = c_scene.n_clusters;
#pragma unroll 1
      for (int c = 0; c < n_clusters; c++) {
        const AxisCluster& C = c_scene.cluster[c];
        const int axis = C.axis;
        const float ok = axis == 0 ? ray.o.x : axis == 1 ? ray.o.y : ray.o.z;
        const float oa = axis == 0 ? ray.o.y : axis == 1 ? ray.o.z : ray.o.x;
        const float ob = axis == 0 ? ray.o.z : axis == 1 ? ray.o.x : ray.o.y;
        const float ik = axis == 0 ? ix : axis == 1 ? iy : iz;
        const float da = axis == 0 ? ray.d.y : axis == 1 ? ray.d.z : ray.d.x;
        const float db = axis == 0 ? ray.d.z : axis == 1 ? ray.d.x : ray.d.y;
        const float t = (C.r_k - ok) * ik;
        const float pa = fmaf(t, da, oa), pb = fmaf(t, db, ob);
        const bool inside = (t > kEpsF) & (t < t_best) & (pa >= C.a_min) & (pa <= C.a_max) & (pb >= C.b_min) & (pb <= C.b_max);
        if (__any_sync(__activemask(), inside)) {
          any |= scan_axis<kAnyHit>(C.first, C.count, ok, oa, ob, ik, da, db, t_best, prim);
          if (kAnyHit && any) return true;
        }
      }
    }
    const int n_groups = c_scene.n_groups;
    for (int g = 0; g < n_groups; g++) {
      const int k = intersect_group(g, ray, t_best, t_best);
      if (k >= 0) {
        prim = k;
        any = true;
        if (kAnyHit) return true;
      }
    }
    return any;
  }

  static __device__ __forceinline__ HitRecord closest(const SceneView&, const Ray& ray, bool active) {
    HitRecord hit;
    if (!active) {
      hit.t = CUDART_NAN_F;
      hit.prim = -1;
      return hit;
    }
    hit.t = CUDART_MAX_NORMAL_F;  // Cast(ray) == Cast(ray, numeric_limits::max), acceleration.h:46-51
    hit.prim = -1;
    scan_triangles<false>(ray, hit.t, hit.prim);
    const int n_tris = c_scene.n_tris;
    const int n_spheres = c_scene.n_spheres;
#pragma unroll 1
    for (int i = 0; i < n_spheres; i++) {
      if (intersect_sphere(c_scene.sphere[i], ray, hit.t, hit.t)) hit.prim = n_tris + i;
    }
    if (hit.prim < 0) hit.t = CUDART_NAN_F;
    return hit;
  }

  // true if anything lies in (kEPS, t_max)
  static __device__ __forceinline__ bool occluded(const SceneView&, const Ray& ray, float t_max, bool active) {
    if (!active) return false;
    float t = t_max;
    int prim = -1;
    if (scan_triangles<true>(ray, t, prim)) return true;
    const int n_spheres = c_scene.n_spheres;
#pragma unroll 1
    for (int i = 0; i < n_spheres; i++) {
      if (intersect_sphere(c_scene.sphere[i], ray, t_max, t)) return true;
    }
    return false;
  }
};

}  // namespace amber_cuda
