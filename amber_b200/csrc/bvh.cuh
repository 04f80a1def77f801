// bvh.cuh -- 4-wide BVH: layout, device traversal.  The host builder is bvh_build.h.
//
// Replaces raytracer::BVH (include/amber/raytracer/acceleration_bvh.h): there a
// pointer-linked binary tree walked recursively; here a flat array of 128-byte
// 4-wide nodes in SoA form walked with a per-lane stack in shared memory.  Tree shape
// is ours (binned SAH, collapsed 2 -> 4); only closest-hit semantics are shared:
// nearest t in (kEPS, t_best), strict `<` update (acceleration_bvh.h:346-351).
//
// Node = 8 float4 (one 128 B line):
//   [0] lo.x of children 0..3   [1] hi.x   [2] lo.y   [3] hi.y   [4] lo.z   [5] hi.z
//   [6] child reference (int bits): >= 0 inner node index; < 0 leaf, ~ref =
//       (first_triangle << 3) | triangle_count; kEmptyChild for an unused lane
//   [7] unused
// Triangles are stored in leaf order as TriIsect (3 float4).
#pragma once

#include "intersect.cuh"

namespace amber_cuda {

constexpr int kBvhStack = 48;
constexpr int kBvhLeafMax = 4;
constexpr int32_t kEmptyChild = 0x7fffffff;

__device__ __forceinline__ int32_t bvh_leaf_first(int32_t ref) { return (~ref) >> 3; }
__device__ __forceinline__ int32_t bvh_leaf_count(int32_t ref) { return (~ref) & 7; }

struct WideBvh {
  // Slab test of one child against [0, t_best]; returns entry distance or +inf.
  // fminf/fmaxf drop NaNs (0 * inf), which only ever makes the test more permissive;
  // boxes are padded at build time so rounding never culls a true hit.
  static __device__ __forceinline__ float child_entry(float lox, float hix, float loy, float hiy, float loz,
                                                      float hiz, const float3& idir, const float3& oidir,
                                                      float t_best) {
    const float tx0 = fmaf(lox, idir.x, -oidir.x), tx1 = fmaf(hix, idir.x, -oidir.x);
    const float ty0 = fmaf(loy, idir.y, -oidir.y), ty1 = fmaf(hiy, idir.y, -oidir.y);
    const float tz0 = fmaf(loz, idir.z, -oidir.z), tz1 = fmaf(hiz, idir.z, -oidir.z);
    const float t_near = fmaxf(fmaxf(fminf(tx0, tx1), fminf(ty0, ty1)), fmaxf(fminf(tz0, tz1), 0.0f));
    const float t_far = fminf(fminf(fmaxf(tx0, tx1), fmaxf(ty0, ty1)), fminf(fmaxf(tz0, tz1), t_best));
    return t_near <= t_far ? t_near : CUDART_INF_F;
  }

  template <bool kAnyHit>
  static __device__ __forceinline__ bool traverse(const SceneView& scene, const Ray& ray, HitRecord& hit) {
    // per-lane stack, column layout: entry k of thread t at stack[k * blockDim + t]
    extern __shared__ int32_t bvh_stack_smem[];
    int32_t* stack = bvh_stack_smem + threadIdx.x;
    const int stride = blockDim.x;
    int sp = 0;

    const float3 idir = v3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    const float3 oidir = v3(ray.o.x * idir.x, ray.o.y * idir.y, ray.o.z * idir.z);
    int32_t current = scene.bvh_root;
    if (current == kEmptyChild) return false;

    while (true) {
      if (current >= 0) {
        const float4* node = scene.bvh_nodes + static_cast<size_t>(current) * 8;
        const float4 lox = __ldg(node + 0), hix = __ldg(node + 1), loy = __ldg(node + 2), hiy = __ldg(node + 3);
        const float4 loz = __ldg(node + 4), hiz = __ldg(node + 5), ref4 = __ldg(node + 6);
        float d0 = child_entry(lox.x, hix.x, loy.x, hiy.x, loz.x, hiz.x, idir, oidir, hit.t);
        float d1 = child_entry(lox.y, hix.y, loy.y, hiy.y, loz.y, hiz.y, idir, oidir, hit.t);
        float d2 = child_entry(lox.z, hix.z, loy.z, hiy.z, loz.z, hiz.z, idir, oidir, hit.t);
        float d3 = child_entry(lox.w, hix.w, loy.w, hiy.w, loz.w, hiz.w, idir, oidir, hit.t);
        int32_t c0 = __float_as_int(ref4.x), c1 = __float_as_int(ref4.y), c2 = __float_as_int(ref4.z),
                c3 = __float_as_int(ref4.w);
        // sort the four (distance, child) pairs ascending: 5-comparator network
#define AMBER_CSWAP(da, ca, db, cb)            \
  if (db < da) {                               \
    const float td = da; da = db; db = td;     \
    const int32_t tc = ca; ca = cb; cb = tc;   \
  }
        AMBER_CSWAP(d0, c0, d1, c1) AMBER_CSWAP(d2, c2, d3, c3) AMBER_CSWAP(d0, c0, d2, c2)
        AMBER_CSWAP(d1, c1, d3, c3) AMBER_CSWAP(d1, c1, d2, c2)
#undef AMBER_CSWAP
        // push far ones first, continue with the nearest
        if (d3 < CUDART_INF_F) { stack[sp * stride] = c3; sp++; }
        if (d2 < CUDART_INF_F) { stack[sp * stride] = c2; sp++; }
        if (d1 < CUDART_INF_F) { stack[sp * stride] = c1; sp++; }
        if (d0 < CUDART_INF_F) {
          current = c0;
          continue;
        }
      } else {
        const int first = bvh_leaf_first(current), count = bvh_leaf_count(current);
        for (int k = 0; k < count; k++) {
          const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + first + k);
          if (intersect_triangle(__ldg(tri), __ldg(tri + 1), __ldg(tri + 2), ray, hit.t, hit.t)) {
            hit.prim = first + k;
            if (kAnyHit) return true;
          }
        }
      }
      if (sp == 0) break;
      sp--;
      current = stack[sp * stride];
    }
    return hit.prim >= 0;
  }

  static __device__ __forceinline__ HitRecord closest(const SceneView& scene, const Ray& ray) {
    // spheres (and any constant-resident triangles) first: shrinks t_best for the walk
    HitRecord hit = ConstList::closest(scene, ray);
    if (hit.prim < 0) hit.t = CUDART_MAX_NORMAL_F;
    const int sphere_hit = hit.prim;
    hit.prim = -1;
    traverse<false>(scene, ray, hit);
    if (hit.prim < 0) hit.prim = sphere_hit >= 0 ? scene.n_tris + sphere_hit : -1;
    if (hit.prim < 0) hit.t = CUDART_NAN_F;
    return hit;
  }

  static __device__ __forceinline__ bool occluded(const SceneView& scene, const Ray& ray, float t_max) {
    if (ConstList::occluded(scene, ray, t_max)) return true;
    HitRecord hit;
    hit.t = t_max;
    hit.prim = -1;
    return traverse<true>(scene, ray, hit);
  }
};

}  // namespace amber_cuda
