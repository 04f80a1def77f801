// bvh.cuh -- 4-wide BVH: layout, device traversal.  The host builder is bvh_build.h.
//
// Replaces raytracer::BVH (include/amber/raytracer/acceleration_bvh.h): there a
// pointer-linked binary tree walked recursively; here a flat array of 128-byte
// 4-wide nodes in SoA form walked with a per-lane stack in shared memory, the lanes of a
// warp kept together by votes (see traverse).  Tree shape
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

constexpr int kBvhStackMax = 160;  // entries per lane the 227 KB of shared memory allow at 256 threads + margin
#ifndef AMBER_BVH_REFILL_LANES
#define AMBER_BVH_REFILL_LANES 20
#endif
#ifndef AMBER_BVH_MIN_DESCEND_LANES
#define AMBER_BVH_MIN_DESCEND_LANES 12
#endif
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

  // One node step of one lane: four slab tests, children sorted by entry distance, the far
  // ones pushed (farthest first); returns the next reference to visit (kEmptyChild = done).
  static __device__ __forceinline__ int32_t node_step(const SceneView& scene, int32_t current, const float3& idir,
                                                      const float3& oidir, float t_best, int32_t* stack, int stride,
                                                      int& sp) {
    const float4* node = scene.bvh_nodes + static_cast<size_t>(current) * 8;
    const float4 lox = __ldg(node + 0), hix = __ldg(node + 1), loy = __ldg(node + 2), hiy = __ldg(node + 3);
    const float4 loz = __ldg(node + 4), hiz = __ldg(node + 5), ref4 = __ldg(node + 6);
    float d0 = child_entry(lox.x, hix.x, loy.x, hiy.x, loz.x, hiz.x, idir, oidir, t_best);
    float d1 = child_entry(lox.y, hix.y, loy.y, hiy.y, loz.y, hiz.y, idir, oidir, t_best);
    float d2 = child_entry(lox.z, hix.z, loy.z, hiy.z, loz.z, hiz.z, idir, oidir, t_best);
    float d3 = child_entry(lox.w, hix.w, loy.w, hiy.w, loz.w, hiz.w, idir, oidir, t_best);
    int32_t c0 = __float_as_int(ref4.x), c1 = __float_as_int(ref4.y), c2 = __float_as_int(ref4.z),
            c3 = __float_as_int(ref4.w);
    // sort the four (distance, child) pairs ascending: 5-comparator network, branch-free
#define AMBER_CSWAP(da, ca, db, cb)            \
  {                                            \
    const bool swap = db < da;                 \
    const float td = swap ? db : da;           \
    const int32_t tc = swap ? cb : ca;         \
    db = swap ? da : db;                       \
    cb = swap ? ca : cb;                       \
    da = td;                                   \
    ca = tc;                                   \
  }
    AMBER_CSWAP(d0, c0, d1, c1) AMBER_CSWAP(d2, c2, d3, c3) AMBER_CSWAP(d0, c0, d2, c2)
    AMBER_CSWAP(d1, c1, d3, c3) AMBER_CSWAP(d1, c1, d2, c2)
#undef AMBER_CSWAP
    if (d3 < CUDART_INF_F) { stack[sp * stride] = c3; sp++; }
    if (d2 < CUDART_INF_F) { stack[sp * stride] = c2; sp++; }
    if (d1 < CUDART_INF_F) { stack[sp * stride] = c1; sp++; }
    if (d0 < CUDART_INF_F) return c0;
    if (sp > 0) {
      sp--;
      return stack[sp * stride];
    }
    return kEmptyChild;
  }

  // Persistent warp-converged traversal with dynamic ray fetch (Aila & Laine's "while-while"
  // with replacement, restated for independent thread scheduling).  A warp owns 32 ray states;
  // the loop alternates three phases, each opened by a warp vote (a reconvergence point):
  //   fetch  : when >= kRefillLanes lanes are idle, they store their finished hits and take the
  //            next rays from a global cursor (one atomic per warp per refill)
  //   descend: lanes holding an inner node do node steps while enough lanes still hold one
  //   leaves : lanes holding a leaf test its <= 4 triangles, then pop
  // Rays of one warp differ wildly in length on an incoherent wavefront (5 of 32 lanes per
  // instruction without replacement, under ncu on the 1.3 M-triangle scene); refilling idle
  // lanes keeps the SIMT width used.
  //   Source: uint32_t count(); bool load(i, Ray&, float& t_best) (false: nothing to trace,
  //   the source has written that item's result itself); void store(i, hit, found).
  template <bool kAnyHit, class Source>
  static __device__ __forceinline__ void trace_persistent(const SceneView& scene, Source& source,
                                                          unsigned int* cursor) {
    extern __shared__ int32_t bvh_stack_smem[];
    int32_t* stack = bvh_stack_smem + threadIdx.x;
    const int stride = blockDim.x;
    constexpr uint32_t kFull = 0xffffffffu;
    constexpr int kRefillLanes = AMBER_BVH_REFILL_LANES;
    constexpr int kMinDescendLanes = AMBER_BVH_MIN_DESCEND_LANES;
    const int lane = threadIdx.x & 31;
    const uint32_t lane_lt = (1u << lane) - 1u;
    const uint32_t n = source.count();

    int sp = 0;
    int32_t current = kEmptyChild;  // kEmptyChild doubles as "idle"
    uint32_t item = 0;
    bool have_item = false, found = false, exhausted = false;
    Ray ray;
    ray.o = v3(0.0f, 0.0f, 0.0f);
    ray.d = v3(0.0f, 0.0f, 1.0f);
    float3 idir = v3(0.0f, 0.0f, 0.0f), oidir = idir;
    HitRecord hit;
    hit.t = 0.0f;
    hit.prim = -1;

    while (true) {
      // ---- fetch ----------------------------------------------------------------------
      const bool idle = current == kEmptyChild;
      const uint32_t idle_mask = __ballot_sync(kFull, idle);
      if (idle_mask == kFull && exhausted) break;
      if (!exhausted && (__popc(idle_mask) >= kRefillLanes || idle_mask == kFull)) {
        if (idle && have_item) {
          source.store(item, hit, found);
          have_item = false;
        }
        const uint32_t wanted = __popc(idle_mask);
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(cursor, wanted);
        base = __shfl_sync(kFull, base, 0);
        exhausted = base + wanted >= n;
        if (idle) {
          const uint32_t mine = base + __popc(idle_mask & lane_lt);
          float t_best = CUDART_MAX_NORMAL_F;
          if (mine < n && source.load(mine, ray, t_best)) {
            item = mine;
            have_item = true;
            found = false;
            hit.t = t_best;
            hit.prim = -1;
            sp = 0;
            idir = v3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
            oidir = v3(ray.o.x * idir.x, ray.o.y * idir.y, ray.o.z * idir.z);
            current = scene.bvh_root;
          }
        }
      }
      // ---- descend --------------------------------------------------------------------
      while (true) {
        const bool inner = static_cast<uint32_t>(current) < static_cast<uint32_t>(kEmptyChild);
        const uint32_t inner_lanes = __ballot_sync(kFull, inner);
        if (inner_lanes == 0) break;
        if (__popc(inner_lanes) < kMinDescendLanes && __any_sync(kFull, current < 0)) break;
        if (inner) current = node_step(scene, current, idir, oidir, hit.t, stack, stride, sp);
      }
      // ---- leaves ---------------------------------------------------------------------
      if (current < 0) {
        const int first = bvh_leaf_first(current), count = bvh_leaf_count(current);
#pragma unroll
        for (int k = 0; k < kBvhLeafMax; k++) {
          if (k < count) {
            const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + first + k);
            if (intersect_triangle(__ldg(tri), __ldg(tri + 1), __ldg(tri + 2), ray, hit.t, hit.t)) {
              hit.prim = first + k;
              found = true;
            }
          }
        }
        if (kAnyHit && found) {
          sp = 0;
          current = kEmptyChild;
        } else if (sp > 0) {
          sp--;
          current = stack[sp * stride];
        } else {
          current = kEmptyChild;
        }
      }
    }
    if (have_item) source.store(item, hit, found);
  }

  static constexpr bool kPersistent = true;
};

}  // namespace amber_cuda
