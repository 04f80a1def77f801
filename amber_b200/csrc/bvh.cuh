// bvh.cuh -- 4-wide BVH: layout, device traversal.  The host builder is bvh_build.h.
//
// Replaces raytracer::BVH (include/amber/raytracer/acceleration_bvh.h): there a
// pointer-linked binary tree walked recursively; here a flat array of 64-byte
// 4-wide nodes with quantised child boxes walked with a per-lane stack in shared memory, the lanes of a
// warp kept together by votes (see traverse).  Tree shape
// is ours (binned SAH, collapsed 2 -> 4); only closest-hit semantics are shared:
// nearest t in (kEPS, t_best), strict `<` update (acceleration_bvh.h:346-351).
//
// Node = 4 float4 = 64 B = two sectors, read with two 256-bit loads.  Incoherent traversal is
// bound by the L1 gather path (every lane reads its own line; measured: lane-gathers cost in
// proportion to the SECTORS they touch, scripts/micro/gather.cu), so the node is as small as four
// children allow: child boxes are 8-bit grid indices on the node's own box,
//   plane = origin + q * scale   (lo rounded down, hi rounded up: conservative; bvh_build.h)
//   [0] origin.xyz, scale.x      [1] scale.yz, x-lo word, x-hi word (byte k = child k)
//   [2] y-lo, y-hi, z-lo, z-hi   [3] child references (int bits): >= 0 inner node index; < 0
//       leaf, ~ref = (first_triangle << 3) | triangle_count; kEmptyChild for an unused lane
// Triangles are stored in leaf order as TriIsect (3 float4).
#pragma once

#include "intersect.cuh"

namespace amber_cuda {

constexpr int kBvhStackMax = 160;  // entries per lane the 227 KB of shared memory allow at 256 threads + margin
// Vote thresholds of trace_persistent, measured on the 1.3 M-triangle scene (closest hit: 8/16
// 544 us, 20/12 592 us per launch; any hit: 20/12 286 us, 8/16 339 us -- shadow rays end early, so
// refilling often costs more than the idle lanes do).
#ifndef AMBER_BVH_REFILL_LANES
#define AMBER_BVH_REFILL_LANES 8
#endif
#ifndef AMBER_BVH_MIN_DESCEND_LANES
#define AMBER_BVH_MIN_DESCEND_LANES 16
#endif
#ifndef AMBER_BVH_REFILL_LANES_ANYHIT
#define AMBER_BVH_REFILL_LANES_ANYHIT 20
#endif
#ifndef AMBER_BVH_MIN_DESCEND_LANES_ANYHIT
#define AMBER_BVH_MIN_DESCEND_LANES_ANYHIT 12
#endif
#ifndef AMBER_BVH_MIN_BLOCKS
#define AMBER_BVH_MIN_BLOCKS 5  // resident 256-thread blocks per SM the traversal kernels are compiled for (<= 51 registers)
#endif
#ifndef AMBER_BVH_POSTPONE
#define AMBER_BVH_POSTPONE 0  // 1: a lane that reaches a leaf parks it and keeps descending (measured: no gain)
#endif
#ifndef AMBER_BVH_PREFETCH_L2
#define AMBER_BVH_PREFETCH_L2 0  // 1: when a node step finds a leaf child in range, its triangles are prefetched into L2
#endif
#ifndef AMBER_BVH_TRI_NO_ALLOCATE
#define AMBER_BVH_TRI_NO_ALLOCATE 0  // 1: triangle records bypass L1 allocation (they are rarely re-read; nodes are)
#endif
constexpr int kBvhLeafMax = 4;
constexpr int32_t kEmptyChild = 0x7fffffff;

// 256-bit read-only load (LDG.E.256, sm_100+): one L1 wavefront brings 32 B of a lane's node where a
// 128-bit load brings 16 -- the traversal is bound by L1 wavefronts (every lane reads its own line).
struct Wide8 {
  float4 a, b;
};
__device__ __forceinline__ Wide8 ldg256(const void* p) {
  Wide8 r;
  asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
      : "l"(p));
  return r;
}

// 128-bit read-only load that does not allocate in L1 (streaming data next to the re-used nodes).
__device__ __forceinline__ float4 ldg_no_allocate(const float4* p) {
  float4 r;
  asm("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
  return r;
}
__device__ __forceinline__ float4 load_triangle_word(const float4* p) {
  return AMBER_BVH_TRI_NO_ALLOCATE ? ldg_no_allocate(p) : __ldg(p);
}
__device__ __forceinline__ void prefetch_l2(const void* p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ int32_t bvh_leaf_first(int32_t ref) { return (~ref) >> 3; }
__device__ __forceinline__ int32_t bvh_leaf_count(int32_t ref) { return (~ref) & 7; }

struct WideBvh {
  // Byte k of a packed plane word as a float: PRMT builds 2^23 + q, the subtraction is exact.
  template <int k>
  static __device__ __forceinline__ float plane_magic(uint32_t word) {  // 2^23 + byte k
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440 | k));
  }
  template <int k>
  static __device__ __forceinline__ float plane_byte(uint32_t word) {
    return __uint_as_float(__byte_perm(word, 0x4B000000u, 0x7440 | k)) - 8388608.0f;
  }

  // Slab test of child k against [0, t_best] in the node's quantised frame: the plane at grid
  // index q lies at ray parameter q * s + o with s = scale * idir, o = (origin - O) * idir per
  // axis; `near`/`far` are the plane words already chosen by the sign of the direction, so no
  // min/max per axis is left.  Returns the entry distance or +inf.  fminf/fmaxf drop NaNs
  // (0 * inf on an axis-parallel ray), which only ever makes the test more permissive; boxes
  // are padded at build time so rounding never culls a true hit.
  template <int k>
  static __device__ __forceinline__ float child_entry(uint32_t near_x, uint32_t far_x, uint32_t near_y, uint32_t far_y,
                                                      uint32_t near_z, uint32_t far_z, const float3& s,
                                                      const float3& o, float t_best, int32_t ref) {
#ifdef AMBER_BVH_SCALAR_DECODE
    const float nx = fmaf(plane_byte<k>(near_x), s.x, o.x), fx = fmaf(plane_byte<k>(far_x), s.x, o.x);
    const float ny = fmaf(plane_byte<k>(near_y), s.y, o.y), fy = fmaf(plane_byte<k>(far_y), s.y, o.y);
    const float nz = fmaf(plane_byte<k>(near_z), s.z, o.z), fz = fmaf(plane_byte<k>(far_z), s.z, o.z);
#else
    // the near and far plane of an axis go through one packed FADD2 + FFMA2 (sm_100)
    const float2 magic = make_float2(-8388608.0f, -8388608.0f);
    const float2 x = __ffma2_rn(__fadd2_rn(make_float2(plane_magic<k>(near_x), plane_magic<k>(far_x)), magic),
                                make_float2(s.x, s.x), make_float2(o.x, o.x));
    const float2 y = __ffma2_rn(__fadd2_rn(make_float2(plane_magic<k>(near_y), plane_magic<k>(far_y)), magic),
                                make_float2(s.y, s.y), make_float2(o.y, o.y));
    const float2 z = __ffma2_rn(__fadd2_rn(make_float2(plane_magic<k>(near_z), plane_magic<k>(far_z)), magic),
                                make_float2(s.z, s.z), make_float2(o.z, o.z));
    const float nx = x.x, fx = x.y, ny = y.x, fy = y.y, nz = z.x, fz = z.y;
#endif
    const float t_near = fmaxf(fmaxf(nx, ny), fmaxf(nz, 0.0f));
    const float t_far = fminf(fminf(fx, fy), fminf(fz, t_best));
    return (t_near <= t_far) & (ref != kEmptyChild) ? t_near : CUDART_INF_F;
  }

  // One node step of one lane: four slab tests, children sorted by entry distance, the far
  // ones pushed (farthest first); returns the next reference to visit (kEmptyChild = done).
  static __device__ __forceinline__ int32_t node_step(const SceneView& scene, int32_t current, const float3& idir,
                                                      const float3& oidir, float t_best, int32_t* stack, int stride,
                                                      int& sp) {
    const float4* node = scene.bvh_nodes + static_cast<size_t>(current) * 4;
    const Wide8 head = ldg256(node), tail = ldg256(node + 2);
    const float3 s = v3(head.a.w * idir.x, head.b.x * idir.y, head.b.y * idir.z);
    const float3 o = v3(fmaf(head.a.x, idir.x, -oidir.x), fmaf(head.a.y, idir.y, -oidir.y),
                        fmaf(head.a.z, idir.z, -oidir.z));
    const uint32_t x_lo = __float_as_uint(head.b.z), x_hi = __float_as_uint(head.b.w);
    const uint32_t y_lo = __float_as_uint(tail.a.x), y_hi = __float_as_uint(tail.a.y);
    const uint32_t z_lo = __float_as_uint(tail.a.z), z_hi = __float_as_uint(tail.a.w);
    const bool rx = idir.x < 0.0f, ry = idir.y < 0.0f, rz = idir.z < 0.0f;
    const uint32_t near_x = rx ? x_hi : x_lo, far_x = rx ? x_lo : x_hi;
    const uint32_t near_y = ry ? y_hi : y_lo, far_y = ry ? y_lo : y_hi;
    const uint32_t near_z = rz ? z_hi : z_lo, far_z = rz ? z_lo : z_hi;
    const float4 ref4 = tail.b;
    int32_t c0 = __float_as_int(ref4.x), c1 = __float_as_int(ref4.y), c2 = __float_as_int(ref4.z),
            c3 = __float_as_int(ref4.w);
    float d0 = child_entry<0>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c0);
    float d1 = child_entry<1>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c1);
    float d2 = child_entry<2>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c2);
    float d3 = child_entry<3>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c3);
    if (AMBER_BVH_PREFETCH_L2) {
      // a leaf child in range will be visited unless a nearer hit cuts the ray short first: start moving its
      // triangles from DRAM to L2 now (the leaf phase's loads are the traversal's longest stalls)
      const int32_t refs[4] = {c0, c1, c2, c3};
      const float dist[4] = {d0, d1, d2, d3};
#pragma unroll
      for (int k = 0; k < 4; k++) {
        if (refs[k] < 0 && dist[k] < CUDART_INF_F) {
          const TriIsect* tri = scene.tri_isect + bvh_leaf_first(refs[k]);
          prefetch_l2(tri);
          if (bvh_leaf_count(refs[k]) > 2) prefetch_l2(tri + 2);
        }
      }
    }
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
  // kCount (profiling mode): node steps and triangle tests are counted per lane and added to
  // counters[0] / counters[1] once per warp at the end -- the traffic floor of the roofline.
  template <bool kAnyHit, bool kCount = false, class Source>
  static __device__ __forceinline__ void trace_persistent(const SceneView& scene, Source& source,
                                                          unsigned int* cursor,
                                                          unsigned long long* counters = nullptr) {
    extern __shared__ int32_t bvh_stack_smem[];
    int32_t* stack = bvh_stack_smem + threadIdx.x;
    const int stride = blockDim.x;
    constexpr uint32_t kFull = 0xffffffffu;
    constexpr int kRefillLanes = kAnyHit ? AMBER_BVH_REFILL_LANES_ANYHIT : AMBER_BVH_REFILL_LANES;
    constexpr int kMinDescendLanes = kAnyHit ? AMBER_BVH_MIN_DESCEND_LANES_ANYHIT : AMBER_BVH_MIN_DESCEND_LANES;
    const int lane = threadIdx.x & 31;
    const uint32_t lane_lt = (1u << lane) - 1u;
    const uint32_t n = source.count();

    int sp = 0;
    int32_t current = kEmptyChild;  // kEmptyChild doubles as "idle"
    // Speculative traversal (Aila & Laine): a lane that reaches a leaf parks it here and goes on with the
    // next stack entry instead of waiting for the warp's leaf phase; the warp tests triangles when enough
    // lanes hold a parked leaf, so a leaf phase runs with many lanes instead of the few that happened to
    // reach a leaf in the same step.  The price: the parked leaf's hit does not shorten the ray for the
    // node steps taken in between.
    int32_t parked = kEmptyChild;
    uint32_t item = 0;
    bool have_item = false, found = false, exhausted = false;
    Ray ray;
    ray.o = v3(0.0f, 0.0f, 0.0f);
    ray.d = v3(0.0f, 0.0f, 1.0f);
    float3 idir = v3(0.0f, 0.0f, 0.0f), oidir = idir;
    HitRecord hit;
    hit.t = 0.0f;
    hit.prim = -1;
    uint32_t node_steps = 0, triangle_tests = 0;

    while (true) {
      // ---- fetch ----------------------------------------------------------------------
      const bool idle = current == kEmptyChild && parked == kEmptyChild;
      const uint32_t idle_mask = __ballot_sync(kFull, idle);
      // (Measured and dropped: a per-warp reserved range that grows after refills that found no ray, so that sweeping a
      // draining pool costs fewer atomics -- the extra bookkeeping cost the steady state 4 %: 51.6 -> 53.9 ms per 16 spp,
      // and a short job gained nothing: 29.7 -> 30.9 ms per 8 spp.)
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
        if (AMBER_BVH_POSTPONE && current < 0 && parked == kEmptyChild) {
          parked = current;
          if (sp > 0) {
            sp--;
            current = stack[sp * stride];
          } else {
            current = kEmptyChild;
          }
        }
        const bool inner = static_cast<uint32_t>(current) < static_cast<uint32_t>(kEmptyChild);
        const uint32_t inner_lanes = __ballot_sync(kFull, inner);
        if (inner_lanes == 0) break;
        const bool has_leaf = AMBER_BVH_POSTPONE ? parked != kEmptyChild : current < 0;
        if (__popc(inner_lanes) < kMinDescendLanes && __any_sync(kFull, has_leaf)) break;
        if (inner) {
          current = node_step(scene, current, idir, oidir, hit.t, stack, stride, sp);
          if (kCount) node_steps++;
        }
      }
      // ---- leaves ---------------------------------------------------------------------
      // (Dealing the (lane, triangle) pairs of a leaf phase out over the whole warp by shuffle --
      // one pass of ~20 busy lanes for four passes of ~5 -- was measured slower: 570 vs 544 us.)
      const int32_t leaf = AMBER_BVH_POSTPONE ? parked : current;
      if (leaf < 0) {
        const int first = bvh_leaf_first(leaf), count = bvh_leaf_count(leaf);
        if (kCount) triangle_tests += count;
#pragma unroll
        for (int k = 0; k < kBvhLeafMax; k++) {
          if (k < count) {
            const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + first + k);
            if (intersect_triangle(load_triangle_word(tri), load_triangle_word(tri + 1), load_triangle_word(tri + 2), ray, hit.t, hit.t)) {
              hit.prim = first + k;
              found = true;
            }
          }
        }
        if (AMBER_BVH_POSTPONE) {
          parked = kEmptyChild;
          if (kAnyHit && found) {
            sp = 0;
            current = kEmptyChild;
          }
        } else if (kAnyHit && found) {
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
    if (kCount) {
      node_steps = __reduce_add_sync(kFull, node_steps);
      triangle_tests = __reduce_add_sync(kFull, triangle_tests);
      if (lane == 0) {
        atomicAdd(&counters[0], static_cast<unsigned long long>(node_steps));
        atomicAdd(&counters[1], static_cast<unsigned long long>(triangle_tests));
      }
    }
  }

  static constexpr bool kPersistent = true;
};

}  // namespace amber_cuda
