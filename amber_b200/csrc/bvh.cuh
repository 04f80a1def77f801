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
//   plane = origin + g(code) * scale   (g: bvh_build.h plane_coordinate; lo rounded down, hi rounded up: conservative)
//   [0] origin.xyz, scale.x      [1] scale.yz, x-lo word, x-hi word (byte k = child k)
//   [2] y-lo, y-hi, z-lo, z-hi   [3] child references (int bits): >= 0 inner node index; < 0
//       leaf, ~ref = (first_triangle << 3) | triangle_count; kEmptyChild for an unused lane
// Triangles are stored in leaf order as TriIsect (3 float4).
#pragma once

#include "intersect.cuh"

namespace amber_cuda {

#ifndef AMBER_BVH_WIDTH
#define AMBER_BVH_WIDTH 4
#endif
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
#ifndef AMBER_BVH_LEAF_PREFETCH
#define AMBER_BVH_LEAF_PREFETCH 0  // 8-wide: 1 = L1 prefetch of a triangle group's later triangles at the start of the leaf phase
#endif
#ifndef AMBER_BVH_LEAF_PAIRS
#define AMBER_BVH_LEAF_PAIRS 0  // 8-wide: 1 = the leaf phase loads two triangles per round
#endif
#ifndef AMBER_BVH_TRI_NO_ALLOCATE
#define AMBER_BVH_TRI_NO_ALLOCATE 0  // 1: triangle records bypass L1 allocation (they are rarely re-read; nodes are)
#endif
// Phase policy of trace_persistent.  0: descend while >= kMinDescendLanes lanes hold an inner node, test leaves as soon as fewer
// do and ANY lane holds a leaf, refill only between the two.  1: one vote per step picks the phase -- refill as soon as
// kRefillLanes lanes are idle, test leaves only once kLeafLanes lanes hold one (or nothing else is left to do), else descend.
#ifndef AMBER_BVH_POLICY
#define AMBER_BVH_POLICY 0
#endif
#ifndef AMBER_BVH_LEAF_LANES
#define AMBER_BVH_LEAF_LANES 12
#endif
#ifndef AMBER_BVH_LEAF_LANES_ANYHIT
#define AMBER_BVH_LEAF_LANES_ANYHIT 12
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

// Plane code -> grid coordinate (bvh_build.h, plane_coordinate): byte k of a packed plane word becomes bits 16-23 of
// the float 0x3F000000 | code << 16.  `magic` = 0x3F000000 must arrive in a REGISTER the compiler cannot see through (a
// word of the node, a constant-bank load): PRMT takes one immediate, and with the selector as the immediate the
// instruction stands alone -- with a known magic the compiler makes THAT the immediate and spends a move on the
// selector in front of every permute (48 moves per 8-wide node step).
template <int k>
__device__ __forceinline__ float plane_float(uint32_t word, uint32_t magic) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(word), "r"(magic), "n"(0x7044 | (k << 8)));
  return __uint_as_float(r);
}

#if AMBER_BVH_WIDTH == 4
struct WideBvh {
  // Slab test of child k against [0, t_best] in the node's quantised frame: the plane with grid
  // coordinate g lies at ray parameter g * s + o with s = scale * idir, o = (origin - O) * idir per
  // axis; `near`/`far` are the plane words already chosen by the sign of the direction, so no
  // min/max per axis is left.  Returns the entry distance or +inf.  fminf/fmaxf drop NaNs
  // (0 * inf on an axis-parallel ray), which only ever makes the test more permissive; boxes
  // are padded at build time so rounding never culls a true hit.
  template <int k>
  static __device__ __forceinline__ float child_entry(uint32_t near_x, uint32_t far_x, uint32_t near_y, uint32_t far_y,
                                                      uint32_t near_z, uint32_t far_z, const float3& s,
                                                      const float3& o, float t_best, int32_t ref, uint32_t magic) {
    // the near and far plane of an axis go through one packed FFMA2 (sm_100)
    const float2 x = __ffma2_rn(make_float2(plane_float<k>(near_x, magic), plane_float<k>(far_x, magic)),
                                make_float2(s.x, s.x), make_float2(o.x, o.x));
    const float2 y = __ffma2_rn(make_float2(plane_float<k>(near_y, magic), plane_float<k>(far_y, magic)),
                                make_float2(s.y, s.y), make_float2(o.y, o.y));
    const float2 z = __ffma2_rn(make_float2(plane_float<k>(near_z, magic), plane_float<k>(far_z, magic)),
                                make_float2(s.z, s.z), make_float2(o.z, o.z));
    const float t_near = fmaxf(fmaxf(x.x, y.x), fmaxf(z.x, 0.0f));
    const float t_far = fminf(fminf(x.y, y.y), fminf(z.y, t_best));
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
    const uint32_t magic = c_scene.bvh_plane_magic;  // (an opaque register: see plane_float)
    float d0 = child_entry<0>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c0, magic);
    float d1 = child_entry<1>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c1, magic);
    float d2 = child_entry<2>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c2, magic);
    float d3 = child_entry<3>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, c3, magic);
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

#if AMBER_BVH_POLICY
    constexpr int kLeafLanes = kAnyHit ? AMBER_BVH_LEAF_LANES_ANYHIT : AMBER_BVH_LEAF_LANES;
    while (true) {
      const bool inner = static_cast<uint32_t>(current) < static_cast<uint32_t>(kEmptyChild);
      const uint32_t inner_mask = __ballot_sync(kFull, inner);
      const uint32_t leaf_mask = __ballot_sync(kFull, current < 0);
      const uint32_t idle_mask = ~(inner_mask | leaf_mask);
      if (idle_mask == kFull && exhausted) break;
      if (!exhausted && (__popc(idle_mask) >= kRefillLanes)) {
        // ---- fetch --------------------------------------------------------------------
        const bool idle = current == kEmptyChild;
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
        continue;
      }
      if (__popc(leaf_mask) >= kLeafLanes || inner_mask == 0) {
        // ---- leaves -------------------------------------------------------------------
        if (current < 0) {
          const int first = bvh_leaf_first(current), count = bvh_leaf_count(current);
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
        continue;
      }
      // ---- descend --------------------------------------------------------------------
      if (inner) {
        current = node_step(scene, current, idir, oidir, hit.t, stack, stride, sp);
        if (kCount) node_steps++;
      }
    }
#else
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
#endif
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

#else
// ---- 8-wide ------------------------------------------------------------------------------------------------
// Node = 6 float4 = 96 B = three sectors, read with three 256-bit loads (host side: bvh_build.h, build8):
//   words 0-2 origin, 3-5 scale                  (child planes = origin + g(code) * scale, 8 bits per plane, as above)
//   word  6   child_base: node index of the first INNER child; inner children are consecutive in slot order
//   word  7   tri_base: index of the first triangle of the first LEAF child; leaf triangles are consecutive
//   words 8-9 meta, byte k = child slot k: inner (1 << 5) | (24 + k); leaf (count one-bits << 5) | offset from
//             tri_base (count <= 3, offset + count <= 24); empty 0
//   word  10  imask: bit k set = slot k holds an inner child          word 11 0x3F000000 (plane_float)
//   words 12-23 planes: x lo [slots 0-3][4-7], x hi, y lo, y hi, z lo, z hi
// One node step yields a 32-bit HIT MASK: bits 0-23 the triangles of the leaf children the ray enters (bit =
// offset from tri_base), bits 24-31 the inner children it enters at bit 24 + (slot ^ octinv), octinv = 7 ^ the
// ray's direction octant -- the builder puts children into the slot that matches their position in the node, so the
// HIGHEST set bit is the nearest child for this ray: no distances are compared, nothing is sorted (Ylitie, Karras,
// Laine, "Efficient incoherent ray traversal on GPUs through compressed wide BVHs", 2017; restated for our node).
// The traversal state of a ray is a NODE GROUP (child_base, imask << 8 | pending inner children) and a TRIANGLE
// GROUP (tri_base, pending triangles); a node step pushes at most ONE stack entry -- the rest of the group it took
// its node from -- so the per-lane stack is `depth` 8-byte entries instead of 3 * depth 4-byte ones.
__device__ __forceinline__ uint32_t sign_extend_s8x4(uint32_t x) {  // each byte: 0xff if its top bit is set, else 0
  uint32_t r;
  asm("prmt.b32 %0, %1, 0, 0x0000ba98;" : "=r"(r) : "r"(x));
  return r;
}

struct WideBvh {
  // Slab test of the child in byte k of the plane words (see the 4-wide child_entry): does the ray enter it
  // within [0, t_best]?  An empty slot has lo = 255 > hi = 0 on every axis and fails.
  template <int k>
  static __device__ __forceinline__ bool child_hit(uint32_t near_x, uint32_t far_x, uint32_t near_y, uint32_t far_y,
                                                   uint32_t near_z, uint32_t far_z, const float3& s, const float3& o,
                                                   float t_best, uint32_t magic) {
    const float2 x = __ffma2_rn(make_float2(plane_float<k>(near_x, magic), plane_float<k>(far_x, magic)),
                                make_float2(s.x, s.x), make_float2(o.x, o.x));
    const float2 y = __ffma2_rn(make_float2(plane_float<k>(near_y, magic), plane_float<k>(far_y, magic)),
                                make_float2(s.y, s.y), make_float2(o.y, o.y));
    const float2 z = __ffma2_rn(make_float2(plane_float<k>(near_z, magic), plane_float<k>(far_z, magic)),
                                make_float2(s.z, s.z), make_float2(o.z, o.z));
    const float t_near = fmaxf(fmaxf(x.x, y.x), fmaxf(z.x, 0.0f));
    const float t_far = fminf(fminf(x.y, y.y), fminf(z.y, t_best));
    return t_near <= t_far;
  }

  // Four children (one word of each plane kind): OR their bits into the hit mask.
  static __device__ __forceinline__ uint32_t four_children(uint32_t lo_x, uint32_t hi_x, uint32_t lo_y, uint32_t hi_y,
                                                           uint32_t lo_z, uint32_t hi_z, bool rx, bool ry, bool rz,
                                                           const float3& s, const float3& o, float t_best, uint32_t meta4,
                                                           uint32_t octinv4, uint32_t magic) {
    const uint32_t near_x = rx ? hi_x : lo_x, far_x = rx ? lo_x : hi_x;
    const uint32_t near_y = ry ? hi_y : lo_y, far_y = ry ? lo_y : hi_y;
    const uint32_t near_z = rz ? hi_z : lo_z, far_z = rz ? lo_z : hi_z;
    // byte-parallel: which children are inner (index >= 24: bits 3 and 4 set), where their bits go, which bits
    const uint32_t is_inner4 = (meta4 & (meta4 << 1)) & 0x10101010u;
    const uint32_t index4 = (meta4 ^ (octinv4 & sign_extend_s8x4(is_inner4 << 3))) & 0x1f1f1f1fu;
    const uint32_t bits4 = (meta4 >> 5) & 0x07070707u;
    uint32_t mask = 0;
    if (child_hit<0>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, magic)) mask |= (bits4 & 0xffu) << (index4 & 0xffu);
    if (child_hit<1>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, magic)) mask |= __byte_perm(bits4, 0, 0x4441) << __byte_perm(index4, 0, 0x4441);
    if (child_hit<2>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, magic)) mask |= __byte_perm(bits4, 0, 0x4442) << __byte_perm(index4, 0, 0x4442);
    if (child_hit<3>(near_x, far_x, near_y, far_y, near_z, far_z, s, o, t_best, magic)) mask |= (bits4 >> 24) << (index4 >> 24);
    return mask;
  }

  // One node step of one lane: the node's hit mask for this ray, split into the two groups.
  static __device__ __forceinline__ void node_step(const SceneView& scene, uint32_t node_index, const float3& idir,
                                                   const float3& oidir, float t_best, uint32_t octinv,
                                                   uint32_t& group_base, uint32_t& group_hits, uint32_t& tri_base,
                                                   uint32_t& tri_mask) {
    const float4* node = scene.bvh_nodes + static_cast<size_t>(node_index) * 6;
    const Wide8 a = ldg256(node), b = ldg256(node + 2), c = ldg256(node + 4);
    const float3 s = v3(a.a.w * idir.x, a.b.x * idir.y, a.b.y * idir.z);
    const float3 o = v3(fmaf(a.a.x, idir.x, -oidir.x), fmaf(a.a.y, idir.y, -oidir.y), fmaf(a.a.z, idir.z, -oidir.z));
    const bool rx = idir.x < 0.0f, ry = idir.y < 0.0f, rz = idir.z < 0.0f;
    const uint32_t octinv4 = octinv * 0x01010101u;
    uint32_t mask = four_children(__float_as_uint(b.b.x), __float_as_uint(b.b.z), __float_as_uint(c.a.x), __float_as_uint(c.a.z),
                                  __float_as_uint(c.b.x), __float_as_uint(c.b.z), rx, ry, rz, s, o, t_best,
                                  __float_as_uint(b.a.x), octinv4, __float_as_uint(b.a.w));
    mask |= four_children(__float_as_uint(b.b.y), __float_as_uint(b.b.w), __float_as_uint(c.a.y), __float_as_uint(c.a.w),
                          __float_as_uint(c.b.y), __float_as_uint(c.b.w), rx, ry, rz, s, o, t_best,
                          __float_as_uint(b.a.y), octinv4, __float_as_uint(b.a.w));
    group_base = __float_as_uint(a.b.z);
    group_hits = (mask >> 24) | (__float_as_uint(b.a.z) << 8);
    tri_base = __float_as_uint(a.b.w);
    tri_mask = mask & 0x00ffffffu;
  }

  static __device__ __forceinline__ uint32_t octant_inverse(const float3& d) {
    return 7u ^ ((d.x < 0.0f ? 1u : 0u) | (d.y < 0.0f ? 2u : 0u) | (d.z < 0.0f ? 4u : 0u));
  }
  // The group that holds only the root (node 0 = "inner child in slot 0 of a parent whose child_base is 0").
  static __device__ __forceinline__ uint32_t root_hits(uint32_t octinv) { return (1u << 8) | (1u << octinv); }
  // Takes the nearest pending child out of a node group; returns its node index.
  static __device__ __forceinline__ uint32_t take_child(uint32_t group_base, uint32_t& group_hits, uint32_t octinv) {
    const uint32_t p = 31u - static_cast<uint32_t>(__clz(static_cast<int>(group_hits & 0xffu)));
    group_hits ^= 1u << p;
    const uint32_t slot = p ^ octinv;
    return group_base + __popc((group_hits >> 8) & ((1u << slot) - 1u));
  }

  // One thread walks one ray on its own (no votes): the BDPT kernel's casts, where lanes of a warp are at
  // different stages of different paths.  `stack`: `depth` entries.
  template <bool kAnyHit>
  static __device__ __forceinline__ bool walk_lane(const SceneView& scene, const Ray& ray, HitRecord& hit, uint2* stack) {
    const float3 idir = v3(1.0f / ray.d.x, 1.0f / ray.d.y, 1.0f / ray.d.z);
    const float3 oidir = v3(ray.o.x * idir.x, ray.o.y * idir.y, ray.o.z * idir.z);
    const uint32_t octinv = octant_inverse(ray.d);
    uint32_t group_base = 0, group_hits = root_hits(octinv), tri_base = 0, tri_mask = 0;
    int sp = 0;
    bool found = false;
    while (true) {
      if (group_hits & 0xffu) {
        const uint32_t node = take_child(group_base, group_hits, octinv);
        if (group_hits & 0xffu) stack[sp++] = make_uint2(group_base, group_hits);
        node_step(scene, node, idir, oidir, hit.t, octinv, group_base, group_hits, tri_base, tri_mask);
      }
      while (tri_mask) {
        const uint32_t k = tri_base + (__ffs(tri_mask) - 1);
        tri_mask &= tri_mask - 1u;
        const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + k);
        if (intersect_triangle(__ldg(tri), __ldg(tri + 1), __ldg(tri + 2), ray, hit.t, hit.t)) {
          hit.prim = static_cast<int32_t>(k);
          found = true;
          if (kAnyHit) return true;
        }
      }
      if ((group_hits & 0xffu) == 0) {
        if (sp == 0) break;
        const uint2 top = stack[--sp];
        group_base = top.x;
        group_hits = top.y;
      }
    }
    return found;
  }

  // Persistent warp-converged traversal with dynamic ray fetch: same three voted phases as the 4-wide version above
  // (fetch / descend / leaves; Source, kCount: see there).  A lane is "inner" while its node group has a pending
  // child and it holds no triangles, "at a leaf" while its triangle group is not empty, idle when both are empty
  // (the stack is popped as soon as a group runs out, so idle implies an empty stack).
  template <bool kAnyHit, bool kCount = false, class Source>
  static __device__ __forceinline__ void trace_persistent(const SceneView& scene, Source& source,
                                                          unsigned int* cursor,
                                                          unsigned long long* counters = nullptr) {
    extern __shared__ uint2 bvh_stack_smem8[];
    uint2* stack = bvh_stack_smem8 + threadIdx.x;
    const int stride = blockDim.x;
    constexpr uint32_t kFull = 0xffffffffu;
    constexpr int kRefillLanes = kAnyHit ? AMBER_BVH_REFILL_LANES_ANYHIT : AMBER_BVH_REFILL_LANES;
    constexpr int kMinDescendLanes = kAnyHit ? AMBER_BVH_MIN_DESCEND_LANES_ANYHIT : AMBER_BVH_MIN_DESCEND_LANES;
    const int lane = threadIdx.x & 31;
    const uint32_t lane_lt = (1u << lane) - 1u;
    const uint32_t n = source.count();

    int sp = 0;
    uint32_t group_base = 0, group_hits = 0, tri_base = 0, tri_mask = 0, octinv = 0;
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
      const bool idle = (group_hits & 0xffu) == 0 && tri_mask == 0;
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
            octinv = octant_inverse(ray.d);
            group_base = 0;
            group_hits = root_hits(octinv);
          }
        }
      }
      // ---- descend --------------------------------------------------------------------
      while (true) {
        const bool inner = tri_mask == 0 && (group_hits & 0xffu) != 0;
        const uint32_t inner_lanes = __ballot_sync(kFull, inner);
        if (inner_lanes == 0) break;
        if (__popc(inner_lanes) < kMinDescendLanes && __any_sync(kFull, tri_mask != 0)) break;
        if (inner) {
          const uint32_t node = take_child(group_base, group_hits, octinv);
          if (group_hits & 0xffu) {
            stack[sp * stride] = make_uint2(group_base, group_hits);
            sp++;
          }
          node_step(scene, node, idir, oidir, hit.t, octinv, group_base, group_hits, tri_base, tri_mask);
          if (kCount) node_steps++;
          if ((group_hits & 0xffu) == 0 && tri_mask == 0 && sp > 0) {
            sp--;
            const uint2 top = stack[sp * stride];
            group_base = top.x;
            group_hits = top.y;
          }
        }
      }
      // ---- leaves ---------------------------------------------------------------------
      if (tri_mask != 0) {
        if (kCount) triangle_tests += __popc(tri_mask);
#if AMBER_BVH_LEAF_PREFETCH
        // the triangles of the group are tested one after the other, each behind its own L2 / DRAM round trip (the
        // kernel's largest stall): ask for all but the first now, so that the round trips overlap
        for (uint32_t rest = tri_mask & (tri_mask - 1u); rest; rest &= rest - 1u) {
          const char* next = reinterpret_cast<const char*>(scene.tri_isect + tri_base + (__ffs(rest) - 1));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(next));
          asm volatile("prefetch.global.L1 [%0];" ::"l"(next + 32));
        }
#endif
#if AMBER_BVH_LEAF_PAIRS
        while (tri_mask) {  // two triangles per round: their six loads are in flight together
          const uint32_t k0 = tri_base + (__ffs(tri_mask) - 1);
          tri_mask &= tri_mask - 1u;
          const bool second = tri_mask != 0;
          const uint32_t k1 = second ? tri_base + (__ffs(tri_mask) - 1) : k0;
          tri_mask &= tri_mask - 1u;
          const float4* t0 = reinterpret_cast<const float4*>(scene.tri_isect + k0);
          const float4* t1 = reinterpret_cast<const float4*>(scene.tri_isect + k1);
          const float4 a0 = __ldg(t0), a1 = __ldg(t0 + 1), a2 = __ldg(t0 + 2);
          const float4 b0 = __ldg(t1), b1 = __ldg(t1 + 1), b2 = __ldg(t1 + 2);
          if (intersect_triangle(a0, a1, a2, ray, hit.t, hit.t)) {
            hit.prim = static_cast<int32_t>(k0);
            found = true;
          }
          if (second && intersect_triangle(b0, b1, b2, ray, hit.t, hit.t)) {
            hit.prim = static_cast<int32_t>(k1);
            found = true;
          }
          if (kAnyHit && found) tri_mask = 0;
        }
#else
        while (tri_mask) {
          const uint32_t k = tri_base + (__ffs(tri_mask) - 1);
          tri_mask &= tri_mask - 1u;
          const float4* tri = reinterpret_cast<const float4*>(scene.tri_isect + k);
          if (intersect_triangle(load_triangle_word(tri), load_triangle_word(tri + 1), load_triangle_word(tri + 2), ray, hit.t, hit.t)) {
            hit.prim = static_cast<int32_t>(k);
            found = true;
            if (kAnyHit) tri_mask = 0;
          }
        }
#endif
        if (kAnyHit && found) {
          sp = 0;
          group_hits = 0;
        } else if ((group_hits & 0xffu) == 0 && sp > 0) {
          sp--;
          const uint2 top = stack[sp * stride];
          group_base = top.x;
          group_hits = top.y;
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

#endif  // AMBER_BVH_WIDTH

}  // namespace amber_cuda
