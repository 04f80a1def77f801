// bvh_build.h -- host-side BVH construction: binned SAH binary tree, collapsed to
// 4-wide 64-byte nodes with 8-bit quantised child boxes (layout: bvh.cuh).
//
// Replaces BVH::BuildBVH / CreateNode / FindSplit / FindSplitAxis /
// SurfaceAreaHeuristic (include/amber/raytracer/acceleration_bvh.h:135-312).  The
// reference sorts the objects four times per node and re-measures up to 45 boxes
// (O(N log^2 N)); this is the usual O(N log N) binned build: 16 bins per axis over the
// CENTROID bounds, cost = C_trav + (N_l A_l + N_r A_r)/A, leaves of at most 4
// triangles.  Subtrees above a size threshold are built on separate std::threads.
// Tree shape does not have to match the reference (closest-hit semantics do).
#pragma once

#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

namespace amber_cuda {
namespace bvh_build {

#ifndef AMBER_BVH_WIDTH
#define AMBER_BVH_WIDTH 4
#endif
constexpr int kWidth = AMBER_BVH_WIDTH;
static_assert(kWidth == 4 || kWidth == 8, "4-wide 64-byte nodes or 8-wide 96-byte nodes");
constexpr int kBins = 16;
#ifndef AMBER_BVH_LEAF_MAX
#define AMBER_BVH_LEAF_MAX (AMBER_BVH_WIDTH == 8 ? 3 : 4)
#endif
constexpr int kLeafMax = AMBER_BVH_LEAF_MAX;  // 8-wide: a node's leaf triangles are 24 bits of its hit mask
static_assert(kLeafMax >= 1 && kLeafMax <= (kWidth == 8 ? 3 : 4), "leaf size the node format can address");
constexpr int32_t kEmpty = 0x7fffffff;

struct Box {
  float lo[3], hi[3];
  void reset() {
    for (int a = 0; a < 3; a++) {
      lo[a] = FLT_MAX;
      hi[a] = -FLT_MAX;
    }
  }
  void grow(const float* p) {
    for (int a = 0; a < 3; a++) {
      lo[a] = std::min(lo[a], p[a]);
      hi[a] = std::max(hi[a], p[a]);
    }
  }
  void grow(const Box& b) {
    for (int a = 0; a < 3; a++) {
      lo[a] = std::min(lo[a], b.lo[a]);
      hi[a] = std::max(hi[a], b.hi[a]);
    }
  }
  float area() const {
    const float dx = hi[0] - lo[0], dy = hi[1] - lo[1], dz = hi[2] - lo[2];
    if (dx < 0 || dy < 0 || dz < 0) return 0.0f;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
  }
};

struct Node2 {
  Box box;
  int32_t left = -1, right = -1;  // children (inner)
  int32_t first = 0, count = 0;   // triangle range (leaf if count > 0)
};

struct Builder {
  const float* verts;  // n * 9
  uint32_t n;
  std::vector<Box> tri_box;
  std::vector<float> centroid;  // n * 3
  std::vector<uint32_t> order;  // permutation: leaf order -> input triangle
  std::vector<Node2> nodes;
  std::atomic<int32_t> node_count{0};

  int32_t alloc_node() { return node_count.fetch_add(1); }

  // Ranges of at least kParallelRange triangles are scanned by several threads (bounds, bins); subtrees are built on
  // their own threads down to kSpawnDepth.  Boxes and counts merge exactly, so the tree does not depend on the
  // number of threads.
  static constexpr uint32_t kParallelRange = 1u << 17;
  static constexpr int kSpawnDepth = 5, kMaxChunks = 32;
  int workers = 1;  // hardware threads to use (build() sets it), at most kMaxChunks

  struct Bins {
    Box box[3][kBins];
    uint32_t count[3][kBins];
    void reset() {
      for (int a = 0; a < 3; a++)
        for (int b = 0; b < kBins; b++) {
          box[a][b].reset();
          count[a][b] = 0;
        }
    }
  };

  // f(chunk, begin, end) over [first, first + count) cut into chunks; returns the number of chunks
  template <class F>
  int for_chunks(uint32_t first, uint32_t count, int depth, F&& f) {
    const int chunks = count >= kParallelRange ? std::max(1, std::min(kMaxChunks, workers >> std::min(depth, 30))) : 1;
    if (chunks == 1) {
      f(0, first, first + count);
      return 1;
    }
    const uint32_t per = (count + chunks - 1) / chunks;
    std::vector<std::thread> pool;
    for (int c = 1; c < chunks; c++)
      pool.emplace_back([&f, c, first, count, per]() {
        f(c, std::min(first + count, first + c * per), std::min(first + count, first + (c + 1) * per));
      });
    f(0, first, std::min(first + count, first + per));
    for (auto& t : pool) t.join();
    return chunks;
  }

  void build_range(int32_t node_index, uint32_t first, uint32_t count, int depth) {
    Node2& node = nodes[node_index];
    Box bounds, cbounds;
    {
      Box part_bounds[kMaxChunks], part_cbounds[kMaxChunks];
      const int chunks = for_chunks(first, count, depth, [&](int c, uint32_t begin, uint32_t end) {
        Box b, cb;
        b.reset();
        cb.reset();
        for (uint32_t i = begin; i < end; i++) {
          b.grow(tri_box[order[i]]);
          cb.grow(&centroid[3 * static_cast<size_t>(order[i])]);
        }
        part_bounds[c] = b;
        part_cbounds[c] = cb;
      });
      bounds = part_bounds[0];
      cbounds = part_cbounds[0];
      for (int c = 1; c < chunks; c++) {
        bounds.grow(part_bounds[c]);
        cbounds.grow(part_cbounds[c]);
      }
    }
    node.box = bounds;
    if (count <= static_cast<uint32_t>(std::min(2, kLeafMax))) {
      node.first = first;
      node.count = count;
      return;
    }
    // binned SAH over the centroid bounds: the three axes are binned in one pass over the range
    float bin_scale[3];
    for (int axis = 0; axis < 3; axis++) {
      const float extent = cbounds.hi[axis] - cbounds.lo[axis];
      bin_scale[axis] = extent > 0 ? kBins / extent : 0.0f;
    }
    std::vector<Bins> part(1);
    {
      const int want = count >= kParallelRange ? std::max(1, std::min(kMaxChunks, workers >> std::min(depth, 30))) : 1;
      part.resize(want);
      for_chunks(first, count, depth, [&](int c, uint32_t begin, uint32_t end) {
        Bins& bins = part[c];
        bins.reset();
        for (uint32_t i = begin; i < end; i++) {
          const uint32_t t = order[i];
          const Box& tb = tri_box[t];
          for (int axis = 0; axis < 3; axis++) {
            int b = static_cast<int>((centroid[3 * static_cast<size_t>(t) + axis] - cbounds.lo[axis]) * bin_scale[axis]);
            b = std::min(kBins - 1, std::max(0, b));
            bins.count[axis][b]++;
            bins.box[axis][b].grow(tb);
          }
        }
      });
      for (size_t c = 1; c < part.size(); c++)
        for (int axis = 0; axis < 3; axis++)
          for (int b = 0; b < kBins; b++) {
            part[0].count[axis][b] += part[c].count[axis][b];
            part[0].box[axis][b].grow(part[c].box[axis][b]);
          }
    }
    float best_cost = FLT_MAX;
    int best_axis = -1, best_bin = -1;
    const float parent_area = std::max(bounds.area(), 1e-30f);
    for (int axis = 0; axis < 3; axis++) {
      const float extent = cbounds.hi[axis] - cbounds.lo[axis];
      if (!(extent > 0)) continue;
      const Box* bin_box = part[0].box[axis];
      const uint32_t* bin_count = part[0].count[axis];
      float right_area[kBins];
      uint32_t right_count[kBins];
      Box acc;
      acc.reset();
      uint32_t cnt = 0;
      for (int b = kBins - 1; b > 0; b--) {
        acc.grow(bin_box[b]);
        cnt += bin_count[b];
        right_area[b] = acc.area();
        right_count[b] = cnt;
      }
      acc.reset();
      cnt = 0;
      for (int b = 0; b < kBins - 1; b++) {
        acc.grow(bin_box[b]);
        cnt += bin_count[b];
        if (cnt == 0 || right_count[b + 1] == 0) continue;
        const float cost = 1.0f + (cnt * acc.area() + right_count[b + 1] * right_area[b + 1]) / parent_area;
        if (cost < best_cost) {
          best_cost = cost;
          best_axis = axis;
          best_bin = b;
        }
      }
    }
    uint32_t mid;
    if (best_axis < 0 || (best_cost >= static_cast<float>(count) && count <= kLeafMax)) {
      if (count <= kLeafMax) {
        node.first = first;
        node.count = count;
        return;
      }
      // no useful SAH split (coincident centroids): split the range in half
      mid = first + count / 2;
    } else {
      const float scale = bin_scale[best_axis];
      const float lo = cbounds.lo[best_axis];
      auto it = std::partition(order.begin() + first, order.begin() + first + count, [&](uint32_t t) {
        int b = static_cast<int>((centroid[3 * t + best_axis] - lo) * scale);
        b = std::min(kBins - 1, std::max(0, b));
        return b <= best_bin;
      });
      mid = static_cast<uint32_t>(it - order.begin());
      if (mid == first || mid == first + count) mid = first + count / 2;
    }
    const int32_t left = alloc_node(), right = alloc_node();
    nodes[node_index].left = left;
    nodes[node_index].right = right;
    const uint32_t n_left = mid - first, n_right = first + count - mid;
    if (depth < kSpawnDepth && count > (1u << 15) && workers > 1) {
      std::thread worker([&, left, first, n_left, depth]() { build_range(left, first, n_left, depth + 1); });
      build_range(right, mid, n_right, depth + 1);
      worker.join();
    } else {
      build_range(left, first, n_left, depth + 1);
      build_range(right, mid, n_right, depth + 1);
    }
  }
};

struct WideBvhHost {
  std::vector<float> nodes;     // n_nodes * kNodeFloats (bvh.cuh)
  std::vector<uint32_t> order;  // leaf order -> input triangle index
  int32_t root = 0;             // 4-wide: 0, or a leaf reference when the whole scene is one leaf; 8-wide: always node 0
  uint32_t n_nodes = 0;
  int depth = 0;                // wide nodes on the longest root-to-leaf chain (sizes the traversal stack)
};

inline int32_t leaf_ref(int32_t first, int32_t count) { return ~((first << 3) | count); }

// pad so that float rounding in the slab test can never cull a true hit: the device evaluates
// t = q * (s * idir) + (o * idir - O * idir) (bvh.cuh node_step), four roundings of relative size 2^-24 on
// terms as large as the ray origin, the node origin and the node extent -- so the slack scales with the
// largest coordinate of the WHOLE scene (`scene_scale`: max |coordinate| of the root box, at least 1), not with
// the child's own coordinates: a child near the coordinate origin of a scene that reaches out to 1000 units
// is decoded with the rounding error of 1000-unit operands.
inline void padded(const Box& b, float scene_scale, float* lo, float* hi) {
  for (int a = 0; a < 3; a++) {
    lo[a] = b.lo[a] - 2e-6f * scene_scale;
    hi[a] = b.hi[a] + 2e-6f * scene_scale;
  }
}

// Plane codes.  Code b (8 bits) stands for the grid coordinate g(b) = the float whose bit pattern is
// 0x3F000000 | b << 16 -- 0.5 + b/256 for b < 128, b/128 from there on (the code's top bit is the exponent's lowest) --
// and the plane lies at origin + g(b) * scale.  The device builds g(b) with ONE byte permute and feeds it to the FMA
// that takes the plane into ray space: no integer-to-float conversion, no subtraction of a magic constant.  The grid is
// uneven (steps of extent/382 in the lower third of the node, extent/191 above), which a box hierarchy does not mind.
constexpr uint32_t kPlaneMagic = 0x3F000000u;

// fmaf as ONE hardware instruction where the CPU has it: without -mfma std::fmaf is a library call that emulates the
// single rounding in software (~100 ns), and the plane search below makes ~200 of them per node.
#if defined(__x86_64__) && defined(__GNUC__)
__attribute__((target("fma"))) inline float fmaf_hardware(float a, float b, float c) { return __builtin_fmaf(a, b, c); }
inline float fused_multiply_add(float a, float b, float c) {
  static const bool hardware = __builtin_cpu_supports("fma");
  return hardware ? fmaf_hardware(a, b, c) : std::fmaf(a, b, c);
}
#else
inline float fused_multiply_add(float a, float b, float c) { return std::fmaf(a, b, c); }
#endif
inline float plane_coordinate(int code) {
  const uint32_t bits = kPlaneMagic | (static_cast<uint32_t>(code) << 16);
  float g;
  std::memcpy(&g, &bits, 4);
  return g;
}

// One axis of up to `width` child boxes: lo codes round down, hi codes round up, checked with the very expression
// the device decodes with (fmaf(g, scale, origin)), so a decoded box always contains the true one.  Unused slots get
// lo = 255 > hi = 0: the slab test fails for every ray.
inline void quantise_axis(const float* lo, const float* hi, const bool* used, int width, float& origin, float& scale,
                          uint8_t* q_lo, uint8_t* q_hi) {
  float bottom = FLT_MAX, top = -FLT_MAX;
  for (int k = 0; k < width; k++) {
    if (!used[k]) continue;
    bottom = std::min(bottom, lo[k]);
    top = std::max(top, hi[k]);
  }
  const float span = plane_coordinate(255) - plane_coordinate(0);
  // The grid spans [bottom - margin, top + margin]; rounding in o, s or the decode can leave it short at an end by an
  // ulp or two of the coordinates, so the margin starts at zero and doubles from one ulp until every box is covered.
  float margin = 0.0f;
  for (bool ok = false; !ok;) {
    ok = true;
    const float s = std::max(((top + margin) - (bottom - margin)) / span, 1e-30f);
    const float o = (bottom - margin) - plane_coordinate(0) * s;
    auto plane = [&](int code) { return fused_multiply_add(plane_coordinate(code), s, o); };
    for (int k = 0; k < width && ok; k++) {
      int a = 255, b = 0;
      if (used[k]) {
        a = 0;
        for (int step = 128; step > 0; step >>= 1)  // largest code whose plane is <= lo (the planes are monotone)
          if (a + step <= 255 && plane(a + step) <= lo[k]) a += step;
        b = 255;
        for (int step = 128; step > 0; step >>= 1)  // smallest code whose plane is >= hi
          if (b - step >= 0 && plane(b - step) >= hi[k]) b -= step;
        if (plane(a) > lo[k] || plane(b) < hi[k]) ok = false;
      }
      q_lo[k] = static_cast<uint8_t>(a);
      q_hi[k] = static_cast<uint8_t>(b);
    }
    origin = o;
    scale = s;
    if (!ok) margin = std::max(2.0f * margin, 1.2e-7f * std::max(std::max(std::fabs(bottom), std::fabs(top)), 1e-30f));
  }
}

// One 64-byte node (bvh.cuh): four children.  lo/hi: [n_kids][3]; unused lanes get ref = kEmpty.
inline void encode_node(const float (*lo)[3], const float (*hi)[3], const int32_t* refs, int n_kids, float* out16) {
  uint32_t words[6] = {0, 0, 0, 0, 0, 0};  // x lo, x hi, y lo, y hi, z lo, z hi; byte k = child k
  float origin[3], scale[3];
  for (int a = 0; a < 3; a++) {
    float axis_lo[4], axis_hi[4];
    bool used[4];
    uint8_t q_lo[4], q_hi[4];
    for (int k = 0; k < 4; k++) {
      used[k] = k < n_kids;
      axis_lo[k] = used[k] ? lo[k][a] : 0.0f;
      axis_hi[k] = used[k] ? hi[k][a] : 0.0f;
    }
    quantise_axis(axis_lo, axis_hi, used, 4, origin[a], scale[a], q_lo, q_hi);
    std::memcpy(&words[2 * a], q_lo, 4);
    std::memcpy(&words[2 * a + 1], q_hi, 4);
  }
  out16[0] = origin[0]; out16[1] = origin[1]; out16[2] = origin[2];
  out16[3] = scale[0]; out16[4] = scale[1]; out16[5] = scale[2];
  std::memcpy(&out16[6], words, sizeof words);
  for (int k = 0; k < 4; k++) {
    const int32_t ref = k < n_kids ? refs[k] : kEmpty;
    std::memcpy(&out16[12 + k], &ref, 4);
  }
}

constexpr int kNodeFloats = kWidth == 8 ? 24 : 16;  // 96 B / 64 B

// ---- 8-wide nodes (layout and traversal: bvh.cuh) ----------------------------------------------------------
// Collapse 2 -> 8.  Children sit in the slot whose sign pattern (bit a set: towards +axis a) best matches the
// direction from the node's centre to theirs, so that `slot ^ (7 ^ ray octant)` orders them front to back without
// any distance sort (Ylitie, Karras, Laine 2017).  A node's inner children are consecutive nodes (child_base + rank
// among the inner slots) and the triangles of its leaf children are consecutive in the triangle order (tri_base +
// 5-bit offset), which is what lets one node address eight children in 16 bytes.
inline WideBvhHost build8(const Builder& b, int32_t root) {
  WideBvhHost out;
  const std::vector<Node2>& n2 = b.nodes;
  float scene_scale = 1.0f;
  for (int a = 0; a < 3; a++)
    scene_scale = std::max(scene_scale, std::max(std::fabs(n2[root].box.lo[a]), std::fabs(n2[root].box.hi[a])));
  struct Pending {
    int32_t node2;
    int32_t wide;
    int depth;
  };
  std::vector<Pending> queue;
  queue.push_back({root, 0, 1});
  out.n_nodes = 1;
  out.order.reserve(b.order.size());
  out.nodes.resize(kNodeFloats);
  for (size_t head = 0; head < queue.size(); head++) {
    const Pending item = queue[head];
    out.depth = std::max(out.depth, item.depth);
    int32_t kids[8];
    int n_kids;
    if (n2[item.node2].count > 0) {  // the whole scene is one leaf: a root with one child
      kids[0] = item.node2;
      n_kids = 1;
    } else {
      kids[0] = n2[item.node2].left;
      kids[1] = n2[item.node2].right;
      n_kids = 2;
      while (n_kids < 8) {  // always open the inner child with the largest surface area
        int best = -1;
        float best_area = -1.0f;
        for (int k = 0; k < n_kids; k++) {
          if (n2[kids[k]].count > 0) continue;
          const float a = n2[kids[k]].box.area();
          if (a > best_area) {
            best_area = a;
            best = k;
          }
        }
        if (best < 0) break;
        const int32_t open = kids[best];
        kids[best] = n2[open].left;
        kids[n_kids++] = n2[open].right;
      }
    }
    // slots: greedy assignment by the largest projection of (child centre - node centre) on the slot's diagonal
    const Box& parent = n2[item.node2].box;
    int slot_kid[8];
    for (int s = 0; s < 8; s++) slot_kid[s] = -1;
    bool placed[8] = {false};
    for (int round = 0; round < n_kids; round++) {
      float best = -FLT_MAX;
      int best_k = -1, best_s = -1;
      for (int k = 0; k < n_kids; k++) {
        if (placed[k]) continue;
        const Box& cb = n2[kids[k]].box;
        for (int s = 0; s < 8; s++) {
          if (slot_kid[s] >= 0) continue;
          float score = 0.0f;
          for (int a = 0; a < 3; a++) {
            const float d = 0.5f * (cb.lo[a] + cb.hi[a]) - 0.5f * (parent.lo[a] + parent.hi[a]);
            score += ((s >> a) & 1) ? d : -d;
          }
          if (score > best) {
            best = score;
            best_k = k;
            best_s = s;
          }
        }
      }
      placed[best_k] = true;
      slot_kid[best_s] = kids[best_k];
    }
    float lo[3][8], hi[3][8];
    bool used[8];
    uint8_t meta[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    uint32_t imask = 0;
    const uint32_t child_base = out.n_nodes, tri_base = static_cast<uint32_t>(out.order.size());
    for (int s = 0; s < 8; s++) {
      used[s] = slot_kid[s] >= 0;
      if (!used[s]) continue;
      const Node2& child = n2[slot_kid[s]];
      float plo[3], phi[3];
      padded(child.box, scene_scale, plo, phi);
      for (int a = 0; a < 3; a++) {
        lo[a][s] = plo[a];
        hi[a][s] = phi[a];
      }
      if (child.count > 0) {  // leaf: `count` one-bits above the 5-bit offset of its first triangle
        const uint32_t offset = static_cast<uint32_t>(out.order.size()) - tri_base;
        meta[s] = static_cast<uint8_t>((((1u << child.count) - 1u) << 5) | offset);
        for (int32_t t = 0; t < child.count; t++) out.order.push_back(b.order[child.first + t]);
      } else {                // inner: one bit at index 24 + slot
        imask |= 1u << s;
        meta[s] = static_cast<uint8_t>((1u << 5) | (24u + s));
        queue.push_back({slot_kid[s], static_cast<int32_t>(out.n_nodes++), item.depth + 1});
      }
    }
    if (out.nodes.size() < static_cast<size_t>(out.n_nodes) * kNodeFloats)
      out.nodes.resize(std::max(out.nodes.size() * 2, static_cast<size_t>(out.n_nodes) * kNodeFloats));
    float* node = &out.nodes[static_cast<size_t>(item.wide) * 24];
    uint32_t words[24];
    uint8_t q[6][8];  // x lo, x hi, y lo, y hi, z lo, z hi
    float origin[3], scale[3];
    for (int a = 0; a < 3; a++) quantise_axis(lo[a], hi[a], used, 8, origin[a], scale[a], q[2 * a], q[2 * a + 1]);
    std::memcpy(&words[0], origin, 12);
    std::memcpy(&words[3], scale, 12);
    words[6] = child_base;
    words[7] = tri_base;
    std::memcpy(&words[8], meta, 8);
    words[10] = imask;
    words[11] = kPlaneMagic;  // read by the device as the second operand of its byte permutes (bvh.cuh, plane_float)
    for (int plane = 0; plane < 6; plane++) std::memcpy(&words[12 + 2 * plane], q[plane], 8);
    std::memcpy(node, words, sizeof words);
  }
  out.nodes.resize(static_cast<size_t>(out.n_nodes) * kNodeFloats);
  out.root = 0;
  return out;
}

inline WideBvhHost build(const float* verts9, uint32_t n) {
  WideBvhHost out;
  Builder b;
  b.verts = verts9;
  b.n = n;
  b.tri_box.resize(n);
  b.centroid.resize(3 * static_cast<size_t>(n));
  b.order.resize(n);
  for (uint32_t i = 0; i < n; i++) {
    const float* v = verts9 + 9 * static_cast<size_t>(i);
    Box box;
    box.reset();
    box.grow(v);
    box.grow(v + 3);
    box.grow(v + 6);
    b.tri_box[i] = box;
    for (int a = 0; a < 3; a++) b.centroid[3 * static_cast<size_t>(i) + a] = 0.5f * (box.lo[a] + box.hi[a]);
    b.order[i] = i;
  }
  b.nodes.resize(std::max<size_t>(1, 2 * static_cast<size_t>(n)));
  b.workers = static_cast<int>(std::max(1u, std::min<unsigned>(Builder::kMaxChunks, std::thread::hardware_concurrency())));
  const int32_t root = b.alloc_node();
  b.build_range(root, 0, n, 0);

  if (kWidth == 8) {
    if (n == 0) {  // a dummy node whose slots are all empty keeps the device pointer valid
      out.nodes.assign(kNodeFloats, 0.0f);
      uint32_t empty_planes[12];
      for (int plane = 0; plane < 6; plane++) empty_planes[2 * plane] = empty_planes[2 * plane + 1] = (plane & 1) ? 0u : 0xffffffffu;
      std::memcpy(&out.nodes[12], empty_planes, sizeof empty_planes);
      std::memcpy(&out.nodes[11], &kPlaneMagic, 4);
      out.n_nodes = 1;
      out.depth = 1;
      return out;
    }
    return build8(b, root);
  }
  out.order = b.order;
  const std::vector<Node2>& n2 = b.nodes;
  if (n2[root].count > 0 || n == 0) {
    out.root = n == 0 ? kEmpty : leaf_ref(n2[root].first, n2[root].count);
    // a dummy node keeps the device pointer valid
    out.nodes.assign(kNodeFloats, 0.0f);
    out.n_nodes = 1;
    return out;
  }
  // collapse 2 -> 4: breadth-first, each wide node adopts up to four descendants,
  // always opening the inner child with the largest surface area
  float scene_scale = 1.0f;
  for (int a = 0; a < 3; a++)
    scene_scale = std::max(scene_scale, std::max(std::fabs(n2[root].box.lo[a]), std::fabs(n2[root].box.hi[a])));
  struct Pending {
    int32_t node2;
    int32_t wide;
  };
  struct Wide4 {
    int32_t kids[4], refs[4];
    int n_kids;
  };
  std::vector<Pending> queue;
  std::vector<Wide4> plan;  // first the shape of the wide tree, serially (node indices are handed out in queue order) ...
  queue.push_back({root, 0});
  out.n_nodes = 1;
  for (size_t head = 0; head < queue.size(); head++) {
    const Pending item = queue[head];
    int32_t kids[4];
    int n_kids = 2;
    kids[0] = n2[item.node2].left;
    kids[1] = n2[item.node2].right;
    while (n_kids < 4) {
      int best = -1;
      float best_area = -1.0f;
      for (int k = 0; k < n_kids; k++) {
        if (n2[kids[k]].count > 0) continue;
        const float a = n2[kids[k]].box.area();
        if (a > best_area) {
          best_area = a;
          best = k;
        }
      }
      if (best < 0) break;
      const int32_t open = kids[best];
      kids[best] = n2[open].left;
      kids[n_kids++] = n2[open].right;
    }
    Wide4 wide;
    wide.n_kids = n_kids;
    for (int k = 0; k < n_kids; k++) {
      const Node2& child = n2[kids[k]];
      wide.kids[k] = kids[k];
      if (child.count > 0) {
        wide.refs[k] = leaf_ref(child.first, child.count);
      } else {
        wide.refs[k] = static_cast<int32_t>(out.n_nodes++);
        queue.push_back({kids[k], wide.refs[k]});
      }
    }
    plan.push_back(wide);  // plan[item.wide]: the queue is in node order
  }
  // ... then the nodes are encoded (the plane search of quantise_axis is most of this phase), independently of each other
  out.nodes.resize(static_cast<size_t>(out.n_nodes) * kNodeFloats);
  const unsigned threads = std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  auto encode_range = [&](size_t begin, size_t end) {
    for (size_t w = begin; w < end; w++) {
      float lo[4][3], hi[4][3];
      for (int k = 0; k < plan[w].n_kids; k++) padded(n2[plan[w].kids[k]].box, scene_scale, lo[k], hi[k]);
      encode_node(lo, hi, plan[w].refs, plan[w].n_kids, &out.nodes[w * kNodeFloats]);
    }
  };
  std::vector<std::thread> pool;
  const size_t per = (plan.size() + threads - 1) / threads;
  for (unsigned t = 1; t < threads && plan.size() > 4096; t++)
    pool.emplace_back(encode_range, std::min(plan.size(), t * per), std::min(plan.size(), (t + 1) * per));
  encode_range(0, pool.empty() ? plan.size() : std::min(plan.size(), per));
  for (auto& t : pool) t.join();
  return out;
}

}  // namespace bvh_build
}  // namespace amber_cuda
