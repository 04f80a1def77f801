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

constexpr int kBins = 16;
constexpr int kLeafMax = 4;
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

  void build_range(int32_t node_index, uint32_t first, uint32_t count, int depth) {
    Node2& node = nodes[node_index];
    Box bounds, cbounds;
    bounds.reset();
    cbounds.reset();
    for (uint32_t i = first; i < first + count; i++) {
      bounds.grow(tri_box[order[i]]);
      cbounds.grow(&centroid[3 * order[i]]);
    }
    node.box = bounds;
    if (count <= 2) {
      node.first = first;
      node.count = count;
      return;
    }
    // binned SAH over the centroid bounds
    float best_cost = FLT_MAX;
    int best_axis = -1, best_bin = -1;
    const float parent_area = std::max(bounds.area(), 1e-30f);
    for (int axis = 0; axis < 3; axis++) {
      const float extent = cbounds.hi[axis] - cbounds.lo[axis];
      if (!(extent > 0)) continue;
      Box bin_box[kBins];
      uint32_t bin_count[kBins] = {0};
      for (auto& b : bin_box) b.reset();
      const float scale = kBins / extent;
      for (uint32_t i = first; i < first + count; i++) {
        const uint32_t t = order[i];
        int b = static_cast<int>((centroid[3 * t + axis] - cbounds.lo[axis]) * scale);
        b = std::min(kBins - 1, std::max(0, b));
        bin_count[b]++;
        bin_box[b].grow(tri_box[t]);
      }
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
      const float extent = cbounds.hi[best_axis] - cbounds.lo[best_axis];
      const float scale = kBins / extent;
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
    if (depth < 3 && count > (1u << 16)) {
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
  std::vector<float> nodes;     // n_nodes * kNodeFloats (4 float4 per node, bvh.cuh)
  std::vector<uint32_t> order;  // leaf order -> input triangle index
  int32_t root = 0;             // 0, or a leaf reference when the whole scene is one leaf
  uint32_t n_nodes = 0;
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

// One 64-byte node (bvh.cuh): the children's boxes quantised to 8 bits per plane on the grid
// origin + q * scale of the node's own box; lo rounds down, hi rounds up, checked with the very
// expression the device decodes with (fmaf(q, scale, origin)), so a decoded box always contains
// the true one.  lo/hi: [n_kids][3]; unused lanes get ref = kEmpty (and lo > hi).
inline void encode_node(const float (*lo)[3], const float (*hi)[3], const int32_t* refs, int n_kids, float* out16) {
  uint32_t words[6] = {0, 0, 0, 0, 0, 0};  // x lo, x hi, y lo, y hi, z lo, z hi; byte k = child k
  float origin[3], scale[3];
  for (int a = 0; a < 3; a++) {
    float o = FLT_MAX, top = -FLT_MAX;
    for (int k = 0; k < n_kids; k++) {
      o = std::min(o, lo[k][a]);
      top = std::max(top, hi[k][a]);
    }
    float s = std::max((top - o) / 255.0f, 1e-30f);
    uint32_t w_lo = 0, w_hi = 0;
    for (bool ok = false; !ok;) {
      ok = true;
      w_lo = w_hi = 0;
      for (int k = 0; k < 4 && ok; k++) {
        int q_lo = 255, q_hi = 0;  // unused lane: empty interval
        if (k < n_kids) {
          q_lo = std::min(255, std::max(0, static_cast<int>(std::floor((lo[k][a] - o) / s))));
          while (q_lo > 0 && std::fmaf(static_cast<float>(q_lo), s, o) > lo[k][a]) q_lo--;
          q_hi = std::min(255, std::max(0, static_cast<int>(std::ceil((hi[k][a] - o) / s))));
          while (q_hi < 255 && std::fmaf(static_cast<float>(q_hi), s, o) < hi[k][a]) q_hi++;
          if (std::fmaf(static_cast<float>(q_hi), s, o) < hi[k][a]) ok = false;  // grid too short: widen it
        }
        w_lo |= static_cast<uint32_t>(q_lo) << (8 * k);
        w_hi |= static_cast<uint32_t>(q_hi) << (8 * k);
      }
      if (!ok) s = std::max(std::nextafter(s, FLT_MAX), s * 1.0001f);
    }
    origin[a] = o;
    scale[a] = s;
    words[2 * a] = w_lo;
    words[2 * a + 1] = w_hi;
  }
  out16[0] = origin[0]; out16[1] = origin[1]; out16[2] = origin[2];
  out16[3] = scale[0]; out16[4] = scale[1]; out16[5] = scale[2];
  std::memcpy(&out16[6], words, sizeof words);
  for (int k = 0; k < 4; k++) {
    const int32_t ref = k < n_kids ? refs[k] : kEmpty;
    std::memcpy(&out16[12 + k], &ref, 4);
  }
}

constexpr int kNodeFloats = 16;  // 64 B

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
  const int32_t root = b.alloc_node();
  b.build_range(root, 0, n, 0);

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
  std::vector<Pending> queue;
  queue.push_back({root, 0});
  out.n_nodes = 1;
  out.nodes.resize(kNodeFloats);
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
    float lo[4][3], hi[4][3];
    int32_t refs[4];
    for (int k = 0; k < n_kids; k++) {
      const Node2& child = n2[kids[k]];
      padded(child.box, scene_scale, lo[k], hi[k]);
      if (child.count > 0) {
        refs[k] = leaf_ref(child.first, child.count);
      } else {
        refs[k] = static_cast<int32_t>(out.n_nodes++);
        queue.push_back({kids[k], refs[k]});
      }
    }
    if (out.nodes.size() < static_cast<size_t>(out.n_nodes) * kNodeFloats)
      out.nodes.resize(std::max(out.nodes.size() * 2, static_cast<size_t>(out.n_nodes) * kNodeFloats));
    encode_node(lo, hi, refs, n_kids, &out.nodes[static_cast<size_t>(item.wide) * kNodeFloats]);
  }
  out.nodes.resize(static_cast<size_t>(out.n_nodes) * kNodeFloats);
  return out;
}

}  // namespace bvh_build
}  // namespace amber_cuda
