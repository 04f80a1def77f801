// CPU check of the host BVH builder (amber_b200/csrc/bvh_build.h): every triangle lands in
// exactly one leaf, every DECODED (quantised) child box contains its subtree, depth fits the
// device stack.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <vector>

#include "../../amber_b200/csrc/bvh_build.h"

using namespace amber_cuda::bvh_build;

struct Item {
  int32_t ref;  // 4-wide: child reference; 8-wide: node index
  float lo[3], hi[3];
  int depth;
};

static int check_leaf(const WideBvhHost& bvh, const float* verts, const Item& it, int first, int count, std::vector<int>& seen) {
  if (count < 1 || count > kLeafMax) { std::printf("bad leaf count %d\n", count); return 1; }
  for (int k = 0; k < count; k++) {
    const uint32_t tri = bvh.order[first + k];
    seen[tri]++;
    for (int v = 0; v < 3; v++)
      for (int a = 0; a < 3; a++) {
        const float p = verts[9 * (size_t)tri + 3 * v + a];
        if (p < it.lo[a] || p > it.hi[a]) { std::printf("vertex outside box\n"); return 1; }
      }
  }
  return 0;
}

// decode as the device does (bvh.cuh): plane = fmaf(g(code), scale, origin); a ray reaches a triangle only through
// every ancestor's box, so the intersection is carried down and the leaf check proves each triangle lies inside ALL
// the boxes on its path (a quantised child box may stick out of its parent's by a grid step; that is harmless)
static Item child_item(const Item& parent, const float* node, const uint8_t* codes /* [6][width] */, int width, int k) {
  Item child;
  child.ref = 0;
  child.depth = parent.depth + 1;
  for (int a = 0; a < 3; a++) {
    child.lo[a] = std::max(parent.lo[a], std::fmaf(plane_coordinate(codes[(2 * a) * width + k]), node[3 + a], node[a]));
    child.hi[a] = std::min(parent.hi[a], std::fmaf(plane_coordinate(codes[(2 * a + 1) * width + k]), node[3 + a], node[a]));
  }
  return child;
}

static int check(const WideBvhHost& bvh, const float* verts, uint32_t n) {
  std::vector<int> seen(n, 0);
  int max_depth = 0;
  std::vector<Item> stack;
  Item root{bvh.root, {-1e30f, -1e30f, -1e30f}, {1e30f, 1e30f, 1e30f}, kWidth == 8 ? 1 : 0};
  stack.push_back(root);
  while (!stack.empty()) {
    Item it = stack.back();
    stack.pop_back();
    if (kWidth == 4 && it.ref == kEmpty) continue;
    if (it.depth > max_depth) max_depth = it.depth;
    if (kWidth == 4 && it.ref < 0) {
      if (check_leaf(bvh, verts, it, (~it.ref) >> 3, (~it.ref) & 7, seen)) return 1;
      continue;
    }
    const float* node = &bvh.nodes[(size_t)it.ref * kNodeFloats];
    if (kWidth == 4) {
      uint8_t codes[6 * 4];
      std::memcpy(codes, &node[6], sizeof codes);
      for (int k = 0; k < 4; k++) {
        Item child = child_item(it, node, codes, 4, k);
        std::memcpy(&child.ref, &node[12 + k], 4);
        stack.push_back(child);
      }
      continue;
    }
    uint32_t words[24];
    std::memcpy(words, node, sizeof words);
    if (words[11] != kPlaneMagic) { std::printf("node without the plane magic\n"); return 1; }
    uint8_t meta[8], codes[6 * 8];
    std::memcpy(meta, &words[8], 8);
    std::memcpy(codes, &words[12], 48);
    uint32_t inner_rank = 0, next_offset = 0;
    for (int k = 0; k < 8; k++) {
      Item child = child_item(it, node, codes, 8, k);
      const bool inner = (words[10] >> k) & 1;
      if (inner) {
        if (meta[k] != ((1u << 5) | (24u + k))) { std::printf("bad inner meta\n"); return 1; }
        child.ref = (int32_t)(words[6] + inner_rank++);
        if ((uint32_t)child.ref >= bvh.n_nodes) { std::printf("child out of range\n"); return 1; }
        stack.push_back(child);
      } else if (meta[k]) {
        const int bits = meta[k] >> 5, offset = meta[k] & 31;
        const int count = bits == 1 ? 1 : bits == 3 ? 2 : bits == 7 ? 3 : -1;
        if ((uint32_t)offset != next_offset || offset + count > 24) { std::printf("bad leaf offset\n"); return 1; }
        next_offset += count;
        if (check_leaf(bvh, verts, child, (int)words[7] + offset, count, seen)) return 1;
      } else {
        for (int a = 0; a < 3; a++)
          if (codes[(2 * a) * 8 + k] != 255 || codes[(2 * a + 1) * 8 + k] != 0) { std::printf("empty slot with a box\n"); return 1; }
      }
    }
  }
  for (uint32_t i = 0; i < n; i++) if (seen[i] != 1) { std::printf("triangle %u seen %d times\n", i, seen[i]); return 1; }
  std::printf("ok n=%u nodes=%u depth=%d\n", n, bvh.n_nodes, max_depth);
  if (kWidth == 8) return max_depth != bvh.depth || max_depth + 1 > 64;
  return 3 * max_depth + 4 > 48 ? 1 : 0;
}

int main(int argc, char** argv) {
  const uint32_t n = argc > 1 ? std::atoi(argv[1]) : 20000;
  std::mt19937 rng(7);
  std::uniform_real_distribution<float> u(-1.f, 1.f), s(-0.02f, 0.02f);
  std::vector<float> verts(9 * (size_t)n);
  for (uint32_t i = 0; i < n; i++) {
    const float c[3] = {u(rng), u(rng), u(rng)};
    for (int v = 0; v < 3; v++) for (int a = 0; a < 3; a++) verts[9 * (size_t)i + 3 * v + a] = c[a] + s(rng);
  }
  // a block of coincident triangles: the SAH has no split for them
  for (uint32_t i = 0; i + 1 < n && i < 40; i++) std::memcpy(&verts[9 * (size_t)(i + 1)], &verts[0], 36);
  if (check(build(verts.data(), n), verts.data(), n)) return 1;
  if (check(build(verts.data(), 3), verts.data(), 3)) return 1;   // single leaf
  if (check(build(verts.data(), 0), verts.data(), 0)) return 1;   // empty
  return 0;
}
