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

static int check(const WideBvhHost& bvh, const float* verts, uint32_t n) {
  std::vector<int> seen(n, 0);
  int max_depth = 0;
  struct Item { int32_t ref; float lo[3], hi[3]; int depth; };
  std::vector<Item> stack;
  Item root{bvh.root, {-1e30f, -1e30f, -1e30f}, {1e30f, 1e30f, 1e30f}, 0};
  stack.push_back(root);
  while (!stack.empty()) {
    Item it = stack.back();
    stack.pop_back();
    if (it.ref == kEmpty) continue;
    if (it.depth > max_depth) max_depth = it.depth;
    if (it.ref < 0) {
      const int first = (~it.ref) >> 3, count = (~it.ref) & 7;
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
      continue;
    }
    // decode as the device does (bvh.cuh): plane = fmaf(q, scale, origin)
    const float* node = &bvh.nodes[(size_t)it.ref * kNodeFloats];
    uint32_t words[6];
    std::memcpy(words, &node[6], sizeof words);
    for (int k = 0; k < 4; k++) {
      Item child;
      std::memcpy(&child.ref, &node[12 + k], 4);
      for (int a = 0; a < 3; a++) {
        child.lo[a] = std::fmaf((float)((words[2 * a] >> (8 * k)) & 0xff), node[3 + a], node[a]);
        child.hi[a] = std::fmaf((float)((words[2 * a + 1] >> (8 * k)) & 0xff), node[3 + a], node[a]);
      }
      child.depth = it.depth + 1;
      // a ray reaches a triangle only through every ancestor's box: carry the intersection down,
      // so the leaf check above proves each triangle lies inside ALL the boxes on its path
      // (a quantised child box may stick out of its parent's by a grid step; that is harmless)
      for (int a = 0; a < 3; a++) {
        child.lo[a] = std::max(child.lo[a], it.lo[a]);
        child.hi[a] = std::min(child.hi[a], it.hi[a]);
      }
      stack.push_back(child);
    }
  }
  for (uint32_t i = 0; i < n; i++) if (seen[i] != 1) { std::printf("triangle %u seen %d times\n", i, seen[i]); return 1; }
  std::printf("ok n=%u nodes=%u depth=%d\n", n, bvh.n_nodes, max_depth);
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
