// bvh_build_gpu.cuh -- the BVH built on the device: linear BVH (Morton order + binary radix tree), collapsed to the same
// 4-wide 64-byte quantised nodes as the host builder's (bvh.cuh), so the traversal kernels do not know which one made
// the tree.  Selected with AMBER_CUDA_BVH_BUILDER=lbvh (amber_cuda.cu, upload_scene); the default stays the host's
// binned SAH build, whose trees trace faster (DESIGN.md section 4.3 has both sets of numbers).
//
// Replaces, like bvh_build.h, BVH::BuildBVH (include/amber/raytracer/acceleration_bvh.h:135-312): tree shape is ours,
// closest-hit semantics are shared.
//   1. triangle boxes, centroids, scene bounds               (one kernel, block reduction + ordered-int atomics)
//   2. 63-bit Morton codes of the centroids, radix sort      (cub::DeviceRadixSort, the one library call)
//   3. binary radix tree over the sorted codes               (Karras 2012: one thread per internal node, duplicates
//                                                             broken by index)
//   4. boxes bottom-up                                        (one thread per leaf, the second child to arrive goes on)
//   5. collapse 2 -> 4 level by level                         (a task per wide node: adopt up to four descendants, always
//                                                             opening the one with the largest surface area; a subtree of
//                                                             <= kLeafMax triangles is a leaf -- its triangles are
//                                                             contiguous in Morton order; children get their node index
//                                                             from an atomic counter) and quantise (the host's
//                                                             quantise_axis, same plane codes, same conservative rounding)
// The sorted triangle order and the nodes go back to the host, which builds the per-triangle arrays as for its own trees.
#pragma once

#include <cub/device/device_radix_sort.cuh>

#include "bvh_build.h"

namespace amber_cuda {
namespace bvh_gpu {

#if AMBER_BVH_WIDTH != 4
// (the device builder writes 4-wide nodes only: with the 8-wide build option the host builds every tree)
inline cudaError_t build(const float*, uint32_t, cudaStream_t, bvh_build::WideBvhHost&) { return cudaErrorNotSupported; }
#else
constexpr int kBuildBlock = 256;
#ifndef AMBER_LBVH_LEAF_MAX
#define AMBER_LBVH_LEAF_MAX 4  // triangles per leaf of the device-built tree (<= bvh_build::kLeafMax)
#endif

struct Box3 {
  float3 lo, hi;
};
__device__ __forceinline__ Box3 box_union(const Box3& a, const Box3& b) {
  Box3 r;
  r.lo = make_float3(fminf(a.lo.x, b.lo.x), fminf(a.lo.y, b.lo.y), fminf(a.lo.z, b.lo.z));
  r.hi = make_float3(fmaxf(a.hi.x, b.hi.x), fmaxf(a.hi.y, b.hi.y), fmaxf(a.hi.z, b.hi.z));
  return r;
}
__device__ __forceinline__ float box_area(const Box3& b) {
  const float dx = b.hi.x - b.lo.x, dy = b.hi.y - b.lo.y, dz = b.hi.z - b.lo.z;
  return 2.0f * (dx * dy + dy * dz + dz * dx);
}

// float <-> unsigned with the same order, for atomicMin / atomicMax
__device__ __forceinline__ uint32_t ordered(float f) {
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__host__ __device__ __forceinline__ float unordered(uint32_t u) {
  const uint32_t v = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  float f;
#ifdef __CUDA_ARCH__
  f = __uint_as_float(v);
#else
  std::memcpy(&f, &v, 4);
#endif
  return f;
}

// bounds[0..2] = min of the centroids, [3..5] = max (ordered encoding)
__global__ void __launch_bounds__(kBuildBlock) prim_bounds_kernel(const float* __restrict__ verts9, uint32_t n, Box3* __restrict__ prim_box,
                                                                  uint32_t* __restrict__ bounds) {
  const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
  float3 c_lo = make_float3(CUDART_INF_F, CUDART_INF_F, CUDART_INF_F), c_hi = make_float3(-CUDART_INF_F, -CUDART_INF_F, -CUDART_INF_F);
  if (i < n) {
    const float* v = verts9 + 9 * static_cast<size_t>(i);
    Box3 b;
    b.lo = make_float3(fminf(fminf(v[0], v[3]), v[6]), fminf(fminf(v[1], v[4]), v[7]), fminf(fminf(v[2], v[5]), v[8]));
    b.hi = make_float3(fmaxf(fmaxf(v[0], v[3]), v[6]), fmaxf(fmaxf(v[1], v[4]), v[7]), fmaxf(fmaxf(v[2], v[5]), v[8]));
    prim_box[i] = b;
    c_lo = c_hi = make_float3(0.5f * (b.lo.x + b.hi.x), 0.5f * (b.lo.y + b.hi.y), 0.5f * (b.lo.z + b.hi.z));
  }
  const float lo[3] = {c_lo.x, c_lo.y, c_lo.z}, hi[3] = {c_hi.x, c_hi.y, c_hi.z};
#pragma unroll
  for (int a = 0; a < 3; a++) {
    float l = lo[a], h = hi[a];
#pragma unroll
    for (int delta = 16; delta > 0; delta >>= 1) {
      l = fminf(l, __shfl_xor_sync(0xffffffffu, l, delta));
      h = fmaxf(h, __shfl_xor_sync(0xffffffffu, h, delta));
    }
    if ((threadIdx.x & 31) == 0) {
      atomicMin(&bounds[a], ordered(l));
      atomicMax(&bounds[3 + a], ordered(h));
    }
  }
}

__device__ __forceinline__ unsigned long long spread21(uint32_t x) {  // 21 bits -> every third bit of 63
  unsigned long long v = x & 0x1fffffull;
  v = (v | (v << 32)) & 0x1f00000000ffffull;
  v = (v | (v << 16)) & 0x1f0000ff0000ffull;
  v = (v | (v << 8)) & 0x100f00f00f00f00full;
  v = (v | (v << 4)) & 0x10c30c30c30c30c3ull;
  v = (v | (v << 2)) & 0x1249249249249249ull;
  return v;
}

__global__ void __launch_bounds__(kBuildBlock) morton_kernel(const Box3* __restrict__ prim_box, uint32_t n, const uint32_t* __restrict__ bounds,
                                                             unsigned long long* __restrict__ codes, uint32_t* __restrict__ ids) {
  const uint32_t i = blockIdx.x * kBuildBlock + threadIdx.x;
  if (i >= n) return;
  const Box3 b = prim_box[i];
  const float c[3] = {0.5f * (b.lo.x + b.hi.x), 0.5f * (b.lo.y + b.hi.y), 0.5f * (b.lo.z + b.hi.z)};
  unsigned long long code = 0;
#pragma unroll
  for (int a = 0; a < 3; a++) {
    const float lo = unordered(bounds[a]), hi = unordered(bounds[3 + a]);
    const float extent = hi - lo;
    const float u = extent > 0.0f ? (c[a] - lo) / extent : 0.0f;
    const uint32_t q = min(2097151u, static_cast<uint32_t>(fmaxf(u, 0.0f) * 2097152.0f));
    code |= spread21(q) << a;
  }
  codes[i] = code;
  ids[i] = i;
}

// Node numbering of the binary tree: internal nodes 0 .. n-2 (0 = root), leaf k = n-1+k (k in sorted order).
struct RadixTree {
  int32_t* parent;  // [2n-1]
  int32_t* left;    // [n-1]
  int32_t* right;   // [n-1]
  uint32_t* first;  // [n-1] range of sorted triangles under an internal node
  uint32_t* last;
  Box3* box;        // [2n-1]
  uint32_t* arrived;  // [n-1] bottom-up fit: children that have reported
};

__device__ __forceinline__ int common_prefix(const unsigned long long* codes, int n, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const unsigned long long a = codes[i], b = codes[j];
  if (a != b) return __clzll(static_cast<long long>(a ^ b));
  return 64 + __clz(i ^ j);  // equal codes: the index breaks the tie
}

__global__ void __launch_bounds__(kBuildBlock) radix_tree_kernel(const unsigned long long* __restrict__ codes, int n, RadixTree tree) {
  const int i = blockIdx.x * kBuildBlock + threadIdx.x;
  if (i >= n - 1) return;
  // Karras, "Maximizing parallelism in the construction of BVHs, octrees, and k-d trees", 2012, fig. 4
  const int d = common_prefix(codes, n, i, i + 1) > common_prefix(codes, n, i, i - 1) ? 1 : -1;
  const int delta_min = common_prefix(codes, n, i, i - d);
  int l_max = 2;
  while (common_prefix(codes, n, i, i + l_max * d) > delta_min) l_max <<= 1;
  int l = 0;
  for (int t = l_max >> 1; t > 0; t >>= 1)
    if (common_prefix(codes, n, i, i + (l + t) * d) > delta_min) l += t;
  const int j = i + l * d;
  const int delta_node = common_prefix(codes, n, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1;; t = (t + 1) >> 1) {
    if (common_prefix(codes, n, i, i + (s + t) * d) > delta_node) s += t;
    if (t == 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int lo = min(i, j), hi = max(i, j);
  const int left = lo == gamma ? (n - 1) + gamma : gamma;
  const int right = hi == gamma + 1 ? (n - 1) + gamma + 1 : gamma + 1;
  tree.left[i] = left;
  tree.right[i] = right;
  tree.first[i] = static_cast<uint32_t>(lo);
  tree.last[i] = static_cast<uint32_t>(hi);
  tree.parent[left] = i;
  tree.parent[right] = i;
  if (i == 0) tree.parent[0] = -1;
}

__global__ void __launch_bounds__(kBuildBlock) fit_kernel(const Box3* __restrict__ prim_box, const uint32_t* __restrict__ ids, int n, RadixTree tree) {
  const int k = blockIdx.x * kBuildBlock + threadIdx.x;
  if (k >= n) return;
  int node = (n - 1) + k;
  tree.box[node] = prim_box[ids[k]];
  __threadfence();
  for (node = tree.parent[node]; node >= 0; node = tree.parent[node]) {
    if (atomicAdd(&tree.arrived[node], 1u) == 0) return;  // the first child to arrive leaves the node to the second
    __threadfence();
    tree.box[node] = box_union(tree.box[tree.left[node]], tree.box[tree.right[node]]);
    __threadfence();
  }
}

// ---- collapse + quantise -------------------------------------------------------------------------------------
__device__ __forceinline__ float plane_coordinate_device(int code) { return __uint_as_float(bvh_build::kPlaneMagic | (static_cast<uint32_t>(code) << 16)); }

// bvh_build.h quantise_axis for four children, on the device: same grid, same search, same conservative rounding.
__device__ void quantise_axis_device(const float* lo, const float* hi, int n_kids, float& origin, float& scale, uint32_t& word_lo,
                                     uint32_t& word_hi) {
  float bottom = CUDART_INF_F, top = -CUDART_INF_F;
  for (int k = 0; k < n_kids; k++) {
    bottom = fminf(bottom, lo[k]);
    top = fmaxf(top, hi[k]);
  }
  const float span = plane_coordinate_device(255) - plane_coordinate_device(0);
  float margin = 0.0f;
  for (bool ok = false; !ok;) {
    ok = true;
    const float s = fmaxf(((top + margin) - (bottom - margin)) / span, 1e-30f);
    const float o = (bottom - margin) - plane_coordinate_device(0) * s;
    word_lo = word_hi = 0;
    for (int k = 0; k < 4 && ok; k++) {
      int a = 255, b = 0;
      if (k < n_kids) {
        a = 0;
        for (int step = 128; step > 0; step >>= 1)
          if (a + step <= 255 && fmaf(plane_coordinate_device(a + step), s, o) <= lo[k]) a += step;
        b = 255;
        for (int step = 128; step > 0; step >>= 1)
          if (b - step >= 0 && fmaf(plane_coordinate_device(b - step), s, o) >= hi[k]) b -= step;
        if (fmaf(plane_coordinate_device(a), s, o) > lo[k] || fmaf(plane_coordinate_device(b), s, o) < hi[k]) ok = false;
      }
      word_lo |= static_cast<uint32_t>(a) << (8 * k);
      word_hi |= static_cast<uint32_t>(b) << (8 * k);
    }
    origin = o;
    scale = s;
    if (!ok) margin = fmaxf(2.0f * margin, 1.2e-7f * fmaxf(fmaxf(fabsf(bottom), fabsf(top)), 1e-30f));
  }
}

struct CollapseTask {
  int32_t node2;  // internal node of the binary tree
  int32_t wide;   // index of the 4-wide node it becomes
};

__global__ void __launch_bounds__(kBuildBlock) collapse_kernel(RadixTree tree, int n, const CollapseTask* __restrict__ tasks, uint32_t n_tasks,
                                                               CollapseTask* __restrict__ next_tasks, uint32_t* __restrict__ counters /* [0] nodes, [1] next tasks */,
                                                               float4* __restrict__ nodes, float scene_scale) {
  const uint32_t t = blockIdx.x * kBuildBlock + threadIdx.x;
  if (t >= n_tasks) return;
  const CollapseTask task = tasks[t];
  auto triangles_under = [&](int node) -> uint32_t { return node >= n - 1 ? 1u : tree.last[node] - tree.first[node] + 1u; };
  auto is_leaf = [&](int node) { return triangles_under(node) <= static_cast<uint32_t>(AMBER_LBVH_LEAF_MAX); };
  int kids[4];
  int n_kids = 2;
  kids[0] = tree.left[task.node2];
  kids[1] = tree.right[task.node2];
  while (n_kids < 4) {
    int best = -1;
    float best_area = -1.0f;
    for (int k = 0; k < n_kids; k++) {
      if (is_leaf(kids[k])) continue;
      const float a = box_area(tree.box[kids[k]]);
      if (a > best_area) {
        best_area = a;
        best = k;
      }
    }
    if (best < 0) break;
    const int open = kids[best];
    kids[best] = tree.left[open];
    kids[n_kids++] = tree.right[open];
  }
  float lo[3][4], hi[3][4];
  int32_t refs[4] = {bvh_build::kEmpty, bvh_build::kEmpty, bvh_build::kEmpty, bvh_build::kEmpty};
  for (int k = 0; k < n_kids; k++) {
    const Box3 b = tree.box[kids[k]];
    const float pad = 2e-6f * scene_scale;  // bvh_build.h, padded()
    lo[0][k] = b.lo.x - pad; lo[1][k] = b.lo.y - pad; lo[2][k] = b.lo.z - pad;
    hi[0][k] = b.hi.x + pad; hi[1][k] = b.hi.y + pad; hi[2][k] = b.hi.z + pad;
    if (is_leaf(kids[k])) {
      const uint32_t first = kids[k] >= n - 1 ? static_cast<uint32_t>(kids[k] - (n - 1)) : tree.first[kids[k]];
      refs[k] = ~static_cast<int32_t>((first << 3) | triangles_under(kids[k]));
    } else {
      refs[k] = static_cast<int32_t>(atomicAdd(&counters[0], 1u));
      next_tasks[atomicAdd(&counters[1], 1u)] = CollapseTask{kids[k], refs[k]};
    }
  }
  float origin[3], scale[3];
  uint32_t words[6];
  for (int a = 0; a < 3; a++) quantise_axis_device(lo[a], hi[a], n_kids, origin[a], scale[a], words[2 * a], words[2 * a + 1]);
  float4* out = nodes + 4 * static_cast<size_t>(task.wide);
  out[0] = make_float4(origin[0], origin[1], origin[2], scale[0]);
  out[1] = make_float4(scale[1], scale[2], __uint_as_float(words[0]), __uint_as_float(words[1]));
  out[2] = make_float4(__uint_as_float(words[2]), __uint_as_float(words[3]), __uint_as_float(words[4]), __uint_as_float(words[5]));
  out[3] = make_float4(__int_as_float(refs[0]), __int_as_float(refs[1]), __int_as_float(refs[2]), __int_as_float(refs[3]));
}

template <class T>
struct Scratch {  // freed when the build returns
  T* ptr = nullptr;
  cudaError_t alloc(size_t count) { return cudaMalloc(&ptr, std::max<size_t>(1, count) * sizeof(T)); }
  ~Scratch() {
    if (ptr) cudaFree(ptr);
  }
};

// Returns cudaSuccess and fills `out` (same contents as bvh_build::build would give for its tree); n must exceed kLeafMax.
inline cudaError_t build(const float* verts9, uint32_t n_tris, cudaStream_t stream, bvh_build::WideBvhHost& out) {
  const int n = static_cast<int>(n_tris);
  const unsigned int grid_n = (n_tris + kBuildBlock - 1) / kBuildBlock;
#define BUILD_TRY(expr)                  \
  do {                                   \
    const cudaError_t e__ = (expr);      \
    if (e__ != cudaSuccess) return e__;  \
  } while (0)
  Scratch<float> d_verts;
  Scratch<Box3> d_prim_box, d_box;
  Scratch<uint32_t> d_bounds, d_ids, d_ids_sorted, d_first, d_last, d_arrived, d_counters;
  Scratch<unsigned long long> d_codes, d_codes_sorted;
  Scratch<int32_t> d_parent, d_left, d_right;
  Scratch<CollapseTask> d_tasks[2];
  Scratch<float4> d_nodes;
  Scratch<unsigned char> d_sort_temp;
  BUILD_TRY(d_verts.alloc(9 * static_cast<size_t>(n_tris)));
  BUILD_TRY(d_prim_box.alloc(n_tris));
  BUILD_TRY(d_box.alloc(2 * static_cast<size_t>(n_tris)));
  BUILD_TRY(d_bounds.alloc(6));
  BUILD_TRY(d_ids.alloc(n_tris));
  BUILD_TRY(d_ids_sorted.alloc(n_tris));
  BUILD_TRY(d_codes.alloc(n_tris));
  BUILD_TRY(d_codes_sorted.alloc(n_tris));
  BUILD_TRY(d_first.alloc(n_tris));
  BUILD_TRY(d_last.alloc(n_tris));
  BUILD_TRY(d_arrived.alloc(n_tris));
  BUILD_TRY(d_counters.alloc(2));
  BUILD_TRY(d_parent.alloc(2 * static_cast<size_t>(n_tris)));
  BUILD_TRY(d_left.alloc(n_tris));
  BUILD_TRY(d_right.alloc(n_tris));
  BUILD_TRY(d_tasks[0].alloc(n_tris));
  BUILD_TRY(d_tasks[1].alloc(n_tris));
  BUILD_TRY(d_nodes.alloc(4 * static_cast<size_t>(n_tris)));  // fewer wide nodes than triangles
  BUILD_TRY(cudaMemcpyAsync(d_verts.ptr, verts9, 9 * static_cast<size_t>(n_tris) * sizeof(float), cudaMemcpyHostToDevice, stream));
  const uint32_t bounds_init[6] = {0xffffffffu, 0xffffffffu, 0xffffffffu, 0u, 0u, 0u};
  BUILD_TRY(cudaMemcpyAsync(d_bounds.ptr, bounds_init, sizeof bounds_init, cudaMemcpyHostToDevice, stream));
  BUILD_TRY(cudaMemsetAsync(d_arrived.ptr, 0, n_tris * sizeof(uint32_t), stream));
  prim_bounds_kernel<<<grid_n, kBuildBlock, 0, stream>>>(d_verts.ptr, n_tris, d_prim_box.ptr, d_bounds.ptr);
  morton_kernel<<<grid_n, kBuildBlock, 0, stream>>>(d_prim_box.ptr, n_tris, d_bounds.ptr, d_codes.ptr, d_ids.ptr);
  size_t temp_bytes = 0;
  BUILD_TRY(cub::DeviceRadixSort::SortPairs(nullptr, temp_bytes, d_codes.ptr, d_codes_sorted.ptr, d_ids.ptr, d_ids_sorted.ptr, n, 0, 63, stream));
  BUILD_TRY(d_sort_temp.alloc(temp_bytes));
  BUILD_TRY(cub::DeviceRadixSort::SortPairs(d_sort_temp.ptr, temp_bytes, d_codes.ptr, d_codes_sorted.ptr, d_ids.ptr, d_ids_sorted.ptr, n, 0, 63, stream));
  RadixTree tree{d_parent.ptr, d_left.ptr, d_right.ptr, d_first.ptr, d_last.ptr, d_box.ptr, d_arrived.ptr};
  radix_tree_kernel<<<grid_n, kBuildBlock, 0, stream>>>(d_codes_sorted.ptr, n, tree);
  fit_kernel<<<grid_n, kBuildBlock, 0, stream>>>(d_prim_box.ptr, d_ids_sorted.ptr, n, tree);
  // the root box gives the padding scale (bvh_build.h, padded())
  Box3 root_box;
  BUILD_TRY(cudaMemcpyAsync(&root_box, d_box.ptr, sizeof(Box3), cudaMemcpyDeviceToHost, stream));
  BUILD_TRY(cudaStreamSynchronize(stream));
  float scene_scale = 1.0f;
  for (float v : {root_box.lo.x, root_box.lo.y, root_box.lo.z, root_box.hi.x, root_box.hi.y, root_box.hi.z})
    scene_scale = std::max(scene_scale, std::fabs(v));
  // collapse, level by level
  const CollapseTask root_task{0, 0};
  BUILD_TRY(cudaMemcpyAsync(d_tasks[0].ptr, &root_task, sizeof root_task, cudaMemcpyHostToDevice, stream));
  uint32_t counters[2] = {1u, 0u};  // wide nodes so far, tasks of the next level
  uint32_t n_tasks = 1;
  for (int level = 0; n_tasks > 0; level++) {
    counters[1] = 0;
    BUILD_TRY(cudaMemcpyAsync(d_counters.ptr, counters, sizeof counters, cudaMemcpyHostToDevice, stream));
    collapse_kernel<<<(n_tasks + kBuildBlock - 1) / kBuildBlock, kBuildBlock, 0, stream>>>(
        tree, n, d_tasks[level & 1].ptr, n_tasks, d_tasks[(level + 1) & 1].ptr, d_counters.ptr, d_nodes.ptr, scene_scale);
    BUILD_TRY(cudaMemcpyAsync(counters, d_counters.ptr, sizeof counters, cudaMemcpyDeviceToHost, stream));
    BUILD_TRY(cudaStreamSynchronize(stream));
    n_tasks = counters[1];
    out.depth = level + 1;
  }
  BUILD_TRY(cudaGetLastError());
  out.n_nodes = counters[0];
  out.root = 0;
  out.nodes.resize(static_cast<size_t>(out.n_nodes) * bvh_build::kNodeFloats);
  out.order.resize(n_tris);
  BUILD_TRY(cudaMemcpyAsync(out.nodes.data(), d_nodes.ptr, out.nodes.size() * sizeof(float), cudaMemcpyDeviceToHost, stream));
  BUILD_TRY(cudaMemcpyAsync(out.order.data(), d_ids_sorted.ptr, n_tris * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  BUILD_TRY(cudaStreamSynchronize(stream));
#undef BUILD_TRY
  return cudaSuccess;
}
#endif  // AMBER_BVH_WIDTH == 4

}  // namespace bvh_gpu
}  // namespace amber_cuda
