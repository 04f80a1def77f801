// amber_cuda.cu -- host side of libamber_cuda.so: the C ABI of include/amber_cuda.h.
//
// Scene preparation (flattening, light table, lens constants, acceleration build)
// runs here on the host in float arithmetic that mirrors the reference constructors
// it replaces; everything per-ray runs in the kernels of kernels.cuh.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <condition_variable>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/amber_cuda.h"
#include "bvh.cuh"
#include "bvh_build.h"
#include "kernels.cuh"
#include "bvh_build_gpu.cuh"

namespace amber_cuda {


namespace detail {

thread_local std::string g_create_error;
std::atomic<uint64_t> g_next_ctx_id{1};
constexpr int kMaxDevices = 64;
// which context's data currently sits in this device's __constant__ banks
std::atomic<uint64_t> g_const_owner[kMaxDevices];
// The scene and the camera of a context live in process-wide __constant__ symbols while its kernels
// run, so contexts that share a device take turns: this lock is held from bind_constants to the last
// synchronisation of a render job or parity call (DeviceTurn below).
std::mutex g_device_turn[kMaxDevices];

template <class T>
struct DeviceBuffer {
  T* ptr = nullptr;
  size_t count = 0;
  cudaError_t resize(size_t n) {
    if (n <= count && ptr) return cudaSuccess;
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    count = 0;
    if (n == 0) return cudaSuccess;
    const cudaError_t err = cudaMalloc(&ptr, n * sizeof(T));
    if (err == cudaSuccess) count = n;
    return err;
  }
  cudaError_t upload(const std::vector<T>& host, cudaStream_t stream) {
    cudaError_t err = resize(host.size());
    if (err != cudaSuccess || host.empty()) return err;
    return cudaMemcpyAsync(ptr, host.data(), host.size() * sizeof(T), cudaMemcpyHostToDevice, stream);
  }
  void release() {
    if (ptr) cudaFree(ptr);
    ptr = nullptr;
    count = 0;
  }
};

}  // namespace detail
}  // namespace amber_cuda

using namespace amber_cuda;
using namespace amber_cuda::detail;

struct amber_cuda_ctx {
  uint64_t id = 0;
  int device = 0;
  cudaStream_t stream = nullptr;
  cudaEvent_t ev_begin = nullptr, ev_end = nullptr;
  std::string error;

  // scene
  bool has_scene = false;
  bool use_bvh = false;
  ConstScene const_scene;
  CameraDev camera;
  bool has_sensor = false;
  SceneView view{};
  std::vector<int32_t> internal_to_prim;
  DeviceBuffer<TriShade> d_tri_shade;
  DeviceBuffer<TriIsect> d_tri_isect;
  DeviceBuffer<float4> d_spheres;
  DeviceBuffer<int32_t> d_tri_material, d_sphere_material, d_internal_to_prim;
  DeviceBuffer<MaterialDev> d_materials;
  DeviceBuffer<uint8_t> d_prim_class;
  DeviceBuffer<LightDev> d_lights;
  DeviceBuffer<float4> d_bvh_nodes;
  uint32_t bvh_nodes = 0;
  uint32_t bvh_stack_entries = 0;  // per-lane traversal stack: 3 * depth + 1
  unsigned int bvh_grid = 0;       // persistent grid: SMs * resident blocks
  size_t bvh_smem = 0;

  // film: W*H float4 sums + one trailing float4 that carries the sample count across an external
  // all-reduce (amber_cuda_film_device)
  DeviceBuffer<float4> d_film;
  DeviceBuffer<float> d_resolved;
  DeviceBuffer<uint8_t> d_resolved8;
  uint64_t samples_done = 0;
  bool samples_in_film = false;  // the trailing element is authoritative (after an external reduce)
  float4* h_tail = nullptr;      // pinned staging for that element

  // path pool
  PathPool pool{};
  DeviceBuffer<float4> d_ray, d_state, d_shadow;
  DeviceBuffer<float2> d_hit;
  DeviceBuffer<uint32_t> d_q_new, d_q_shadow;
  // wavefront BDPT (bdpt.cuh, BdptPool): allocated by the first such render
  DeviceBuffer<float4> d_bdpt_state;
  DeviceBuffer<BdptEvent> d_bdpt_events;
  DeviceBuffer<uint32_t> d_bdpt_queue;
  DeviceBuffer<uint2> d_bdpt_evaluate;
  DeviceBuffer<float4> d_bdpt_evaluate_ray;
  DeviceBuffer<float2> d_bdpt_evaluate_hit;
  DeviceBuffer<Control> d_ctl;
  Control* h_ctl = nullptr;  // pinned

  // scratch for the host-buffer entry points
  DeviceBuffer<float> d_scratch_f[6];
  DeviceBuffer<int32_t> d_scratch_i[2];
  DeviceBuffer<float4> d_scratch_4[3];

  // statistics
  amber_cuda_stats stats{};
  bool profiling = false;

  // render queue: amber_cuda_render_async hands jobs to this context's worker thread
  std::thread worker;
  std::mutex queue_mutex;
  std::condition_variable queue_cv, idle_cv;
  std::deque<amber_cuda_render_params> jobs;
  int running = 0;    // calls taken by the worker whose paths have not all been started yet
  bool busy = false;  // the worker is inside a render (its wavefront may still be draining)
  bool jobs_may_follow = false;
  std::deque<uint64_t> absorbed_ends;  // end ordinal (samples per pixel) of each call in the running wavefront
  bool quit = false;
  int job_rc = AMBER_CUDA_OK;  // first failure of the queue, reported by the next wait
  std::string job_error;
};

namespace amber_host {

int fail(amber_cuda_ctx* ctx, int code, const char* fmt, ...) {
  char buffer[512];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buffer, sizeof buffer, fmt, ap);
  va_end(ap);
  if (ctx) ctx->error = buffer; else g_create_error = buffer;
  return code;
}

#define CUDA_TRY(ctx, expr)                                                                   \
  do {                                                                                        \
    const cudaError_t err__ = (expr);                                                         \
    if (err__ != cudaSuccess)                                                                 \
      return fail(ctx, AMBER_CUDA_ECUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(err__), \
                  __FILE__, __LINE__);                                                        \
  } while (0)

// Holds a device's turn (see g_device_turn) for the length of a scope.
struct DeviceTurn {
  std::unique_lock<std::mutex> lock;
  explicit DeviceTurn(const amber_cuda_ctx* ctx) {
    if (ctx->device >= 0 && ctx->device < kMaxDevices) lock = std::unique_lock<std::mutex>(g_device_turn[ctx->device]);
  }
};

void worker_main(amber_cuda_ctx* ctx);
int wait_idle(amber_cuda_ctx* ctx, int max_pending = 0);

// AMBER_CUDA_TRACE=1 in the environment: one line per queue event on stderr (microseconds since load)
inline bool trace_on() {
  static const bool on = std::getenv("AMBER_CUDA_TRACE") != nullptr;
  return on;
}
inline void trace(const amber_cuda_ctx* ctx, const char* fmt, ...) {
  if (!trace_on()) return;
  static const auto t0 = std::chrono::steady_clock::now();
  char buffer[256];
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(buffer, sizeof buffer, fmt, ap);
  va_end(ap);
  std::fprintf(stderr, "[amber_cuda %10.0f us ctx %llu] %s\n",
               std::chrono::duration<double, std::micro>(std::chrono::steady_clock::now() - t0).count(),
               static_cast<unsigned long long>(ctx->id), buffer);
}

inline unsigned int blocks_for(uint64_t n, int block = kBlock) {
  return static_cast<unsigned int>((n + block - 1) / block);
}

// --- float3 host helpers: every operation rounds to float, as the reference's
// Vector3<float> arithmetic does (built without FMA, CMakeLists.txt:5) ------------
struct F3 {
  float x, y, z;
};
inline F3 sub(F3 a, F3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline F3 crossf(F3 u, F3 v) {
  return {u.y * v.z - u.z * v.y, u.z * v.x - u.x * v.z, u.x * v.y - u.y * v.x};
}
inline float dotf(F3 a, F3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float lengthf(F3 a) { return std::sqrt(dotf(a, a)); }
inline F3 normalizef(F3 a) {
  const float l = lengthf(a);
  return {a.x / l, a.y / l, a.z / l};
}

void make_triangle(const float* v, TriIsect& isect, TriShade& shade) {
  const F3 v0{v[0], v[1], v[2]}, v1{v[3], v[4], v[5]}, v2{v[6], v[7], v[8]};
  const F3 e1 = sub(v1, v0), e2 = sub(v2, v0);
  const F3 normal = normalizef(crossf(e1, e2));  // primitive_triangle.cc:59-68
  shade.v0_nx = make_float4(v0.x, v0.y, v0.z, normal.x);
  shade.e1_ny = make_float4(e1.x, e1.y, e1.z, normal.y);
  shade.e2_nz = make_float4(e2.x, e2.y, e2.z, normal.z);
  // plane form from the float edges, evaluated in double and rounded once
  const double a[3] = {e1.x, e1.y, e1.z}, b[3] = {e2.x, e2.y, e2.z};
  const double n[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  const double nn = n[0] * n[0] + n[1] * n[1] + n[2] * n[2];
  const double nu[3] = {(b[1] * n[2] - b[2] * n[1]) / nn, (b[2] * n[0] - b[0] * n[2]) / nn,
                        (b[0] * n[1] - b[1] * n[0]) / nn};
  const double nv[3] = {(n[1] * a[2] - n[2] * a[1]) / nn, (n[2] * a[0] - n[0] * a[2]) / nn,
                        (n[0] * a[1] - n[1] * a[0]) / nn};
  isect.n_v0x = make_float4((float)n[0], (float)n[1], (float)n[2], v0.x);
  isect.nu_v0y = make_float4((float)nu[0], (float)nu[1], (float)nu[2], v0.y);
  isect.nv_v0z = make_float4((float)nv[0], (float)nv[1], (float)nv[2], v0.z);
}


// Batches the triangles of a small scene by supporting plane for the constant-memory list.
// `tri_ids` (creation-order ids) is permuted so that each group is contiguous: the planes
// perpendicular to x, then y, then z (AxisEntry: no dot products), then the other groups, each in
// order of their first member; inside a group the parallelogram pairs (earlier triangle first, so a
// hit on the shared diagonal goes to it, as in a creation-order scan), then the lone
// triangles.  All geometry is evaluated in double from the float vertices.
struct ConstListHost {
  std::vector<AxisEntry> axis_entries;
  std::vector<uint8_t> axis_tri;
  int32_t n_axis_entries[3] = {0, 0, 0};
  std::vector<AxisCluster> clusters;
  std::vector<PlaneGroup> groups;
  std::vector<PlaneEntry> entries;
};

void group_by_plane(const amber_cuda_primitive* prims, std::vector<uint32_t>& tri_ids, ConstListHost& out) {
  struct D3 { double x, y, z; };
  auto sub = [](D3 a, D3 b) { return D3{a.x - b.x, a.y - b.y, a.z - b.z}; };
  auto cross = [](D3 a, D3 b) { return D3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; };
  auto dot = [](D3 a, D3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; };
  auto vertex = [&](uint32_t id, int k) {
    const float* d = prims[id].data + 3 * k;
    return D3{d[0], d[1], d[2]};
  };
  struct Group { D3 anchor, unit_n, n; double scale; std::vector<uint32_t> members; };
  std::vector<Group> found;
  for (uint32_t id : tri_ids) {
    const D3 v0 = vertex(id, 0), v1 = vertex(id, 1), v2 = vertex(id, 2);
    const D3 n = cross(sub(v1, v0), sub(v2, v0));
    const double len = std::sqrt(dot(n, n));
    int home = -1;
    if (len > 0) {
      const D3 unit{n.x / len, n.y / len, n.z / len};
      for (size_t g = 0; g < found.size() && home < 0; g++) {
        const Group& G = found[g];
        const D3 c = cross(unit, G.unit_n);
        if (dot(c, c) > 1e-14) continue;  // not parallel
        const double tol = 1e-7 * G.scale;
        if (std::fabs(dot(G.unit_n, sub(v0, G.anchor))) <= tol && std::fabs(dot(G.unit_n, sub(v1, G.anchor))) <= tol &&
            std::fabs(dot(G.unit_n, sub(v2, G.anchor))) <= tol)
          home = static_cast<int>(g);
      }
      if (home < 0) {
        const double scale = std::max({1.0, std::fabs(v0.x), std::fabs(v0.y), std::fabs(v0.z)});
        found.push_back({v0, unit, n, scale, {}});
        home = static_cast<int>(found.size()) - 1;
      }
    } else {  // degenerate triangle: a group of its own; its NaN plane rejects every ray, as the
      found.push_back({v0, D3{0, 0, 0}, n, 1.0, {}});  // reference's det == 0 does (hit.h:90-94)
      home = static_cast<int>(found.size()) - 1;
    }
    found[home].members.push_back(id);
  }
  auto same = [](D3 a, D3 b) { return a.x == b.x && a.y == b.y && a.z == b.z; };
  auto comp = [](D3 a, int k) { return k == 0 ? a.x : k == 1 ? a.y : a.z; };
  struct Corner { D3 r, e1, e2; };  // u = nu.(x - r), v = nv.(x - r) with nu, nv from e1, e2
  auto edge_vectors = [&](const Corner& c, D3& nu, D3& nv) {
    const D3 n = cross(c.e1, c.e2);
    const double nn = dot(n, n);
    nu = cross(c.e2, n);
    nv = cross(n, c.e1);
    nu = D3{nu.x / nn, nu.y / nn, nu.z / nn};
    nv = D3{nv.x / nn, nv.y / nn, nv.z / nn};
  };
  struct Batch { int axis; std::vector<Corner> quads, tris; std::vector<uint32_t> quad_ids, lone_ids; };
  std::vector<Batch> batches;
  for (const Group& G : found) {
    // pair up triangles that form a parallelogram: A = (r, p, q), B = {p, q, p + q - r}
    const size_t m = G.members.size();
    std::vector<int> partner(m, -1);
    Batch batch;
    batch.axis = -1;
    for (int k = 0; k < 3 && batch.axis < 0; k++) {  // every vertex has the anchor's k-coordinate exactly
      bool flat = G.scale > 0 && comp(G.unit_n, k) != 0;
      for (size_t i = 0; i < m && flat; i++)
        for (int c = 0; c < 3; c++) flat = flat && comp(vertex(G.members[i], c), k) == comp(G.anchor, k);
      if (flat) batch.axis = k;
    }
    for (size_t a = 0; a < m; a++) {
      if (partner[a] >= 0) continue;
      const D3 A[3] = {vertex(G.members[a], 0), vertex(G.members[a], 1), vertex(G.members[a], 2)};
      for (size_t b = a + 1; b < m && partner[a] < 0; b++) {
        if (partner[b] >= 0) continue;
        const D3 B[3] = {vertex(G.members[b], 0), vertex(G.members[b], 1), vertex(G.members[b], 2)};
        for (int lone = 0; lone < 3 && partner[a] < 0; lone++) {  // A's corner off the shared edge
          const D3 r = A[lone], p = A[(lone + 1) % 3], q = A[(lone + 2) % 3];
          const D3 fourth{p.x + q.x - r.x, p.y + q.y - r.y, p.z + q.z - r.z};
          int shared = 0, other = -1;
          for (int k = 0; k < 3; k++) {
            if (same(B[k], p) || same(B[k], q)) shared++; else other = k;
          }
          if (shared != 2 || other < 0) continue;
          const D3 diff = sub(B[other], fourth);
          if (std::sqrt(dot(diff, diff)) > 1e-6 * G.scale) continue;
          partner[a] = static_cast<int>(b);
          partner[b] = static_cast<int>(a);
          batch.quads.push_back({r, sub(p, r), sub(q, r)});
          batch.quad_ids.push_back(G.members[a]);
          batch.quad_ids.push_back(G.members[b]);
        }
      }
      if (partner[a] < 0) {
        batch.tris.push_back({A[0], sub(A[1], A[0]), sub(A[2], A[0])});
        batch.lone_ids.push_back(G.members[a]);
      }
    }
    batches.push_back(std::move(batch));
  }

  // emit: entries perpendicular to x, y, z (AxisEntry), the clustered planes (three or more entries
  // in one plane: AxisCluster), then everything else (PlaneGroup)
  tri_ids.clear();
  auto clustered = [&](const Batch& B) { return B.quads.size() + B.tris.size() >= 3; };
  size_t n_clustered = 0;
  for (const Batch& B : batches) n_clustered += B.axis >= 0 && clustered(B);
  const bool use_clusters = n_clustered <= static_cast<size_t>(kMaxAxisClusters);
  for (int pass = 0; pass < 4; pass++) {  // pass 0-2: unclustered entries of axis x, y, z; pass 3: clusters
    for (size_t g = 0; g < found.size(); g++) {
      const Batch& B = batches[g];
      if (B.axis < 0) continue;
      const bool in_cluster = use_clusters && clustered(B);
      if (pass < 3 ? (in_cluster || B.axis != pass) : !in_cluster) continue;
      const int axis = B.axis;
      const int ka = (axis + 1) % 3, kb = (axis + 2) % 3;
      AxisCluster cluster{};
      cluster.first = static_cast<int32_t>(out.axis_entries.size());
      cluster.axis = axis;
      cluster.r_k = (float)comp(found[g].anchor, axis);
      double lo_a = 1e300, hi_a = -1e300, lo_b = 1e300, hi_b = -1e300;
      uint32_t tri = static_cast<uint32_t>(tri_ids.size());
      for (int is_tri = 0; is_tri < 2; is_tri++) {
        for (const Corner& c : is_tri ? B.tris : B.quads) {
          D3 nu, nv;
          edge_vectors(c, nu, nv);
          for (int corner = 0; corner < 4; corner++) {  // r, r + e1, r + e2, r + e1 + e2 (a superset for a triangle)
            const double wa = comp(c.r, ka) + (corner & 1 ? comp(c.e1, ka) : 0) + (corner & 2 ? comp(c.e2, ka) : 0);
            const double wb = comp(c.r, kb) + (corner & 1 ? comp(c.e1, kb) : 0) + (corner & 2 ? comp(c.e2, kb) : 0);
            lo_a = std::min(lo_a, wa); hi_a = std::max(hi_a, wa);
            lo_b = std::min(lo_b, wb); hi_b = std::max(hi_b, wb);
          }
          AxisEntry e;
          e.r_k = (float)comp(c.r, axis);
          e.r_a = (float)comp(c.r, ka);
          e.r_b = (float)comp(c.r, kb);
          e.nu_a = (float)comp(nu, ka);
          e.nu_b = (float)comp(nu, kb);
          e.nv_a = (float)comp(nv, ka);
          e.nv_b = (float)comp(nv, kb);
          e.limit = is_tri ? 1.0f : 2.0f;
          out.axis_entries.push_back(e);
          out.axis_tri.push_back(static_cast<uint8_t>(tri));
          if (pass < 3) out.n_axis_entries[axis]++;
          tri += is_tri ? 1 : 2;
        }
      }
      if (pass == 3) {
        // the rectangle is widened by far more than the rounding of the fast plane point
        const double slack = 1e-4 * std::max({hi_a - lo_a, hi_b - lo_b, 1e-3});
        cluster.count = static_cast<int32_t>(out.axis_entries.size()) - cluster.first;
        cluster.a_min = (float)(lo_a - slack); cluster.a_max = (float)(hi_a + slack);
        cluster.b_min = (float)(lo_b - slack); cluster.b_max = (float)(hi_b + slack);
        out.clusters.push_back(cluster);
      }
      tri_ids.insert(tri_ids.end(), B.quad_ids.begin(), B.quad_ids.end());
      tri_ids.insert(tri_ids.end(), B.lone_ids.begin(), B.lone_ids.end());
    }
  }
  for (size_t g = 0; g < found.size(); g++) {
    const Batch& B = batches[g];
    if (B.axis >= 0) continue;
    const Group& G = found[g];
    PlaneGroup group;
    const int32_t firsts = static_cast<int32_t>(out.entries.size()) | (static_cast<int32_t>(tri_ids.size()) << 16);
    const int32_t counts = static_cast<int32_t>(B.quads.size()) | (static_cast<int32_t>(B.tris.size()) << 16);
    group.n_gx = make_float4((float)G.n.x, (float)G.n.y, (float)G.n.z, (float)G.anchor.x);
    group.g_range = make_float4((float)G.anchor.y, (float)G.anchor.z, __builtin_bit_cast(float, firsts),
                                __builtin_bit_cast(float, counts));
    out.groups.push_back(group);
    for (const std::vector<Corner>* list : {&B.quads, &B.tris}) {
      for (const Corner& c : *list) {
        D3 nu, nv;
        edge_vectors(c, nu, nv);
        const D3 rel = sub(c.r, G.anchor);
        PlaneEntry e;
        e.nu_cu = make_float4((float)nu.x, (float)nu.y, (float)nu.z, (float)-dot(nu, rel));
        e.nv_cv = make_float4((float)nv.x, (float)nv.y, (float)nv.z, (float)-dot(nv, rel));
        out.entries.push_back(e);
      }
    }
    tri_ids.insert(tri_ids.end(), B.quad_ids.begin(), B.quad_ids.end());
    tri_ids.insert(tri_ids.end(), B.lone_ids.begin(), B.lone_ids.end());
  }
}

// BasicThin::BasicThin, src/amber/scene/lens_thin.cc:32-57
int make_lens(amber_cuda_ctx* ctx, const amber_cuda_lens& in, LensDev& lens) {
  if (in.kind != AMBER_CUDA_LENS_THIN && in.kind != AMBER_CUDA_LENS_PINHOLE)
    return fail(ctx, AMBER_CUDA_EINVAL, "unknown lens kind %d", in.kind);
  if (in.kind == AMBER_CUDA_LENS_PINHOLE && !(in.sensor_distance > 0))
    return fail(ctx, AMBER_CUDA_EINVAL, "pinhole lens: sensor_distance must be positive");
  if (in.kind == AMBER_CUDA_LENS_THIN && (in.n_blades < 1 || in.n_blades > kMaxBlades))
    return fail(ctx, AMBER_CUDA_EINVAL, "lens.n_blades must be in [1, %d]", kMaxBlades);
  lens.kind = in.kind;
  const float* t = in.transform;
  lens.origin[0] = t[3]; lens.origin[1] = t[7]; lens.origin[2] = t[11];
  const float e11 = t[0], e12 = t[1], e13 = t[2], e21 = t[4], e22 = t[5], e23 = t[6], e31 = t[8],
              e32 = t[9], e33 = t[10];
  const float g[9] = {e11, e12, e13, e21, e22, e23, e31, e32, e33};
  std::memcpy(lens.global, g, sizeof g);
  // Matrix3::Inverse, include/amber/prelude/matrix3.h:110-145
  const float det = e11 * (e22 * e33 - e23 * e32) - e21 * (e12 * e33 - e13 * e32) + e31 * (e12 * e23 - e13 * e22);
  if (det == 0.0f) return fail(ctx, AMBER_CUDA_EINVAL, "lens transform is singular");
  const float dinv = 1 / det;
  const float inv[9] = {+dinv * (e22 * e33 - e23 * e32), -dinv * (e12 * e33 - e13 * e32), +dinv * (e12 * e23 - e13 * e22),
                        -dinv * (e21 * e33 - e23 * e31), +dinv * (e11 * e33 - e13 * e31), -dinv * (e11 * e23 - e13 * e21),
                        +dinv * (e21 * e32 - e22 * e31), -dinv * (e11 * e32 - e12 * e31), +dinv * (e11 * e22 - e12 * e21)};
  std::memcpy(lens.local, inv, sizeof inv);
  if (in.kind == AMBER_CUDA_LENS_PINHOLE) {  // BasicPinhole::BasicPinhole, lens_pinhole.cc:31-42
    lens.sensor_distance = in.sensor_distance;
    lens.n_blades = 0;
    lens.p_area = 1.0f;  // kDiracDelta
    return AMBER_CUDA_OK;
  }
  lens.focus_distance = in.focus_distance;
  lens.sensor_distance = 1 / (1 / in.focal_length - 1 / in.focus_distance);
  lens.focus_over_sensor = -lens.focus_distance / lens.sensor_distance;
  lens.n_blades = in.n_blades;
  const float pi = kPiF;
  const float nb = static_cast<float>(in.n_blades);
  auto rim = [&](int i, float* out) {
    const float angle = 2 * pi / nb * static_cast<float>(i);
    const F3 p{in.radius * std::cos(angle), in.radius * std::sin(angle), in.radius * 0.0f};
    out[0] = lens.origin[0] + (g[0] * p.x + g[1] * p.y + g[2] * p.z);
    out[1] = lens.origin[1] + (g[3] * p.x + g[4] * p.y + g[5] * p.z);
    out[2] = lens.origin[2] + (g[6] * p.x + g[7] * p.y + g[8] * p.z);
  };
  for (int i = 0; i < in.n_blades; i++) {
    float* b = lens.blades[i];
    rim(i, b);
    rim(i + 1, b + 3);
    b[6] = lens.origin[0]; b[7] = lens.origin[1]; b[8] = lens.origin[2];
    const F3 v0{b[0], b[1], b[2]}, v1{b[3], b[4], b[5]}, v2{b[6], b[7], b[8]};
    const F3 n = normalizef(crossf(sub(v1, v0), sub(v2, v0)));
    b[9] = n.x; b[10] = n.y; b[11] = n.z;
  }
  const float* b0 = lens.blades[0];
  const F3 v0{b0[0], b0[1], b0[2]}, v1{b0[3], b0[4], b0[5]}, v2{b0[6], b0[7], b0[8]};
  const float area = lengthf(crossf(sub(v1, v0), sub(v2, v0))) / 2;  // Triangle::SurfaceArea
  lens.p_area = 1.0f / (area * nb);
  return AMBER_CUDA_OK;
}

int bind_constants(amber_cuda_ctx* ctx) {
  if (ctx->device < 64 && g_const_owner[ctx->device].load() == ctx->id) return AMBER_CUDA_OK;
  CUDA_TRY(ctx, cudaMemcpyToSymbolAsync(c_scene, &ctx->const_scene, sizeof(ConstScene), 0,
                                        cudaMemcpyHostToDevice, ctx->stream));
  CUDA_TRY(ctx, cudaMemcpyToSymbolAsync(c_camera, &ctx->camera, sizeof(CameraDev), 0,
                                        cudaMemcpyHostToDevice, ctx->stream));
  if (ctx->device < 64) g_const_owner[ctx->device].store(ctx->id);
  return AMBER_CUDA_OK;
}

void invalidate_constants(amber_cuda_ctx* ctx) {
  if (ctx->device < 64 && g_const_owner[ctx->device].load() == ctx->id) g_const_owner[ctx->device].store(0);
}

int ensure_pool(amber_cuda_ctx* ctx, uint32_t slots) {
  if (ctx->pool.n_slots == slots && ctx->d_ray.ptr) return AMBER_CUDA_OK;
  CUDA_TRY(ctx, ctx->d_ray.resize(2 * static_cast<size_t>(slots)));
  CUDA_TRY(ctx, ctx->d_state.resize(2 * static_cast<size_t>(slots)));
  CUDA_TRY(ctx, ctx->d_hit.resize(slots));
  CUDA_TRY(ctx, ctx->d_q_new.resize(slots));
  CUDA_TRY(ctx, ctx->d_q_shadow.resize(slots));
  CUDA_TRY(ctx, ctx->d_shadow.resize(2 * static_cast<size_t>(slots)));
  ctx->pool.ray = ctx->d_ray.ptr;
  ctx->pool.state = ctx->d_state.ptr;
  ctx->pool.hit = ctx->d_hit.ptr;
  ctx->pool.q_new = ctx->d_q_new.ptr;
  ctx->pool.shadow = ctx->d_shadow.ptr;
  ctx->pool.q_shadow = ctx->d_q_shadow.ptr;
  ctx->pool.n_slots = slots;
  return AMBER_CUDA_OK;
}

// f(begin, end) over [0, n) on the host's threads (large scenes: the per-triangle loops of upload_scene)
template <class F>
void parallel_ranges(size_t n, F&& f) {
  const size_t threads = n < (1u << 16) ? 1 : std::max(1u, std::min(32u, std::thread::hardware_concurrency()));
  if (threads == 1) {
    f(static_cast<size_t>(0), n);
    return;
  }
  const size_t per = (n + threads - 1) / threads;
  std::vector<std::thread> pool;
  for (size_t t = 1; t < threads; t++) pool.emplace_back([&f, t, per, n]() { f(std::min(n, t * per), std::min(n, (t + 1) * per)); });
  f(static_cast<size_t>(0), std::min(n, per));
  for (auto& t : pool) t.join();
}

// Depth of the flattened 4-wide tree (root = depth 1); a node step pushes at most 3 children,
// so 3 * depth + 1 stack entries per lane can never overflow.
int wide_bvh_depth(const bvh_build::WideBvhHost& bvh) {
  if (bvh.root < 0 || bvh.root == bvh_build::kEmpty || bvh.n_nodes == 0) return 0;
  std::vector<std::pair<int32_t, int>> todo{{bvh.root, 1}};
  int depth = 0;
  while (!todo.empty()) {
    const auto [node, d] = todo.back();
    todo.pop_back();
    depth = std::max(depth, d);
    for (int c = 0; c < 4 && bvh_build::kWidth == 4; c++) {
      int32_t ref;
      std::memcpy(&ref, &bvh.nodes[static_cast<size_t>(node) * bvh_build::kNodeFloats + 12 + c], sizeof ref);
      if (ref >= 0 && ref != bvh_build::kEmpty) todo.push_back({ref, d + 1});
    }
  }
  return depth;
}

// Persistent BVH kernels: stack size from the tree depth, grid = SMs x resident blocks.
int configure_bvh_launch(amber_cuda_ctx* ctx, const bvh_build::WideBvhHost& bvh) {
  // 4-wide: a node step pushes up to three 4-byte child references; 8-wide: at most one 8-byte node group
  const int depth = bvh_build::kWidth == 8 ? bvh.depth : wide_bvh_depth(bvh);
  const size_t entry_bytes = bvh_build::kWidth == 8 ? sizeof(uint2) : sizeof(int32_t);
  const uint32_t entry_limit = bvh_build::kWidth == 8 ? 64u : static_cast<uint32_t>(kBvhStackMax);  // 8-wide: bdpt.cuh kLaneStackMax
  ctx->bvh_stack_entries = static_cast<uint32_t>(bvh_build::kWidth == 8 ? depth + 1 : 3 * depth + 1);
  if (ctx->bvh_stack_entries > entry_limit)
    return fail(ctx, AMBER_CUDA_EINVAL, "BVH depth %d needs %u stack entries per lane (limit %u)", depth,
                ctx->bvh_stack_entries, entry_limit);
  ctx->bvh_smem = static_cast<size_t>(ctx->bvh_stack_entries) * kBlock * entry_bytes;
  cudaDeviceProp prop;
  CUDA_TRY(ctx, cudaGetDeviceProperties(&prop, ctx->device));
  const void* kernels[] = {reinterpret_cast<const void*>(&extend_kernel<WideBvh>),
                           reinterpret_cast<const void*>(&connect_kernel<WideBvh>),
                           reinterpret_cast<const void*>(&cast_bvh_kernel),
                           reinterpret_cast<const void*>(&extend_bvh_counting_kernel),
                           reinterpret_cast<const void*>(&connect_bvh_counting_kernel),
                           reinterpret_cast<const void*>(&bdpt_trace_kernel),
                           reinterpret_cast<const void*>(&bdpt_trace_evaluate_kernel)};
  int resident = 1 << 30;
  for (const void* k : kernels) {
    // the limit is per function and process-wide: always the maximum, so that a later upload of a shallower
    // tree in another context cannot lower it under a launch of this one
    CUDA_TRY(ctx, cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       static_cast<int>(kBvhStackMax * kBlock * sizeof(int32_t))));
    int blocks = 0;
    CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks, k, kBlock, ctx->bvh_smem));
    resident = std::min(resident, blocks);
  }
  if (resident < 1) return fail(ctx, AMBER_CUDA_ECUDA, "BVH kernels do not fit on an SM (%zu B of shared memory)", ctx->bvh_smem);
#ifdef AMBER_BVH_CARVEOUT
  // what the stacks do not need of the SM's 228 KB is L1 for nodes and triangles: ask for exactly that split
  const int carveout = static_cast<int>(std::min<size_t>(100, (ctx->bvh_smem + 1024) * static_cast<size_t>(resident) * 100 / (228 * 1024) + 1));
  for (const void* k : kernels) CUDA_TRY(ctx, cudaFuncSetAttribute(k, cudaFuncAttributePreferredSharedMemoryCarveout, carveout));
#endif
  ctx->bvh_grid = static_cast<unsigned int>(prop.multiProcessorCount * resident);
  return AMBER_CUDA_OK;
}

}  // namespace amber_host
using namespace amber_host;

// =====================================================================================
//                                    C ABI
// =====================================================================================
extern "C" {

int amber_cuda_abi_version(void) { return AMBER_CUDA_ABI_VERSION; }

const char* amber_cuda_last_error(const amber_cuda_ctx* ctx) {
  return ctx ? ctx->error.c_str() : g_create_error.c_str();
}

int amber_cuda_create(amber_cuda_ctx** out, int device) {
  if (!out) return fail(nullptr, AMBER_CUDA_EINVAL, "out is NULL");
  *out = nullptr;
  int n_devices = 0;
  if (cudaGetDeviceCount(&n_devices) != cudaSuccess || n_devices == 0) {
    cudaGetLastError();
    return fail(nullptr, AMBER_CUDA_ENODEV, "no CUDA device (libamber_cuda has no CPU fallback)");
  }
  if (device < 0 || device >= n_devices)
    return fail(nullptr, AMBER_CUDA_EINVAL, "device %d out of range [0, %d)", device, n_devices);
  amber_cuda_ctx* ctx = new amber_cuda_ctx();
  ctx->id = g_next_ctx_id.fetch_add(1);
  ctx->device = device;
  cudaError_t err = cudaSetDevice(device);
  if (err == cudaSuccess) err = cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking);
  if (err == cudaSuccess) err = cudaEventCreate(&ctx->ev_begin);
  if (err == cudaSuccess) err = cudaEventCreate(&ctx->ev_end);
  if (err == cudaSuccess) err = cudaMallocHost(&ctx->h_ctl, 2 * sizeof(Control));  // two poll slots
  if (err == cudaSuccess) err = cudaMallocHost(&ctx->h_tail, sizeof(float4));
  if (err == cudaSuccess) err = ctx->d_ctl.resize(1);
  if (err == cudaSuccess) err = cudaMemsetAsync(ctx->d_ctl.ptr, 0, sizeof(Control), ctx->stream);
  if (err != cudaSuccess) {
    fail(nullptr, AMBER_CUDA_ECUDA, "context creation failed: %s", cudaGetErrorString(err));
    if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
    if (ctx->h_tail) cudaFreeHost(ctx->h_tail);
    delete ctx;
    return AMBER_CUDA_ECUDA;
  }
  ctx->worker = std::thread(worker_main, ctx);
  *out = ctx;
  return AMBER_CUDA_OK;
}

int amber_cuda_destroy(amber_cuda_ctx* ctx) {
  if (!ctx) return AMBER_CUDA_OK;
  wait_idle(ctx);
  {
    std::lock_guard<std::mutex> guard(ctx->queue_mutex);
    ctx->quit = true;
  }
  ctx->queue_cv.notify_all();
  if (ctx->worker.joinable()) ctx->worker.join();
  cudaSetDevice(ctx->device);
  cudaStreamSynchronize(ctx->stream);
  invalidate_constants(ctx);
  ctx->d_tri_shade.release(); ctx->d_tri_isect.release(); ctx->d_spheres.release();
  ctx->d_tri_material.release(); ctx->d_sphere_material.release(); ctx->d_internal_to_prim.release();
  ctx->d_materials.release(); ctx->d_lights.release(); ctx->d_bvh_nodes.release();
  ctx->d_film.release(); ctx->d_resolved.release(); ctx->d_resolved8.release();
  ctx->d_ray.release(); ctx->d_state.release(); ctx->d_hit.release(); ctx->d_shadow.release();
  ctx->d_bdpt_state.release(); ctx->d_bdpt_events.release(); ctx->d_bdpt_queue.release(); ctx->d_bdpt_evaluate.release();
  ctx->d_bdpt_evaluate_ray.release(); ctx->d_bdpt_evaluate_hit.release();
  ctx->d_q_new.release(); ctx->d_q_shadow.release(); ctx->d_ctl.release(); ctx->d_prim_class.release();
  for (auto& b : ctx->d_scratch_f) b.release();
  for (auto& b : ctx->d_scratch_i) b.release();
  for (auto& b : ctx->d_scratch_4) b.release();
  if (ctx->h_ctl) cudaFreeHost(ctx->h_ctl);
  if (ctx->h_tail) cudaFreeHost(ctx->h_tail);
  if (ctx->ev_begin) cudaEventDestroy(ctx->ev_begin);
  if (ctx->ev_end) cudaEventDestroy(ctx->ev_end);
  if (ctx->stream) cudaStreamDestroy(ctx->stream);
  delete ctx;
  return AMBER_CUDA_OK;
}

int amber_cuda_upload_scene(amber_cuda_ctx* ctx, const amber_cuda_scene_desc* scene) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!scene || (scene->n_primitives && !scene->primitives) || !scene->materials || scene->n_materials == 0)
    return fail(ctx, AMBER_CUDA_EINVAL, "scene description is incomplete");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const auto build_t0 = std::chrono::steady_clock::now();

  // ---- materials ---------------------------------------------------------------
  std::vector<MaterialDev> materials(scene->n_materials);
  for (uint32_t i = 0; i < scene->n_materials; i++) {
    const amber_cuda_material& in = scene->materials[i];
    if (in.kind < 0 || in.kind > AMBER_CUDA_MAT_EYE)
      return fail(ctx, AMBER_CUDA_EINVAL, "material %u: unknown kind %d", i, in.kind);
    MaterialDev m{};
    m.kind = in.kind;
    m.r = in.rgb[0]; m.g = in.rgb[1]; m.b = in.rgb[2];
    m.param = in.param;
    if (in.kind == AMBER_CUDA_MAT_REFRACTION || in.kind == AMBER_CUDA_MAT_EYE) m.r = m.g = m.b = 1.0f;
    if (in.kind == AMBER_CUDA_MAT_REFRACTION) {
      // BasicRefraction::Fresnel, material_refraction.cc:265-269: std::pow(float, int) is double
      const float q = (in.param - 1) / (in.param + 1);
      m.r0 = static_cast<float>(std::pow(static_cast<double>(q), 2));
    }
    m.nee = (in.kind == AMBER_CUDA_MAT_LAMBERTIAN || in.kind == AMBER_CUDA_MAT_PHONG) ? 1 : 0;
    materials[i] = m;
  }

  // ---- primitives: triangles first, then spheres (internal order) ---------------
  std::vector<uint32_t> tri_ids, sphere_ids;
  for (uint32_t i = 0; i < scene->n_primitives; i++) {
    const amber_cuda_primitive& p = scene->primitives[i];
    if (p.material < 0 || static_cast<uint32_t>(p.material) >= scene->n_materials)
      return fail(ctx, AMBER_CUDA_EINVAL, "primitive %u: material %d out of range", i, p.material);
    if (p.kind == AMBER_CUDA_PRIM_TRIANGLE) tri_ids.push_back(i);
    else if (p.kind == AMBER_CUDA_PRIM_SPHERE) sphere_ids.push_back(i);
    else return fail(ctx, AMBER_CUDA_EINVAL, "primitive %u: unknown kind %d", i, p.kind);
  }
  if (sphere_ids.size() > static_cast<size_t>(kMaxConstSpheres))
    return fail(ctx, AMBER_CUDA_EINVAL, "at most %d spheres are supported", kMaxConstSpheres);
  if (tri_ids.size() >= (1u << 28)) return fail(ctx, AMBER_CUDA_EINVAL, "too many triangles");
  const bool use_bvh = tri_ids.size() > static_cast<size_t>(kMaxConstTris);

  bvh_build::WideBvhHost bvh;
  if (use_bvh) {
    std::vector<float> verts(tri_ids.size() * 9);
    parallel_ranges(tri_ids.size(), [&](size_t begin, size_t end) {
      for (size_t k = begin; k < end; k++) std::memcpy(&verts[9 * k], scene->primitives[tri_ids[k]].data, 9 * sizeof(float));
    });
    // AMBER_CUDA_BVH_BUILDER=lbvh: the tree is built on the device (bvh_build_gpu.cuh): tens of milliseconds instead
    // of half a second, for a tree that traces slower -- worth it when the render is shorter than the host build.
    const char* builder = std::getenv("AMBER_CUDA_BVH_BUILDER");
    const auto bvh_t0 = std::chrono::steady_clock::now();
    bool built = false;
    if (bvh_build::kWidth == 4 && builder && std::strcmp(builder, "lbvh") == 0 && tri_ids.size() > 64) {
      built = bvh_gpu::build(verts.data(), static_cast<uint32_t>(tri_ids.size()), ctx->stream, bvh) == cudaSuccess;
      if (!built) {
        cudaGetLastError();
        bvh = bvh_build::WideBvhHost();
      }
    }
    if (!built) bvh = bvh_build::build(verts.data(), static_cast<uint32_t>(tri_ids.size()));
    trace(ctx, "upload_scene: %zu triangles, BVH of %u nodes by the %s builder in %.1f ms", tri_ids.size(), bvh.n_nodes,
          built ? "device (lbvh)" : "host (binned SAH)",
          std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - bvh_t0).count());
    std::vector<uint32_t> reordered(tri_ids.size());
    for (size_t k = 0; k < tri_ids.size(); k++) reordered[k] = tri_ids[bvh.order[k]];
    tri_ids.swap(reordered);
  }

  // constant-memory list: batch the triangles by supporting plane (scene_device.cuh, PlaneGroup)
  ConstListHost const_list;
  if (!use_bvh) group_by_plane(scene->primitives, tri_ids, const_list);

  const size_t n_tris = tri_ids.size(), n_spheres = sphere_ids.size();
  std::vector<TriIsect> tri_isect(n_tris);
  std::vector<TriShade> tri_shade(n_tris);
  std::vector<int32_t> tri_material(n_tris), sphere_material(n_spheres);
  std::vector<float4> spheres(n_spheres);
  ctx->internal_to_prim.assign(n_tris + n_spheres, 0);
  parallel_ranges(n_tris, [&](size_t begin, size_t end) {
    for (size_t k = begin; k < end; k++) {
      const amber_cuda_primitive& p = scene->primitives[tri_ids[k]];
      make_triangle(p.data, tri_isect[k], tri_shade[k]);
      tri_material[k] = p.material;
      ctx->internal_to_prim[k] = static_cast<int32_t>(tri_ids[k]);
    }
  });
  for (size_t k = 0; k < n_spheres; k++) {
    const amber_cuda_primitive& p = scene->primitives[sphere_ids[k]];
    spheres[k] = make_float4(p.data[0], p.data[1], p.data[2], p.data[3]);
    sphere_material[k] = p.material;
    ctx->internal_to_prim[n_tris + k] = static_cast<int32_t>(sphere_ids[k]);
  }

  // hit class of every primitive (internal order): what `shade` sorts a tile by
  std::vector<uint8_t> prim_class(n_tris + n_spheres);
  parallel_ranges(n_tris, [&](size_t begin, size_t end) {
    for (size_t k = begin; k < end; k++) prim_class[k] = static_cast<uint8_t>(materials[tri_material[k]].kind * 2);
  });
  for (size_t k = 0; k < n_spheres; k++)
    prim_class[n_tris + k] = static_cast<uint8_t>(materials[sphere_material[k]].kind * 2 + 1);

  // ---- lights: emitting objects, LightSet::LightSet (light_set.h:61-81) ---------
  // internal index: triangles [0, n_tris), spheres n_tris + k
  struct HostLight {
    size_t prim;
    float power, irradiance_sum;
  };
  std::vector<HostLight> host_lights;
  auto original_id = [&](size_t internal) {
    return internal < n_tris ? tri_ids[internal] : sphere_ids[internal - n_tris];
  };
  for (size_t k = 0; k < n_tris + n_spheres; k++) {
    const MaterialDev& m = materials[k < n_tris ? tri_material[k] : sphere_material[k - n_tris]];
    if (m.kind != AMBER_CUDA_MAT_DIFFUSE_LIGHT) continue;
    float area;
    if (k < n_tris) {
      const TriShade& t = tri_shade[k];
      const F3 e1{t.e1_ny.x, t.e1_ny.y, t.e1_ny.z}, e2{t.e2_nz.x, t.e2_nz.y, t.e2_nz.z};
      area = lengthf(crossf(e1, e2)) / 2;  // Triangle::SurfaceArea, primitive_triangle.cc:130-134
    } else {
      const float radius = spheres[k - n_tris].w;
      area = 4 * kPiF * radius * radius;  // Sphere::SurfaceArea, primitive_sphere.cc:109-113
    }
    // DiffuseLight::Irradiance = radiance * pi; Power = Sum(area * irradiance), object.h:98-103
    const float ir = m.r * kPiF, ig = m.g * kPiF, ib = m.b * kPiF;
    const float power = area * ir + area * ig + area * ib;
    if (power > 0) host_lights.push_back({k, power, ir + ig + ib});
  }
  // ascending power; equal powers in creation order (the BVH permutes `tri_ids`)
  std::stable_sort(host_lights.begin(), host_lights.end(), [&](const HostLight& a, const HostLight& b) {
    return a.power < b.power || (a.power == b.power && original_id(a.prim) < original_id(b.prim));
  });
  std::vector<LightDev> lights(host_lights.size());
  float total_power = 0;
  for (size_t i = 0; i < host_lights.size(); i++) total_power += host_lights[i].power;
  float cumulative = 0;
  for (size_t i = 0; i < host_lights.size(); i++) {
    const size_t k = host_lights[i].prim;
    const int32_t material_index = k < n_tris ? tri_material[k] : sphere_material[k - n_tris];
    const MaterialDev& m = materials[material_index];
    cumulative += host_lights[i].power;
    const float pdf_area = host_lights[i].irradiance_sum / total_power;  // LightSet::PDFArea, light_set.h:106-111
    LightDev& l = lights[i];
    std::memset(&l, 0, sizeof l);
    if (k < n_tris) {
      const TriShade& t = tri_shade[k];
      const amber_cuda_primitive& p = scene->primitives[tri_ids[k]];
      l.v0 = make_float4(t.v0_nx.x, t.v0_nx.y, t.v0_nx.z, cumulative);
      l.v1 = make_float4(p.data[3], p.data[4], p.data[5], 0.0f);
      l.v2 = make_float4(p.data[6], p.data[7], p.data[8], 0.0f);
      l.n = make_float4(t.v0_nx.w, t.e1_ny.w, t.e2_nz.w, 0.0f);
    } else {
      const float4 sphere = spheres[k - n_tris];
      l.v0 = make_float4(sphere.x, sphere.y, sphere.z, cumulative);
      l.shape.x = sphere.w;
    }
    // weight of a light ray, LightSet::GenerateRay: Irradiance / PDFArea (light_set.h:101)
    l.v1.w = m.r * kPiF / pdf_area;
    l.v2.w = m.g * kPiF / pdf_area;
    l.n.w = m.b * kPiF / pdf_area;
    l.le_over_pdf = make_float4(m.r / pdf_area, m.g / pdf_area, m.b / pdf_area,
                                __builtin_bit_cast(float, material_index));
  }

  // ---- lens --------------------------------------------------------------------
  LensDev lens{};
  const int lens_rc = make_lens(ctx, scene->lens, lens);
  if (lens_rc != AMBER_CUDA_OK) return lens_rc;

  // ---- commit ------------------------------------------------------------------
  // from here on the context's scene is in flux: it counts as "no scene" until the last fallible call
  // has succeeded, so a failed upload can never leave new constants next to old device pointers
  const auto upload_t0 = std::chrono::steady_clock::now();
  ctx->has_scene = false;
  ctx->view = SceneView{};
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  invalidate_constants(ctx);
  ctx->use_bvh = use_bvh;
  std::memset(&ctx->const_scene, 0, sizeof(ConstScene));
  ctx->const_scene.n_tris = use_bvh ? 0 : static_cast<int32_t>(n_tris);
  ctx->const_scene.n_spheres = static_cast<int32_t>(n_spheres);
  ctx->const_scene.bvh_plane_magic = bvh_build::kPlaneMagic;
  ctx->const_scene.n_groups = static_cast<int32_t>(const_list.groups.size());
  for (int axis = 0; axis < 3; axis++) ctx->const_scene.n_axis_entries[axis] = const_list.n_axis_entries[axis];
  ctx->const_scene.n_clusters = static_cast<int32_t>(const_list.clusters.size());
  for (size_t c = 0; c < const_list.clusters.size(); c++) ctx->const_scene.cluster[c] = const_list.clusters[c];
  for (size_t k = 0; k < const_list.axis_entries.size(); k++) {
    ctx->const_scene.axis_entry[k] = const_list.axis_entries[k];
    ctx->const_scene.axis_tri[k] = const_list.axis_tri[k];
  }
  {  // pair the unclustered entries of each axis for the packed scan (scene_device.cuh, AxisPair)
    size_t entry = 0, pair = 0;
    for (int axis = 0; axis < 3; axis++) {
      const size_t n = static_cast<size_t>(const_list.n_axis_entries[axis]);
      for (size_t k = 0; k < n; k += 2, pair++) {
        const AxisEntry& a = const_list.axis_entries[entry + k];
        AxisEntry b{};
        b.limit = -1.0f;
        if (k + 1 < n) b = const_list.axis_entries[entry + k + 1];
        AxisPair& P = ctx->const_scene.axis_pair[pair];
        P.r_k = make_float2(a.r_k, b.r_k);
        P.neg_r_a = make_float2(-a.r_a, -b.r_a);
        P.neg_r_b = make_float2(-a.r_b, -b.r_b);
        P.nu_a = make_float2(a.nu_a, b.nu_a);
        P.nu_b = make_float2(a.nu_b, b.nu_b);
        P.nv_a = make_float2(a.nv_a, b.nv_a);
        P.nv_b = make_float2(a.nv_b, b.nv_b);
        P.limit = make_float2(a.limit, b.limit);
      }
      entry += n;
    }
  }
  for (size_t g = 0; g < const_list.groups.size(); g++) ctx->const_scene.group[g] = const_list.groups[g];
  for (size_t k = 0; k < const_list.entries.size(); k++) ctx->const_scene.entry[k] = const_list.entries[k];
  for (size_t k = 0; k < n_spheres; k++) ctx->const_scene.sphere[k] = spheres[k];
  ctx->camera.lens = lens;

  CUDA_TRY(ctx, ctx->d_tri_shade.upload(tri_shade, ctx->stream));
  CUDA_TRY(ctx, ctx->d_tri_isect.upload(tri_isect, ctx->stream));
  CUDA_TRY(ctx, ctx->d_spheres.upload(spheres, ctx->stream));
  CUDA_TRY(ctx, ctx->d_tri_material.upload(tri_material, ctx->stream));
  CUDA_TRY(ctx, ctx->d_sphere_material.upload(sphere_material, ctx->stream));
  CUDA_TRY(ctx, ctx->d_internal_to_prim.upload(ctx->internal_to_prim, ctx->stream));
  CUDA_TRY(ctx, ctx->d_materials.upload(materials, ctx->stream));
  CUDA_TRY(ctx, ctx->d_lights.upload(lights, ctx->stream));
  CUDA_TRY(ctx, ctx->d_prim_class.upload(prim_class, ctx->stream));
  if (use_bvh) {
    std::vector<float4> nodes(bvh.nodes.size() / 4);
    std::memcpy(nodes.data(), bvh.nodes.data(), bvh.nodes.size() * sizeof(float));
    CUDA_TRY(ctx, ctx->d_bvh_nodes.upload(nodes, ctx->stream));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));

  SceneView& view = ctx->view;
  view.tri_shade = ctx->d_tri_shade.ptr;
  view.spheres = ctx->d_spheres.ptr;
  view.tri_material = ctx->d_tri_material.ptr;
  view.sphere_material = ctx->d_sphere_material.ptr;
  view.materials = ctx->d_materials.ptr;
  view.lights = ctx->d_lights.ptr;
  view.prim_class = ctx->d_prim_class.ptr;
  view.n_tris = static_cast<int32_t>(n_tris);
  view.n_spheres = static_cast<int32_t>(n_spheres);
  view.n_lights = static_cast<int32_t>(lights.size());
  view.total_power = total_power;
  view.bvh_nodes = ctx->d_bvh_nodes.ptr;
  view.tri_isect = ctx->d_tri_isect.ptr;
  view.bvh_root = use_bvh ? bvh.root : kEmptyChild;
  ctx->bvh_nodes = use_bvh ? bvh.n_nodes : 0;
  if (use_bvh) {
    const int rc = configure_bvh_launch(ctx, bvh);
    if (rc != AMBER_CUDA_OK) return rc;
  }
  const auto upload_t1 = std::chrono::steady_clock::now();
  ctx->stats.scene_build_ms = std::chrono::duration<double, std::milli>(upload_t0 - build_t0).count();
  ctx->stats.scene_upload_ms = std::chrono::duration<double, std::milli>(upload_t1 - upload_t0).count();
  ctx->has_scene = true;
  return AMBER_CUDA_OK;
}

int amber_cuda_set_sensor(amber_cuda_ctx* ctx, uint32_t width, uint32_t height, float film_width,
                          float film_height) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (width == 0 || height == 0 || !(film_width > 0) || !(film_height > 0))
    return fail(ctx, AMBER_CUDA_EINVAL, "bad sensor %ux%u, film %g x %g", width, height, film_width, film_height);
  if (static_cast<uint64_t>(width) * height >= (1ull << 31))
    return fail(ctx, AMBER_CUDA_EINVAL, "sensor too large");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  invalidate_constants(ctx);
  SensorDev& s = ctx->camera.sensor;
  s.width = width;
  s.height = height;
  s.film_width = film_width;
  s.film_height = film_height;
  // sensor.Size() / sensor.SceneArea(): pixel count (integer) over a float product
  s.size_over_area = static_cast<float>(static_cast<uint64_t>(width) * height) / (film_width * film_height);
  s.inv_area = 1 / (film_width * film_height);
  ctx->has_sensor = true;
  CUDA_TRY(ctx, ctx->d_film.resize(static_cast<size_t>(width) * height + 1));  // + the sample-count element
  return amber_cuda_film_clear(ctx);
}

int amber_cuda_film_clear(amber_cuda_ctx* ctx) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_sensor) return fail(ctx, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t n = static_cast<size_t>(ctx->camera.sensor.width) * ctx->camera.sensor.height + 1;
  ctx->samples_in_film = false;
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_film.ptr, 0, n * sizeof(float4), ctx->stream));
  CUDA_TRY(ctx, cudaMemsetAsync(ctx->d_ctl.ptr, 0, sizeof(Control), ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctx->samples_done = 0;
  return AMBER_CUDA_OK;
}

}  // extern "C"

namespace amber_host {

__global__ void control_begin_kernel(Control* ctl, unsigned long long total_paths, unsigned int slots,
                                     unsigned int n_pixels) {
  ctl->next_path = 0;
  ctl->base_pixel = 0;
  ctl->base_sample = 0;
  ctl->n_pixels = n_pixels;
  ctl->total_paths = total_paths;
  ctl->n_new = slots;
  ctl->n_marked = slots;  // pool_reset_kernel empties every slot
  ctl->n_shadow = 0;
  ctl->n_alive = 0;
  ctl->done = 0;
  ctl->raygen_blocks_done = 0;
  for (int k = 0; k < 4; k++) ctl->bvh_work[k] = 0;
}

// A queued render call that continues the running one is absorbed into its wavefront: more paths to
// start, nothing else changes.  `done` is cleared, so a poll copy taken after this kernel is current.
__global__ void control_extend_kernel(Control* ctl, unsigned long long more_paths) {
  ctl->total_paths += more_paths;
  ctl->done = 0;
}

constexpr int kStepsPerPoll = 8;
// sixteen waves of shade's 512-slot tiles at 8 resident blocks per SM on 148 SMs
constexpr uint32_t kDefaultPoolPaths = 148u * 4u * 1024u * 8u;

struct KernelTimer {
  // profiling mode: one event pair per launch, summed per kernel class after the render
  std::vector<cudaEvent_t> events;
  std::vector<int> kinds;
  size_t used = 0;
  cudaEvent_t next() {
    if (used == events.size()) {
      cudaEvent_t e;
      cudaEventCreate(&e);
      events.push_back(e);
    }
    return events[used++];
  }
  ~KernelTimer() {
    for (cudaEvent_t e : events) cudaEventDestroy(e);
  }
};

void launch_extend_counting(const RenderArgs& args, unsigned int grid, size_t smem, cudaStream_t stream) {
  extend_bvh_counting_kernel<<<grid, kBlock, smem, stream>>>(args);
}
void launch_connect_counting(const RenderArgs& args, unsigned int grid, size_t smem, cudaStream_t stream) {
  connect_bvh_counting_kernel<<<grid, kBlock, smem, stream>>>(args);
}

uint32_t effective_stride(const amber_cuda_render_params& p) { return p.sample_stride ? p.sample_stride : 1u; }

// Does `next` continue `current` (the call whose wavefront is running, which ends at ordinal `ordinals_done`)?
bool continues(const amber_cuda_render_params& current, uint64_t ordinals_done, const amber_cuda_render_params& next) {
  return next.estimator == current.estimator && next.seed == current.seed && next.max_depth == current.max_depth &&
         next.pool_paths == current.pool_paths && next.pixel_begin == current.pixel_begin &&
         next.pixel_count == current.pixel_count && effective_stride(next) == effective_stride(current) &&
         next.sample_begin == current.sample_begin + ordinals_done * effective_stride(current) && next.sample_count > 0;
}

template <class Accel>
int render_loop(amber_cuda_ctx* ctx, const RenderArgs& args, const amber_cuda_render_params& first, bool nee,
                size_t dynamic_smem, uint64_t* ordinals_rendered) {
  const unsigned int grid = blocks_for(args.pool.n_slots);
  const unsigned int shade_grid = blocks_for(args.pool.n_slots, kShadeTile);
  const unsigned int trace_grid = Accel::kPersistent ? ctx->bvh_grid : grid;
  // raygen and the list connect are grid-stride loops over queues whose length only the device knows:
  // one wave of resident blocks (kQueueBlocks per SM), never more than the pool
  int sm_count = 148;
  cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device);
  const unsigned int queue_grid = std::min<unsigned int>(grid, static_cast<unsigned int>(sm_count) * static_cast<unsigned int>(kQueueBlocks));
  const unsigned int connect_grid = Accel::kPersistent ? ctx->bvh_grid : queue_grid;
  const bool light_tracing = args.estimator == AMBER_CUDA_ESTIMATOR_LIGHT_TRACING;
  cudaStream_t stream = ctx->stream;
  KernelTimer timer;
  const bool profile = ctx->profiling;
  double ms_by_kind[4] = {0, 0, 0, 0};  // raygen, extend, shade, connect
  uint64_t extend_launches = 0;
  auto timed = [&](int kind, auto&& launch) {
    if (profile) {
      cudaEventRecord(timer.next(), stream);
      launch();
      cudaEventRecord(timer.next(), stream);
      timer.kinds.push_back(kind);
    } else {
      launch();
    }
  };
  cudaEvent_t poll_event[2];
  Control* poll_copy = ctx->h_ctl;  // pinned, one slot per poll in flight
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&poll_event[0], cudaEventDisableTiming));
  CUDA_TRY(ctx, cudaEventCreateWithFlags(&poll_event[1], cudaEventDisableTiming));
  int rc = AMBER_CUDA_OK;
  bool done = false;
  uint64_t ordinals = first.sample_count;  // samples per pixel this wavefront has been asked for so far
  int polls_valid_from = 0;                // poll copies of earlier batches predate the last extension
  for (int batch = 0; !done; batch++) {
    // absorb queued calls that continue this one: the pool keeps running instead of draining
    for (;;) {
      uint64_t more = 0;
      {
        std::lock_guard<std::mutex> guard(ctx->queue_mutex);
        if (!ctx->jobs.empty() && continues(first, ordinals, ctx->jobs.front()) &&
            first.sample_begin + (ordinals + ctx->jobs.front().sample_count) * effective_stride(first) < (1ull << 32)) {
          more = ctx->jobs.front().sample_count;
          ctx->jobs.pop_front();
          ctx->running++;
          ctx->absorbed_ends.push_back(ordinals + more);
        }
      }
      if (more == 0) break;
      trace(ctx, "batch %d: absorbed a call of %llu samples (wavefront now %llu)", batch, (unsigned long long)more,
            (unsigned long long)(ordinals + more));
      control_extend_kernel<<<1, 1, 0, stream>>>(args.ctl, more * args.n_pixels);
      ctx->stats.kernel_launches++;
      ctx->stats.chained_renders++;
      ordinals += more;
      polls_valid_from = batch;
    }
    for (int step = 0; step < kStepsPerPoll; step++) {
      timed(0, [&] {
        const unsigned int raygen_grid = AMBER_RAYGEN_LOOP ? std::min<unsigned int>(grid, static_cast<unsigned int>(sm_count) * AMBER_RAYGEN_BLOCKS) : grid;
        if (light_tracing) raygen_kernel<true><<<raygen_grid, kBlock, 0, stream>>>(args);
        else raygen_kernel<false><<<raygen_grid, kBlock, 0, stream>>>(args);
      });
#if !AMBER_FUSE_BOOKKEEP
      bookkeep_kernel<<<1, 1, 0, stream>>>(args.ctl);
#endif
      timed(1, [&] {
        if (AMBER_FUSE_EXTEND && !Accel::kPersistent) return;  // experiment: raygen and shade found the hits
        if (profile && Accel::kPersistent) launch_extend_counting(args, trace_grid, dynamic_smem, stream);
        else extend_kernel<Accel><<<trace_grid, kBlock, dynamic_smem, stream>>>(args);
      });
      timed(2, [&] {
        if (light_tracing) shade_kernel<true><<<shade_grid, kShadeBlock, 0, stream>>>(args);
        else shade_kernel<false><<<shade_grid, kShadeBlock, 0, stream>>>(args);
      });
      if (nee) timed(3, [&] {
        if (profile && Accel::kPersistent) launch_connect_counting(args, connect_grid, dynamic_smem, stream);
        else connect_kernel<Accel><<<connect_grid, kBlock, dynamic_smem, stream>>>(args);
      });
      ctx->stats.kernel_launches += (nee ? 4 : 3) + (AMBER_FUSE_BOOKKEEP ? 0 : 1);
      extend_launches++;
    }
    // poll: the copy of batch b is read while batch b+1 is already queued
    if (batch > 0) {
      cudaError_t err = cudaEventSynchronize(poll_event[(batch - 1) & 1]);
      if (err == cudaSuccess) err = cudaGetLastError();  // launch failures surface here, once per poll
      if (err != cudaSuccess) {
        rc = fail(ctx, AMBER_CUDA_ECUDA, "render failed: %s", cudaGetErrorString(err));
        break;
      }
      const Control& seen = poll_copy[(batch - 1) & 1];
      if (seen.done && batch - 1 >= polls_valid_from) done = true;
      // calls whose paths have all been started no longer count as pending (amber_cuda_wait_pending):
      // the caller may queue the next one while this wavefront is still full
      bool released = false;
      {
        std::lock_guard<std::mutex> guard(ctx->queue_mutex);
        const uint64_t started = seen.next_path / args.n_pixels;
        while (!ctx->absorbed_ends.empty() && ctx->absorbed_ends.front() <= started) {
          ctx->absorbed_ends.pop_front();
          ctx->running--;
          released = true;
        }
      }
      if (released) ctx->idle_cv.notify_all();
      trace(ctx, "batch %d: poll of batch %d: next_path %llu of %llu, alive %u, done %u%s", batch, batch - 1,
            (unsigned long long)seen.next_path, (unsigned long long)seen.total_paths, seen.n_alive, seen.done,
            released ? ", released" : "");
    }
    if (!done) {
      cudaMemcpyAsync(&poll_copy[batch & 1], args.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream);
      cudaEventRecord(poll_event[batch & 1], stream);
    }
  }
  cudaEventDestroy(poll_event[0]);
  cudaEventDestroy(poll_event[1]);
  *ordinals_rendered = ordinals;
  if (rc != AMBER_CUDA_OK) return rc;
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  if (profile) {
    for (size_t k = 0; k < timer.kinds.size(); k++) {
      float ms = 0;
      cudaEventElapsedTime(&ms, timer.events[2 * k], timer.events[2 * k + 1]);
      ms_by_kind[timer.kinds[k]] += ms;
    }
    if (trace_on()) {  // AMBER_CUDA_TRACE + profiling: the last iterations kernel by kernel (the drain tail of a render)
      const size_t per_iteration = nee ? 4 : 3;
      const size_t iterations = timer.kinds.size() / per_iteration, from = iterations > 48 ? iterations - 48 : 0;
      for (size_t it = from; it < iterations; it++) {
        float us[4] = {0, 0, 0, 0};
        for (size_t k = it * per_iteration; k < (it + 1) * per_iteration; k++) {
          float ms = 0;
          cudaEventElapsedTime(&ms, timer.events[2 * k], timer.events[2 * k + 1]);
          us[timer.kinds[k]] = 1e3f * ms;
        }
        float whole = 0;
        if (it + 1 < iterations) cudaEventElapsedTime(&whole, timer.events[2 * it * per_iteration], timer.events[2 * (it + 1) * per_iteration]);
        trace(ctx, "iteration %zu: raygen %.1f extend %.1f shade %.1f connect %.1f us, start to next start %.1f us", it, us[0], us[1],
              us[2], us[3], 1e3f * whole);
      }
    }
  }
  ctx->stats.raygen_ms = ms_by_kind[0];
  ctx->stats.extend_ms = ms_by_kind[1];
  ctx->stats.shade_ms = ms_by_kind[2];
  ctx->stats.connect_ms = ms_by_kind[3];
  ctx->stats.extend_launches = extend_launches;
  return AMBER_CUDA_OK;
}

// Bidirectional path tracing as a wavefront (bdpt.cuh, second half): a pool of pixel samples in flight, one subpath step
// per item per iteration, connections from a queue.
// Pool: 2 M items = 5.4 GB of events + 200 MB of state (512 k / 1 M / 2 M / 4 M items: 601 / 682 / 726 / 739 M pixel samples/s
// on the Cornell box at 64 spp).  AMBER_CUDA_BDPT_POOL overrides it (experiments).
constexpr uint32_t kBdptPoolItems = 1u << 21;
// candidates that pass the probe, per pool item and iteration: ~0.25 on the Cornell box; beyond the capacity they are
// evaluated in place
constexpr uint32_t kBdptEvaluatePerItem = 2;
constexpr int AMBER_CUDA_ENOMEM_POOL = -1000;  // internal: the wavefront's pool could not be allocated
uint32_t bdpt_pool_items() {
  const char* text = std::getenv("AMBER_CUDA_BDPT_POOL");
  const long long wanted = text ? std::atoll(text) : 0;
  return wanted >= 1024 && wanted <= (1ll << 24) ? static_cast<uint32_t>(wanted) : kBdptPoolItems;
}
// Which implementation: the wavefront (825 against 470 M pixel samples/s on the Cornell box, 165 against 131 on the
// 1.3 M-triangle scene) unless AMBER_CUDA_BDPT=megakernel asks for round 1's one-kernel version (tests compare the two).
bool bdpt_wavefront_selected() {
  const char* choice = std::getenv("AMBER_CUDA_BDPT");
  return !(choice && std::strcmp(choice, "megakernel") == 0);
}
template <class Accel>
int render_bdpt_wavefront(amber_cuda_ctx* ctx, const amber_cuda_render_params* params, const RenderArgs& args,
                          uint64_t total_pairs, int cap) {
  const uint64_t wanted = std::min<uint64_t>(bdpt_pool_items(), total_pairs);
  const uint32_t n_items = static_cast<uint32_t>((wanted + kBdptGrowBlock - 1) / kBdptGrowBlock * kBdptGrowBlock);
  if (ctx->d_bdpt_state.resize(static_cast<size_t>(kBdptStateWords) * n_items) != cudaSuccess ||
      ctx->d_bdpt_events.resize(static_cast<size_t>(2 * kBdptMaxEvents) * n_items) != cudaSuccess ||
      ctx->d_bdpt_queue.resize(n_items) != cudaSuccess ||
      ctx->d_bdpt_evaluate.resize(static_cast<size_t>(kBdptEvaluatePerItem) * n_items) != cudaSuccess ||
      (Accel::kPersistent && (ctx->d_bdpt_evaluate_ray.resize(2 * static_cast<size_t>(kBdptEvaluatePerItem) * n_items) != cudaSuccess ||
                              ctx->d_bdpt_evaluate_hit.resize(static_cast<size_t>(kBdptEvaluatePerItem) * n_items) != cudaSuccess))) {
    // no room for the pool (2.7 KB per item): the caller runs the one-kernel version, which needs none
    cudaGetLastError();
    ctx->d_bdpt_state.release();
    ctx->d_bdpt_events.release();
    ctx->d_bdpt_queue.release();
    ctx->d_bdpt_evaluate.release();
    ctx->d_bdpt_evaluate_ray.release();
    ctx->d_bdpt_evaluate_hit.release();
    return AMBER_CUDA_ENOMEM_POOL;
  }
  BdptPool pool;
  // (the arrays may be larger than this render needs: the state words are n_items apart, whatever the allocation)
  pool.state = ctx->d_bdpt_state.ptr;
  pool.events = ctx->d_bdpt_events.ptr;
  pool.q_connect = ctx->d_bdpt_queue.ptr;
  pool.q_evaluate = ctx->d_bdpt_evaluate.ptr;
  pool.evaluate_ray = ctx->d_bdpt_evaluate_ray.ptr;
  pool.evaluate_hit = ctx->d_bdpt_evaluate_hit.ptr;
  pool.evaluate_capacity = kBdptEvaluatePerItem * n_items;
  pool.n_items = n_items;
  cudaStream_t stream = ctx->stream;
  int sm_count = 148, resident = 1;
  CUDA_TRY(ctx, cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device));
  CUDA_TRY(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, bdpt_connect_kernel<Accel>, kBdptConnectBlock, 0));
  const unsigned int connect_grid = static_cast<unsigned int>(sm_count * std::max(1, resident));
  const unsigned int grow_grid = n_items / kBdptGrowBlock;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_begin, stream));
  // every item starts in the "take a pair" phase: all-zero ids
  CUDA_TRY(ctx, cudaMemsetAsync(pool.state + 3 * static_cast<size_t>(n_items), 0, sizeof(float4) * n_items, stream));
  CUDA_TRY(ctx, cudaMemsetAsync(&ctx->d_ctl.ptr->cursor_bdpt, 0, sizeof(unsigned long long), stream));
  CUDA_TRY(ctx, cudaMemsetAsync(&ctx->d_ctl.ptr->bdpt_n_connect[0], 0, 6 * sizeof(unsigned int), stream));
  Control* seen = ctx->h_ctl;
  const uint64_t max_iterations = 8 * ((total_pairs + n_items - 1) / n_items + 1) * (2ull * kBdptMaxEvents + 2) + 64;
  for (uint64_t iteration = 0;;) {
    for (int step = 0; step < kStepsPerPoll; step++, iteration++) {
      const uint32_t parity = static_cast<uint32_t>(iteration & 1u);
      if constexpr (Accel::kPersistent) {
        bdpt_begin_kernel<<<grow_grid, kBdptGrowBlock, 0, stream>>>(args, pool, total_pairs);
        bdpt_trace_kernel<<<ctx->bvh_grid, kBlock, ctx->bvh_smem, stream>>>(args, pool, cap);
        bdpt_event_kernel<<<grow_grid, kBdptGrowBlock, 0, stream>>>(args, pool, cap, parity);
        ctx->stats.kernel_launches += 2;
      } else {
        bdpt_grow_kernel<Accel><<<grow_grid, kBdptGrowBlock, 0, stream>>>(args, pool, total_pairs, cap, parity);
      }
      bdpt_connect_kernel<Accel><<<connect_grid, kBdptConnectBlock, 0, stream>>>(args, pool, parity);
      if constexpr (Accel::kPersistent) {
        bdpt_trace_evaluate_kernel<<<ctx->bvh_grid, kBlock, ctx->bvh_smem, stream>>>(args, pool);
        ctx->stats.kernel_launches++;
      }
      bdpt_evaluate_kernel<Accel><<<connect_grid, kBdptConnectBlock, 0, stream>>>(args, pool);
      ctx->stats.kernel_launches += 3;
    }
    CUDA_TRY(ctx, cudaMemcpyAsync(seen, args.ctl, sizeof(Control), cudaMemcpyDeviceToHost, stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(stream));
    CUDA_TRY(ctx, cudaGetLastError());
    if (seen->bdpt_done >= n_items) break;
    // every iteration advances every unfinished item by one step, and a pair takes at most 2 * 16 of them: a render
    // that is still going far beyond that has a broken invariant -- fail instead of spinning
    if (iteration > max_iterations)
      return fail(ctx, AMBER_CUDA_ECUDA, "bidirectional wavefront did not finish within %llu iterations (%u of %u items done)",
                  static_cast<unsigned long long>(max_iterations), seen->bdpt_done, n_items);
  }
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_end, stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_end));
  float ms = 0;
  CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
  ctx->stats.last_render_ms = ms;
  ctx->samples_done += params->sample_count;
  return AMBER_CUDA_OK;
}

// Bidirectional path tracing (bdpt.cuh): one kernel, one thread per (pixel, sample).
int render_bdpt(amber_cuda_ctx* ctx, const amber_cuda_render_params* params, uint32_t n_pixels, uint64_t total_pairs) {
  int rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;
  RenderArgs args{};
  args.scene = ctx->view;
  args.ctl = ctx->d_ctl.ptr;
  args.film = ctx->d_film.ptr;
  args.seed = params->seed;
  args.sample_begin = params->sample_begin;
  args.sample_stride = effective_stride(*params);
  args.n_pixels = n_pixels;
  args.estimator = params->estimator;
  args.inv_n_pixels = 1.0f / static_cast<float>(n_pixels);
  const int cap = params->max_depth ? static_cast<int>(params->max_depth) : kBdptMaxEvents;
  if (bdpt_wavefront_selected()) {
    rc = ctx->use_bvh ? render_bdpt_wavefront<WideBvh>(ctx, params, args, total_pairs, cap)
                      : render_bdpt_wavefront<ConstList>(ctx, params, args, total_pairs, cap);
    if (rc != AMBER_CUDA_ENOMEM_POOL) return rc;
  }
  // persistent warps: one resident wave, lanes fetch pair indices from ctl->cursor_bdpt
  int resident = 0, sm_count = 148;
  CUDA_TRY(ctx, cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, ctx->device));
  CUDA_TRY(ctx, ctx->use_bvh
                    ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, bdpt_kernel<WideBvh>, BdptLaunch<WideBvh>::kThreads, 0)
                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&resident, bdpt_kernel<ConstList>, BdptLaunch<ConstList>::kThreads, 0));
  const uint64_t threads = ctx->use_bvh ? BdptLaunch<WideBvh>::kThreads : BdptLaunch<ConstList>::kThreads;
  const uint64_t blocks_needed = (total_pairs + threads - 1) / threads;
  const unsigned int grid = static_cast<unsigned int>(
      std::min<uint64_t>(blocks_needed, static_cast<uint64_t>(sm_count) * std::max(1, resident)));
  cudaStream_t stream = ctx->stream;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_begin, stream));
  CUDA_TRY(ctx, cudaMemsetAsync(&ctx->d_ctl.ptr->cursor_bdpt, 0, sizeof(unsigned long long), stream));
  if (ctx->use_bvh) {
    bdpt_kernel<WideBvh><<<grid, BdptLaunch<WideBvh>::kThreads, 0, stream>>>(args, total_pairs, cap);
  } else {
    bdpt_kernel<ConstList><<<grid, BdptLaunch<ConstList>::kThreads, 0, stream>>>(args, total_pairs, cap);
  }
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_end, stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_end));
  CUDA_TRY(ctx, cudaGetLastError());
  float ms = 0;
  CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
  ctx->stats.last_render_ms = ms;
  ctx->samples_done += params->sample_count;
  return AMBER_CUDA_OK;
}

// Argument checks of a render call: done on the caller's thread, so that a bad call fails at once.
int validate_render(amber_cuda_ctx* ctx, const amber_cuda_render_params* params) {
  if (!params) return fail(ctx, AMBER_CUDA_EINVAL, "params is NULL");
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (!ctx->has_sensor) return fail(ctx, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  if (params->estimator != AMBER_CUDA_ESTIMATOR_REFERENCE && params->estimator != AMBER_CUDA_ESTIMATOR_NEE &&
      params->estimator != AMBER_CUDA_ESTIMATOR_LIGHT_TRACING && params->estimator != AMBER_CUDA_ESTIMATOR_BDPT)
    return fail(ctx, AMBER_CUDA_EINVAL, "unknown estimator %d", params->estimator);
  if (params->estimator == AMBER_CUDA_ESTIMATOR_BDPT && params->max_depth > static_cast<uint32_t>(kBdptMaxEvents))
    return fail(ctx, AMBER_CUDA_EINVAL, "bidirectional path tracing: max_depth (events per subpath) must be <= %d",
                kBdptMaxEvents);
  const bool from_lights = params->estimator == AMBER_CUDA_ESTIMATOR_LIGHT_TRACING || params->estimator == AMBER_CUDA_ESTIMATOR_BDPT;
  if (from_lights) {
    // LightSet::GenerateRay on an empty set is undefined in the reference (light_set.h:87-92)
    if (ctx->view.n_lights == 0)
      return fail(ctx, AMBER_CUDA_EINVAL, "light / bidirectional path tracing needs at least one emitter");
  }
  const uint64_t n_frame = static_cast<uint64_t>(ctx->camera.sensor.width) * ctx->camera.sensor.height;
  if (params->pixel_count != 0) {
    if (from_lights) return fail(ctx, AMBER_CUDA_EINVAL, "a pixel window needs a path tracing estimator");
    if (static_cast<uint64_t>(params->pixel_begin) + params->pixel_count > n_frame)
      return fail(ctx, AMBER_CUDA_EINVAL, "pixel window [%u, +%u) leaves the %llu-pixel frame", params->pixel_begin,
                  params->pixel_count, static_cast<unsigned long long>(n_frame));
  } else if (params->pixel_begin != 0) {
    return fail(ctx, AMBER_CUDA_EINVAL, "pixel_begin without pixel_count");
  }
  if (params->sample_begin + static_cast<uint64_t>(params->sample_count) * effective_stride(*params) >= (1ull << 32))
    return fail(ctx, AMBER_CUDA_EINVAL, "sample indices must stay below 2^32");
  return AMBER_CUDA_OK;
}

// One render call (and every queued call it absorbs) on the context's device; runs on the worker thread.
int run_render(amber_cuda_ctx* ctx, const amber_cuda_render_params& job) {
  const amber_cuda_render_params* params = &job;
  if (params->sample_count == 0) return AMBER_CUDA_OK;
  trace(ctx, "job start: samples [%llu, +%u) stride %u", (unsigned long long)job.sample_begin, job.sample_count,
        effective_stride(job));
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  DeviceTurn turn(ctx);
  if (ctx->samples_in_film) {  // an external reduce left the count in the film: take it back first
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_tail, ctx->d_film.ptr + ctx->camera.sensor.width * ctx->camera.sensor.height,
                                  sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->samples_done = static_cast<uint64_t>(ctx->h_tail->x) + (static_cast<uint64_t>(ctx->h_tail->y) << 16) +
                        (static_cast<uint64_t>(ctx->h_tail->z) << 32);
    ctx->samples_in_film = false;
  }

  const uint32_t n_frame = ctx->camera.sensor.width * ctx->camera.sensor.height;
  const uint32_t n_pixels = params->pixel_count ? params->pixel_count : n_frame;
  const uint64_t total_paths = static_cast<uint64_t>(n_pixels) * params->sample_count;
  if (params->estimator == AMBER_CUDA_ESTIMATOR_BDPT) return render_bdpt(ctx, params, n_pixels, total_paths);
  uint64_t slots = params->pool_paths ? params->pool_paths : kDefaultPoolPaths;
  // a queue of small calls renders as ONE wavefront: then the pool is sized for it, not for the first call
  if (!ctx->jobs_may_follow) slots = std::min<uint64_t>(slots, total_paths);
  slots = std::max<uint64_t>(kBlock, (slots + kBlock - 1) / kBlock * kBlock);
  int rc = ensure_pool(ctx, static_cast<uint32_t>(slots));
  if (rc != AMBER_CUDA_OK) return rc;
  rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;

  RenderArgs args{};
  args.pool = ctx->pool;
  args.scene = ctx->view;
  args.ctl = ctx->d_ctl.ptr;
  args.film = ctx->d_film.ptr;
  args.seed = params->seed;
  args.sample_begin = params->sample_begin;
  args.sample_stride = effective_stride(*params);
  args.n_pixels = n_pixels;
  args.pixel_begin = params->pixel_begin;
  args.max_depth = params->max_depth ? std::min<uint32_t>(params->max_depth, 1u << 16) : (1u << 16);
  args.estimator = params->estimator;
  args.inv_n_pixels = 1.0f / static_cast<float>(n_frame);

  cudaStream_t stream = ctx->stream;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_begin, stream));
  control_begin_kernel<<<1, 1, 0, stream>>>(args.ctl, total_paths, args.pool.n_slots, n_pixels);
  pool_reset_kernel<<<blocks_for(args.pool.n_slots), kBlock, 0, stream>>>(args);
  ctx->stats.kernel_launches += 2;
  const bool nee = params->estimator == AMBER_CUDA_ESTIMATOR_NEE && ctx->view.n_lights > 0;
  uint64_t ordinals = 0;
  if (ctx->use_bvh) {
    rc = render_loop<WideBvh>(ctx, args, *params, nee, ctx->bvh_smem, &ordinals);
  } else {
    rc = render_loop<ConstList>(ctx, args, *params, nee, 0, &ordinals);
  }
  if (rc != AMBER_CUDA_OK) return rc;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_end, stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_end));
  CUDA_TRY(ctx, cudaGetLastError());
  float ms = 0;
  CUDA_TRY(ctx, cudaEventElapsedTime(&ms, ctx->ev_begin, ctx->ev_end));
  ctx->stats.last_render_ms = ms;
  ctx->samples_done += ordinals;
  trace(ctx, "job end: %llu samples per pixel in %.3f ms of device time, %llu iterations", (unsigned long long)ordinals, ms,
        (unsigned long long)ctx->stats.extend_launches);
  return AMBER_CUDA_OK;
}

void worker_main(amber_cuda_ctx* ctx) {
  for (;;) {
    amber_cuda_render_params job;
    {
      std::unique_lock<std::mutex> lock(ctx->queue_mutex);
      ctx->queue_cv.wait(lock, [&] { return ctx->quit || !ctx->jobs.empty(); });
      if (ctx->jobs.empty()) return;  // quit
      job = ctx->jobs.front();
      ctx->jobs.pop_front();
      ctx->running = 1;
      ctx->busy = true;
      ctx->absorbed_ends.assign(1, job.sample_count);
      ctx->jobs_may_follow = !ctx->jobs.empty();
    }
    int rc = AMBER_CUDA_OK;
    {
      std::lock_guard<std::mutex> guard(ctx->queue_mutex);
      rc = ctx->job_rc;
    }
    if (rc == AMBER_CUDA_OK) rc = run_render(ctx, job);  // after a failure the rest of the queue is dropped
    {
      std::lock_guard<std::mutex> guard(ctx->queue_mutex);
      if (rc != AMBER_CUDA_OK && ctx->job_rc == AMBER_CUDA_OK) {
        ctx->job_rc = rc;
        ctx->job_error = ctx->error;
      }
      ctx->running = 0;
      ctx->busy = false;
      ctx->absorbed_ends.clear();
    }
    ctx->idle_cv.notify_all();
  }
}

// Every entry point but render_async / pending starts here: the queue is drained first.
int wait_idle(amber_cuda_ctx* ctx, int max_pending) {
  std::unique_lock<std::mutex> lock(ctx->queue_mutex);
  if (max_pending > 0) {
    ctx->idle_cv.wait(lock, [&] { return static_cast<int>(ctx->jobs.size()) + ctx->running <= max_pending; });
  } else {
    ctx->idle_cv.wait(lock, [&] { return ctx->jobs.empty() && !ctx->busy; });
  }
  return AMBER_CUDA_OK;
}

// The film's sample count after the work queued on the stream: the host's, or (after an external
// reduce) the one carried by the film's trailing element.  Synchronises the stream.
int film_sample_count(amber_cuda_ctx* ctx, uint64_t* samples_done) {
  if (ctx->samples_in_film) {
    const size_t n_pixels = static_cast<size_t>(ctx->camera.sensor.width) * ctx->camera.sensor.height;
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_tail, ctx->d_film.ptr + n_pixels, sizeof(float4), cudaMemcpyDeviceToHost, ctx->stream));
  }
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  if (ctx->samples_in_film)
    ctx->samples_done = static_cast<uint64_t>(ctx->h_tail->x) + (static_cast<uint64_t>(ctx->h_tail->y) << 16) +
                        (static_cast<uint64_t>(ctx->h_tail->z) << 32);
  if (samples_done) *samples_done = ctx->samples_done;
  return AMBER_CUDA_OK;
}

}  // namespace amber_host

using namespace amber_host;

extern "C" {

int amber_cuda_render_async(amber_cuda_ctx* ctx, const amber_cuda_render_params* params) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  {
    // scene and sensor only change while the queue is idle (upload_scene / set_sensor wait), so the
    // checks below see the state the job will run with
    std::lock_guard<std::mutex> guard(ctx->queue_mutex);
    if (ctx->job_rc != AMBER_CUDA_OK) {
      ctx->error = ctx->job_error;
      return ctx->job_rc;  // sticky until amber_cuda_wait reports it
    }
  }
  const int rc = validate_render(ctx, params);
  if (rc != AMBER_CUDA_OK) return rc;
  if (params->sample_count == 0) return AMBER_CUDA_OK;
  {
    std::lock_guard<std::mutex> guard(ctx->queue_mutex);
    ctx->jobs.push_back(*params);
  }
  ctx->queue_cv.notify_one();
  return AMBER_CUDA_OK;
}

int amber_cuda_wait(amber_cuda_ctx* ctx) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  wait_idle(ctx);
  std::lock_guard<std::mutex> guard(ctx->queue_mutex);
  const int rc = ctx->job_rc;
  if (rc != AMBER_CUDA_OK) ctx->error = ctx->job_error;
  ctx->job_rc = AMBER_CUDA_OK;
  return rc;
}

int amber_cuda_pending(amber_cuda_ctx* ctx) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  std::lock_guard<std::mutex> guard(ctx->queue_mutex);
  return static_cast<int>(ctx->jobs.size()) + ctx->running;
}

int amber_cuda_wait_pending(amber_cuda_ctx* ctx, int max_pending) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (max_pending <= 0) return amber_cuda_wait(ctx);
  wait_idle(ctx, max_pending);
  return AMBER_CUDA_OK;
}

int amber_cuda_render(amber_cuda_ctx* ctx, const amber_cuda_render_params* params) {
  const int rc = amber_cuda_render_async(ctx, params);
  if (rc != AMBER_CUDA_OK) return rc;
  return amber_cuda_wait(ctx);
}

int amber_cuda_resolve(amber_cuda_ctx* ctx, float* host_rgb, uint64_t* samples_done) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_sensor) return fail(ctx, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  if (!host_rgb) return fail(ctx, AMBER_CUDA_EINVAL, "host_rgb is NULL");
  const int wait_rc = amber_cuda_wait(ctx);
  if (wait_rc != AMBER_CUDA_OK) return wait_rc;
  trace(ctx, "resolve: queue drained");
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const uint32_t n_pixels = ctx->camera.sensor.width * ctx->camera.sensor.height;
  CUDA_TRY(ctx, ctx->d_resolved.resize(static_cast<size_t>(n_pixels) * 3));
  const float inv = ctx->samples_done ? 1.0f / static_cast<float>(ctx->samples_done) : 0.0f;
  resolve_kernel<<<blocks_for(n_pixels), kBlock, 0, ctx->stream>>>(ctx->d_film.ptr, n_pixels, inv,
                                                                  ctx->samples_in_film ? 1 : 0, ctx->d_resolved.ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(host_rgb, ctx->d_resolved.ptr, static_cast<size_t>(n_pixels) * 3 * sizeof(float),
                                cudaMemcpyDeviceToHost, ctx->stream));
  trace(ctx, "resolve: kernel and copy of %zu bytes queued", static_cast<size_t>(n_pixels) * 3 * sizeof(float));
  const int count_rc = film_sample_count(ctx, samples_done);  // synchronises the stream
  if (count_rc != AMBER_CUDA_OK) return count_rc;
  trace(ctx, "resolve: done");
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_resolve_ldr(amber_cuda_ctx* ctx, uint8_t* host_rgb8, int flip_x, uint64_t* samples_done) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_sensor) return fail(ctx, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  if (!host_rgb8) return fail(ctx, AMBER_CUDA_EINVAL, "host_rgb8 is NULL");
  const int wait_rc = amber_cuda_wait(ctx);
  if (wait_rc != AMBER_CUDA_OK) return wait_rc;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const uint32_t width = ctx->camera.sensor.width, height = ctx->camera.sensor.height;
  const size_t n_pixels = static_cast<size_t>(width) * height;
  CUDA_TRY(ctx, ctx->d_resolved8.resize(n_pixels * 3));
  const float inv = ctx->samples_done ? 1.0f / static_cast<float>(ctx->samples_done) : 0.0f;
  resolve_tonemap_kernel<<<blocks_for(n_pixels), kBlock, 0, ctx->stream>>>(ctx->d_film.ptr, width, height, inv,
                                                                          ctx->samples_in_film ? 1 : 0, flip_x ? 1 : 0,
                                                                          ctx->d_resolved8.ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(host_rgb8, ctx->d_resolved8.ptr, n_pixels * 3, cudaMemcpyDeviceToHost, ctx->stream));
  const int count_rc = film_sample_count(ctx, samples_done);  // synchronises the stream
  if (count_rc != AMBER_CUDA_OK) return count_rc;
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_film_device(amber_cuda_ctx* ctx, void** device_ptr, uint64_t* n_floats, uint64_t* samples_done) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_sensor) return fail(ctx, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  const int wait_rc = amber_cuda_wait(ctx);
  if (wait_rc != AMBER_CUDA_OK) return wait_rc;
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  const size_t n_pixels = static_cast<size_t>(ctx->camera.sensor.width) * ctx->camera.sensor.height;
  if (!ctx->samples_in_film) {  // publish the host's count in the trailing element: it travels with the film
    *ctx->h_tail = make_float4(static_cast<float>(ctx->samples_done & 0xffffu), static_cast<float>((ctx->samples_done >> 16) & 0xffffu),
                               static_cast<float>(ctx->samples_done >> 32), 0.0f);
    CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_film.ptr + n_pixels, ctx->h_tail, sizeof(float4), cudaMemcpyHostToDevice, ctx->stream));
    CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  }
  if (device_ptr) *device_ptr = ctx->d_film.ptr;
  if (n_floats) *n_floats = static_cast<uint64_t>(n_pixels + 1) * 4;
  if (samples_done) *samples_done = ctx->samples_done;
  return AMBER_CUDA_OK;
}

int amber_cuda_film_set_samples(amber_cuda_ctx* ctx, uint64_t samples_done) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  wait_idle(ctx);
  if (samples_done == AMBER_CUDA_SAMPLES_IN_FILM) {
    ctx->samples_in_film = true;
  } else {
    ctx->samples_in_film = false;
    ctx->samples_done = samples_done;
  }
  return AMBER_CUDA_OK;
}

int amber_cuda_film_reduce(amber_cuda_ctx* const* ctxs, int n) {
  if (!ctxs || n < 1 || !ctxs[0]) return AMBER_CUDA_EINVAL;
  amber_cuda_ctx* dst = ctxs[0];
  if (n - 1 > kMaxReduceSources) return fail(dst, AMBER_CUDA_EINVAL, "at most %d films can be summed into one", kMaxReduceSources);
  if (!dst->has_sensor) return fail(dst, AMBER_CUDA_ESTATE, "set_sensor has not been called");
  const SensorDev& sensor = dst->camera.sensor;
  const uint32_t n_pixels = sensor.width * sensor.height;
  for (int k = 0; k < n; k++) {
    if (!ctxs[k]) return fail(dst, AMBER_CUDA_EINVAL, "context %d is NULL", k);
    const int rc = amber_cuda_wait(ctxs[k]);
    if (rc != AMBER_CUDA_OK) {
      if (k > 0) dst->error = ctxs[k]->error;
      return rc;
    }
    if (!ctxs[k]->has_sensor || ctxs[k]->camera.sensor.width != sensor.width || ctxs[k]->camera.sensor.height != sensor.height)
      return fail(dst, AMBER_CUDA_EINVAL, "context %d has a different sensor", k);
    if (ctxs[k]->samples_in_film) return fail(dst, AMBER_CUDA_ESTATE, "context %d holds an externally reduced film", k);
    for (int j = 0; j < k; j++)
      if (ctxs[j] == ctxs[k]) return fail(dst, AMBER_CUDA_EINVAL, "context %d is listed twice", k);
  }
  if (n == 1) return AMBER_CUDA_OK;
  CUDA_TRY(dst, cudaSetDevice(dst->device));
  FilmSources sources{};
  std::vector<float4*> staged;  // films of devices without peer access, copied next to dst
  auto release = [&] { for (float4* p : staged) cudaFree(p); };
  for (int k = 1; k < n; k++) {
    const amber_cuda_ctx* src = ctxs[k];
    const float4* film = src->d_film.ptr;
    if (src->device != dst->device) {
      int can = 0;
      CUDA_TRY(dst, cudaDeviceCanAccessPeer(&can, dst->device, src->device));
      if (can) {
        const cudaError_t err = cudaDeviceEnablePeerAccess(src->device, 0);
        if (err == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError(); else if (err != cudaSuccess) can = 0;
      }
      if (!can) {
        cudaGetLastError();
        float4* copy = nullptr;
        if (cudaMalloc(&copy, static_cast<size_t>(n_pixels) * sizeof(float4)) != cudaSuccess) {
          release();
          return fail(dst, AMBER_CUDA_ECUDA, "film_reduce: no memory to stage the film of device %d", src->device);
        }
        staged.push_back(copy);
        const cudaError_t err = cudaMemcpyPeerAsync(copy, dst->device, film, src->device,
                                                    static_cast<size_t>(n_pixels) * sizeof(float4), dst->stream);
        if (err != cudaSuccess) {
          release();
          return fail(dst, AMBER_CUDA_ECUDA, "film_reduce: peer copy from device %d failed: %s", src->device, cudaGetErrorString(err));
        }
        film = copy;
      }
    }
    sources.film[sources.n++] = film;
  }
  int sm_count = 148;
  cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dst->device);
  const unsigned int grid = std::min<unsigned int>(blocks_for(n_pixels), static_cast<unsigned int>(sm_count) * 8u);
  film_sum_kernel<<<grid, kBlock, 0, dst->stream>>>(dst->d_film.ptr, sources, n_pixels);
  dst->stats.kernel_launches++;
  const cudaError_t err = cudaStreamSynchronize(dst->stream);
  release();
  if (err != cudaSuccess) return fail(dst, AMBER_CUDA_ECUDA, "film_reduce failed: %s", cudaGetErrorString(err));
  CUDA_TRY(dst, cudaGetLastError());
  for (int k = 1; k < n; k++) dst->samples_done += ctxs[k]->samples_done;
  return AMBER_CUDA_OK;
}

// the caller holds the device's turn
static int cast_device_locked(amber_cuda_ctx* ctx, const void* ray_origin4, const void* ray_direction4, uint64_t n,
                              void* hit4, float* elapsed_ms) {
  int rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;
  const float4* o = static_cast<const float4*>(ray_origin4);
  const float4* d = static_cast<const float4*>(ray_direction4);
  float4* h = static_cast<float4*>(hit4);
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_begin, ctx->stream));
  if (ctx->use_bvh) {
    if (n >= (1ull << 32)) return fail(ctx, AMBER_CUDA_EINVAL, "at most 2^32 - 1 rays per call");
    CUDA_TRY(ctx, cudaMemsetAsync(&ctx->d_ctl.ptr->cursor_cast, 0, sizeof(unsigned int), ctx->stream));
    cast_bvh_kernel<<<ctx->bvh_grid, kBlock, ctx->bvh_smem, ctx->stream>>>(ctx->view, o, d, static_cast<uint32_t>(n), h,
                                                                         &ctx->d_ctl.ptr->cursor_cast);
  } else {
    cast_kernel<ConstList><<<blocks_for(n), kBlock, 0, ctx->stream>>>(ctx->view, o, d, n, h);
  }
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaEventRecord(ctx->ev_end, ctx->stream));
  CUDA_TRY(ctx, cudaEventSynchronize(ctx->ev_end));
  CUDA_TRY(ctx, cudaGetLastError());
  if (elapsed_ms) CUDA_TRY(ctx, cudaEventElapsedTime(elapsed_ms, ctx->ev_begin, ctx->ev_end));
  return AMBER_CUDA_OK;
}

int amber_cuda_cast_device(amber_cuda_ctx* ctx, const void* ray_origin4, const void* ray_direction4, uint64_t n,
                           void* hit4, float* elapsed_ms) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (n == 0) return AMBER_CUDA_OK;
  if (!ray_origin4 || !ray_direction4 || !hit4) return fail(ctx, AMBER_CUDA_EINVAL, "NULL buffer");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  DeviceTurn turn(ctx);
  return cast_device_locked(ctx, ray_origin4, ray_direction4, n, hit4, elapsed_ms);
}

int amber_cuda_cast(amber_cuda_ctx* ctx, const float* origins, const float* directions, uint64_t n,
                    int32_t* prim, float* t) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (n == 0) return AMBER_CUDA_OK;
  if (!origins || !directions || !prim || !t) return fail(ctx, AMBER_CUDA_EINVAL, "NULL buffer");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  DeviceTurn turn(ctx);
  cudaStream_t stream = ctx->stream;
  CUDA_TRY(ctx, ctx->d_scratch_f[0].resize(3 * n));
  CUDA_TRY(ctx, ctx->d_scratch_f[1].resize(3 * n));
  CUDA_TRY(ctx, ctx->d_scratch_f[2].resize(n));
  CUDA_TRY(ctx, ctx->d_scratch_i[0].resize(n));
  for (auto& b : ctx->d_scratch_4) CUDA_TRY(ctx, b.resize(n));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_scratch_f[0].ptr, origins, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->d_scratch_f[1].ptr, directions, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  pack_rays_kernel<<<blocks_for(n), kBlock, 0, stream>>>(ctx->d_scratch_f[0].ptr, ctx->d_scratch_f[1].ptr, n,
                                                        ctx->d_scratch_4[0].ptr, ctx->d_scratch_4[1].ptr);
  const int rc = cast_device_locked(ctx, ctx->d_scratch_4[0].ptr, ctx->d_scratch_4[1].ptr, n,
                                    ctx->d_scratch_4[2].ptr, nullptr);
  if (rc != AMBER_CUDA_OK) return rc;
  unpack_hits_kernel<<<blocks_for(n), kBlock, 0, stream>>>(ctx->d_scratch_4[2].ptr, n, ctx->d_internal_to_prim.ptr,
                                                          ctx->d_scratch_i[0].ptr, ctx->d_scratch_f[2].ptr);
  ctx->stats.kernel_launches += 2;
  CUDA_TRY(ctx, cudaMemcpyAsync(prim, ctx->d_scratch_i[0].ptr, n * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(t, ctx->d_scratch_f[2].ptr, n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

}  // extern "C"

// ---- amber_cuda_cast_static: closest hit without any allocation ---------------------------------------
// Scratch of the entry point below: the packed rays (origin, direction as float4) and the float4 hit records
// of up to AMBER_CUDA_CAST_STATIC_MAX rays, in a file-scope device array instead of cudaMalloc'ed buffers:
// 3 x 16 B x 65 536 = 3 MiB, held by every GPU that loads the library until the process exits.  In host code
// the array's NAME is a host-side stand-in, not a device address; the address the kernels get comes from
// cudaGetSymbolAddress, asked once per device with that device current (static_scratch_for).
static __device__ __align__(256) float4 g_cast_static_scratch[3 * AMBER_CUDA_CAST_STATIC_MAX];

namespace amber_host {

std::mutex g_static_scratch_mutex;
float4* g_static_scratch_ptr[kMaxDevices] = {};

// The device address of g_cast_static_scratch on the context's device (which must be current).
int static_scratch_for(amber_cuda_ctx* ctx, float4** out) {
  if (ctx->device < 0 || ctx->device >= kMaxDevices) return fail(ctx, AMBER_CUDA_EINVAL, "device %d out of range", ctx->device);
  std::lock_guard<std::mutex> guard(g_static_scratch_mutex);
  if (!g_static_scratch_ptr[ctx->device]) {
    void* address = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&address, g_cast_static_scratch);
    if (err != cudaSuccess || !address)
      return fail(ctx, AMBER_CUDA_ECUDA, "cudaGetSymbolAddress(g_cast_static_scratch) failed on device %d: %s", ctx->device,
                  cudaGetErrorString(err));
    g_static_scratch_ptr[ctx->device] = static_cast<float4*>(address);
  }
  *out = g_static_scratch_ptr[ctx->device];
  return AMBER_CUDA_OK;
}

}  // namespace amber_host

extern "C" {

int amber_cuda_cast_static(amber_cuda_ctx* ctx, const float* device_origins3, const float* device_directions3, uint64_t n,
                           int32_t* device_prim, float* device_t) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (n > static_cast<uint64_t>(AMBER_CUDA_CAST_STATIC_MAX))
    return fail(ctx, AMBER_CUDA_EINVAL, "amber_cuda_cast_static: %llu rays exceed the static scratch of %d rays "
                "(AMBER_CUDA_CAST_STATIC_MAX); split the call or use amber_cuda_cast",
                static_cast<unsigned long long>(n), AMBER_CUDA_CAST_STATIC_MAX);
  if (n == 0) return AMBER_CUDA_OK;
  if (!device_origins3 || !device_directions3 || !device_prim || !device_t) return fail(ctx, AMBER_CUDA_EINVAL, "NULL buffer");
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  float4* scratch = nullptr;
  int rc = static_scratch_for(ctx, &scratch);
  if (rc != AMBER_CUDA_OK) return rc;
  DeviceTurn turn(ctx);
  rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;
  // every element the kernels read is written first by this call (rays by pack, hits by the cast), so the
  // previous call's contents need no clearing; the BVH walk's ray cursor is the one thing that starts at zero
  float4* o4 = scratch;
  float4* d4 = scratch + AMBER_CUDA_CAST_STATIC_MAX;
  float4* h4 = scratch + 2 * AMBER_CUDA_CAST_STATIC_MAX;
  cudaStream_t stream = ctx->stream;
  pack_rays_kernel<<<blocks_for(n), kBlock, 0, stream>>>(device_origins3, device_directions3, n, o4, d4);
  if (ctx->use_bvh) {
    CUDA_TRY(ctx, cudaMemsetAsync(&ctx->d_ctl.ptr->cursor_cast, 0, sizeof(unsigned int), stream));
    cast_bvh_kernel<<<ctx->bvh_grid, kBlock, ctx->bvh_smem, stream>>>(ctx->view, o4, d4, static_cast<uint32_t>(n), h4,
                                                                    &ctx->d_ctl.ptr->cursor_cast);
  } else {
    cast_kernel<ConstList><<<blocks_for(n), kBlock, 0, stream>>>(ctx->view, o4, d4, n, h4);
  }
  unpack_hits_kernel<<<blocks_for(n), kBlock, 0, stream>>>(h4, n, ctx->d_internal_to_prim.ptr, device_prim, device_t);
  ctx->stats.kernel_launches += 3;
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

static int sample_scatter(amber_cuda_ctx* ctx, bool importance, int32_t material, uint64_t n, const float* normal3,
                          const float* dir_out3, const float* xi8, float* dir_in3, float* weight3,
                          int32_t* draws) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (material < 0 || static_cast<size_t>(material) >= ctx->d_materials.count)
    return fail(ctx, AMBER_CUDA_EINVAL, "material %d out of range", material);
  if (n == 0) return AMBER_CUDA_OK;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  cudaStream_t stream = ctx->stream;
  auto& f = ctx->d_scratch_f;
  CUDA_TRY(ctx, f[0].resize(3 * n)); CUDA_TRY(ctx, f[1].resize(3 * n)); CUDA_TRY(ctx, f[2].resize(8 * n));
  CUDA_TRY(ctx, f[3].resize(3 * n)); CUDA_TRY(ctx, f[4].resize(3 * n));
  CUDA_TRY(ctx, ctx->d_scratch_i[0].resize(n));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[0].ptr, normal3, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[1].ptr, dir_out3, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[2].ptr, xi8, 8 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  sample_light_kernel<<<blocks_for(n), kBlock, 0, stream>>>(ctx->view, material, importance, n, f[0].ptr, f[1].ptr,
                                                           f[2].ptr, f[3].ptr, f[4].ptr, ctx->d_scratch_i[0].ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(dir_in3, f[3].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(weight3, f[4].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(draws, ctx->d_scratch_i[0].ptr, n * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_sample_light(amber_cuda_ctx* ctx, int32_t material, uint64_t n, const float* normal3,
                            const float* dir_out3, const float* xi8, float* dir_in3, float* weight3,
                            int32_t* draws) {
  return sample_scatter(ctx, false, material, n, normal3, dir_out3, xi8, dir_in3, weight3, draws);
}

int amber_cuda_sample_importance(amber_cuda_ctx* ctx, int32_t material, uint64_t n, const float* normal3,
                                 const float* dir_out3, const float* xi8, float* dir_in3, float* weight3,
                                 int32_t* draws) {
  return sample_scatter(ctx, true, material, n, normal3, dir_out3, xi8, dir_in3, weight3, draws);
}

int amber_cuda_light_rays(amber_cuda_ctx* ctx, uint64_t n, const float* xi5, float* origin3, float* direction3,
                          float* weight3) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene) return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene has not been called");
  if (ctx->view.n_lights == 0) return fail(ctx, AMBER_CUDA_EINVAL, "the scene has no emitting triangle");
  if (n == 0) return AMBER_CUDA_OK;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  cudaStream_t stream = ctx->stream;
  auto& f = ctx->d_scratch_f;
  CUDA_TRY(ctx, f[0].resize(5 * n));
  for (int k = 1; k <= 3; k++) CUDA_TRY(ctx, f[k].resize(3 * n));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[0].ptr, xi5, 5 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  light_rays_kernel<<<blocks_for(n), kBlock, 0, stream>>>(ctx->view, n, f[0].ptr, f[1].ptr, f[2].ptr, f[3].ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(origin3, f[1].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(direction3, f[2].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(weight3, f[3].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_lens_response(amber_cuda_ctx* ctx, uint64_t n, const float* position3, const float* dir_out3,
                             int32_t* pixel, float* value) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene || !ctx->has_sensor)
    return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene and set_sensor must be called first");
  if (n == 0) return AMBER_CUDA_OK;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  DeviceTurn turn(ctx);
  int rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;
  cudaStream_t stream = ctx->stream;
  auto& f = ctx->d_scratch_f;
  CUDA_TRY(ctx, f[0].resize(3 * n)); CUDA_TRY(ctx, f[1].resize(3 * n)); CUDA_TRY(ctx, f[2].resize(n));
  CUDA_TRY(ctx, ctx->d_scratch_i[0].resize(n));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[0].ptr, position3, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[1].ptr, dir_out3, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  lens_response_kernel<<<blocks_for(n), kBlock, 0, stream>>>(n, f[0].ptr, f[1].ptr, ctx->d_scratch_i[0].ptr, f[2].ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(pixel, ctx->d_scratch_i[0].ptr, n * sizeof(int32_t), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(value, f[2].ptr, n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_normalize(amber_cuda_ctx* ctx, uint64_t n, const float* vector3, const float* denominator,
                         float* unit3) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!vector3 || !unit3) return fail(ctx, AMBER_CUDA_EINVAL, "vector3 / unit3 is NULL");
  if (n == 0) return AMBER_CUDA_OK;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  cudaStream_t stream = ctx->stream;
  auto& f = ctx->d_scratch_f;
  CUDA_TRY(ctx, f[0].resize(3 * n)); CUDA_TRY(ctx, f[1].resize(3 * n));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[0].ptr, vector3, 3 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  if (denominator) {
    CUDA_TRY(ctx, f[2].resize(n));
    CUDA_TRY(ctx, cudaMemcpyAsync(f[2].ptr, denominator, n * sizeof(float), cudaMemcpyHostToDevice, stream));
  }
  normalize_kernel<<<blocks_for(n), kBlock, 0, stream>>>(n, f[0].ptr, denominator ? f[2].ptr : nullptr, f[1].ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(unit3, f[1].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_eye_rays(amber_cuda_ctx* ctx, uint64_t n, const uint32_t* px, const uint32_t* py, const float* xi5,
                        float* origin3, float* direction3, float* weight) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  if (!ctx->has_scene || !ctx->has_sensor)
    return fail(ctx, AMBER_CUDA_ESTATE, "upload_scene and set_sensor must be called first");
  if (n == 0) return AMBER_CUDA_OK;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  DeviceTurn turn(ctx);
  int rc = bind_constants(ctx);
  if (rc != AMBER_CUDA_OK) return rc;
  cudaStream_t stream = ctx->stream;
  auto& f = ctx->d_scratch_f;
  CUDA_TRY(ctx, f[0].resize(5 * n)); CUDA_TRY(ctx, f[1].resize(3 * n)); CUDA_TRY(ctx, f[2].resize(3 * n));
  CUDA_TRY(ctx, f[3].resize(n));
  CUDA_TRY(ctx, ctx->d_scratch_i[0].resize(n)); CUDA_TRY(ctx, ctx->d_scratch_i[1].resize(n));
  uint32_t* d_px = reinterpret_cast<uint32_t*>(ctx->d_scratch_i[0].ptr);
  uint32_t* d_py = reinterpret_cast<uint32_t*>(ctx->d_scratch_i[1].ptr);
  CUDA_TRY(ctx, cudaMemcpyAsync(d_px, px, n * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(d_py, py, n * sizeof(uint32_t), cudaMemcpyHostToDevice, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(f[0].ptr, xi5, 5 * n * sizeof(float), cudaMemcpyHostToDevice, stream));
  eye_rays_kernel<<<blocks_for(n), kBlock, 0, stream>>>(n, d_px, d_py, f[0].ptr, f[1].ptr, f[2].ptr, f[3].ptr);
  ctx->stats.kernel_launches++;
  CUDA_TRY(ctx, cudaMemcpyAsync(origin3, f[1].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(direction3, f[2].ptr, 3 * n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaMemcpyAsync(weight, f[3].ptr, n * sizeof(float), cudaMemcpyDeviceToHost, stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(stream));
  CUDA_TRY(ctx, cudaGetLastError());
  return AMBER_CUDA_OK;
}

int amber_cuda_get_stats(amber_cuda_ctx* ctx, amber_cuda_stats* out) {
  if (!ctx || !out) return AMBER_CUDA_EINVAL;
  wait_idle(ctx);
  CUDA_TRY(ctx, cudaSetDevice(ctx->device));
  Control ctl;
  CUDA_TRY(ctx, cudaMemcpyAsync(ctx->h_ctl, ctx->d_ctl.ptr, sizeof(Control), cudaMemcpyDeviceToHost, ctx->stream));
  CUDA_TRY(ctx, cudaStreamSynchronize(ctx->stream));
  ctl = *ctx->h_ctl;
  ctx->stats.paths = ctl.paths;
  ctx->stats.closest_rays = ctl.closest_rays;
  ctx->stats.shadow_rays = ctl.shadow_rays;
  ctx->stats.truncated_paths = ctl.truncated;
  ctx->stats.phong_capped = ctl.phong_capped;
  ctx->stats.light_splats = ctl.splats;
  ctx->stats.bvh_node_visits = ctl.bvh_work[0];
  ctx->stats.bvh_triangle_tests = ctl.bvh_work[1];
  ctx->stats.bvh_node_visits_connect = ctl.bvh_work[2];
  ctx->stats.bvh_triangle_tests_connect = ctl.bvh_work[3];
  ctx->stats.accel = ctx->use_bvh ? 1u : 0u;
  ctx->stats.bvh_node_bytes = static_cast<uint32_t>(bvh_build::kNodeFloats * sizeof(float));
  ctx->stats.bvh_triangle_bytes = static_cast<uint32_t>(sizeof(TriIsect));
  ctx->stats.bvh_nodes = ctx->bvh_nodes;
  *out = ctx->stats;
  return AMBER_CUDA_OK;
}

int amber_cuda_set_profiling(amber_cuda_ctx* ctx, int profile) {
  if (!ctx) return AMBER_CUDA_EINVAL;
  wait_idle(ctx);
  ctx->profiling = profile != 0;
  return AMBER_CUDA_OK;
}

}  // extern "C"
