// scene_device.cuh -- device-side scene layout.
//
// Two acceleration layouts sit behind the same closest-hit / any-hit interface:
//  * ConstList (this file): scenes of up to kMaxConstTris triangles + kMaxConstSpheres
//    spheres live entirely in __constant__ memory and are tested warp-uniformly in
//    one loop -- every lane reads the same primitive, so operands come through the
//    uniform datapath and no lane ever waits on a gather.  This is the layout the
//    built-in Cornell box (25 objects) uses; the reference itself ends up testing all
//    25 objects per ray there (its SAH tree collapses to one leaf).
//  * WideBvh (bvh.cuh): larger scenes, 4-wide 64-byte quantised nodes in global memory.
#pragma once

#include <cstdint>

#include "vec.cuh"

namespace amber_cuda {

constexpr int kMaxConstTris = 128;
constexpr int kMaxConstSpheres = 32;
constexpr int kMaxBlades = 16;

// float just below the reference's `kEPS = 1e-6L` (include/amber/constants.h:26):
// (float)1e-6 = 9.99999997e-7 < 1e-6L, so for a float t, `t < kEPS` (reject,
// src/amber/scene/primitive_triangle.cc:122-125) <=> t <= kEpsF and `t > kEPS`
// (accept, src/amber/scene/primitive_sphere.cc:90,98) <=> t > kEpsF.
constexpr float kEpsF = 1e-6f;
constexpr float kRussianRoulette = 0.9375f;  // include/amber/constants.h:28
constexpr float kPiF = 3.14159274101257324f;  // static_cast<float>(kPI)

// Triangle in intersection form.  With n = E1 x E2, T = o - v0:
//   t = -(n.T)/(n.d),  h = T + t d,  u = h.nu,  v = h.nv
// where nu = (E2 x n)/|n|^2 and nv = (n x E1)/|n|^2.  Algebraically the test of
// src/amber/scene/primitive_triangle.cc:97-128 (det = -n.d, Q.E2 = n.T); the edge
// tests use the same inclusive bounds.
struct TriIsect {
  float4 n_v0x;   // n.xyz,  v0.x
  float4 nu_v0y;  // nu.xyz, v0.y
  float4 nv_v0z;  // nv.xyz, v0.z
};

// Triangle in shading form: position = v0 + u E1 + v E2 (primitive_triangle.cc:127),
// stored geometric normal = normalize((v1-v0) x (v2-v0)) (primitive_triangle.cc:59-68).
struct TriShade {
  float4 v0_nx;  // v0.xyz, normal.x
  float4 e1_ny;  // E1.xyz, normal.y
  float4 e2_nz;  // E2.xyz, normal.z
};

struct alignas(16) MaterialDev {
  int32_t kind;
  float r, g, b;
  float param;  // Phong exponent / ior
  float r0;     // Refraction: Fresnel reflectance at normal incidence (material_refraction.cc:265-269)
  int32_t nee;  // 1 if the surface is connectable: SurfaceType::Diffuse (Lambertian, Phong)
  int32_t pad;
};

// Emitting object, sorted by power as include/amber/scene/light_set.h:61-81: a triangle (sampled by
// next-event estimation, light tracing and BDPT) or a sphere (light tracing and BDPT only; the path
// tracers count a sphere's emission on every hit).
struct alignas(16) LightDev {
  float4 v0;  // triangle: v0 xyz | sphere: centre xyz ; w: cumulative power
  float4 v1;  // triangle: v1 xyz ; w: light-ray weight r (Irradiance / PDFArea)
  float4 v2;  // triangle: v2 xyz ; w: light-ray weight g
  float4 n;   // triangle: normal xyz ; w: light-ray weight b
  float4 le_over_pdf;  // radiance / pdf_area (rgb), material index (int bits)
  float4 shape;        // x: radius of a sphere, 0 for a triangle
};

struct LensDev {
  float origin[3];
  float global[9];  // row-major lens-local -> world (Matrix3 part of the transform)
  float local[9];   // inverse
  float focus_distance, sensor_distance, p_area;
  float focus_over_sensor;  // -focus_distance / sensor_distance, lens_thin.cc:92
  int32_t kind;             // AMBER_CUDA_LENS_*: 1 = BasicPinhole (src/amber/scene/lens_pinhole.cc)
  int32_t n_blades;
  float blades[kMaxBlades][12];  // v0 v1 v2 normal
};

struct SensorDev {
  uint32_t width, height;
  float film_width, film_height;
  float size_over_area;  // sensor.Size() / sensor.SceneArea(), lens_thin.cc:145
  float inv_area;        // 1 / sensor.SceneArea(), lens_pinhole.cc:103
};

// Constant-memory list, triangles batched by supporting plane.  Triangles that lie in one
// plane (the two halves of a wall, the blades of the aperture) share ONE ray/plane
// intersection.  With g = the group's anchor vertex, n = its (unnormalised) normal,
// T = o - g:   t = -(n.T)/(n.d),  h = T + t d.
// Inside a group, two triangles that form a parallelogram (r, p, q | q, p + q - r, p) are ONE
// entry: s = ns.h + cs, r = nr.h + cr are its edge coordinates from corner r, the hit is
// inside iff s, r in [0, 1] and belongs to the first triangle iff s + r <= 1.  A lone
// triangle is an entry with the usual u, v >= 0, u + v <= 1.  Either way an entry is two
// float4: (nu, cu), (nv, cv) with cu = -nu.(v0 - g), i.e. u = nu.h + cu.
struct PlaneGroup {
  float4 n_gx;     // n.xyz, g.x
  float4 g_range;  // g.y, g.z, (first entry | first triangle << 16), (quads | lone triangles << 16)
};
struct PlaneEntry {
  float4 nu_cu;  // nu.xyz, cu
  float4 nv_cv;  // nv.xyz, cv
};

// Triangles whose plane is perpendicular to a coordinate axis k (walls, floors, table tops: most
// of an architectural scene) need no dot products.  With a, b the other two axes and r the entry's
// corner:   t = (r_k - o_k) / d_k,   h = (o_a - r_a + t d_a, o_b - r_b + t d_b),
//           s = nu_a h_a + nu_b h_b,   q = nv_a h_a + nv_b h_b      (1/d once per ray).
// One entry is a parallelogram (two triangles, limit = 2) or a lone triangle (limit = 1); the hit
// is inside iff s, q in [0, 1] and s + q <= limit, and belongs to triangle `tri` + (s + q > 1).
// Entries are self-contained (no per-plane state), so the scan is one flat loop per axis.
struct AxisEntry {
  float r_k, r_a;
  float r_b, nu_a;
  float nu_b, nv_a;
  float nv_b, limit;
};

// Two consecutive unclustered entries of one axis, field by field as float2, so that the scan evaluates
// both with packed FADD2 / FMUL2 / FFMA2 (sm_100: one instruction, two lanes of fp32).  r_a and r_b are
// stored negated (h = t d + (o + (-r))).  An odd entry count is padded with an entry nothing can hit
// (nu = nv = 0, limit = -1: s + q = 0 > limit).
struct AxisPair {
  float2 r_k, neg_r_a, neg_r_b, nu_a, nu_b, nv_a, nv_b, limit;
};

// Three or more axis entries in ONE plane (the blades of the aperture, a tessellated wall) form a
// cluster: the ray/plane point is tested against the cluster's bounding rectangle first and the
// entries are scanned only if some lane of the warp lands inside it.  The Cornell box's aperture fan
// (6 entries, reached by 3 rays in 1e5) is skipped this way.
constexpr int kMaxAxisClusters = 16;
struct AxisCluster {
  float r_k, a_min, a_max, b_min;
  float b_max;
  int32_t first, count, axis;  // entries axis_entry[first, first + count)
};
struct ConstScene {
  int32_t n_tris, n_spheres, n_groups;
  int32_t n_axis_entries[3];  // unclustered entries perpendicular to x, y, z: axis_entry[] holds them in that order
  int32_t n_clusters;         // clustered entries follow in axis_entry[]
  uint32_t bvh_plane_magic;   // 0x3F000000, read by the 4-wide BVH decode so that it sits in a register (bvh.cuh, plane_float)
  int32_t pad;
  AxisCluster cluster[kMaxAxisClusters];
  AxisEntry axis_entry[kMaxConstTris];
  AxisPair axis_pair[kMaxConstTris / 2 + 3];  // the unclustered entries again, paired per axis (x, y, z in order)
  uint8_t axis_tri[kMaxConstTris];  // first triangle (internal index) of each entry
  PlaneGroup group[kMaxConstTris];
  PlaneEntry entry[kMaxConstTris];  // group by group: quads, then lone triangles
  float4 sphere[kMaxConstSpheres];  // centre xyz, radius
};

// Everything a lane indexes on its own (shading data, materials, lights) lives in
// global memory and is read through the L1/read-only path; only data that a whole warp
// reads at the same address sits in __constant__ memory.
struct SceneView {
  const TriShade* tri_shade;      // [n_tris], internal triangle order
  const float4* spheres;          // [n_spheres]
  const int32_t* tri_material;    // [n_tris]
  const int32_t* sphere_material; // [n_spheres]
  const MaterialDev* materials;
  const LightDev* lights;         // sorted by power, cumulative power in v0.w
  const uint8_t* prim_class;      // [n_tris + n_spheres] material kind * 2 + (sphere ? 1 : 0)
  int32_t n_tris, n_spheres, n_lights;
  float total_power;
  // WideBvh only
  const float4* bvh_nodes;
  const TriIsect* tri_isect;
  int32_t bvh_root;  // 0 (node 0), or a leaf reference if the whole mesh is one leaf
};

struct CameraDev {
  LensDev lens;
  SensorDev sensor;
};

// Hit record written by `extend`, one float2 per path slot:
//   t    (NaN on a miss, include/amber/prelude/hit.h:62-66)
//   prim internal primitive index (int bits; -1 on a miss); triangles first, then spheres.
// Barycentrics are not carried: `shade` recomputes them once per hit with the reference's
// exact operation sequence (kernels.cuh, surface_point).
struct HitRecord {
  float t;
  int32_t prim;
};

}  // namespace amber_cuda
