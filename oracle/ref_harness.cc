// ref_harness.cc -- TEST INFRASTRUCTURE (oracle/).  Not part of the product.
//
// extern "C" entry points over the UNMODIFIED reference (compiled from
// /root/reference by oracle/Makefile into oracle/_ref/libamber_ref.so).  Used to
//   (1) pin the CPU restatement oracle/amber_oracle.c,
//   (2) generate the golden fixtures under tests/golden/ (scripts/make_golden.py),
//   (3) serve as the timed CPU baseline (bench.py --impl reference).
// Nothing here is reference source: it only CALLS the reference's public factories
// and interfaces.  The scene arrives as the same flat description the CUDA library
// takes (include/amber_cuda.h) and is rebuilt with the reference's own factories
// (MakeTriangle / MakeSphere / MakeLambertian / ... / MakeThinLens,
// scene::Scene<RGB>::Create<BVH>), so both sides render the identical scene.
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <mutex>
#include <thread>
#include <tuple>
#include <vector>

#include "amber/etude/cornel_box.h"
#include "amber/postprocess/filmic.h"
#include "amber/postprocess/gamma.h"
#include "amber/prelude/hit.h"
#include "amber/prelude/image.h"
#include "amber/prelude/matrix4.h"
#include "amber/prelude/pixel.h"
#include "amber/prelude/ray.h"
#include "amber/prelude/sampling.h"
#include "amber/prelude/vector3.h"
#include "amber/raytracer/acceleration_bvh.h"
#include "amber/raytracer/acceleration_list.h"
#include "amber/rendering/algorithm_bdpt.h"
#include "amber/rendering/algorithm_lt.h"
#include "amber/rendering/algorithm_pt.h"
#include "amber/rendering/bidirectional_path_sampling.h"
#include "amber/rendering/multiple_importance_sampling.h"
#include "amber/rendering/context.h"
#include "amber/rendering/leading.h"
#include "amber/rendering/scatter.h"
#include "amber/rendering/sensor.h"
#include "amber/scene/lens_pinhole.h"
#include "amber/scene/lens_thin.h"
#include "amber/scene/material_diffuse_light.h"
#include "amber/scene/material_eye.h"
#include "amber/scene/material_lambertian.h"
#include "amber/scene/material_phong.h"
#include "amber/scene/material_refraction.h"
#include "amber/scene/material_specular.h"
#include "amber/scene/primitive_sphere.h"
#include "amber/scene/primitive_triangle.h"
#include "amber/scene/scene.h"

#include "../include/amber_cuda.h"
#include "philox_stream.h"

using namespace amber;
using postprocess::RGB;
using scene::AABB;
using scene::Hit;
using scene::Matrix4;
using scene::Ray;
using scene::UnitVector3;
using scene::Vector3;

namespace {

// Context with the semantics of src/amber/cli/context.cc:34-60 (counter + mutex).
struct HarnessContext : rendering::Context {
  std::size_t n_threads, n_iterations, started = 0;
  mutable std::mutex mutex;
  HarnessContext(std::size_t threads, std::size_t iterations)
      : n_threads(threads), n_iterations(iterations) {}
  std::size_t ThreadCount() const noexcept { return n_threads; }
  std::size_t IterationCount() const noexcept {
    std::lock_guard<std::mutex> lock(mutex);
    return started;
  }
  bool Iterate() noexcept {
    std::lock_guard<std::mutex> lock(mutex);
    if (started >= n_iterations) return false;
    started++;
    return true;
  }
};

// Sampler::operator() is virtual (include/amber/prelude/sampling.h:35-41): feed the
// reference code from the device's Philox stream, or from a scripted list.
struct StreamSampler : prelude::Sampler {
  philox_stream stream;
  bool camera = true;  // camera vertex: sequential slots; surface vertex: BSDF-sampling slots
  result_type operator()() {
    return static_cast<result_type>(camera ? philox_stream_next_camera(&stream)
                                           : philox_stream_next_bsdf(&stream));
  }
  // The Russian-roulette uniform has a fixed slot of its own (philox_stream.h).
  rendering::real_type RussianRoulette() { return philox_stream_rr(&stream); }
};

struct ScriptedSampler : prelude::Sampler {
  const float* xi = nullptr;
  int n = 0, used = 0;
  result_type operator()() {
    // past the scripted list: a Weyl sequence, identical in the oracle and the kernels
    const float v = used < n ? xi[used]
                             : static_cast<float>((static_cast<uint32_t>(used) * 2654435769u) >> 8) * (1.0f / 16777216.0f);
    used++;
    return static_cast<result_type>(v);
  }
};

// Id-carrying object for the reference's acceleration templates
// (include/amber/raytracer/acceleration_bvh.h:38-49 takes any type with
// Center/BoundingBox/Intersect).
struct IdObject {
  const scene::Primitive* primitive;
  int id;
  const Vector3 Center() const noexcept { return primitive->Center(); }
  const AABB BoundingBox() const noexcept { return primitive->BoundingBox(); }
  Hit Intersect(const Ray& ray) const noexcept { return primitive->Intersect(ray); }
};
using IdAcceleration = raytracer::Acceleration<scene::real_type, IdObject>;

// Counts Cast calls (virtual, include/amber/scene/scene.h:60-61).
std::atomic<unsigned long long> g_casts{0};
struct CountingScene : scene::Scene<RGB> {
  using Base = scene::Scene<RGB>;
  explicit CountingScene(Base&& base) : Base(std::move(base)) {}
  std::tuple<rendering::ObjectPointer, scene::Hit> Cast(const scene::Ray& ray) const {
    thread_local unsigned long long local = 0;
    if (++local == 4096) {
      g_casts.fetch_add(local, std::memory_order_relaxed);
      local = 0;
    }
    return Base::Cast(ray);
  }
};

std::unique_ptr<scene::Material<RGB>> MakeMaterial(const amber_cuda_material& m) {
  const RGB rgb(m.rgb[0], m.rgb[1], m.rgb[2]);
  switch (m.kind) {
    case AMBER_CUDA_MAT_LAMBERTIAN: return scene::MakeLambertian(rgb);
    case AMBER_CUDA_MAT_PHONG: return scene::MakePhong(rgb, m.param);
    case AMBER_CUDA_MAT_SPECULAR: return scene::MakeSpecular(rgb);
    case AMBER_CUDA_MAT_REFRACTION: return scene::MakeRefraction<RGB>(m.param);
    case AMBER_CUDA_MAT_DIFFUSE_LIGHT: return scene::MakeDiffuseLight(rgb);
    case AMBER_CUDA_MAT_EYE: return scene::MakeEye<RGB>();
  }
  return nullptr;
}

std::unique_ptr<scene::Primitive> MakePrimitive(const amber_cuda_primitive& p) {
  const float* d = p.data;
  if (p.kind == AMBER_CUDA_PRIM_TRIANGLE) {
    return scene::MakeTriangle(Vector3(d[0], d[1], d[2]), Vector3(d[3], d[4], d[5]),
                               Vector3(d[6], d[7], d[8]));
  }
  return scene::MakeSphere(Vector3(d[0], d[1], d[2]), d[3]);
}

struct RefScene {
  // for ref_cast: every flat primitive (aperture triangles included), ids = flat index
  std::vector<std::unique_ptr<scene::Primitive>> id_primitives;
  std::unique_ptr<raytracer::BVH<scene::real_type, IdObject>> id_bvh;
  std::unique_ptr<raytracer::List<scene::real_type, IdObject>> id_list;
  // the reference scene proper
  std::unique_ptr<CountingScene> scene;
};

Matrix4 LensTransform(const amber_cuda_lens& lens) {
  const float* t = lens.transform;
  return Matrix4(t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7], t[8], t[9], t[10], t[11], t[12],
                 t[13], t[14], t[15]);
}

rendering::Sensor MakeSensor(uint32_t w, uint32_t h, float film_w, float film_h) {
  return rendering::Sensor(w, h, film_w, film_h);
}

void ImageToFloats(const prelude::Image<RGB>& image, uint32_t w, uint32_t h, float* rgb) {
  for (uint32_t y = 0; y < h; y++) {
    for (uint32_t x = 0; x < w; x++) {
      const auto& value = image[prelude::Pixel(x, y)];
      float* out = rgb + (static_cast<size_t>(x) + static_cast<size_t>(y) * w) * 3;
      out[0] = value.X();
      out[1] = value.Y();
      out[2] = value.Z();
    }
  }
}

}  // namespace

extern "C" {

// Build the reference scene from the flat description.  Primitives whose material is
// EYE are the lens aperture: the reference takes those from lens->ApertureObjects()
// (src/amber/etude/cornel_box.cc:62-64), so they are not re-created for the scene
// proper -- only for the id-carrying cast structures.
void* ref_scene_create(const amber_cuda_scene_desc* desc) {
  auto out = std::make_unique<RefScene>();

  std::vector<IdObject> ids_a, ids_b;
  for (uint32_t i = 0; i < desc->n_primitives; i++) {
    out->id_primitives.emplace_back(MakePrimitive(desc->primitives[i]));
    ids_a.push_back({out->id_primitives.back().get(), static_cast<int>(i)});
    ids_b.push_back({out->id_primitives.back().get(), static_cast<int>(i)});
  }
  out->id_bvh = std::make_unique<raytracer::BVH<scene::real_type, IdObject>>(std::move(ids_a));
  out->id_list = std::make_unique<raytracer::List<scene::real_type, IdObject>>(std::move(ids_b));

  std::vector<std::unique_ptr<scene::Primitive>> primitives;
  std::vector<std::unique_ptr<scene::Material<RGB>>> materials;
  std::vector<scene::Object<RGB>> objects;
  auto lens = desc->lens.kind == AMBER_CUDA_LENS_PINHOLE
                  ? scene::MakePinholeLens<RGB>(LensTransform(desc->lens), desc->lens.sensor_distance)
                  : scene::MakeThinLens<RGB>(LensTransform(desc->lens), desc->lens.focal_length,
                                             desc->lens.focus_distance, desc->lens.radius,
                                             static_cast<std::size_t>(desc->lens.n_blades));
  bool has_aperture = false;
  for (uint32_t i = 0; i < desc->n_primitives; i++) {
    if (desc->materials[desc->primitives[i].material].kind == AMBER_CUDA_MAT_EYE) {
      has_aperture = true;
    }
  }
  // The pinhole's aperture object is a degenerate triangle (lens_pinhole.cc:38) that no ray can
  // hit (its Intersect returns t = NaN); it is in the reference scene as a user of the reference
  // would put it there, and absent from the flat description.
  if (desc->lens.kind == AMBER_CUDA_LENS_PINHOLE) has_aperture = true;
  if (has_aperture) {
    for (const auto& object : lens->ApertureObjects()) objects.emplace_back(*object);
  }
  for (uint32_t i = 0; i < desc->n_materials; i++) {
    materials.emplace_back(MakeMaterial(desc->materials[i]));
  }
  for (uint32_t i = 0; i < desc->n_primitives; i++) {
    const auto& p = desc->primitives[i];
    if (desc->materials[p.material].kind == AMBER_CUDA_MAT_EYE) continue;
    primitives.emplace_back(MakePrimitive(p));
    objects.emplace_back(primitives.back().get(), materials[p.material].get());
  }
  out->scene = std::make_unique<CountingScene>(
      scene::Scene<RGB>::Create<raytracer::BVH<scene::real_type, scene::Object<RGB>>>(
          std::move(primitives), std::move(materials), std::move(objects), std::move(lens)));
  return out.release();
}

// The reference's own built-in scene, exactly as src/amber/cli/application.cc:69-73
// constructs it.  (No id structures: ref_cast is not available on it.)
void* ref_scene_create_cornel_box(void) {
  auto out = std::make_unique<RefScene>();
  out->scene = std::make_unique<CountingScene>(etude::CornelBox(0.050, 0.050, 6));
  return out.release();
}

void ref_scene_destroy(void* handle) { delete static_cast<RefScene*>(handle); }

// Closest hit through the reference's BVH (accel = 0) or List (accel = 1).
int ref_cast(void* handle, int accel, uint64_t n, const float* origins, const float* directions,
             int32_t* prim, float* t, int threads) {
  auto* ref = static_cast<RefScene*>(handle);
  if (!ref->id_bvh) return -1;
  const IdAcceleration& acceleration =
      accel == 0 ? static_cast<const IdAcceleration&>(*ref->id_bvh)
                 : static_cast<const IdAcceleration&>(*ref->id_list);
  if (threads < 1) threads = 1;
  std::vector<std::thread> pool;
  for (int worker = 0; worker < threads; worker++) {
    pool.emplace_back([&, worker]() {
      for (uint64_t i = worker; i < n; i += threads) {
        const Ray ray(Vector3(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]),
                      UnitVector3(directions[3 * i], directions[3 * i + 1], directions[3 * i + 2]));
        Hit hit;
        const IdObject* object;
        std::tie(hit, object) = acceleration.Cast(ray);
        prim[i] = object ? object->id : -1;
        t[i] = hit.Distance();
      }
    });
  }
  for (auto& thread : pool) thread.join();
  return 0;
}

// Material::SampleLight with scripted uniforms (8 per sample).
int ref_sample_light(const amber_cuda_material* material, uint64_t n, const float* normal,
                     const float* dir_out, const float* xi8, float* dir_in, float* weight,
                     int32_t* draws) {
  const auto m = MakeMaterial(*material);
  if (!m) return -1;
  for (uint64_t i = 0; i < n; i++) {
    ScriptedSampler sampler;
    sampler.xi = xi8 + 8 * i;
    sampler.n = 8;
    const UnitVector3 nn(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
    const UnitVector3 wo(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]);
    const auto scatter = m->SampleLight(nn, wo, sampler);
    dir_in[3 * i] = scatter.DirectionIn().X();
    dir_in[3 * i + 1] = scatter.DirectionIn().Y();
    dir_in[3 * i + 2] = scatter.DirectionIn().Z();
    weight[3 * i] = scatter.Weight().X();
    weight[3 * i + 1] = scatter.Weight().Y();
    weight[3 * i + 2] = scatter.Weight().Z();
    draws[i] = sampler.used;
  }
  return 0;
}

// Material::BSDF and Material::Radiance (for the NEE connection term).
int ref_bsdf(const amber_cuda_material* material, uint64_t n, const float* normal,
             const float* dir_out, const float* dir_in, float* bsdf, float* radiance) {
  const auto m = MakeMaterial(*material);
  if (!m) return -1;
  for (uint64_t i = 0; i < n; i++) {
    const UnitVector3 nn(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
    const UnitVector3 wo(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]);
    const UnitVector3 wi(dir_in[3 * i], dir_in[3 * i + 1], dir_in[3 * i + 2]);
    const auto f = m->BSDF(nn, wo, wi);
    const auto le = m->Radiance(nn, wo);
    bsdf[3 * i] = f.X(); bsdf[3 * i + 1] = f.Y(); bsdf[3 * i + 2] = f.Z();
    radiance[3 * i] = le.X(); radiance[3 * i + 1] = le.Y(); radiance[3 * i + 2] = le.Z();
  }
  return 0;
}

// Triangle::SampleSurfacePoint / SurfaceArea (light point sampling for NEE).
int ref_triangle_sample(const float* v9, uint64_t n, const float* xi2, float* point, float* normal,
                        float* area) {
  const auto tri = scene::MakeTriangle(Vector3(v9[0], v9[1], v9[2]), Vector3(v9[3], v9[4], v9[5]),
                                       Vector3(v9[6], v9[7], v9[8]));
  *area = tri->SurfaceArea();
  for (uint64_t i = 0; i < n; i++) {
    ScriptedSampler sampler;
    sampler.xi = xi2 + 2 * i;
    sampler.n = 2;
    const auto ray = tri->SampleSurfacePoint(sampler);
    point[3 * i] = ray.Origin().X(); point[3 * i + 1] = ray.Origin().Y(); point[3 * i + 2] = ray.Origin().Z();
    normal[3 * i] = ray.Direction().X(); normal[3 * i + 1] = ray.Direction().Y(); normal[3 * i + 2] = ray.Direction().Z();
  }
  return 0;
}

// scene.GenerateEyeRay on a pixel-bound sensor with scripted uniforms (5 per ray).
int ref_eye_rays(void* handle, uint32_t w, uint32_t h, float film_w, float film_h, uint64_t n,
                 const uint32_t* px, const uint32_t* py, const float* xi5, float* origin,
                 float* direction, float* weight) {
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  for (uint64_t i = 0; i < n; i++) {
    ScriptedSampler sampler;
    sampler.xi = xi5 + 5 * i;
    sampler.n = 5;
    const auto generated =
        ref->scene->GenerateEyeRay(sensor.Bind(prelude::Pixel(px[i], py[i])), sampler);
    const auto& leading = std::get<rendering::Leading<RGB>>(generated);
    const auto ray = leading.Ray();
    origin[3 * i] = ray.Origin().X(); origin[3 * i + 1] = ray.Origin().Y(); origin[3 * i + 2] = ray.Origin().Z();
    direction[3 * i] = ray.Direction().X(); direction[3 * i + 1] = ray.Direction().Y(); direction[3 * i + 2] = ray.Direction().Z();
    weight[i] = leading.Weight().X();
  }
  return 0;
}

// Unit checks for light tracing (row N2): Material::SampleImportance, Scene::Response
// (lens + sensor), Scene::GenerateLightRay -- each with scripted uniforms.
int ref_sample_importance(const amber_cuda_material* material, uint64_t n, const float* normal,
                          const float* dir_out, const float* xi8, float* dir_in, float* weight,
                          int32_t* draws) {
  const auto m = MakeMaterial(*material);
  if (!m) return -1;
  for (uint64_t i = 0; i < n; i++) {
    ScriptedSampler sampler;
    sampler.xi = xi8 + 8 * i;
    sampler.n = 8;
    const UnitVector3 nn(normal[3 * i], normal[3 * i + 1], normal[3 * i + 2]);
    const UnitVector3 wo(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]);
    const auto scatter = m->SampleImportance(nn, wo, sampler);
    dir_in[3 * i] = scatter.DirectionIn().X();
    dir_in[3 * i + 1] = scatter.DirectionIn().Y();
    dir_in[3 * i + 2] = scatter.DirectionIn().Z();
    weight[3 * i] = scatter.Weight().X();
    weight[3 * i + 1] = scatter.Weight().Y();
    weight[3 * i + 2] = scatter.Weight().Z();
    if (draws) draws[i] = sampler.used;
  }
  return 0;
}

int ref_lens_response(void* handle, uint32_t w, uint32_t h, float film_w, float film_h, uint64_t n,
                      const float* position, const float* dir_out, int32_t* pixel, float* value) {
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  const rendering::Scene<RGB>& scene = *ref->scene;
  for (uint64_t i = 0; i < n; i++) {
    const Vector3 p(position[3 * i], position[3 * i + 1], position[3 * i + 2]);
    const UnitVector3 d(dir_out[3 * i], dir_out[3 * i + 1], dir_out[3 * i + 2]);
    const auto response = scene.Response(sensor, p, d);
    if (response.Pixel()) {
      pixel[i] = static_cast<int32_t>(response.Pixel().X() + response.Pixel().Y() * w);
      value[i] = response.Value().X();
    } else {
      pixel[i] = -1;
      value[i] = 0;
    }
  }
  return 0;
}

int ref_light_rays(void* handle, uint64_t n, const float* xi5, float* origin, float* direction,
                   float* weight) {
  auto* ref = static_cast<RefScene*>(handle);
  const rendering::Scene<RGB>& scene = *ref->scene;
  for (uint64_t i = 0; i < n; i++) {
    ScriptedSampler sampler;
    sampler.xi = xi5 + 5 * i;
    sampler.n = 5;
    const auto generated = scene.GenerateLightRay(sampler);
    const auto& leading = std::get<rendering::Leading<RGB>>(generated);
    const auto ray = leading.Ray();
    origin[3 * i] = ray.Origin().X(); origin[3 * i + 1] = ray.Origin().Y(); origin[3 * i + 2] = ray.Origin().Z();
    direction[3 * i] = ray.Direction().X(); direction[3 * i + 1] = ray.Direction().Y(); direction[3 * i + 2] = ray.Direction().Z();
    weight[3 * i] = leading.Weight().X(); weight[3 * i + 1] = leading.Weight().Y(); weight[3 * i + 2] = leading.Weight().Z();
  }
  return 0;
}

// The reference's own path tracer (algo = 0, MakeRGBPathTracing), bidirectional path
// tracer (algo = 1) or light tracer (algo = 2, MakeRGBLightTracing), unmodified, `threads`
// std::threads, `spp` iterations.  Engines are seeded from std::random_device inside the reference => not reproducible.
int ref_render(void* handle, int algo, uint32_t w, uint32_t h, float film_w, float film_h,
               uint32_t spp, uint32_t threads, float* rgb, uint64_t* casts, double* seconds) {
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  HarnessContext context(threads, spp);
  auto algorithm = algo == 0   ? rendering::MakeRGBPathTracing()
                   : algo == 1 ? rendering::MakeRGBBidirectionalPathTracing()
                               : rendering::MakeRGBLightTracing();
  g_casts = 0;
  const auto begin = std::chrono::steady_clock::now();
  const auto image = algorithm->Render(*ref->scene, sensor, context);
  const auto end = std::chrono::steady_clock::now();
  if (seconds) *seconds = std::chrono::duration<double>(end - begin).count();
  if (casts) *casts = g_casts.load();  // granularity 4096 per thread
  if (rgb) ImageToFloats(image, w, h, rgb);
  return 0;
}

// Reproducible reference render: the loop of src/amber/rendering/algorithm_pt.cc:125-160
// restated over the reference's own Scene API, every uniform drawn from the device's
// Philox stream for (pixel, sample).  max_depth caps the number of rays per path (the
// reference has no cap).  stats: [0] rays cast, [1] truncated paths.
int ref_render_stream(void* handle, uint32_t w, uint32_t h, float film_w, float film_h,
                      uint64_t sample_begin, uint32_t sample_count, uint64_t seed,
                      uint32_t max_depth, uint32_t threads, float* rgb, uint64_t* stats) {
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  const rendering::Scene<RGB>& scene = *ref->scene;
  if (max_depth == 0) max_depth = 1u << 16;
  if (threads < 1) threads = 1;
  std::atomic<uint64_t> total_rays{0}, total_truncated{0};
  std::vector<std::thread> pool;
  for (uint32_t worker = 0; worker < threads; worker++) {
    pool.emplace_back([&, worker]() {
      StreamSampler sampler;
      uint64_t rays = 0, truncated = 0;
      for (uint32_t y = worker; y < h; y += threads) {
        for (uint32_t x = 0; x < w; x++) {
          double sum[3] = {0, 0, 0};
          for (uint64_t s = sample_begin; s < sample_begin + sample_count; s++) {
            philox_stream_reset(&sampler.stream, seed, x + y * w, s);
            sampler.camera = true;
            auto generated = scene.GenerateEyeRay(sensor.Bind(prelude::Pixel(x, y)), sampler);
            auto& object = std::get<rendering::ObjectPointer>(generated);
            auto& leading = std::get<rendering::Leading<RGB>>(generated);
            RGB measurement(0);
            auto ray = leading.Ray();
            auto weight = leading.Weight();
            for (uint32_t depth = 1;; depth++) {
              rendering::Hit hit;
              philox_stream_begin_vertex(&sampler.stream, depth);
              sampler.camera = false;
              std::tie(object, hit) = scene.Cast(ray);
              rays++;
              if (!hit) break;
              measurement += weight * scene.Radiance(object, hit.Normal(), -ray.Direction());
              const auto scatter =
                  scene.SampleLight(object, hit.Normal(), -ray.Direction(), sampler);
              const auto p = std::min<rendering::real_type>(kRussianRoulette, Max(scatter.Weight()));
              if (sampler.RussianRoulette() >= p) break;
              if (depth >= max_depth) {
                truncated++;
                break;
              }
              ray = rendering::Ray(hit.Position(), scatter.DirectionIn());
              weight *= scatter.Weight() / p;
            }
            sum[0] += measurement.X();
            sum[1] += measurement.Y();
            sum[2] += measurement.Z();
          }
          float* out = rgb + (static_cast<size_t>(x) + static_cast<size_t>(y) * w) * 3;
          for (int c = 0; c < 3; c++) out[c] = static_cast<float>(sum[c] / sample_count);
        }
      }
      total_rays += rays;
      total_truncated += truncated;
    });
  }
  for (auto& thread : pool) thread.join();
  if (stats) {
    stats[0] = total_rays;
    stats[1] = total_truncated;
  }
  return 0;
}

// Reproducible reference LIGHT TRACING: the loop of src/amber/rendering/algorithm_lt.cc:124-163
// restated over the reference's own Scene API.  One iteration = w*h light paths
// (algorithm_lt.cc:112-122); path i of sample s draws every uniform from the device's Philox
// stream keyed (i, s): vertex 0 is the light (choice, point u, v, direction: five sequential
// slots, as the camera's), vertex k >= 1 the k-th surface interaction.  rgb = mean over the
// samples, as ParallelMean returns it.  stats: [0] rays cast, [1] truncated paths, [2] splats.
int ref_render_lt_stream(void* handle, uint32_t w, uint32_t h, float film_w, float film_h,
                         uint64_t sample_begin, uint32_t sample_count, uint64_t seed,
                         uint32_t max_depth, uint32_t threads, float* rgb, uint64_t* stats) {
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  const rendering::Scene<RGB>& scene = *ref->scene;
  if (max_depth == 0) max_depth = 1u << 16;
  if (threads < 1) threads = 1;
  const uint64_t n_pixels = static_cast<uint64_t>(w) * h;
  std::atomic<uint64_t> total_rays{0}, total_truncated{0}, total_splats{0};
  std::vector<std::vector<double>> films(threads, std::vector<double>(n_pixels * 3, 0.0));
  std::vector<std::thread> pool;
  for (uint32_t worker = 0; worker < threads; worker++) {
    pool.emplace_back([&, worker]() {
      StreamSampler sampler;
      std::vector<double>& film = films[worker];
      uint64_t rays = 0, truncated = 0, splats = 0;
      for (uint64_t s = sample_begin; s < sample_begin + sample_count; s++) {
        for (uint64_t i = worker; i < n_pixels; i += threads) {
          philox_stream_reset(&sampler.stream, seed, static_cast<uint32_t>(i), s);
          sampler.camera = true;
          auto generated = scene.GenerateLightRay(sampler);
          auto& object = std::get<rendering::ObjectPointer>(generated);
          auto& leading = std::get<rendering::Leading<RGB>>(generated);
          auto ray = leading.Ray();
          auto weight = leading.Weight();
          for (uint32_t depth = 1;; depth++) {
            rendering::Hit hit;
            philox_stream_begin_vertex(&sampler.stream, depth);
            sampler.camera = false;
            std::tie(object, hit) = scene.Cast(ray);
            rays++;
            if (!hit) break;
            if (scene.Surface(object) == rendering::SurfaceType::Eye) {
              const auto response = scene.Response(sensor, hit.Position(), -ray.Direction());
              if (response.Pixel()) {
                const RGB value = weight * response.Value() / n_pixels;  // image.Size()
                const auto pixel = response.Pixel();
                double* out = film.data() + (static_cast<size_t>(pixel.X()) + static_cast<size_t>(pixel.Y()) * w) * 3;
                out[0] += value.X();
                out[1] += value.Y();
                out[2] += value.Z();
                splats++;
              }
            }
            const auto scatter =
                scene.SampleImportance(object, hit.Normal(), -ray.Direction(), sampler);
            const auto p = std::min<rendering::real_type>(kRussianRoulette, Max(scatter.Weight()));
            if (sampler.RussianRoulette() >= p) break;
            if (depth >= max_depth) {
              truncated++;
              break;
            }
            ray = rendering::Ray(hit.Position(), scatter.DirectionIn());
            weight *= scatter.Weight() / p;
          }
        }
      }
      total_rays += rays;
      total_truncated += truncated;
      total_splats += splats;
    });
  }
  for (auto& thread : pool) thread.join();
  for (uint64_t k = 0; k < n_pixels * 3; k++) {
    double sum = 0;
    for (uint32_t worker = 0; worker < threads; worker++) sum += films[worker][k];
    rgb[k] = static_cast<float>(sum / sample_count);
  }
  if (stats) {
    stats[0] = total_rays;
    stats[1] = total_truncated;
    stats[2] = total_splats;
  }
  return 0;
}

// Reproducible reference BIDIRECTIONAL PATH TRACING: BidirectionalPathTracing::Thread::operator()
// and ::Render (src/amber/rendering/algorithm_bdpt.cc:116-148) restated -- a dozen lines -- over the
// reference's OWN GenerateLightSubpath / GenerateEyeSubpath / Combine
// (include/amber/rendering/bidirectional_path_sampling.h:726-1068, called unmodified).  Each
// subpath draws from its own sequential Philox stream (philox_stream.h).  max_vertices caps
// the events per subpath (the generators' own s / t arguments; 0 = no cap, as the reference).
// stats: [0] light-subpath events, [1] eye-subpath events, [2] light-image splats.
int ref_render_bdpt_stream(void* handle, uint32_t w, uint32_t h, float film_w, float film_h,
                           uint64_t sample_begin, uint32_t sample_count, uint64_t seed,
                           uint32_t max_vertices, uint32_t threads, float* rgb, uint64_t* stats) {
  using rendering::path_size_type;
  auto* ref = static_cast<RefScene*>(handle);
  const auto sensor = MakeSensor(w, h, film_w, film_h);
  const rendering::Scene<RGB>& scene = *ref->scene;
  const path_size_type cap = max_vertices ? max_vertices : std::numeric_limits<path_size_type>::max();
  if (threads < 1) threads = 1;
  const uint64_t n_pixels = static_cast<uint64_t>(w) * h;
  std::atomic<uint64_t> total_light{0}, total_eye{0}, total_splats{0};
  std::vector<std::vector<double>> films(threads, std::vector<double>(n_pixels * 3, 0.0));
  std::vector<std::thread> pool;
  for (uint32_t worker = 0; worker < threads; worker++) {
    pool.emplace_back([&, worker]() {
      StreamSampler sampler;
      sampler.camera = true;  // sequential slots
      rendering::Subpath<RGB> light_path, eye_path;
      rendering::BidirectionalPathSamplingBuffer<RGB> buffer;
      prelude::SparseImage<RGB> light_image;
      std::vector<double>& film = films[worker];
      uint64_t n_light = 0, n_eye = 0, n_splats = 0;
      for (uint64_t s = sample_begin; s < sample_begin + sample_count; s++) {
        for (uint64_t i = worker; i < n_pixels; i += threads) {
          const prelude::Pixel pixel(i % w, i / w);
          const auto bound = sensor.Bind(pixel);
          light_image.Clear();
          philox_stream_reset(&sampler.stream, seed, static_cast<uint32_t>(i), s);
          philox_stream_begin_sequence(&sampler.stream, PHILOX_SUBPATH_LIGHT);
          rendering::GenerateLightSubpath(scene, sampler, light_path, cap);
          philox_stream_begin_sequence(&sampler.stream, PHILOX_SUBPATH_EYE);
          rendering::GenerateEyeSubpath(scene, bound, sampler, eye_path, cap);
          const RGB measurement = rendering::Combine(scene, bound, light_path, eye_path,
                                                     rendering::BalanceHeuristic(), buffer, light_image);
          n_light += light_path.Size();
          n_eye += eye_path.Size();
          double* out = film.data() + i * 3;
          out[0] += measurement.X();
          out[1] += measurement.Y();
          out[2] += measurement.Z();
          for (const auto& splat : light_image) {
            const auto p = splat.Pixel();
            double* to = film.data() + (static_cast<size_t>(p.X()) + static_cast<size_t>(p.Y()) * w) * 3;
            to[0] += splat.Value().X();
            to[1] += splat.Value().Y();
            to[2] += splat.Value().Z();
            n_splats++;
          }
        }
      }
      total_light += n_light;
      total_eye += n_eye;
      total_splats += n_splats;
    });
  }
  for (auto& thread : pool) thread.join();
  for (uint64_t k = 0; k < n_pixels * 3; k++) {
    double sum = 0;
    for (uint32_t worker = 0; worker < threads; worker++) sum += films[worker][k];
    rgb[k] = static_cast<float>(sum / sample_count);
  }
  if (stats) {
    stats[0] = total_light;
    stats[1] = total_eye;
    stats[2] = total_splats;
  }
  return 0;
}


// The reference's post-process as src/amber/cli/application.cc:98-106 applies it:
// Gamma()(Filmic()(image)).  rgb: w*h*3 floats (x + y*w), ldr: the same layout, 8 bit.
int ref_tonemap(uint32_t w, uint32_t h, const float* rgb, uint8_t* ldr) {
  postprocess::HDRImage image(w, h);
  for (uint32_t y = 0; y < h; y++) {
    for (uint32_t x = 0; x < w; x++) {
      const float* in = rgb + (static_cast<size_t>(x) + static_cast<size_t>(y) * w) * 3;
      image[prelude::Pixel(x, y)] = postprocess::HDR(in[0], in[1], in[2]);
    }
  }
  const auto out = postprocess::Gamma()(postprocess::Filmic()(image));
  for (uint32_t y = 0; y < h; y++) {
    for (uint32_t x = 0; x < w; x++) {
      const auto& value = out[prelude::Pixel(x, y)];
      uint8_t* o = ldr + (static_cast<size_t>(x) + static_cast<size_t>(y) * w) * 3;
      o[0] = value.X();
      o[1] = value.Y();
      o[2] = value.Z();
    }
  }
  return 0;
}

}  // extern "C"
