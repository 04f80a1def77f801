// adapter.hpp -- C++ host side of the drop-in: amber's own scene / algorithm API over the
// C ABI of include/amber_cuda.h.
//
// Compiled against the REFERENCE's headers (-I<amber>/include); nothing here restates
// reference arithmetic.  It provides, in amber's namespaces and naming style,
//
//   amber::cuda::FlatSceneBuilder        the reference's factory verbs (MakeTriangle, MakeSphere,
//                                        MakeLambertian, MakePhong, MakeSpecular, MakeRefraction,
//                                        MakeDiffuseLight, MakeThinLens, MakePinholeLens); builds the real reference
//                                        scene (scene::Scene<RGB>::Create<BVH>,
//                                        include/amber/scene/scene.h:169-193) AND records the flat
//                                        description the GPU needs, in creation order
//   amber::cuda::Scene                   : scene::Scene<RGB>, FlatSceneProvider -- usable by every
//                                        CPU algorithm of the reference unchanged
//   amber::cuda::MakeRGBCudaPathTracing  std::unique_ptr<rendering::Algorithm<RGB>>, the factory
//                                        convention of include/amber/rendering/algorithm_pt.h:30-31
//   amber::cuda::MakeRGBCudaLightTracing the same for the light tracer (algorithm_lt.h:29-30)
//   amber::cuda::MakeRGBCudaBidirectionalPathTracing   and for BDPT (algorithm_bdpt.h)
//
// Why a builder: scene::Object<R> keeps its primitive and material private
// (include/amber/scene/object.h:93-96) and the concrete classes are .cc-local, so geometry
// cannot be read back out of an arbitrary rendering::Scene.  Render() on a scene that is not
// a FlatSceneProvider throws std::invalid_argument -- there is no CPU fallback.
//
// Error convention: the reference uses exceptions only (include/amber/prelude/image.h:73-75);
// every non-zero C-ABI return is rethrown as std::runtime_error(amber_cuda_last_error()).
#pragma once

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>
#include <tuple>
#include <utility>
#include <vector>

#include "amber/prelude/matrix4.h"
#include "amber/raytracer/acceleration_bvh.h"
#include "amber/rendering/algorithm.h"
#include "amber/rendering/context.h"
#include "amber/rendering/sensor.h"
#include "amber/scene/lens_pinhole.h"
#include "amber/scene/lens_thin.h"
#include "amber/scene/material_diffuse_light.h"
#include "amber/scene/material_lambertian.h"
#include "amber/scene/material_phong.h"
#include "amber/scene/material_refraction.h"
#include "amber/scene/material_specular.h"
#include "amber/scene/primitive_sphere.h"
#include "amber/scene/primitive_triangle.h"
#include "amber/scene/scene.h"

#include "../amber_cuda.h"

namespace amber {
namespace cuda {

using rendering::RGB;
using scene::Matrix4;
using scene::real_type;
using scene::Vector3;

/** Anything that can hand the GPU path its flat scene description. */
class FlatSceneProvider
{
public:
  virtual ~FlatSceneProvider() {}
  virtual const amber_cuda_scene_desc& FlatDescription() const noexcept = 0;
};

/** The reference scene plus its flat description. */
class Scene
: public scene::Scene<RGB>
, public FlatSceneProvider
{
public:
  Scene(
    scene::Scene<RGB>&& base,
    std::vector<amber_cuda_primitive>&& primitives,
    std::vector<amber_cuda_material>&& materials,
    const amber_cuda_lens& lens
  )
  : scene::Scene<RGB>(std::move(base))
  , primitives_(std::move(primitives))
  , materials_(std::move(materials))
  {
    desc_.n_primitives = static_cast<uint32_t>(primitives_.size());
    desc_.primitives = primitives_.data();
    desc_.n_materials = static_cast<uint32_t>(materials_.size());
    desc_.materials = materials_.data();
    desc_.lens = lens;
  }

  const amber_cuda_scene_desc& FlatDescription() const noexcept { return desc_; }

private:
  std::vector<amber_cuda_primitive> primitives_;
  std::vector<amber_cuda_material> materials_;
  amber_cuda_scene_desc desc_;
};

/** A sensor that remembers its film size exactly (rendering::Sensor only exposes the area). */
class Sensor
: public rendering::Sensor
{
public:
  Sensor(
    rendering::pixel_size_type pixel_width,
    rendering::pixel_size_type pixel_height,
    real_type scene_width,
    real_type scene_height
  ) noexcept
  : rendering::Sensor(pixel_width, pixel_height, scene_width, scene_height)
  , scene_width_(scene_width)
  , scene_height_(scene_height)
  {}

  real_type SceneWidth() const noexcept { return scene_width_; }
  real_type SceneHeight() const noexcept { return scene_height_; }

private:
  real_type scene_width_, scene_height_;
};

/** Scene construction with the reference's factory verbs. */
class FlatSceneBuilder
{
public:
  using MaterialId = int32_t;

  MaterialId MakeLambertian(const RGB& kd)
  { return AddMaterial(scene::MakeLambertian(kd), AMBER_CUDA_MAT_LAMBERTIAN, kd, 0); }
  MaterialId MakePhong(const RGB& ks, real_type exponent)
  { return AddMaterial(scene::MakePhong(ks, exponent), AMBER_CUDA_MAT_PHONG, ks, exponent); }
  MaterialId MakeSpecular(const RGB& ks)
  { return AddMaterial(scene::MakeSpecular(ks), AMBER_CUDA_MAT_SPECULAR, ks, 0); }
  MaterialId MakeRefraction(real_type ior)
  { return AddMaterial(scene::MakeRefraction<RGB>(ior), AMBER_CUDA_MAT_REFRACTION, RGB(1), ior); }
  MaterialId MakeDiffuseLight(const RGB& radiance)
  { return AddMaterial(scene::MakeDiffuseLight(radiance), AMBER_CUDA_MAT_DIFFUSE_LIGHT, radiance, 0); }

  void MakeTriangle(const Vector3& v0, const Vector3& v1, const Vector3& v2, MaterialId material)
  {
    CheckMaterial(material);
    primitives_.emplace_back(scene::MakeTriangle(v0, v1, v2));
    objects_.emplace_back(primitives_.back().get(), materials_[material].get());
    amber_cuda_primitive p;
    p.kind = AMBER_CUDA_PRIM_TRIANGLE;
    p.material = material;
    const Vector3 v[3] = {v0, v1, v2};
    for (int k = 0; k < 3; k++) {
      p.data[3 * k + 0] = v[k].X();
      p.data[3 * k + 1] = v[k].Y();
      p.data[3 * k + 2] = v[k].Z();
    }
    flat_primitives_.push_back(p);
  }

  void MakeSphere(const Vector3& center, real_type radius, MaterialId material)
  {
    CheckMaterial(material);
    primitives_.emplace_back(scene::MakeSphere(center, radius));
    objects_.emplace_back(primitives_.back().get(), materials_[material].get());
    amber_cuda_primitive p;
    std::memset(&p, 0, sizeof p);
    p.kind = AMBER_CUDA_PRIM_SPHERE;
    p.material = material;
    p.data[0] = center.X();
    p.data[1] = center.Y();
    p.data[2] = center.Z();
    p.data[3] = radius;
    flat_primitives_.push_back(p);
  }

  /** MakeThinLens (include/amber/scene/lens_thin.h:30-38).  transform: row-major 4x4 in the
   *  argument order of the Matrix4 constructor.  The aperture blades become scene objects, as
   *  src/amber/etude/cornel_box.cc:62-64 does. */
  void MakeThinLens(
    const real_type (&transform)[16],
    real_type focal_length,
    real_type focus_distance,
    real_type radius,
    std::size_t n_blades
  )
  {
    const real_type* t = transform;
    lens_ = scene::MakeThinLens<RGB>(
      Matrix4(t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7],
              t[8], t[9], t[10], t[11], t[12], t[13], t[14], t[15]),
      focal_length, focus_distance, radius, n_blades);
    flat_lens_ = amber_cuda_lens();
    std::memcpy(flat_lens_.transform, transform, sizeof flat_lens_.transform);
    flat_lens_.kind = AMBER_CUDA_LENS_THIN;
    flat_lens_.focal_length = focal_length;
    flat_lens_.focus_distance = focus_distance;
    flat_lens_.radius = radius;
    flat_lens_.n_blades = static_cast<int32_t>(n_blades);
  }

  /** MakePinholeLens (include/amber/scene/lens_pinhole.h:32-37).  Its aperture object -- a
   *  degenerate triangle at the lens origin (src/amber/scene/lens_pinhole.cc:38) that no ray can
   *  hit -- goes into the reference scene like any lens's, and is left out of the flat
   *  description. */
  void MakePinholeLens(const real_type (&transform)[16], real_type sensor_distance)
  {
    const real_type* t = transform;
    lens_ = scene::MakePinholeLens<RGB>(
      Matrix4(t[0], t[1], t[2], t[3], t[4], t[5], t[6], t[7],
              t[8], t[9], t[10], t[11], t[12], t[13], t[14], t[15]),
      sensor_distance);
    flat_lens_ = amber_cuda_lens();
    std::memcpy(flat_lens_.transform, transform, sizeof flat_lens_.transform);
    flat_lens_.kind = AMBER_CUDA_LENS_PINHOLE;
    flat_lens_.sensor_distance = sensor_distance;
  }

  /** scene::Scene<RGB>::Create<Acceleration> + the flat description. */
  template <typename Acceleration = raytracer::BVH<real_type, scene::Object<RGB>>>
  std::unique_ptr<Scene> Build();

private:
  std::vector<std::unique_ptr<scene::Primitive>> primitives_;
  std::vector<std::unique_ptr<scene::Material<RGB>>> materials_;
  std::vector<scene::Object<RGB>> objects_;
  std::unique_ptr<scene::Lens<RGB>> lens_;
  std::vector<amber_cuda_primitive> flat_primitives_;
  std::vector<amber_cuda_material> flat_materials_;
  amber_cuda_lens flat_lens_ = {};

  MaterialId AddMaterial(std::unique_ptr<scene::Material<RGB>>&& material, int32_t kind, const RGB& rgb, real_type param)
  {
    materials_.emplace_back(std::move(material));
    amber_cuda_material m;
    m.kind = kind;
    m.rgb[0] = rgb.X();
    m.rgb[1] = rgb.Y();
    m.rgb[2] = rgb.Z();
    m.param = param;
    flat_materials_.push_back(m);
    return static_cast<MaterialId>(flat_materials_.size() - 1);
  }

  void CheckMaterial(MaterialId material) const
  {
    if (material < 0 || static_cast<std::size_t>(material) >= materials_.size()) {
      throw std::out_of_range("FlatSceneBuilder: unknown material id");
    }
  }
};

namespace detail {

/** Feeds fixed uniforms to reference code (Sampler::operator() is virtual,
 *  include/amber/prelude/sampling.h:35-41). */
class ScriptedSampler
: public prelude::Sampler
{
public:
  ScriptedSampler(result_type a, result_type b) noexcept : values_{a, b}, used_(0) {}
  result_type operator()() { return values_[used_++ & 1]; }

private:
  result_type values_[2];
  unsigned used_;
};

/** The three vertices of a reference triangle object, read back exactly through its public
 *  SampleSurfacePoint: (u, v) = (0, 0) yields v0, (0, 1) flips to (1, 0) and yields v1,
 *  (1, 0) flips to (0, 1) and yields v2 (src/amber/scene/primitive_triangle.cc:136-150;
 *  a weight of exactly 1 or 0 leaves the float coordinates untouched). */
inline void
TriangleVertices(const scene::Object<RGB>& object, float (&out)[9])
{
  const double uv[3][2] = {{0, 0}, {0, 1}, {1, 0}};
  for (int k = 0; k < 3; k++) {
    ScriptedSampler sampler(uv[k][0], uv[k][1]);
    const auto point = object.SampleSurfacePoint(sampler).Origin();
    out[3 * k + 0] = point.X();
    out[3 * k + 1] = point.Y();
    out[3 * k + 2] = point.Z();
  }
}

}

template <typename Acceleration>
std::unique_ptr<Scene>
FlatSceneBuilder::Build()
{
  if (!lens_) {
    throw std::logic_error("FlatSceneBuilder: MakeThinLens / MakePinholeLens has not been called");
  }

  // Aperture blades first, as src/amber/etude/cornel_box.cc:62-64 emplaces them; in the flat
  // description they are triangles with the pass-through Eye material
  // (include/amber/scene/material_eye.h:146-155).
  std::vector<scene::Object<RGB>> objects;
  std::vector<amber_cuda_primitive> flat_primitives;
  std::vector<amber_cuda_material> flat_materials(flat_materials_);
  amber_cuda_material eye;
  eye.kind = AMBER_CUDA_MAT_EYE;
  eye.rgb[0] = eye.rgb[1] = eye.rgb[2] = 1;
  eye.param = 0;
  flat_materials.push_back(eye);
  const int32_t eye_id = static_cast<int32_t>(flat_materials.size() - 1);
  for (const auto& object : lens_->ApertureObjects()) {
    objects.emplace_back(*object);
    if (flat_lens_.kind == AMBER_CUDA_LENS_PINHOLE) continue;
    amber_cuda_primitive p;
    p.kind = AMBER_CUDA_PRIM_TRIANGLE;
    p.material = eye_id;
    detail::TriangleVertices(*object, p.data);
    flat_primitives.push_back(p);
  }
  objects.insert(objects.end(), objects_.begin(), objects_.end());
  flat_primitives.insert(flat_primitives.end(), flat_primitives_.begin(), flat_primitives_.end());

  auto scene = std::make_unique<Scene>(
    scene::Scene<RGB>::template Create<Acceleration>(
      std::move(primitives_), std::move(materials_), std::move(objects), std::move(lens_)),
    std::move(flat_primitives), std::move(flat_materials), flat_lens_);
  objects_.clear();
  flat_primitives_.clear();
  flat_materials_.clear();
  return scene;
}

/** Options of the GPU path tracer. */
struct CudaPathTracingOptions
{
  int estimator = AMBER_CUDA_ESTIMATOR_NEE;  // AMBER_CUDA_ESTIMATOR_REFERENCE: algorithm_pt.cc sample for sample
  uint32_t max_depth = 0;                    // 0 = 2^16; the reference has no cap (algorithm_pt.cc:137)
  uint64_t seed = 1;
  std::vector<int> devices = {0};            // samples are dealt to these GPUs, which render concurrently
  uint32_t batch_spp = 64;                   // Context::Iterate() calls consumed per queued render call (raised for small
                                             // frames so that a call holds at least min_batch_paths camera paths)
  uint64_t min_batch_paths = 1ull << 24;     // ~3 ms of work: keeps expiry responsive and the pool full
  uint32_t pool_paths = 0;
  uint32_t queue_depth = 3;                  // render calls queued ahead per device (they run as one wavefront)
};

namespace detail {

/** RAII over one amber_cuda_ctx; rethrows C-ABI failures as exceptions. */
class Device
{
public:
  explicit Device(int device) : ctx_(nullptr)
  {
    if (amber_cuda_create(&ctx_, device) != AMBER_CUDA_OK) {
      throw std::runtime_error(std::string("amber_cuda_create: ") + amber_cuda_last_error(nullptr));
    }
  }
  ~Device() { amber_cuda_destroy(ctx_); }
  Device(const Device&) = delete;
  Device& operator=(const Device&) = delete;

  void Check(int rc) const
  {
    if (rc != AMBER_CUDA_OK) {
      throw std::runtime_error(std::string("amber_cuda: ") + amber_cuda_last_error(ctx_));
    }
  }

  amber_cuda_ctx* Get() const noexcept { return ctx_; }

private:
  amber_cuda_ctx* ctx_;
};

}

/** rendering::Algorithm<RGB> for `--algorithm pt` on the GPU.
 *
 *  Render() keeps the reference's contract (include/amber/rendering/algorithm.h:31-45,
 *  src/amber/rendering/algorithm_pt.cc:82-95): one successful context.Iterate() is one sample per
 *  pixel over the whole frame, so --spp, --time expiry and SIGINT (src/amber/cli/application.cc:131-147)
 *  keep working and a partial mean is returned on expiry; the result is the mean image, row-major
 *  x + y * W.
 *
 *  Iterations are consumed in batches of batch_spp; batch b is QUEUED on device b mod G
 *  (amber_cuda_render_async) at most queue_depth calls ahead of that device, so the G devices
 *  render concurrently -- as the reference's worker threads do (src/amber/prelude/parallel.cc:29-40,
 *  include/amber/rendering/parallel.h:55-68) -- and the batches of one device run as ONE wavefront
 *  (the path pool is not drained between them).  Device d renders the interleaved global sample
 *  indices {d, d + G, d + 2G, ...}.  At the end the films are summed into the first device's over
 *  peer memory (amber_cuda_film_reduce, the counterpart of Accumulator::operator+=,
 *  include/amber/prelude/accumulator.h:136-166) and resolved once.
 *
 *  Unlike the reference's algorithm objects this one keeps state between calls: contexts, the
 *  uploaded scene and the path pool stay on the devices, and a scene is uploaded again only when
 *  Render() is handed a different one. */
class CudaPathTracing
: public rendering::Algorithm<RGB>
{
public:
  explicit CudaPathTracing(const CudaPathTracingOptions& options) : options_(options)
  {
    if (options_.devices.empty()) {
      throw std::invalid_argument("CudaPathTracing: no device given");
    }
    if (options_.batch_spp == 0) {
      options_.batch_spp = 1;
    }
    if (options_.queue_depth == 0) {
      options_.queue_depth = 1;
    }
  }

  const rendering::Image<RGB>
  Render(
    const rendering::Scene<RGB>& scene,
    const rendering::Sensor& sensor,
    rendering::Context& context
  )
  {
    const auto provider = dynamic_cast<const FlatSceneProvider*>(&scene);
    if (!provider) {
      throw std::invalid_argument(
        "CudaPathTracing: the scene was not built by amber::cuda::FlatSceneBuilder "
        "(geometry cannot be read out of a rendering::Scene; there is no CPU fallback)");
    }

    auto image = sensor.CreateImage<RGB>();
    const uint32_t width = static_cast<uint32_t>(image.Width());
    const uint32_t height = static_cast<uint32_t>(image.Height());
    float film_width, film_height;
    if (const auto exact = dynamic_cast<const Sensor*>(&sensor)) {
      film_width = exact->SceneWidth();
      film_height = exact->SceneHeight();
    } else {
      // rendering::Sensor exposes only the film area; the aspect is that of the pixel grid
      // (src/amber/cli/application.cc:89-94)
      film_width = std::sqrt(sensor.SceneArea() * static_cast<float>(width) / static_cast<float>(height));
      film_height = sensor.SceneArea() / film_width;
    }

    // contexts live as long as the algorithm object; the scene is uploaded when it changes
    if (devices_.empty()) {
      for (const auto index : options_.devices) {
        devices_.emplace_back(std::make_unique<detail::Device>(index));
      }
      scene_key_ = SceneKey();
    }
    const auto key = MakeSceneKey(provider->FlatDescription());
    for (const auto& device : devices_) {
      if (!(key == scene_key_)) {
        device->Check(amber_cuda_upload_scene(device->Get(), &provider->FlatDescription()));
      }
      device->Check(amber_cuda_set_sensor(device->Get(), width, height, film_width, film_height));  // clears the film
    }
    scene_key_ = key;

    // consume Iterate() in batches; batch b is queued on device b mod G
    const std::size_t n_devices = devices_.size();
    std::vector<uint64_t> handed(n_devices, 0);  // samples per pixel handed to each device
    const uint64_t n_pixels = static_cast<uint64_t>(width) * height;
    const uint32_t batch_spp = static_cast<uint32_t>(std::min<uint64_t>(
      1u << 20, std::max<uint64_t>(options_.batch_spp, (options_.min_batch_paths + n_pixels - 1) / n_pixels)));
    for (std::size_t batch = 0;; batch++) {
      uint32_t count = 0;
      while (count < batch_spp && context.Iterate()) {
        count++;
      }
      if (count == 0) {
        break;
      }
      const std::size_t d = batch % n_devices;
      const auto& device = *devices_[d];
      device.Check(amber_cuda_wait_pending(device.Get(), static_cast<int>(options_.queue_depth) - 1));
      amber_cuda_render_params params = amber_cuda_render_params();
      params.sample_begin = d + n_devices * handed[d];
      params.sample_stride = static_cast<uint32_t>(n_devices);
      params.sample_count = count;
      params.max_depth = options_.max_depth;
      params.seed = options_.seed;
      params.estimator = options_.estimator;
      params.pool_paths = options_.pool_paths;
      device.Check(amber_cuda_render_async(device.Get(), &params));
      handed[d] += count;
    }
    for (const auto& device : devices_) {
      device->Check(amber_cuda_wait(device->Get()));
    }

    // sum of the films over peer memory, one resolve: mean = sum / total samples
    std::vector<amber_cuda_ctx*> contexts;
    for (const auto& device : devices_) {
      contexts.push_back(device->Get());
    }
    devices_.front()->Check(amber_cuda_film_reduce(contexts.data(), static_cast<int>(contexts.size())));
    std::vector<float> mean(static_cast<std::size_t>(width) * height * 3);
    uint64_t total = 0;
    devices_.front()->Check(amber_cuda_resolve(devices_.front()->Get(), mean.data(), &total));
    if (total > 0) {
      for (uint32_t y = 0; y < height; y++) {
        for (uint32_t x = 0; x < width; x++) {
          const std::size_t i = (static_cast<std::size_t>(x) + static_cast<std::size_t>(y) * width) * 3;
          image[rendering::Pixel(x, y)] = RGB(mean[i], mean[i + 1], mean[i + 2]);
        }
      }
    }
    // counters of the whole job: summed over the devices
    amber_cuda_get_stats(devices_.front()->Get(), &last_stats_);
    for (std::size_t d = 1; d < n_devices; d++) {
      amber_cuda_stats other;
      amber_cuda_get_stats(devices_[d]->Get(), &other);
      last_stats_.paths += other.paths;
      last_stats_.closest_rays += other.closest_rays;
      last_stats_.shadow_rays += other.shadow_rays;
      last_stats_.truncated_paths += other.truncated_paths;
      last_stats_.chained_renders += other.chained_renders;
    }
    return image;
  }

  const amber_cuda_stats& LastStats() const noexcept { return last_stats_; }

private:
  /** Identity of an uploaded scene: where its arrays live, how long they are, and a sparse
   *  checksum of their bytes (every 257th word), so a rebuilt scene at a recycled address is
   *  noticed without hashing 60 MB per Render(). */
  struct SceneKey
  {
    const void* primitives = nullptr;
    const void* materials = nullptr;
    uint32_t n_primitives = 0, n_materials = 0;
    uint64_t checksum = 0;
    bool operator==(const SceneKey& o) const noexcept
    {
      return primitives == o.primitives && materials == o.materials && n_primitives == o.n_primitives &&
             n_materials == o.n_materials && checksum == o.checksum;
    }
  };

  static SceneKey MakeSceneKey(const amber_cuda_scene_desc& d)
  {
    SceneKey key;
    key.primitives = d.primitives;
    key.materials = d.materials;
    key.n_primitives = d.n_primitives;
    key.n_materials = d.n_materials;
    uint64_t h = 1469598103934665603ull;
    auto mix = [&](const void* data, std::size_t bytes, std::size_t step) {
      const auto words = static_cast<const unsigned char*>(data);
      for (std::size_t i = 0; i + 4 <= bytes; i += 4 * step) {
        uint32_t w;
        std::memcpy(&w, words + i, 4);
        h = (h ^ w) * 1099511628211ull;
      }
    };
    mix(d.primitives, static_cast<std::size_t>(d.n_primitives) * sizeof(amber_cuda_primitive), 257);
    mix(d.materials, static_cast<std::size_t>(d.n_materials) * sizeof(amber_cuda_material), 1);
    mix(&d.lens, sizeof d.lens, 1);
    key.checksum = h;
    return key;
  }

  CudaPathTracingOptions options_;
  std::vector<std::unique_ptr<detail::Device>> devices_;
  SceneKey scene_key_;
  amber_cuda_stats last_stats_ = {};
};

/** Factory in the convention of include/amber/rendering/algorithm_pt.h:30-31; this is the
 *  function src/amber/cli/algorithm_factory.cc:35-79 would register under "pt-cuda". */
inline std::unique_ptr<rendering::Algorithm<RGB>>
MakeRGBCudaPathTracing(const CudaPathTracingOptions& options = CudaPathTracingOptions())
{
  return std::make_unique<CudaPathTracing>(options);
}

/** Row N3: the bidirectional path tracer (src/amber/rendering/algorithm_bdpt.cc:82-148 over
 *  include/amber/rendering/bidirectional_path_sampling.h), in the convention of
 *  include/amber/rendering/algorithm_bdpt.h -- what algorithm_factory.cc would register under
 *  "bdpt-cuda".  options.max_depth is the number of events per subpath (0 = 16, the maximum). */
inline std::unique_ptr<rendering::Algorithm<RGB>>
MakeRGBCudaBidirectionalPathTracing(CudaPathTracingOptions options = CudaPathTracingOptions())
{
  options.estimator = AMBER_CUDA_ESTIMATOR_BDPT;
  return std::make_unique<CudaPathTracing>(options);
}

/** Row N2: the light tracer (src/amber/rendering/algorithm_lt.cc:82-163) behind the same
 *  interface, in the convention of include/amber/rendering/algorithm_lt.h:29-30 -- what
 *  algorithm_factory.cc would register under "lt-cuda".  One Context::Iterate() is W*H light
 *  paths, as LightTracing::Thread::operator() traces them; the estimator field of the options
 *  is overridden. */
inline std::unique_ptr<rendering::Algorithm<RGB>>
MakeRGBCudaLightTracing(CudaPathTracingOptions options = CudaPathTracingOptions())
{
  options.estimator = AMBER_CUDA_ESTIMATOR_LIGHT_TRACING;
  return std::make_unique<CudaPathTracing>(options);
}

}
}
