// adapter_demo.cc -- the drop-in, exercised the way a maintainer of the reference would use it.
//
// Builds the Cornell box of src/amber/etude/cornel_box.cc:38-204 through
// amber::cuda::FlatSceneBuilder (the reference's factory verbs), then renders the SAME scene
// object twice through the SAME interface, rendering::Algorithm<RGB>::Render(scene, sensor, context):
//   1. the reference's own CPU path tracer, MakeRGBPathTracing()   (unmodified reference code)
//   2. the GPU path tracer,                 MakeRGBCudaPathTracing()
// and compares the two images statistically (mean luminance, 4x4 block means).
//
// Built by `make -C oracle adapter_demo` into oracle/_ref/ (it links the compiled reference).
// usage: adapter_demo [size=128] [spp=64] [threads=8] [devices=1]
//        adapter_demo --dump <file>     (flat scene description only; needs no GPU)
//        adapter_demo --throughput <size> <spp> <devices> [batch_spp=64] [repeats=3]
//            the GPU path tracer alone through Algorithm<RGB>::Render on `devices` GPUs, timed per call
//            (the first call uploads the scene and allocates the pools; later ones reuse them), and --
//            with more than one device -- compared with the image of ONE device, which holds the same
//            samples (device d renders the indices d, d + G, ...)
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <mutex>
#include <string>

#include "amber/etude/cornel_box.h"
#include "amber/rendering/algorithm_bdpt.h"
#include "amber/rendering/algorithm_lt.h"
#include "amber/rendering/algorithm_pt.h"

#include "amber_cuda/adapter.hpp"

using namespace amber;
using rendering::RGB;
using scene::Vector3;

namespace {

// cli::Context semantics, src/amber/cli/context.cc:34-67
class DemoContext : public rendering::Context
{
public:
  DemoContext(std::size_t threads, std::size_t iterations) : threads_(threads), iterations_(iterations) {}
  std::size_t ThreadCount() const noexcept { return threads_; }
  std::size_t IterationCount() const noexcept
  {
    std::lock_guard<std::mutex> lock(mutex_);
    return started_;
  }
  bool Iterate() noexcept
  {
    std::lock_guard<std::mutex> lock(mutex_);
    if (started_ >= iterations_) return false;
    started_++;
    return true;
  }

private:
  std::size_t threads_, iterations_, started_ = 0;
  mutable std::mutex mutex_;
};

std::unique_ptr<cuda::Scene> CornelBox()
{
  cuda::FlatSceneBuilder b;
  const scene::real_type transform[16] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 4, 0, 0, 0, 1};
  b.MakeThinLens(transform, 0.050, 4.0, 0.050, 6);
  auto quad = [&](cuda::FlatSceneBuilder::MaterialId m, Vector3 a0, Vector3 a1, Vector3 a2, Vector3 b0, Vector3 b1, Vector3 b2) {
    b.MakeTriangle(a0, a1, a2, m);
    b.MakeTriangle(b0, b1, b2, m);
  };
  quad(b.MakeDiffuseLight(RGB(1e11, 1e11, 1e11)),
       Vector3(0.01, 0.99, 0.01), Vector3(-0.01, 0.99, 0.01), Vector3(-0.01, 0.99, -0.01),
       Vector3(-0.01, 0.99, -0.01), Vector3(0.01, 0.99, -0.01), Vector3(0.01, 0.99, 0.01));
  quad(b.MakeLambertian(RGB(.5, 0, 0)),
       Vector3(-1, 1, 1), Vector3(-1, -1, 1), Vector3(-1, -1, -1), Vector3(-1, -1, -1), Vector3(-1, 1, -1), Vector3(-1, 1, 1));
  quad(b.MakeLambertian(RGB(0, .5, 0)),
       Vector3(1, 1, 1), Vector3(1, 1, -1), Vector3(1, -1, -1), Vector3(1, -1, -1), Vector3(1, -1, 1), Vector3(1, 1, 1));
  quad(b.MakePhong(RGB(.95, .95, .95), 256),
       Vector3(1, 1, -1), Vector3(-1, 1, -1), Vector3(-1, -1, -1), Vector3(-1, -1, -1), Vector3(1, -1, -1), Vector3(1, 1, -1));
  quad(b.MakeLambertian(RGB(.5, .5, .5)),
       Vector3(1, -1, 1), Vector3(1, -1, -1), Vector3(-1, -1, -1), Vector3(-1, -1, -1), Vector3(-1, -1, 1), Vector3(1, -1, 1));
  quad(b.MakeLambertian(RGB(.5, .5, .5)),
       Vector3(1, 1, 1), Vector3(-1, 1, 1), Vector3(-1, 1, -1), Vector3(-1, 1, -1), Vector3(1, 1, -1), Vector3(1, 1, 1));
  const auto water = b.MakeRefraction(1.333);
  quad(water, Vector3(1, -0.5, 1), Vector3(1, -0.5, -1), Vector3(-1, -0.5, -1),
       Vector3(-1, -0.5, -1), Vector3(-1, -0.5, 1), Vector3(1, -0.5, 1));
  quad(water, Vector3(1, -0.5, 1), Vector3(-1, -0.5, 1), Vector3(-1, -1, 1),
       Vector3(-1, -1, 1), Vector3(1, -1, 1), Vector3(1, -0.5, 1));
  b.MakeSphere(Vector3(0.4, -0.6, -0.5), 0.4, b.MakeLambertian(RGB(.5, .5, .5)));
  b.MakeSphere(Vector3(-0.4, -0.7, 0.1), 0.3, b.MakeSpecular(RGB(.95, .95, .95)));
  b.MakeSphere(Vector3(0.1, -0.8, 0.6), 0.2, b.MakeRefraction(1.125));
  return b.Build();
}

struct Summary { double mean; std::vector<double> blocks; };

Summary Summarise(const rendering::Image<RGB>& image, unsigned size)
{
  const unsigned nb = 4, bs = size / nb;
  Summary s{0, std::vector<double>(nb * nb, 0.0)};
  for (unsigned y = 0; y < size; y++) {
    for (unsigned x = 0; x < size; x++) {
      const auto& v = image[rendering::Pixel(x, y)];
      const double lum = 0.2126 * v.X() + 0.7152 * v.Y() + 0.0722 * v.Z();
      s.mean += lum;
      s.blocks[(y / bs) * nb + (x / bs)] += lum;
    }
  }
  s.mean /= double(size) * size;
  for (auto& b : s.blocks) b /= double(bs) * bs;
  return s;
}

}  // namespace

int main(int argc, char** argv)
{
  const unsigned size = argc > 1 ? std::atoi(argv[1]) : 128;
  const unsigned spp = argc > 2 ? std::atoi(argv[2]) : 64;
  const unsigned threads = argc > 3 ? std::atoi(argv[3]) : 8;
  const unsigned n_devices = argc > 4 ? std::atoi(argv[4]) : 1;

  const auto scene = CornelBox();
  if (argc > 2 && std::string(argv[1]) == "--dump") {
    // flat description as the GPU receives it: per primitive kind, 9 floats, then its material
    // (kind, rgb, param) -- compared with the Python builder's Cornell box by tests/test_adapter.py
    const auto& d = scene->FlatDescription();
    std::FILE* f = std::fopen(argv[2], "wb");
    if (!f) return 4;
    for (uint32_t i = 0; i < d.n_primitives; i++) {
      const auto& p = d.primitives[i];
      const auto& m = d.materials[p.material];
      std::fwrite(&p.kind, 4, 1, f);
      std::fwrite(p.data, 4, 9, f);
      std::fwrite(&m.kind, 4, 1, f);
      std::fwrite(m.rgb, 4, 3, f);
      std::fwrite(&m.param, 4, 1, f);
    }
    std::fwrite(&d.lens, sizeof d.lens, 1, f);
    std::fclose(f);
    return 0;
  }
  if (argc > 4 && std::string(argv[1]) == "--throughput") {
    const unsigned t_size = std::atoi(argv[2]), t_spp = std::atoi(argv[3]), t_devices = std::atoi(argv[4]);
    const unsigned batch_spp = argc > 5 ? std::atoi(argv[5]) : 64, repeats = argc > 6 ? std::atoi(argv[6]) : 3;
    const cuda::Sensor t_sensor(t_size, t_size, 0.036f, 0.036f);
    auto run = [&](unsigned n_dev, unsigned n_repeats, bool print) {
      cuda::CudaPathTracingOptions options;
      options.devices.clear();
      for (unsigned d = 0; d < n_dev; d++) options.devices.push_back(int(d));
      options.max_depth = 32;
      options.batch_spp = batch_spp;
      cuda::CudaPathTracing algorithm(options);
      rendering::Image<RGB> image(t_size, t_size);
      for (unsigned r = 0; r < n_repeats; r++) {
        DemoContext context(1, t_spp);
        const auto t0 = std::chrono::steady_clock::now();
        image = algorithm.Render(*scene, t_sensor, context);
        const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        const auto& s = algorithm.LastStats();
        if (print)
          std::printf("{\"through\": \"Algorithm<RGB>::Render\", \"devices\": %u, \"size\": %u, \"spp\": %u, \"batch_spp\": %u, "
                      "\"call\": %u, \"seconds\": %.4f, \"Mrays_per_s\": %.1f, \"Mpaths_per_s\": %.1f, \"chained_renders\": %llu}\n",
                      n_dev, t_size, t_spp, batch_spp, r, seconds, double(s.closest_rays + s.shadow_rays) / seconds / 1e6,
                      double(s.paths) / seconds / 1e6, static_cast<unsigned long long>(s.chained_renders));
      }
      return image;
    };
    try {
      const auto image = run(t_devices, repeats, true);
      if (t_devices > 1) {
        const auto single = run(1, 1, false);
        double worst = 0, sum = 0;
        for (unsigned y = 0; y < t_size; y++) {
          for (unsigned x = 0; x < t_size; x++) {
            const auto& a = image[rendering::Pixel(x, y)];
            const auto& b = single[rendering::Pixel(x, y)];
            const double la = a.X() + a.Y() + a.Z(), lb = b.X() + b.Y() + b.Z();
            worst = std::max(worst, std::fabs(la - lb) / std::max(1e-30, std::max(la, lb)));
            sum += lb;
          }
        }
        std::printf("%u devices vs 1 device, same samples: worst relative pixel difference %.3g (mean %.6g)\n", t_devices,
                    worst, sum / (double(t_size) * t_size));
        if (!(worst < 1e-3)) {
          std::printf("adapter_demo: MISMATCH\n");
          return 1;
        }
      }
    } catch (const std::runtime_error& error) {
      std::printf("pt-cuda failed loudly (no CPU fallback): %s\n", error.what());
      return 3;
    }
    std::printf("adapter_demo: OK\n");
    return 0;
  }
  const cuda::Sensor sensor(size, size, 0.036f, 0.036f * size / size);
  std::printf("scene: %u flat primitives, %u materials\n", scene->FlatDescription().n_primitives,
              scene->FlatDescription().n_materials);

  auto timed = [&](rendering::Algorithm<RGB>& algorithm, unsigned n_threads, unsigned iterations, const char* name) {
    DemoContext context(n_threads, iterations);
    const auto t0 = std::chrono::steady_clock::now();
    auto image = algorithm.Render(*scene, sensor, context);
    const double seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
    std::printf("%-28s %4u spp in %8.3f s  (%.2f Mpaths/s)\n", name, iterations, seconds,
                double(size) * size * iterations / seconds / 1e6);
    return image;
  };

  // 1. the reference's CPU path tracer on the builder's scene
  const auto cpu_algorithm = rendering::MakeRGBPathTracing();
  const auto cpu_image = timed(*cpu_algorithm, threads, spp, "reference pt (CPU)");
  const auto cpu = Summarise(cpu_image, size);

  // 2. the GPU path tracer behind the same interface
  cuda::CudaPathTracingOptions options;
  options.devices.clear();
  for (unsigned d = 0; d < n_devices; d++) options.devices.push_back(int(d));
  std::unique_ptr<rendering::Algorithm<RGB>> gpu_algorithm;
  rendering::Image<RGB> gpu_image(size, size);
  try {
    options.estimator = AMBER_CUDA_ESTIMATOR_REFERENCE;
    gpu_algorithm = cuda::MakeRGBCudaPathTracing(options);
    gpu_image = timed(*gpu_algorithm, 1, spp * 4, "pt-cuda, reference estimator");
  } catch (const std::runtime_error& error) {
    std::printf("pt-cuda failed loudly (no CPU fallback): %s\n", error.what());
    return 3;
  }
  const auto gpu = Summarise(gpu_image, size);
  options.estimator = AMBER_CUDA_ESTIMATOR_NEE;
  gpu_algorithm = cuda::MakeRGBCudaPathTracing(options);
  const auto nee = Summarise(timed(*gpu_algorithm, 1, spp * 4, "pt-cuda, NEE estimator"), size);

  // row N2: the light tracer behind the same interface.  In this scene a light path finds the
  // aperture three times in 1e5, so the CPU reference is run only to show the call; the GPU
  // gets enough paths for its image mean to be compared with the path tracer's.
  const auto cpu_lt = rendering::MakeRGBLightTracing();
  timed(*cpu_lt, threads, std::max(1u, spp / 16), "reference lt (CPU)");
  gpu_algorithm = cuda::MakeRGBCudaLightTracing(options);
  const auto lt = Summarise(timed(*gpu_algorithm, 1, spp * 64, "lt-cuda"), size);

  // row N3: bidirectional path tracing behind the same interface, CPU reference beside it
  const auto cpu_bdpt = rendering::MakeRGBBidirectionalPathTracing();
  const auto bdpt_cpu = Summarise(timed(*cpu_bdpt, threads, std::max(1u, spp / 4), "reference bdpt (CPU)"), size);
  gpu_algorithm = cuda::MakeRGBCudaBidirectionalPathTracing(options);
  const auto bdpt_gpu = Summarise(timed(*gpu_algorithm, 1, spp, "bdpt-cuda"), size);
  std::printf("mean luminance: reference bdpt %.6g | bdpt-cuda %.6g\n", bdpt_cpu.mean, bdpt_gpu.mean);

  // a reference scene that was NOT built by the builder must be refused, not rendered on the CPU
  bool refused = false;
  try {
    const auto builtin = etude::CornelBox(0.050, 0.050, 6);  // src/amber/cli/application.cc:69-73
    DemoContext context(1, 1);
    gpu_algorithm->Render(builtin, sensor, context);
  } catch (const std::invalid_argument& error) {
    refused = true;
    std::printf("built-in etude::CornelBox refused as expected: %s\n", error.what());
  }
  if (!refused) {
    std::printf("adapter_demo: a non-flat scene was not refused\n");
    return 2;
  }

  // The reference estimator has no light sampling: a pixel is zero unless a BSDF-sampled ray
  // happens to hit the 0.02 x 0.02 light, so only large aggregates are comparable -- the overall
  // mean and the means of 4x4 image blocks (median relative deviation over the 16 blocks).
  auto median_rel = [](const Summary& a, const Summary& b) {
    std::vector<double> rel;
    for (std::size_t i = 0; i < a.blocks.size(); i++) {
      const double denom = 0.5 * (a.blocks[i] + b.blocks[i]);
      if (denom > 0) rel.push_back(std::fabs(a.blocks[i] - b.blocks[i]) / denom);
    }
    std::sort(rel.begin(), rel.end());
    return rel.empty() ? 0.0 : rel[rel.size() / 2];
  };
  auto rel_mean = [](const Summary& a, const Summary& b) { return std::fabs(a.mean - b.mean) / (0.5 * (a.mean + b.mean)); };
  std::printf("mean luminance: reference CPU %.6g | pt-cuda reference-estimator %.6g | pt-cuda NEE %.6g\n",
              cpu.mean, gpu.mean, nee.mean);
  std::printf("median relative deviation of 4x4 block means: CPU vs cuda(ref) %.4f, cuda(ref) vs cuda(NEE) %.4f\n",
              median_rel(cpu, gpu), median_rel(gpu, nee));
  std::printf("mean luminance: lt-cuda %.6g (vs pt-cuda NEE: %.3f relative)\n", lt.mean, rel_mean(lt, nee));
  const bool ok = rel_mean(cpu, gpu) < 0.2 && rel_mean(gpu, nee) < 0.1 && median_rel(gpu, nee) < 0.25 &&
                  rel_mean(lt, nee) < 0.3;
  std::printf("%s\n", ok ? "adapter_demo: OK" : "adapter_demo: MISMATCH");
  return ok ? 0 : 1;
}
