// philox.cuh -- per-path counter-based random stream.
//
// Replaces the reference's per-thread mt19937_64 behind Sampler::operator()
// (include/amber/prelude/sampling.h:148-161, src/amber/rendering/algorithm_pt.cc:104).
//
// Philox4x32-10, key = (seed_lo, seed_hi), counter = (pixel, sample, vertex, block).
// A uniform is addressed by (vertex, slot) = word slot%4 of block slot/4, mapped to [0,1)
// as (word >> 8) * 2^-24.  vertex 0 is the camera (slots in call order: blade, u, v,
// film y, film x); vertex k >= 1 is the k-th surface interaction with FIXED slots:
//   0 Russian roulette | 1,2 first BSDF pair | 3 light choice | 4,5 point on the light |
//   8+i BSDF uniform 2+i (Phong rejection retries).
// Fixed slots mean every lane of a warp evaluates block 0 at the same instruction (no
// divergence on "is my buffer empty"), the reference estimator needs ONE Philox block per
// bounce, and a path is resumable from (seed, pixel, sample, vertex) alone -- nothing but
// the bounce count is carried between kernels, and the image does not depend on how
// samples are spread over GPUs.
#pragma once

#include <cstdint>

namespace amber_cuda {

__device__ __forceinline__ uint4 philox4x32_10(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t c2,
                                               uint32_t c3) {
#pragma unroll
  for (int round = 0; round < 10; round++) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
    c0 = hi1 ^ c1 ^ k0;
    c1 = lo1;
    c2 = hi0 ^ c3 ^ k1;
    c3 = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float philox_uniform(uint32_t bits) {
  return static_cast<float>(bits >> 8) * (1.0f / 16777216.0f);
}

// Uniforms of one surface vertex.
struct VertexRng {
  uint32_t key0, key1, pixel, sample, vertex;
  uint4 block0, retry;
  uint32_t bsdf_i;

  __device__ __forceinline__ void init(uint64_t seed, uint32_t pixel_index, uint32_t sample_index,
                                       uint32_t vertex_index) {
    key0 = static_cast<uint32_t>(seed);
    key1 = static_cast<uint32_t>(seed >> 32);
    pixel = pixel_index;
    sample = sample_index;
    vertex = vertex_index;
    block0 = philox4x32_10(key0, key1, pixel, sample, vertex, 0);
    bsdf_i = 0;
  }
  __device__ __forceinline__ float rr() const { return philox_uniform(block0.x); }
  __device__ __forceinline__ float light_choice() const { return philox_uniform(block0.w); }
  __device__ __forceinline__ void light_point(float& u, float& v) const {
    const uint4 b = philox4x32_10(key0, key1, pixel, sample, vertex, 1);
    u = philox_uniform(b.x);
    v = philox_uniform(b.y);
  }
  // next BSDF-sampling uniform, in the reference's call order
  __device__ __forceinline__ float next() {
    const uint32_t i = bsdf_i++;
    if (i == 0) return philox_uniform(block0.y);
    if (i == 1) return philox_uniform(block0.z);
    const uint32_t j = i - 2;
    if ((j & 3u) == 0) retry = philox4x32_10(key0, key1, pixel, sample, vertex, 2 + (j >> 2));
    const uint32_t w = j & 3u;
    return philox_uniform(w == 0 ? retry.x : w == 1 ? retry.y : w == 2 ? retry.z : retry.w);
  }
};

// Uniforms of the camera vertex (five, in call order).
struct CameraRng {
  uint4 block0;
  uint32_t block1_x;
  uint32_t i;
  __device__ __forceinline__ void init(uint64_t seed, uint32_t pixel, uint32_t sample) {
    const uint32_t k0 = static_cast<uint32_t>(seed), k1 = static_cast<uint32_t>(seed >> 32);
    block0 = philox4x32_10(k0, k1, pixel, sample, 0, 0);
    block1_x = philox4x32_10(k0, k1, pixel, sample, 0, 1).x;
    i = 0;
  }
  __device__ __forceinline__ float next() {
    const uint32_t k = i++;
    return philox_uniform(k == 0 ? block0.x : k == 1 ? block0.y : k == 2 ? block0.z : k == 3 ? block0.w : block1_x);
  }
};

}  // namespace amber_cuda
