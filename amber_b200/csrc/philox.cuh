// philox.cuh -- per-path counter-based random stream.
//
// Replaces the reference's per-thread mt19937_64 behind Sampler::operator()
// (include/amber/prelude/sampling.h:148-161, src/amber/rendering/algorithm_pt.cc:104).
// Philox4x32-10, key = (seed_lo, seed_hi), counter = (pixel, sample_lo, sample_hi,
// block); draw j of a path is word j%4 of block j/4 mapped to [0,1) as
// (word >> 8) * 2^-24.  A path is fully described by (seed, pixel, sample) and the
// number of draws it has consumed, so it can be suspended in one kernel and resumed
// in the next, and the image does not depend on how samples are spread over GPUs.
#pragma once

#include <cstdint>

namespace amber_cuda {

struct Philox {
  uint32_t key0, key1;
  uint32_t pixel, sample_lo, sample_hi;
  uint32_t w0, w1, w2, w3;
  uint32_t draws;

  __device__ __forceinline__ void block(uint32_t index) {
    uint32_t c0 = pixel, c1 = sample_lo, c2 = sample_hi, c3 = index;
    uint32_t k0 = key0, k1 = key1;
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
    w0 = c0; w1 = c1; w2 = c2; w3 = c3;
  }

  // Resume a path that has already consumed `consumed` uniforms.
  __device__ __forceinline__ void init(uint64_t seed, uint32_t pixel_index, uint64_t sample,
                                       uint32_t consumed) {
    key0 = static_cast<uint32_t>(seed);
    key1 = static_cast<uint32_t>(seed >> 32);
    pixel = pixel_index;
    sample_lo = static_cast<uint32_t>(sample);
    sample_hi = static_cast<uint32_t>(sample >> 32);
    draws = consumed;
    if (consumed & 3u) block(consumed >> 2);
  }

  __device__ __forceinline__ float next() {
    const uint32_t word = draws & 3u;
    if (word == 0) block(draws >> 2);
    draws++;
    const uint32_t bits = word == 0 ? w0 : word == 1 ? w1 : word == 2 ? w2 : w3;
    return static_cast<float>(bits >> 8) * (1.0f / 16777216.0f);
  }
};

}  // namespace amber_cuda
