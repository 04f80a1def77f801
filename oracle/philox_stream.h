/*
 * philox_stream.h -- TEST INFRASTRUCTURE (oracle/).  Not part of the product.
 *
 * The per-path random stream the CUDA kernels use, restated in plain C so the CPU
 * oracle and the reference harness can replay exactly the uniforms a device path
 * consumed.  Spec (DESIGN.md "Random stream"):
 *
 *   Philox4x32-10 (Salmon et al., SC'11), key = (seed_lo, seed_hi),
 *   counter = (pixel_index, sample_index, vertex, block).  A uniform is addressed by
 *   (vertex, slot): word slot%4 of block slot/4, mapped to [0,1) as (word >> 8) * 2^-24.
 *
 *   vertex 0 is the camera: slots 0.. are consumed in call order (blade, u, v, film y,
 *   film x).  vertex k >= 1 is the k-th surface interaction, with FIXED slots so that every
 *   lane of a warp evaluates the same block at the same place:
 *     slot 0      Russian roulette
 *     slot 1, 2   the first two BSDF-sampling uniforms
 *     slot 3      light choice        \  consumed only by the NEE estimator
 *     slot 4, 5   point on the light  /
 *     slot 8 + i  BSDF-sampling uniform 2 + i (Phong rejection retries)
 *
 * The reference draws `static_cast<float>(uniform_real_distribution<double>(0,1))` from a
 * per-thread mt19937_64 (include/amber/prelude/sampling.h:148-161,
 * src/amber/rendering/algorithm_pt.cc:104); which engine feeds Sampler::operator() is not
 * part of its numerics, so a Sampler subclass fed by this stream drives the unmodified
 * reference code along exactly the device's paths.
 */
#ifndef AMBER_ORACLE_PHILOX_STREAM_H_
#define AMBER_ORACLE_PHILOX_STREAM_H_

#include <stdint.h>

typedef struct philox_stream {
  uint32_t key[2];
  uint32_t pixel, sample, vertex;
  uint32_t words[4];
  uint32_t cached_block; /* block index held in words[], 0xffffffff = none */
  uint32_t camera_i;     /* next camera slot (vertex 0) */
  uint32_t bsdf_i;       /* next BSDF-sampling uniform of the current vertex */
  uint32_t draws;        /* total uniforms consumed (statistics) */
} philox_stream;

static inline void philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]) {
  uint32_t c0 = ctr[0], c1 = ctr[1], c2 = ctr[2], c3 = ctr[3];
  uint32_t k0 = key[0], k1 = key[1];
  for (int round = 0; round < 10; round++) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0;
    const uint32_t n1 = (uint32_t)p1;
    const uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1;
    const uint32_t n3 = (uint32_t)p0;
    c0 = n0; c1 = n1; c2 = n2; c3 = n3;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
  out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}

static inline void philox_stream_begin_vertex(philox_stream* s, uint32_t vertex) {
  s->vertex = vertex;
  s->cached_block = 0xffffffffu;
  s->bsdf_i = 0;
}

static inline void philox_stream_reset(philox_stream* s, uint64_t seed, uint32_t pixel,
                                       uint64_t sample) {
  s->key[0] = (uint32_t)seed;
  s->key[1] = (uint32_t)(seed >> 32);
  s->pixel = pixel;
  s->sample = (uint32_t)sample;
  s->camera_i = 0;
  s->draws = 0;
  philox_stream_begin_vertex(s, 0);
}

static inline float philox_stream_slot(philox_stream* s, uint32_t slot) {
  const uint32_t block = slot >> 2;
  if (block != s->cached_block) {
    const uint32_t ctr[4] = {s->pixel, s->sample, s->vertex, block};
    philox4x32_10(ctr, s->key, s->words);
    s->cached_block = block;
  }
  s->draws++;
  return (float)(s->words[slot & 3u] >> 8) * (1.0f / 16777216.0f);
}

/* Bidirectional path tracing (row N3) draws each SUBPATH from one purely sequential stream: the
 * counter's vertex word is kPhiloxSubpath | subpath (0 = light, 1 = eye) and uniform i is word
 * i % 4 of block i / 4 -- the reference's subpath generators interleave BSDF and Russian-roulette
 * draws in one call order, and the device's per-thread kernel replays exactly that order. */
#define PHILOX_SUBPATH_LIGHT 0x40000000u
#define PHILOX_SUBPATH_EYE 0x40000001u
static inline void philox_stream_begin_sequence(philox_stream* s, uint32_t id) {
  philox_stream_begin_vertex(s, id);
  s->camera_i = 0;
}

/* camera vertex: sequential slots */
static inline float philox_stream_next_camera(philox_stream* s) { return philox_stream_slot(s, s->camera_i++); }
/* surface vertex */
static inline float philox_stream_rr(philox_stream* s) { return philox_stream_slot(s, 0); }
static inline float philox_stream_light(philox_stream* s, uint32_t i) { return philox_stream_slot(s, 3 + i); }
static inline float philox_stream_next_bsdf(philox_stream* s) {
  const uint32_t i = s->bsdf_i++;
  return philox_stream_slot(s, i < 2 ? 1 + i : 8 + (i - 2));
}

#endif /* AMBER_ORACLE_PHILOX_STREAM_H_ */
