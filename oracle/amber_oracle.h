/*
 * amber_oracle.h -- TEST INFRASTRUCTURE.  Not part of the product.
 *
 * Plain-C restatement of amber's path-tracing hot path (the algorithm of
 * /root/reference, cited function by function in amber_oracle.c), taking the same
 * flat scene description as the CUDA library.  It is the checker for the GPU path:
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may call it.
 *
 * Pinning: tests/test_oracle_pinned.py requires this restatement to reproduce the
 * UNMODIFIED reference (oracle/_ref/libamber_ref.so) -- closest hits, BSDF samples,
 * camera rays and whole rendered images bit for bit on the same random stream -- and
 * checks it against the golden fixtures under tests/golden/ generated from that
 * reference (scripts/make_golden.py), so the oracle stays pinned where /root/reference
 * is absent.  The reference has no tests or golden vectors of its own.
 */
#ifndef AMBER_ORACLE_H_
#define AMBER_ORACLE_H_

#include <stdint.h>

#include "../include/amber_cuda.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct oracle_scene oracle_scene;

oracle_scene* oracle_scene_create(const amber_cuda_scene_desc* desc);
void oracle_scene_destroy(oracle_scene* scene);

/* closest hit, brute force in creation order (raytracer::List semantics) */
void oracle_cast(const oracle_scene* scene, uint64_t n, const float* origins, const float* directions,
                 int32_t* prim, float* t, int threads);

void oracle_sample_light(const amber_cuda_material* material, uint64_t n, const float* normal,
                         const float* dir_out, const float* xi8, float* dir_in, float* weight,
                         int32_t* draws);
void oracle_bsdf(const amber_cuda_material* material, uint64_t n, const float* normal,
                 const float* dir_out, const float* dir_in, float* bsdf, float* radiance);
void oracle_eye_rays(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                     uint64_t n, const uint32_t* px, const uint32_t* py, const float* xi5,
                     float* origin, float* direction, float* weight);

/* Path tracer on the device's Philox stream.  estimator: AMBER_CUDA_ESTIMATOR_*.
 * stats[0] closest rays, [1] shadow rays, [2] truncated paths, [3] paths. */
void oracle_render(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                   uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_depth,
                   int estimator, int threads, float* rgb, uint64_t* stats);

/* Light tracing (src/amber/rendering/algorithm_lt.cc:124-163) on the same stream: one sample =
 * w*h light paths keyed (path index, sample).  stats[0] closest rays, [1] splats,
 * [2] truncated paths, [3] paths. */
void oracle_render_lt(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                      uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_depth,
                      int threads, float* rgb, uint64_t* stats);
void oracle_sample_importance(const amber_cuda_material* material, uint64_t n, const float* normal,
                              const float* dir_out, const float* xi8, float* dir_in, float* weight,
                              int32_t* draws);
void oracle_lens_response(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                          uint64_t n, const float* position, const float* dir_out, int32_t* pixel,
                          float* value);
void oracle_light_rays(const oracle_scene* scene, uint64_t n, const float* xi5, float* origin,
                       float* direction, float* weight);

/* Bidirectional path tracing (include/amber/rendering/bidirectional_path_sampling.h,
 * src/amber/rendering/algorithm_bdpt.cc:116-148), each subpath on its sequential stream.
 * max_vertices: events per subpath, 1..64 (0 = 64).  stats[0] light events, [1] eye events,
 * [2] light-image splats. */
void oracle_render_bdpt(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                        uint64_t sample_begin, uint32_t sample_count, uint64_t seed, uint32_t max_vertices,
                        int threads, float* rgb, uint64_t* stats);

/* Filmic (exposure 16) + Gamma (2.2) + 8-bit truncation of n_values floats (any layout). */
void oracle_tonemap(uint64_t n_values, const float* hdr, uint8_t* ldr);

/* test hooks: the Philox stream itself, and the fixed ray set of the intersection-parity
 * test (n camera rays followed by n secondary rays). */
void oracle_philox4x32_10(const uint32_t ctr[4], const uint32_t key[2], uint32_t out[4]);
void oracle_stream(uint64_t seed, uint32_t pixel, uint64_t sample, uint32_t n, float* out);
void oracle_ray_set(const oracle_scene* scene, uint32_t w, uint32_t h, float film_w, float film_h,
                    uint64_t seed, uint64_t n, float* origins, float* directions);

#ifdef __cplusplus
}
#endif
#endif
