// Microbenchmark: dependent random gathers of fixed-size records, one record per lane per step
// (the access pattern of incoherent BVH traversal).  Which record size / load width does the L1
// gather path of sm_100a prefer?   nvcc -gencode arch=compute_100a,code=sm_100a -O3 gather.cu -o gather
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

struct Wide8 { float4 a, b; };
__device__ __forceinline__ Wide8 ldg256(const void* p) {
  Wide8 r;
  asm volatile("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
      : "=f"(r.a.x), "=f"(r.a.y), "=f"(r.a.z), "=f"(r.a.w), "=f"(r.b.x), "=f"(r.b.y), "=f"(r.b.z), "=f"(r.b.w)
      : "l"(p));
  return r;
}
__device__ __forceinline__ uint32_t mix(uint32_t x) {
  x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16; return x;
}
__device__ __forceinline__ uint32_t bits(float4 v) {
  return __float_as_uint(v.x) ^ __float_as_uint(v.y) ^ __float_as_uint(v.z) ^ __float_as_uint(v.w);
}

// MODE: bytes per record and load width.  n_rec records of REC bytes.
template <int REC, int LOADS128, int LOADS256>
__global__ void __launch_bounds__(256) gather(const char* table, uint32_t n_rec, int steps, uint32_t* out, int active_lanes) {
  extern __shared__ int dummy[];
  uint32_t state = mix(blockIdx.x * 256 + threadIdx.x + 1);
  const bool on = (threadIdx.x & 31) < active_lanes;
  for (int s = 0; s < steps; s++) {
    if (on) {
      const uint32_t i = mix(state + s) % n_rec;
      const char* rec = table + static_cast<size_t>(i) * REC;
      uint32_t acc = 0;
#pragma unroll
      for (int k = 0; k < LOADS128; k++) acc ^= bits(__ldg(reinterpret_cast<const float4*>(rec) + k));
#pragma unroll
      for (int k = 0; k < LOADS256; k++) { Wide8 w = ldg256(rec + 32 * k); acc ^= bits(w.a) ^ bits(w.b); }
      state ^= acc;
    }
  }
  if (state == 0x12345678u) out[0] = state;
}

template <int REC, int L128, int L256>
void run(const char* name, const char* table, size_t table_bytes, uint32_t* out, int lanes) {
  const uint32_t n_rec = static_cast<uint32_t>(table_bytes / REC);
  const int steps = 400, grid = 148 * 4, smem = 40 * 1024;
  cudaFuncSetAttribute(gather<REC, L128, L256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  gather<REC, L128, L256><<<grid, 256, smem>>>(table, n_rec, steps, out, lanes);
  cudaEventRecord(a);
  gather<REC, L128, L256><<<grid, 256, smem>>>(table, n_rec, steps, out, lanes);
  cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double recs = double(grid) * 256 * lanes / 32 * steps;
  printf("%-28s table %4zu MB lanes %2d: %7.2f Grec/s  %7.1f GB/s useful  %6.1f ns/step/warp\n", name, table_bytes >> 20, lanes,
         recs / ms / 1e6, recs * (L128 * 16 + L256 * 32) / ms / 1e6, ms * 1e6 / steps);
}

int main() {
  const size_t max_bytes = 96u << 20;
  char* table; cudaMalloc(&table, max_bytes); cudaMemset(table, 1, max_bytes);
  uint32_t* out; cudaMalloc(&out, 4);
  for (size_t mb : {8, 36, 96}) {
    const size_t bytes = mb << 20;
    for (int lanes : {32, 16}) {
      run<128, 7, 0>("128B rec, 7xLDG.128", table, bytes, out, lanes);
      run<128, 0, 4>("128B rec, 4xLDG.256", table, bytes, out, lanes);
      run<128, 0, 3>("128B rec, 3xLDG.256", table, bytes, out, lanes);
      run<96, 0, 3>("96B rec, 3xLDG.256", table, bytes, out, lanes);
      run<64, 4, 0>("64B rec, 4xLDG.128", table, bytes, out, lanes);
      run<64, 0, 2>("64B rec, 2xLDG.256", table, bytes, out, lanes);
      run<32, 2, 0>("32B rec, 2xLDG.128", table, bytes, out, lanes);
      run<32, 0, 1>("32B rec, 1xLDG.256", table, bytes, out, lanes);
      run<48, 3, 0>("48B rec, 3xLDG.128", table, bytes, out, lanes);
      run<16, 1, 0>("16B rec, 1xLDG.128", table, bytes, out, lanes);
    }
  }
  return 0;
}
