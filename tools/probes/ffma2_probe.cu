// Throughput of packed fp32 (FFMA2) vs scalar FFMA on sm_100a: same FLOP count, 8 independent chains per thread.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
__device__ __forceinline__ u64 pk(float a, float b) { float2 t = make_float2(a, b); return *reinterpret_cast<u64*>(&t); }
template <int PACKED>
__global__ void k(float* out, int iters, float s) {
  float a[16];
  for (int i = 0; i < 16; ++i) a[i] = threadIdx.x * 0.001f + i;
  if (PACKED) {
    u64 v[8], m = pk(s, s), c = pk(0.5f, 0.25f);
    for (int i = 0; i < 8; ++i) v[i] = pk(a[2 * i], a[2 * i + 1]);
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 8; ++i) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(v[i]) : "l"(m), "l"(c));
    float acc = 0;
    for (int i = 0; i < 8; ++i) { float2 t = *reinterpret_cast<float2*>(&v[i]); acc += t.x + t.y; }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  } else {
    for (int it = 0; it < iters; ++it)
#pragma unroll
      for (int i = 0; i < 16; ++i) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(a[i]) : "f"(s), "f"(0.5f));
    float acc = 0;
    for (int i = 0; i < 16; ++i) acc += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
  }
}
int main() {
  float* out; cudaMalloc(&out, 148 * 512 * 4);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000;
  for (int p = 0; p < 2; ++p) {
    for (int rep = 0; rep < 2; ++rep) {
      cudaEventRecord(e0);
      if (p) k<1><<<148, 512>>>(out, iters, 0.999f); else k<0><<<148, 512>>>(out, iters, 0.999f);
      cudaEventRecord(e1); cudaEventSynchronize(e1);
      float ms; cudaEventElapsedTime(&ms, e0, e1);
      double fma = 148.0 * 512 * 16 * iters;
      if (rep) printf("%s: %.3f ms, %.1f TFLOP/s, %.2f G warp-instr/s per SM\n", p ? "FFMA2" : "FFMA ", ms, 2 * fma / ms / 1e9,
                      (p ? fma / 2 : fma) / 32 / ms / 1e6 / 148);
    }
  }
  return 0;
}
