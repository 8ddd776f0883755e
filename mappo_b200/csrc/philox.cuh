// Philox4x32-10 (Salmon et al. SC'11), device side.  CPU statement: oracle/philox.py.
#pragma once
#include <stdint.h>

namespace mappo {

struct u32x4 { uint32_t x, y, z, w; };

__host__ __device__ __forceinline__ u32x4 philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3,
                                                        uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint64_t p0 = (uint64_t)0xD2511F53u * c0;
    const uint64_t p1 = (uint64_t)0xCD9E8D57u * c2;
    const uint32_t hi0 = (uint32_t)(p0 >> 32), lo0 = (uint32_t)p0;
    const uint32_t hi1 = (uint32_t)(p1 >> 32), lo1 = (uint32_t)p1;
    const uint32_t n0 = hi1 ^ c1 ^ k0, n2 = hi0 ^ c3 ^ k1;
    c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
    k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
  }
  u32x4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
  return o;
}

__host__ __device__ __forceinline__ uint32_t lane_of(const u32x4& v, int lane) {
  return lane == 0 ? v.x : lane == 1 ? v.y : lane == 2 ? v.z : v.w;
}

// draw i of the (env, episode) stream: block i/4, lane i%4
__host__ __device__ __forceinline__ uint32_t philox_draw(uint32_t env_lo, uint32_t episode, int i,
                                                         uint32_t k0, uint32_t k1) {
  const u32x4 v = philox4x32_10(env_lo, episode, (uint32_t)(i >> 2), 0u, k0, k1);
  return lane_of(v, i & 3);
}

// uint32 -> double in (0,1): (x + 0.5) * 2^-32   (exact)
__host__ __device__ __forceinline__ double u32_to_unit(uint32_t x) {
  return ((double)x + 0.5) * (1.0 / 4294967296.0);
}
// uint32 -> double in (-1,1)
__host__ __device__ __forceinline__ double u32_to_sym(uint32_t x) { return 2.0 * u32_to_unit(x) - 1.0; }
// uint32 -> float in (0,1) with 24 bits: ((x >> 8) + 0.5) * 2^-24  (exact in f32)
__host__ __device__ __forceinline__ float u32_to_unit_f32(uint32_t x) {
  return ((float)(x >> 8) + 0.5f) * (1.0f / 16777216.0f);
}

}  // namespace mappo
