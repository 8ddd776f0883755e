// Pieces shared by the tensor-core policy kernels (mlp_tc.cu: the fused 64-wide network; mlp_wide.cu: the first layer
// of a wide-input network as two streaming GEMMs): tile constants, the fp16-pair split, operand-tile stores in the
// chunk-major layout of umma.cuh, and the 4-thread row exchange.
#pragma once
#include "mlp_rows.cuh"
#include "umma.cuh"
#include <cuda_fp16.h>

namespace mappo {
namespace tc {

constexpr int kThreads = 512, kBM = 128, kNP = 2;
constexpr float kLoScale = 4096.0f, kLoInv = 1.0f / 4096.0f;            // the lo halves are stored times 2^12
constexpr uint32_t kRowChunk = kBM * 16;                                 // bytes of one 8-column chunk of a 128-row tile
constexpr uint32_t kWChunk = 64 * 16;                                    // ... of a 64-row weight tile

// ---- small device helpers -----------------------------------------------------------------------------
__device__ __forceinline__ void bar_named(int id, int nthreads) { asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory"); }

// tanh in fp32: 1 - 2 / (exp(2|x|) + 1) with ex2.approx / rcp.approx (absolute error ~1.5e-7)
__device__ __forceinline__ float tanh_fast(float x) {
  const float ax = fabsf(x);
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(ax * 2.885390081777927f));   // 2 * log2(e)
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
  return copysignf(fmaf(-2.0f, r, 1.0f), x);
}
// the same from a pre-scaled argument y = 2 log2(e) x
constexpr float kTanhScale = 2.885390081777927f;
__device__ __forceinline__ float tanh_scaled(float y) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fabsf(y)));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
  return copysignf(fmaf(-2.0f, r, 1.0f), y);
}
// the same with the reciprocal taken on the FMA pipe (three Newton steps from the exponent-flip estimate): used for
// every other element of the activation loops, whose pace is otherwise set by the 2 MUFU operations per tanh
__device__ __forceinline__ float tanh_scaled_fma(float y) {
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(fminf(fabsf(y), 64.0f)));
  const float d = e + 1.0f;
  float r = __int_as_float(0x7EF311C7 - __float_as_int(d));
  r = fmaf(r, fmaf(-d, r, 1.0f), r);
  r = fmaf(r, fmaf(-d, r, 1.0f), r);
  r = fmaf(r, fmaf(-d, r, 1.0f), r);
  return copysignf(fmaf(-2.0f, r, 1.0f), y);
}
__device__ __forceinline__ float act_tc(float z, int relu) { return relu ? fmaxf(z, 0.0f) : tanh_fast(z); }

// 8 consecutive fp32 -> one 16-byte chunk of fp16 `hi` and one of (scaled) `lo`
__device__ __forceinline__ void split8(const float* v, uint4& hi, uint4& lo) {
  uint32_t h[4], l[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const __half2 hh = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
    const float2 hf = __half22float2(hh);
    const __half2 ll = __floats2half2_rn((v[2 * i] - hf.x) * kLoScale, (v[2 * i + 1] - hf.y) * kLoScale);
    h[i] = *reinterpret_cast<const uint32_t*>(&hh);
    l[i] = *reinterpret_cast<const uint32_t*>(&ll);
  }
  hi = make_uint4(h[0], h[1], h[2], h[3]);
  lo = make_uint4(l[0], l[1], l[2], l[3]);
}

// store this thread's 16 columns [16*cq, 16*cq+16) of row r into a 128-row, 64-column operand tile (both pieces)
__device__ __forceinline__ void store_cols16(unsigned char* tile, int r, int cq, const float* v) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint4 hi, lo;
    split8(v + 8 * h, hi, lo);
    const uint32_t off = (uint32_t)(2 * cq + h) * kRowChunk + (uint32_t)r * 16;
    *reinterpret_cast<uint4*>(tile + off) = hi;
    *reinterpret_cast<uint4*>(tile + 8 * kRowChunk + off) = lo;
  }
}

// sum (a, b) over the 4 threads that share a row; `slot` alternates between two exchange buffers
__device__ __forceinline__ float2 row_sum2(float a, float b, float2* xch, int& slot, int cq, int r, int q4) {
  float2* buf = xch + slot * (4 * kBM);
  buf[cq * kBM + r] = make_float2(a, b);
  bar_named(1 + q4, 128);
  const float2 p0 = buf[r], p1 = buf[kBM + r], p2 = buf[2 * kBM + r], p3 = buf[3 * kBM + r];
  slot ^= 1;
  return make_float2((p0.x + p1.x) + (p2.x + p3.x), (p0.y + p1.y) + (p2.y + p3.y));
}

// ---- uniformly scaled operand pairs (mlp_tc.cu) ----------------------------------------------------------
// x' = S x with S a power of two fixed per operand class; hi = fp16(x'), lo = fp16(x' - hi): both halves carry the SAME
// scale, so hi*hi + hi*lo + lo*hi accumulate in ONE fp32 accumulator (scaled by the product of the two operand scales,
// undone for free in the epilogue's first FFMA).  S keeps lo out of the fp16 subnormals for the magnitudes that matter
// (absolute resolution 2^-25 / S) and bounds the range to |x| < 65504 / S.
constexpr float kSN = 64.0f;      // LayerNorm outputs (|n| <= sqrt(63))
constexpr float kSX = 16.0f;      // input rows (normalised: |x| <= sqrt(in_dim); raw: saturated at +-4094)
constexpr float kSW = 64.0f;      // folded weights
constexpr float kSD = 16.0f;      // per-sample gradients (dz2, dz1, dlogits)

// two scaled fp32 -> packed fp16 hi pair and packed fp16 residual pair: F2FP, 2 x FHFMA (x' - hi, exact), F2FP
__device__ __forceinline__ void split2u(float a, float b, uint32_t& hi, uint32_t& lo) {
  const __half2 hh = __floats2half2_rn(a, b);
  hi = *reinterpret_cast<const uint32_t*>(&hh);
  float da, db;
  asm("{\n\t.reg .b16 l, h, m;\n\tmov.b32 {l, h}, %2;\n\tmov.b16 m, 0xBC00;\n\t"
      "fma.rn.f32.f16 %0, l, m, %3;\n\tfma.rn.f32.f16 %1, h, m, %4;\n\t}"
      : "=f"(da), "=f"(db) : "r"(hi), "f"(a), "f"(b));
  const __half2 ll = __floats2half2_rn(da, db);
  lo = *reinterpret_cast<const uint32_t*>(&ll);
}
__device__ __forceinline__ void split8u(const float* v, uint4& hi, uint4& lo) {
  split2u(v[0], v[1], hi.x, lo.x);
  split2u(v[2], v[3], hi.y, lo.y);
  split2u(v[4], v[5], hi.z, lo.z);
  split2u(v[6], v[7], hi.w, lo.w);
}
// this thread's 16 (scaled) columns of row r -> hi piece at `tile`, lo piece `lo_off` bytes behind it
__device__ __forceinline__ void store_cols16u(unsigned char* tile, uint32_t lo_off, int r, int cq, const float* v) {
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    uint4 hi, lo;
    split8u(v + 8 * h, hi, lo);
    const uint32_t off = (uint32_t)(2 * cq + h) * kRowChunk + (uint32_t)r * 16;
    *reinterpret_cast<uint4*>(tile + off) = hi;
    *reinterpret_cast<uint4*>(tile + lo_off + off) = lo;
  }
}
// one (scaled) fp32 value -> (hi, lo) fp16 at (r, c) of an R-row chunk-major tile
__device__ __forceinline__ void put_pair_u(unsigned char* tile, uint32_t piece, int R, int r, int c, float v) {
  const __half hi = __float2half_rn(v);
  const __half lo = __float2half_rn(v - __half2float(hi));
  const uint32_t off = umma::tile_off16(R, r, c);
  *reinterpret_cast<__half*>(tile + off) = hi;
  *reinterpret_cast<__half*>(tile + piece + off) = lo;
}

// one fp32 value -> (hi, lo) fp16 at (r, c) of an R-row chunk-major tile
__device__ __forceinline__ void put_pair(unsigned char* tile, uint32_t piece, int R, int r, int c, float v) {
  const __half hi = __float2half_rn(v);
  const __half lo = __float2half_rn((v - __half2float(hi)) * kLoScale);
  const uint32_t off = umma::tile_off16(R, r, c);
  *reinterpret_cast<__half*>(tile + off) = hi;
  *reinterpret_cast<__half*>(tile + piece + off) = lo;
}

}  // namespace tc
}  // namespace mappo
