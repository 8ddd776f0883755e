// fp32 FMA tile primitives shared by the register-tiled policy kernels (mlp.cu: feed-forward networks; mlp_gru.cu:
// recurrent networks).  Thread (ty, tx) of a 16-wide thread row owns 4 output columns 4 tx .. 4 tx + 3 of rows
// ty + NTY * i of a tile; tiles live in shared memory with row stride kLD.
#pragma once
#include "mlp_rows.cuh"

namespace mappo {

__device__ __forceinline__ float act_fwd(float z, int relu) { return relu ? fmaxf(z, 0.0f) : tanhf(z); }

// sum over the 16 lanes that share a row (half-warp: lanes differ in tx only)
__device__ __forceinline__ float hw_sum(float v) {
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  return v;
}

// acc[i][c] += sum_k A[row(i)][k] * B[k][4tx + c],  row(i) = ty + NTY*i,  k < 4*K4
template <int RPT, int NTY>
__device__ __forceinline__ void gemm_tile(const float* __restrict__ A, int lda, int K4, const float* __restrict__ B,
                                          float (&acc)[RPT][4], int ty, int tx) {
#pragma unroll 2
  for (int k4 = 0; k4 < K4; ++k4) {
    const float4 b0 = *reinterpret_cast<const float4*>(B + (4 * k4 + 0) * 64 + 4 * tx);
    const float4 b1 = *reinterpret_cast<const float4*>(B + (4 * k4 + 1) * 64 + 4 * tx);
    const float4 b2 = *reinterpret_cast<const float4*>(B + (4 * k4 + 2) * 64 + 4 * tx);
    const float4 b3 = *reinterpret_cast<const float4*>(B + (4 * k4 + 3) * 64 + 4 * tx);
#pragma unroll
    for (int i = 0; i < RPT; ++i) {
      const float4 a = *reinterpret_cast<const float4*>(A + (ty + NTY * i) * lda + 4 * k4);
      acc[i][0] = fmaf(a.x, b0.x, acc[i][0]); acc[i][1] = fmaf(a.x, b0.y, acc[i][1]);
      acc[i][2] = fmaf(a.x, b0.z, acc[i][2]); acc[i][3] = fmaf(a.x, b0.w, acc[i][3]);
      acc[i][0] = fmaf(a.y, b1.x, acc[i][0]); acc[i][1] = fmaf(a.y, b1.y, acc[i][1]);
      acc[i][2] = fmaf(a.y, b1.z, acc[i][2]); acc[i][3] = fmaf(a.y, b1.w, acc[i][3]);
      acc[i][0] = fmaf(a.z, b2.x, acc[i][0]); acc[i][1] = fmaf(a.z, b2.y, acc[i][1]);
      acc[i][2] = fmaf(a.z, b2.z, acc[i][2]); acc[i][3] = fmaf(a.z, b2.w, acc[i][3]);
      acc[i][0] = fmaf(a.w, b3.x, acc[i][0]); acc[i][1] = fmaf(a.w, b3.y, acc[i][1]);
      acc[i][2] = fmaf(a.w, b3.z, acc[i][2]); acc[i][3] = fmaf(a.w, b3.w, acc[i][3]);
    }
  }
}

// z = acc + bias -> t = act(z) -> n = LayerNorm(t) (no affine); writes n to dst and (mean, rstd) per row.
template <int RPT, int NTY>
__device__ __forceinline__ void act_ln_store(float (&acc)[RPT][4], const float* __restrict__ bias, int relu,
                                             float* __restrict__ dst, float* __restrict__ mean_out,
                                             float* __restrict__ rstd_out, int ty, int tx) {
  const float4 b = *reinterpret_cast<const float4*>(bias + 4 * tx);
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = ty + NTY * i;
    float t0 = act_fwd(acc[i][0] + b.x, relu), t1 = act_fwd(acc[i][1] + b.y, relu);
    float t2 = act_fwd(acc[i][2] + b.z, relu), t3 = act_fwd(acc[i][3] + b.w, relu);
    const float mean = hw_sum((t0 + t1) + (t2 + t3)) * (1.0f / 64.0f);
    t0 -= mean; t1 -= mean; t2 -= mean; t3 -= mean;
    const float var = hw_sum((t0 * t0 + t1 * t1) + (t2 * t2 + t3 * t3)) * (1.0f / 64.0f);
    const float rstd = 1.0f / sqrtf(var + kLnEps);
    *reinterpret_cast<float4*>(dst + r * kLD + 4 * tx) = make_float4(t0 * rstd, t1 * rstd, t2 * rstd, t3 * rstd);
    if (tx == 0) { mean_out[r] = mean; rstd_out[r] = rstd; }
  }
}

// dn (in acc) -> dz, through LayerNorm (no affine) and the activation; overwrites the n tile with dz.
// Caller guarantees nobody else still reads `tile` (sync before) and syncs after.
template <int RPT, int NTY>
__device__ __forceinline__ void ln_act_bwd_store(float (&acc)[RPT][4], int relu, float* __restrict__ tile,
                                                 const float* __restrict__ mean_in, const float* __restrict__ rstd_in,
                                                 int ty, int tx) {
#pragma unroll
  for (int i = 0; i < RPT; ++i) {
    const int r = ty + NTY * i;
    const float4 n = *reinterpret_cast<const float4*>(tile + r * kLD + 4 * tx);
    const float s1 = hw_sum((acc[i][0] + acc[i][1]) + (acc[i][2] + acc[i][3])) * (1.0f / 64.0f);
    const float s2 = hw_sum((acc[i][0] * n.x + acc[i][1] * n.y) + (acc[i][2] * n.z + acc[i][3] * n.w)) * (1.0f / 64.0f);
    const float mean = mean_in[r], rstd = rstd_in[r];
    const float sd = 1.0f / rstd;
    // activation derivative from the activation output t = n*sd + mean.  ReLU needs "t == 0" exactly: the
    // forward wrote n = (t - mean) * rstd, so t == 0 <=> n == (0 - mean) * rstd bit for bit.
    const float nz = (0.0f - mean) * rstd;
    float4 da;
    if (relu) {
      da.x = n.x != nz ? 1.0f : 0.0f; da.y = n.y != nz ? 1.0f : 0.0f;
      da.z = n.z != nz ? 1.0f : 0.0f; da.w = n.w != nz ? 1.0f : 0.0f;
    } else {
      const float t0 = fmaf(n.x, sd, mean), t1 = fmaf(n.y, sd, mean), t2 = fmaf(n.z, sd, mean), t3 = fmaf(n.w, sd, mean);
      da.x = 1.0f - t0 * t0; da.y = 1.0f - t1 * t1; da.z = 1.0f - t2 * t2; da.w = 1.0f - t3 * t3;
    }
    float4 dz;
    dz.x = rstd * (acc[i][0] - s1 - n.x * s2) * da.x;
    dz.y = rstd * (acc[i][1] - s1 - n.y * s2) * da.y;
    dz.z = rstd * (acc[i][2] - s1 - n.z * s2) * da.z;
    dz.w = rstd * (acc[i][3] - s1 - n.w * s2) * da.w;
    *reinterpret_cast<float4*>(tile + r * kLD + 4 * tx) = dz;
  }
}

// G[jj][c] += sum_r dz[r][j0 + jj] * n[r][4 tk + c];  db[jj] += sum_r dz[r][j0 + jj] (kept by tk == 0)
template <int BM, int JT>
__device__ __forceinline__ void wgrad_tile(const float* __restrict__ dz, const float* __restrict__ n,
                                           float (&G)[JT][4], float (&db)[JT], int j0, int tk, bool k_active) {
  if (!k_active) return;
#pragma unroll 4
  for (int r = 0; r < BM; ++r) {
    const float4 nv = *reinterpret_cast<const float4*>(n + r * kLD + 4 * tk);
    float d[JT];
#pragma unroll
    for (int q = 0; q < JT / 4; ++q) {
      const float4 v = *reinterpret_cast<const float4*>(dz + r * kLD + j0 + 4 * q);
      d[4 * q] = v.x; d[4 * q + 1] = v.y; d[4 * q + 2] = v.z; d[4 * q + 3] = v.w;
    }
#pragma unroll
    for (int jj = 0; jj < JT; ++jj) {
      G[jj][0] = fmaf(d[jj], nv.x, G[jj][0]); G[jj][1] = fmaf(d[jj], nv.y, G[jj][1]);
      G[jj][2] = fmaf(d[jj], nv.z, G[jj][2]); G[jj][3] = fmaf(d[jj], nv.w, G[jj][3]);
      if (tk == 0) db[jj] += d[jj];
    }
  }
}

// logits of one row from the n2 tile: l[a] = bhf[a] + sum_k n2[r][k] Whf[a][k]
__device__ __forceinline__ void row_logits(const float* __restrict__ n2row, const float* __restrict__ Whf,
                                           const float* __restrict__ bhf, int A, float* l) {
  for (int q = 0; q < A; ++q) l[q] = bhf[q];
#pragma unroll 4
  for (int k4 = 0; k4 < 16; ++k4) {
    const float4 n = *reinterpret_cast<const float4*>(n2row + 4 * k4);
    for (int q = 0; q < A; ++q) {
      const float4 w = *reinterpret_cast<const float4*>(Whf + q * 64 + 4 * k4);
      l[q] = fmaf(n.x, w.x, fmaf(n.y, w.y, fmaf(n.z, w.z, fmaf(n.w, w.w, l[q]))));
    }
  }
}

}  // namespace mappo
