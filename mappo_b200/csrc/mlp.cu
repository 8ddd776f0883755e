// Fused MLP actor / critic kernels (fp32 FMA path): one persistent CTA per SM walks tiles of BM rows.
//   rollout : LayerNorm(obs) -> fc1 -> act -> LN -> fc2 -> act -> LN -> head -> {sample, log-prob | value}
//   update  : the same forward, the PPO loss per row, and the whole backward, all inside the tile, so
//             activations never touch HBM; weight gradients accumulate in registers across the CTA's tiles
//             and leave the kernel once, as one partial row per CTA (summed in a fixed order by optim.cu).
// Reference: mlp.py:6-59, act.py:41-162, distributions.py:14-68, r_actor_critic.py:43-107,146-165,
//            r_mappo.py:52-136 (losses), utils/util.py:23-29 (huber/mse).
// See net.cuh for the LayerNorm-affine folding used here.
#include "mlp_fma.cuh"
#include <stdlib.h>
#include <string.h>

namespace mappo {
namespace tc {   // mlp_tc.cu: the tcgen05 tensor-core tiles
bool supports(const mappo_net_config* cfg);
int grid_for(long long rows);
long long wide_workspace_bytes(long long rows);      // mlp_wide.cu
long long wide_packed_bytes(int in_dim, long long rows);
bool wide_pack_ok(const MlpArgs& a);
int launch_wide_pack(const MlpArgs& a, void* workspace, void* packed, cudaStream_t stream, const char* who);
template <int MODE> int launch(MlpArgs& a, const mappo_net_config* cfg, cudaStream_t stream, const char* who);
}  // namespace tc
}  // namespace mappo

namespace mappo {

// ---- shared memory map (floats) --------------------------------------------------------------------
template <int BM>
struct Smem {
  static constexpr int W1t = 0;                       // [64][64]  W1t[k][j] = W1[j][k] g0[k]   (k < Kp)
  static constexpr int W2t = W1t + 64 * 64;           // [64][64]  W2t[k][j] = W2[j][k] g1[k]
  static constexpr int W2f = W2t + 64 * 64;           // [64][64]  W2f[j][k] = W2[j][k] g1[k]
  static constexpr int Whf = W2f + 64 * 64;           // [16][64]  Whf[a][k] = Wh[a][k] g2[k]
  static constexpr int b1f = Whf + kMaxA * 64;        // [64]
  static constexpr int b2f = b1f + 64;                // [64]
  static constexpr int bhf = b2f + 64;                // [16]
  static constexpr int T0 = bhf + kMaxA;              // [BM][kLD] n0
  static constexpr int T1 = T0 + BM * kLD;            // [BM][kLD] n1, later dz1
  static constexpr int T2 = T1 + BM * kLD;            // [BM][kLD] n2, later dz2
  static constexpr int mean1 = T2 + BM * kLD;         // [BM]
  static constexpr int rstd1 = mean1 + BM;
  static constexpr int mean2 = rstd1 + BM;
  static constexpr int rstd2 = mean2 + BM;
  static constexpr int dlog = rstd2 + BM;             // [BM][kDL]
  static constexpr int total = dlog + BM * kDL;
  static constexpr size_t bytes = sizeof(float) * (size_t)total;
};

// ---- prologue: fold the LayerNorm affines into the following Linear and stage weights in shared memory
template <int BM, bool UPDATE>
__device__ __forceinline__ void stage_weights(float* sm, const MlpArgs& a) {
  using S = Smem<BM>;
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  const int D = L.in_dim, Kp = a.Kp, nt = blockDim.x, tid = threadIdx.x;
  for (int e = tid; e < 64 * 64; e += nt) {
    const int j = e & 63, k = e >> 6;
    float v = 0.0f;
    if (k < D) v = P[L.w1 + j * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
    if (k < Kp) sm[S::W1t + k * 64 + j] = v;
    const float w2 = P[L.w2 + j * 64 + k] * P[L.g1 + k];
    sm[S::W2t + k * 64 + j] = w2;
    if (UPDATE) sm[S::W2f + j * 64 + k] = w2;
  }
  for (int e = tid; e < kMaxA * 64; e += nt) {
    const int aidx = e >> 6, k = e & 63;
    float v = 0.0f;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    if (h < L.n_heads) v = P[L.wh[h] + (aidx - off) * 64 + k] * P[L.g2 + k];
    sm[S::Whf + aidx * 64 + k] = v;
  }
  if (tid < 64) {
    const int j = tid;
    float s1 = P[L.b1 + j];
    if (L.use_fn) for (int k = 0; k < D; ++k) s1 = fmaf(P[L.w1 + j * D + k], P[L.fn_b + k], s1);
    sm[S::b1f + j] = s1;
    float s2 = P[L.b2 + j];
    for (int k = 0; k < 64; ++k) s2 = fmaf(P[L.w2 + j * 64 + k], P[L.be1 + k], s2);
    sm[S::b2f + j] = s2;
  } else if (tid < 64 + kMaxA) {
    const int aidx = tid - 64;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    float s = 0.0f;
    if (h < L.n_heads) {
      s = P[L.bh[h] + (aidx - off)];
      for (int k = 0; k < 64; ++k) s = fmaf(P[L.wh[h] + (aidx - off) * 64 + k], P[L.be2 + k], s);
    }
    sm[S::bhf + aidx] = s;
  }
}

// ---- tile input: gather rows, LayerNorm over the in_dim features (mlp.py:44-45,56-57), zero padding to Kp
template <int BM>
__device__ __forceinline__ void load_input_tile(float* __restrict__ T0, const MlpArgs& a, long long tile) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int D = a.L.in_dim, Kp = a.Kp;
  for (int r = warp; r < BM; r += nw) {
    const long long g = tile * BM + r;
    float v0 = 0.0f, v1 = 0.0f;
    if (g < a.rows) {
      const long long src = a.index ? a.index[g] / a.x_div : g;
      const float* __restrict__ row = a.x + src * (long long)a.ld;
      if (lane < D) v0 = row[lane];
      if (lane + 32 < D) v1 = row[lane + 32];
    }
    if (a.L.use_fn) {
      float s = v0 + v1;
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mean = s / (float)D;
      const float d0 = lane < D ? v0 - mean : 0.0f, d1 = lane + 32 < D ? v1 - mean : 0.0f;
      float q = d0 * d0 + d1 * d1;
      for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float rstd = 1.0f / sqrtf(q / (float)D + kLnEps);
      v0 = d0 * rstd; v1 = d1 * rstd;
    }
    if (lane < Kp) T0[r * kLD + lane] = v0;
    if (lane + 32 < Kp) T0[r * kLD + lane + 32] = v1;
  }
}

// ---- the kernel ------------------------------------------------------------------------------------
// BM rows per tile, NT threads.  Thread (ty, tx): tx = tid % 16 owns 4 output columns, ty = tid / 16 owns rows
// ty + NTY*i.  MODE: 0 rollout actor, 1 rollout critic, 2 update actor, 3 update critic.
template <int BM, int NT, int MODE>
__global__ void __launch_bounds__(NT, 1) mlp_kernel(const MlpArgs a) {
  constexpr bool UPDATE = (MODE >= 2);
  constexpr bool CRITIC = (MODE == 1 || MODE == 3);
  constexpr int NTY = NT / 16, RPT = BM / NTY, JT = 64 / NTY, NQ = (kMaxA * 64 + NT - 1) / NT;
  static_assert(RPT * NTY == BM && JT * NTY == 64 && JT % 4 == 0, "tile shape");
  using S = Smem<BM>;
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const NetLayout& L = a.L;
  const int A = L.a_total, relu = L.use_relu, K4 = a.Kp >> 2;

  stage_weights<BM, UPDATE>(sm, a);

  // persistent gradient accumulators (UPDATE): this thread's 4(j) x 4(k) (or 8 x 4) blocks of G1, G2, its head
  // outputs o = tid + NT*q of G_h[a][k], and the bias sums (kept by tk == 0 / k == 0 threads).
  float G1[JT][4], G2[JT][4], db1[JT], db2[JT], gh[NQ], dbh[NQ];
  double lacc[4] = {0.0, 0.0, 0.0, 0.0};
  if (UPDATE) {
#pragma unroll
    for (int j = 0; j < JT; ++j) {
      db1[j] = db2[j] = 0.0f;
#pragma unroll
      for (int c = 0; c < 4; ++c) G1[j][c] = G2[j][c] = 0.0f;
    }
#pragma unroll
    for (int q = 0; q < NQ; ++q) gh[q] = dbh[q] = 0.0f;
  }
  const int j0 = ty * JT, tk = tx;                       // wgrad block of this thread
  const bool k1_active = (4 * tk < a.Kp);

  for (long long tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
    __syncthreads();                                     // previous tile's readers of T0/T1/T2 are done (and weights staged)
    load_input_tile<BM>(sm + S::T0, a, tile);
    __syncthreads();
    float acc[RPT][4];
    // fc1
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    gemm_tile<RPT, NTY>(sm + S::T0, kLD, K4, sm + S::W1t, acc, ty, tx);
    act_ln_store<RPT, NTY>(acc, sm + S::b1f, relu, sm + S::T1, sm + S::mean1, sm + S::rstd1, ty, tx);
    __syncthreads();
    // fc2
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    gemm_tile<RPT, NTY>(sm + S::T1, kLD, 16, sm + S::W2t, acc, ty, tx);
    act_ln_store<RPT, NTY>(acc, sm + S::b2f, relu, sm + S::T2, sm + S::mean2, sm + S::rstd2, ty, tx);
    __syncthreads();
    // head + per-row loss / sampling: one thread per row
    if (tid < BM) {
      const long long g = tile * BM + tid;
      float l[kMaxA];
      if (g < a.rows) {
        row_logits(sm + S::T2 + tid * kLD, sm + S::Whf, sm + S::bhf, A, l);
        const long long src = a.index ? a.index[g] : g;
        if (MODE == 0) actor_rollout_row(a, g, l);
        if (MODE == 1) for (int m = 0; m < a.group; ++m) a.values_out[g * a.group + m] = l[0];
        if (MODE == 2) actor_loss_row(a, src, l, sm + S::dlog + tid * kDL, lacc);
        if (MODE == 3) {
          const float dv = critic_loss_row(a, src, l[0], lacc);
          float* dl = sm + S::dlog + tid * kDL;
          dl[0] = dv;
          for (int q = 1; q < kMaxA; ++q) dl[q] = 0.0f;
        }
      } else if (UPDATE) {
        float* dl = sm + S::dlog + tid * kDL;
        for (int q = 0; q < kMaxA; ++q) dl[q] = 0.0f;
      }
    }
    if (!UPDATE) continue;
    __syncthreads();
    // head weight gradient: G_h[a][k] += sum_r dlog[r][a] n2[r][k]; output o = a*64 + k
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const int o = tid + NT * q;
      if (o < A * 64) {
        const int aa = o >> 6, k = o & 63;
        float s = 0.0f, sb = 0.0f;
        for (int r = 0; r < BM; ++r) {
          const float d = sm[S::dlog + r * kDL + aa];
          s = fmaf(d, sm[S::T2 + r * kLD + k], s);
          sb += d;
        }
        gh[q] += s;
        if (k == 0) dbh[q] += sb;
      }
    }
    // dn2 = dlog . Whf
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    for (int aa = 0; aa < A; ++aa) {
      const float4 w = *reinterpret_cast<const float4*>(sm + S::Whf + aa * 64 + 4 * tx);
#pragma unroll
      for (int i = 0; i < RPT; ++i) {
        const float d = sm[S::dlog + (ty + NTY * i) * kDL + aa];
        acc[i][0] = fmaf(d, w.x, acc[i][0]); acc[i][1] = fmaf(d, w.y, acc[i][1]);
        acc[i][2] = fmaf(d, w.z, acc[i][2]); acc[i][3] = fmaf(d, w.w, acc[i][3]);
      }
    }
    __syncthreads();                                     // head wgrad finished reading n2
    ln_act_bwd_store<RPT, NTY>(acc, relu, sm + S::T2, sm + S::mean2, sm + S::rstd2, ty, tx);   // T2 := dz2
    __syncthreads();
    // fc2 weight gradient and dgrad
    wgrad_tile<BM, JT>(sm + S::T2, sm + S::T1, G2, db2, j0, tk, true);
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    gemm_tile<RPT, NTY>(sm + S::T2, kLD, 16, sm + S::W2f, acc, ty, tx);                      // dn1 = dz2 . W2'
    __syncthreads();                                     // wgrad2 finished reading n1
    ln_act_bwd_store<RPT, NTY>(acc, relu, sm + S::T1, sm + S::mean1, sm + S::rstd1, ty, tx);   // T1 := dz1
    __syncthreads();
    // fc1 weight gradient (no dgrad into the observations)
    wgrad_tile<BM, JT>(sm + S::T1, sm + S::T0, G1, db1, j0, tk, k1_active);
  }

  if (!UPDATE) return;
  // ---- epilogue: one partial row per CTA, in the flat parameter layout (G in the W slots, db in the b slots)
  float* __restrict__ gp = a.grad_partials + (size_t)blockIdx.x * L.total;
  for (int e = tid; e < L.total; e += NT) gp[e] = 0.0f;   // LayerNorm-affine slots stay zero (filled by optim.cu)
  __syncthreads();
  const int D = L.in_dim;
#pragma unroll
  for (int jj = 0; jj < JT; ++jj) {
    const int j = j0 + jj;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int k = 4 * tk + c;
      if (k < D) gp[L.w1 + j * D + k] = G1[jj][c];
      gp[L.w2 + j * 64 + k] = G2[jj][c];
    }
    if (tk == 0) { gp[L.b1 + j] = db1[jj]; gp[L.b2 + j] = db2[jj]; }
  }
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const int o = tid + NT * q;
    if (o < A * 64) {
      const int aa = o >> 6, k = o & 63;
      int h = 0, off = 0;
      while (aa >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
      gp[L.wh[h] + (aa - off) * 64 + k] = gh[q];
      if (k == 0) gp[L.bh[h] + (aa - off)] = dbh[q];
    }
  }
  // loss sums: block reduction in double
  __shared__ double red[4][32];
  for (int q = 0; q < 4; ++q) {
    double v = lacc[q];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) red[q][tid >> 5] = v;
  }
  __syncthreads();
  if (tid < 4) {
    double v = 0.0;
    for (int w = 0; w < NT / 32; ++w) v += red[tid][w];
    a.loss_partials[(size_t)blockIdx.x * 4 + tid] = v;
  }
  (void)CRITIC;
}

// ---- layer_N > 1 (mlp.py:20-28: fc2 = layer_N hidden blocks Linear -> act -> LayerNorm after fc1) -----------------------
// The same fp32 tile algorithm with NL hidden blocks: tiles T[0 .. NL+1] (input, n after fc1, n after each block), one
// folded weight pair per block, gradients of every block in registers.  BM = 64 rows per tile keeps NL = 3 inside
// 227 KB of shared memory.  Off the tensor-core path on purpose: the reference's MPE scripts all run layer_N = 1.
template <int BM, int NL>
struct SmemDeep {
  static constexpr int W1t = 0;                                   // [64][64]
  static constexpr int Wt = W1t + 64 * 64;                        // [NL][64][64]  Wt[l][k][j] = W_l[j][k] g_prev[k]
  static constexpr int Wf = Wt + NL * 64 * 64;                    // [NL][64][64]  Wf[l][j][k] (dgrad operand)
  static constexpr int Whf = Wf + NL * 64 * 64;                   // [16][64]
  static constexpr int b1f = Whf + kMaxA * 64;                    // [64]
  static constexpr int bf = b1f + 64;                             // [NL][64]
  static constexpr int bhf = bf + NL * 64;                        // [16]
  static constexpr int T = bhf + kMaxA;                           // [NL + 2][BM][kLD]
  static constexpr int mean = T + (NL + 2) * BM * kLD;            // [NL + 1][BM]
  static constexpr int rstd = mean + (NL + 1) * BM;               // [NL + 1][BM]
  static constexpr int dlog = rstd + (NL + 1) * BM;               // [BM][kDL]
  static constexpr int total = dlog + BM * kDL;
  static constexpr size_t bytes = sizeof(float) * (size_t)total;
};

template <int BM, int NL, bool UPDATE>
__device__ __forceinline__ void stage_weights_deep(float* sm, const MlpArgs& a) {
  using S = SmemDeep<BM, NL>;
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  const int D = L.in_dim, Kp = a.Kp, nt = blockDim.x, tid = threadIdx.x;
  for (int e = tid; e < 64 * 64; e += nt) {
    const int j = e & 63, k = e >> 6;
    float v = 0.0f;
    if (k < D) v = P[L.w1 + j * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
    if (k < Kp) sm[S::W1t + k * 64 + j] = v;
    for (int l = 0; l < NL; ++l) {
      const int gprev = l == 0 ? L.g1 : layer_g(L, l - 1);
      const float w = P[layer_w(L, l) + j * 64 + k] * P[gprev + k];
      sm[S::Wt + l * 4096 + k * 64 + j] = w;
      if (UPDATE) sm[S::Wf + l * 4096 + j * 64 + k] = w;
    }
  }
  const int glast = layer_g(L, NL - 1), belast = layer_be(L, NL - 1);
  for (int e = tid; e < kMaxA * 64; e += nt) {
    const int aidx = e >> 6, k = e & 63;
    float v = 0.0f;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    if (h < L.n_heads) v = P[L.wh[h] + (aidx - off) * 64 + k] * P[glast + k];
    sm[S::Whf + aidx * 64 + k] = v;
  }
  if (tid < 64) {
    const int j = tid;
    float s1 = P[L.b1 + j];
    if (L.use_fn) for (int k = 0; k < D; ++k) s1 = fmaf(P[L.w1 + j * D + k], P[L.fn_b + k], s1);
    sm[S::b1f + j] = s1;
    for (int l = 0; l < NL; ++l) {
      const int beprev = l == 0 ? L.be1 : layer_be(L, l - 1);
      float s2 = P[layer_b(L, l) + j];
      for (int k = 0; k < 64; ++k) s2 = fmaf(P[layer_w(L, l) + j * 64 + k], P[beprev + k], s2);
      sm[S::bf + l * 64 + j] = s2;
    }
  } else if (tid < 64 + kMaxA) {
    const int aidx = tid - 64;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    float s = 0.0f;
    if (h < L.n_heads) {
      s = P[L.bh[h] + (aidx - off)];
      for (int k = 0; k < 64; ++k) s = fmaf(P[L.wh[h] + (aidx - off) * 64 + k], P[belast + k], s);
    }
    sm[S::bhf + aidx] = s;
  }
}

template <int BM, int NT, int MODE, int NL>
__global__ void __launch_bounds__(NT, 1) mlp_deep_kernel(const MlpArgs a) {
  constexpr bool UPDATE = (MODE >= 2);
  constexpr int NTY = NT / 16, RPT = BM / NTY, JT = 64 / NTY, NQ = (kMaxA * 64 + NT - 1) / NT;
  static_assert(RPT * NTY == BM && JT * NTY == 64 && JT % 4 == 0, "tile shape");
  using S = SmemDeep<BM, NL>;
  extern __shared__ __align__(16) float sm[];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const NetLayout& L = a.L;
  const int A = L.a_total, relu = L.use_relu, K4 = a.Kp >> 2;
  auto tile_of = [&](int i) { return sm + S::T + i * (BM * kLD); };
  auto mean_of = [&](int i) { return sm + S::mean + i * BM; };
  auto rstd_of = [&](int i) { return sm + S::rstd + i * BM; };

  stage_weights_deep<BM, NL, UPDATE>(sm, a);

  float G1[JT][4], db1[JT], Gx[NL][JT][4], dbx[NL][JT], gh[NQ], dbh[NQ];
  double lacc[4] = {0.0, 0.0, 0.0, 0.0};
  if (UPDATE) {
#pragma unroll
    for (int j = 0; j < JT; ++j) {
      db1[j] = 0.0f;
#pragma unroll
      for (int c = 0; c < 4; ++c) G1[j][c] = 0.0f;
#pragma unroll
      for (int l = 0; l < NL; ++l) {
        dbx[l][j] = 0.0f;
#pragma unroll
        for (int c = 0; c < 4; ++c) Gx[l][j][c] = 0.0f;
      }
    }
#pragma unroll
    for (int q = 0; q < NQ; ++q) gh[q] = dbh[q] = 0.0f;
  }
  const int j0 = ty * JT, tk = tx;
  const bool k1_active = (4 * tk < a.Kp);

  for (long long tile = blockIdx.x; tile < a.ntiles; tile += gridDim.x) {
    __syncthreads();
    load_input_tile<BM>(tile_of(0), a, tile);
    __syncthreads();
    float acc[RPT][4];
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    gemm_tile<RPT, NTY>(tile_of(0), kLD, K4, sm + S::W1t, acc, ty, tx);
    act_ln_store<RPT, NTY>(acc, sm + S::b1f, relu, tile_of(1), mean_of(0), rstd_of(0), ty, tx);
    __syncthreads();
#pragma unroll
    for (int l = 0; l < NL; ++l) {                       // hidden blocks (mlp.py:26-28)
#pragma unroll
      for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
      gemm_tile<RPT, NTY>(tile_of(1 + l), kLD, 16, sm + S::Wt + l * 4096, acc, ty, tx);
      act_ln_store<RPT, NTY>(acc, sm + S::bf + l * 64, relu, tile_of(2 + l), mean_of(1 + l), rstd_of(1 + l), ty, tx);
      __syncthreads();
    }
    float* Tlast = tile_of(NL + 1);
    if (tid < BM) {
      const long long g = tile * BM + tid;
      float l[kMaxA];
      if (g < a.rows) {
        row_logits(Tlast + tid * kLD, sm + S::Whf, sm + S::bhf, A, l);
        const long long src = a.index ? a.index[g] : g;
        if (MODE == 0) actor_rollout_row(a, g, l);
        if (MODE == 1) for (int m = 0; m < a.group; ++m) a.values_out[g * a.group + m] = l[0];
        if (MODE == 2) actor_loss_row(a, src, l, sm + S::dlog + tid * kDL, lacc);
        if (MODE == 3) {
          const float dv = critic_loss_row(a, src, l[0], lacc);
          float* dl = sm + S::dlog + tid * kDL;
          dl[0] = dv;
          for (int q = 1; q < kMaxA; ++q) dl[q] = 0.0f;
        }
      } else if (UPDATE) {
        float* dl = sm + S::dlog + tid * kDL;
        for (int q = 0; q < kMaxA; ++q) dl[q] = 0.0f;
      }
    }
    if (!UPDATE) continue;
    __syncthreads();
    // head weight gradient and dn of the last block
#pragma unroll
    for (int q = 0; q < NQ; ++q) {
      const int o = tid + NT * q;
      if (o < A * 64) {
        const int aa = o >> 6, k = o & 63;
        float s = 0.0f, sb = 0.0f;
        for (int r = 0; r < BM; ++r) {
          const float d = sm[S::dlog + r * kDL + aa];
          s = fmaf(d, Tlast[r * kLD + k], s);
          sb += d;
        }
        gh[q] += s;
        if (k == 0) dbh[q] += sb;
      }
    }
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    for (int aa = 0; aa < A; ++aa) {
      const float4 w = *reinterpret_cast<const float4*>(sm + S::Whf + aa * 64 + 4 * tx);
#pragma unroll
      for (int i = 0; i < RPT; ++i) {
        const float d = sm[S::dlog + (ty + NTY * i) * kDL + aa];
        acc[i][0] = fmaf(d, w.x, acc[i][0]); acc[i][1] = fmaf(d, w.y, acc[i][1]);
        acc[i][2] = fmaf(d, w.z, acc[i][2]); acc[i][3] = fmaf(d, w.w, acc[i][3]);
      }
    }
    __syncthreads();                                     // head wgrad finished reading the last n tile
    ln_act_bwd_store<RPT, NTY>(acc, relu, Tlast, mean_of(NL), rstd_of(NL), ty, tx);     // T[NL+1] := dz of block NL-1
    __syncthreads();
#pragma unroll
    for (int l = NL - 1; l >= 0; --l) {
      // block l: weight gradient from (dz_l, its input n) and the data gradient into that input
      wgrad_tile<BM, JT>(tile_of(2 + l), tile_of(1 + l), Gx[l], dbx[l], j0, tk, true);
#pragma unroll
      for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
      gemm_tile<RPT, NTY>(tile_of(2 + l), kLD, 16, sm + S::Wf + l * 4096, acc, ty, tx);
      __syncthreads();                                   // the weight gradient finished reading the input tile
      ln_act_bwd_store<RPT, NTY>(acc, relu, tile_of(1 + l), mean_of(l), rstd_of(l), ty, tx);   // := dz of the producer
      __syncthreads();
    }
    wgrad_tile<BM, JT>(tile_of(1), tile_of(0), G1, db1, j0, tk, k1_active);             // fc1 (no dgrad into the inputs)
  }

  if (!UPDATE) return;
  float* __restrict__ gp = a.grad_partials + (size_t)blockIdx.x * L.total;
  for (int e = tid; e < L.total; e += NT) gp[e] = 0.0f;
  __syncthreads();
  const int D = L.in_dim;
#pragma unroll
  for (int jj = 0; jj < JT; ++jj) {
    const int j = j0 + jj;
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int k = 4 * tk + c;
      if (k < D) gp[L.w1 + j * D + k] = G1[jj][c];
#pragma unroll
      for (int l = 0; l < NL; ++l) gp[layer_w(L, l) + j * 64 + k] = Gx[l][jj][c];
    }
    if (tk == 0) {
      gp[L.b1 + j] = db1[jj];
#pragma unroll
      for (int l = 0; l < NL; ++l) gp[layer_b(L, l) + j] = dbx[l][jj];
    }
  }
#pragma unroll
  for (int q = 0; q < NQ; ++q) {
    const int o = tid + NT * q;
    if (o < A * 64) {
      const int aa = o >> 6, k = o & 63;
      int h = 0, off = 0;
      while (aa >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
      gp[L.wh[h] + (aa - off) * 64 + k] = gh[q];
      if (k == 0) gp[L.bh[h] + (aa - off)] = dbh[q];
    }
  }
  __shared__ double red[4][32];
  for (int q = 0; q < 4; ++q) {
    double v = lacc[q];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if ((tid & 31) == 0) red[q][tid >> 5] = v;
  }
  __syncthreads();
  if (tid < 4) {
    double v = 0.0;
    for (int w = 0; w < NT / 32; ++w) v += red[tid][w];
    a.loss_partials[(size_t)blockIdx.x * 4 + tid] = v;
  }
}

constexpr int kDeepBM = 64;

static int blocks_for_deep(long long rows) {
  const long long ntiles = (rows + kDeepBM - 1) / kDeepBM;
  int dev = 0, sms = 148;
  if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  return (int)(ntiles < sms ? (ntiles < 1 ? 1 : ntiles) : sms);
}

template <int MODE, int NL>
static int launch_deep_nl(MlpArgs& a, cudaStream_t stream, const char* who) {
  using S = SmemDeep<kDeepBM, NL>;
  auto kern = mlp_deep_kernel<kDeepBM, 256, MODE, NL>;
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)S::bytes));
    attr_set = true;
  }
  kern<<<blocks_for_deep(a.rows), 256, S::bytes, stream>>>(a);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

template <int MODE>
static int launch_deep(MlpArgs& a, const mappo_net_config* cfg, cudaStream_t stream, const char* who) {
  a.L = make_layout(*cfg);
  if (a.x_div < 1) a.x_div = 1;
  a.Kp = (cfg->in_dim + 3) & ~3;
  a.ntiles = (a.rows + kDeepBM - 1) / kDeepBM;
  if (cfg->layer_N == 2) return launch_deep_nl<MODE, 2>(a, stream, who);
  return launch_deep_nl<MODE, 3>(a, stream, who);
}

// ---- host side -------------------------------------------------------------------------------------
constexpr int kBM = 128, kNT = 256;

static int device_sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
}

static int blocks_for(long long rows) {
  const long long ntiles = (rows + kBM - 1) / kBM;
  const int sms = device_sm_count();
  return (int)(ntiles < sms ? (ntiles < 1 ? 1 : ntiles) : sms);
}

// MAPPO_MLP_IMPL=fma selects the fp32 FMA tiles of this file (kept for A/B measurements); default: tensor cores
static bool use_tc(const mappo_net_config* cfg) {
  static int forced_fma = -1;
  if (forced_fma < 0) {
    const char* e = getenv("MAPPO_MLP_IMPL");
    forced_fma = (e && strcmp(e, "fma") == 0) ? 1 : 0;
  }
  return !forced_fma && tc::supports(cfg);
}

template <int MODE>
static int launch_mlp(MlpArgs& a, const mappo_net_config* cfg, cudaStream_t stream, const char* who) {
  if (int rc = check_net(cfg, who)) return rc;
  if (cfg->recurrent) MAPPO_UNSUPPORTED("%s: recurrent (GRU) networks are not built yet", who);
  MAPPO_REQUIRE(a.params && a.x && a.rows > 0 && a.ld >= cfg->in_dim, "%s: bad input arguments", who);
  if (cfg->layer_N > 1) return launch_deep<MODE>(a, cfg, stream, who);      // extra hidden blocks: fp32 tile kernel
  if (use_tc(cfg)) return tc::launch<MODE>(a, cfg, stream, who);
  if (cfg->in_dim > kMaxIn) MAPPO_UNSUPPORTED("%s: in_dim %d > %d is built on the tensor-core path only", who, cfg->in_dim, kMaxIn);
  a.L = make_layout(*cfg);
  if (a.x_div < 1) a.x_div = 1;
  a.Kp = (cfg->in_dim + 3) & ~3;
  a.ntiles = (a.rows + kBM - 1) / kBM;
  const int grid = blocks_for(a.rows);
  auto kern = mlp_kernel<kBM, kNT, MODE>;
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)Smem<kBM>::bytes));
    attr_set = true;
  }
  kern<<<grid, kNT, Smem<kBM>::bytes, stream>>>(a);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

}  // namespace mappo

using namespace mappo;

extern "C" long long mappo_net_param_count(const mappo_net_config* cfg) {
  if (!cfg) return -1;
  return make_layout(*cfg).total;
}

extern "C" int mappo_update_num_blocks(const mappo_net_config* cfg, long long rows) {
  if (cfg && cfg->layer_N > 1) return blocks_for_deep(rows);
  if (use_tc(cfg) || (cfg && cfg->recurrent)) return 2 * tc::grid_for(rows);   // two partial rows per CTA (hi / lo operand halves)
  return blocks_for(rows);
}

extern "C" int mappo_mlp_kernel_path(const mappo_net_config* cfg) {
  if (!cfg || check_net(cfg, "mappo_mlp_kernel_path")) return -1;
  if (cfg->recurrent) return 3;
  if (cfg->layer_N > 1 || !use_tc(cfg)) return 0;
  return cfg->in_dim >= kWideFrom ? 2 : 1;
}

extern "C" long long mappo_mlp_workspace_bytes(const mappo_net_config* cfg, long long rows) {
  if (!cfg || rows < 0) return -1;
  return (cfg->in_dim >= kWideFrom && use_tc(cfg)) ? tc::wide_workspace_bytes(rows) : 0;
}

extern "C" int mappo_actor_forward(const mappo_net_config* cfg, const float* params, const float* obs, int obs_ld,
                                   long long rows, const float* available_actions, int deterministic, uint64_t seed,
                                   uint64_t step, long long row_offset, const float* given_actions,
                                   float* actions_out, float* logp_out, float* entropy_out,
                                   const int64_t* step_offset, void* workspace, void* stream) {
  MAPPO_REQUIRE(logp_out && (actions_out || given_actions), "actor_forward: null outputs");
  MAPPO_REQUIRE(!cfg || !cfg->is_critic, "actor_forward: critic config");
  MlpArgs a{};
  a.params = params; a.x = obs; a.ld = obs_ld; a.rows = rows; a.avail = available_actions;
  a.deterministic = deterministic; a.k0 = (uint32_t)(seed & 0xFFFFFFFFull); a.step = (uint32_t)step;
  a.step_dev = step_offset;
  a.row_offset = row_offset; a.actions_out = actions_out; a.logp_out = logp_out;
  a.given_actions = given_actions; a.entropy_out = entropy_out; a.workspace = workspace;
  return launch_mlp<0>(a, cfg, (cudaStream_t)stream, "actor_forward");
}

extern "C" int mappo_critic_forward(const mappo_net_config* cfg, const float* params, const float* share_obs,
                                    int obs_ld, long long rows, int group, float* values_out, void* workspace,
                                    void* stream) {
  MAPPO_REQUIRE(values_out && group >= 1, "critic_forward: bad outputs");
  MAPPO_REQUIRE(!cfg || cfg->is_critic, "critic_forward: actor config");
  MlpArgs a{};
  a.params = params; a.x = share_obs; a.ld = obs_ld; a.rows = rows; a.group = group; a.values_out = values_out;
  a.workspace = workspace;
  return launch_mlp<1>(a, cfg, (cudaStream_t)stream, "critic_forward");
}

extern "C" int mappo_actor_update(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                                  const float* obs, int obs_ld, long long rows, const int64_t* index,
                                  const float* actions, const float* old_logp, const float* adv,
                                  const float* active_masks, const float* available_actions, float* grad_partials,
                                  double* loss_partials, void* workspace, void* stream) {
  MAPPO_REQUIRE(loss && actions && old_logp && adv && grad_partials && loss_partials, "actor_update: null pointer");
  MAPPO_REQUIRE(!cfg || !cfg->is_critic, "actor_update: critic config");
  MlpArgs a{};
  a.params = params; a.x = obs; a.ld = obs_ld; a.rows = rows; a.index = index; a.actions = actions;
  a.old_logp = old_logp; a.adv = adv; a.active = active_masks; a.avail = available_actions; a.loss = *loss;
  a.grad_partials = grad_partials; a.loss_partials = loss_partials; a.workspace = workspace;
  return launch_mlp<2>(a, cfg, (cudaStream_t)stream, "actor_update");
}

extern "C" int mappo_critic_update(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                                   const float* share_obs, int obs_ld, long long rows, int group,
                                   const int64_t* index, int share_rows_div, const float* value_preds,
                                   const float* returns,
                                   const float* active_masks, const float* vn_stats, float* grad_partials,
                                   double* loss_partials, void* workspace, const void* packed_inputs, void* stream) {
  MAPPO_REQUIRE(loss && value_preds && returns && grad_partials && loss_partials && group >= 1,
                "critic_update: null pointer");
  MAPPO_REQUIRE(!cfg || cfg->is_critic, "critic_update: actor config");
  MAPPO_REQUIRE(!(index && group != 1), "critic_update: a row gather needs group == 1");
  MAPPO_REQUIRE(share_rows_div >= 1, "critic_update: share_rows_div");
  MAPPO_REQUIRE(!loss->use_value_norm || vn_stats, "critic_update: use_value_norm without statistics");
  MlpArgs a{};
  a.params = params; a.x = share_obs; a.ld = obs_ld; a.rows = rows; a.group = group; a.index = index;
  a.x_div = share_rows_div;
  a.value_preds = value_preds; a.returns = returns; a.active = active_masks; a.vn_stats = vn_stats; a.loss = *loss;
  a.grad_partials = grad_partials; a.loss_partials = loss_partials; a.workspace = workspace;
  a.packed = packed_inputs;
  MAPPO_REQUIRE(!packed_inputs || (!index && share_rows_div == 1 && cfg->in_dim >= kWideFrom && use_tc(cfg)),
                "critic_update: packed inputs need the wide first layer and no row gather");
  return launch_mlp<3>(a, cfg, (cudaStream_t)stream, "critic_update");
}

extern "C" long long mappo_critic_packed_bytes(const mappo_net_config* cfg, long long rows) {
  if (!cfg || rows < 0) return -1;
  if (!(cfg->in_dim >= kWideFrom && use_tc(cfg))) return 0;
  MlpArgs a{};
  a.L = make_layout(*cfg); a.x_div = 1;
  return tc::wide_pack_ok(a) ? tc::wide_packed_bytes(cfg->in_dim, rows) : 0;
}

extern "C" int mappo_critic_pack_inputs(const mappo_net_config* cfg, const float* share_obs, int obs_ld, long long rows,
                                        void* packed, void* workspace, void* stream) {
  if (int rc = check_net(cfg, "critic_pack_inputs")) return rc;
  MAPPO_REQUIRE(cfg->is_critic && cfg->in_dim >= kWideFrom && use_tc(cfg), "critic_pack_inputs: only the wide first layer packs its inputs");
  MAPPO_REQUIRE(share_obs && packed && workspace && rows > 0 && obs_ld >= cfg->in_dim, "critic_pack_inputs: bad arguments");
  MlpArgs a{};
  a.L = make_layout(*cfg); a.x = share_obs; a.ld = obs_ld; a.rows = rows; a.x_div = 1;
  return tc::launch_wide_pack(a, workspace, packed, (cudaStream_t)stream, "critic_pack_inputs");
}
