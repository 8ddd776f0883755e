// Fused MLP actor / critic kernels on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a only.
//
// One persistent CTA per SM (512 threads) walks tiles of 128 rows.  Every dense contraction of the tile -- fc1, fc2,
// the data gradient through fc2, and the three weight gradients -- is a `tcgen05.mma.kind::f16` on fp16 operand
// PAIRS: each fp32 operand x is scaled by a power of two S fixed per operand class (mlp_tc_common.cuh: 64 for
// LayerNorm outputs and weights, 16 for input rows and per-sample gradients) and carried as hi = fp16(S x) and
// lo = fp16(S x - hi) -- 22 significant bits, an absolute resolution of 2^-25 / S, valid for |x| < 65504 / S.  Both
// halves carry the same scale, so the products hi*hi + hi*lo + lo*hi accumulate in ONE fp32 tensor-memory
// accumulator whose power-of-two scale the epilogue removes inside its first FFMA; the weight gradients stack the
// hi and lo halves of A along M and of B along N (four equally scaled blocks, added when they leave the kernel).
// Operands live in shared memory in the chunk-major layout of umma.cuh, where the same bytes serve as the K-major
// operand of the forward / data-gradient GEMMs and as the MN-major operand of the weight-gradient GEMMs, so every
// activation is written to shared memory exactly once.  Everything between the GEMMs (bias, tanh / ReLU,
// LayerNorm and their derivatives, the categorical head, the PPO losses) runs in fp32 registers: thread
// (row r, column quarter c) owns 16 of the 64 hidden columns of its row and reads them straight out of TMEM.
// Weight gradients accumulate in TMEM across all tiles of the CTA and leave the kernel once.
//
// Reference: mlp.py:6-59, act.py:41-162, distributions.py:14-68, r_actor_critic.py:43-107,146-165,
//            r_mappo.py:52-136, utils/util.py:23-29.  LayerNorm-affine folding: net.cuh.
#include "mlp_tc_common.cuh"

namespace mappo {
namespace tc {

constexpr int kTmemCols = 512;
constexpr int kRowpGroup = 4;                       // largest critic `group` whose loss inputs arrive by bulk copy
constexpr float kRawMax = 4094.0f;                  // 65504 / kSX, rounded down to an fp16-exact value after scaling
// TMEM column map.  The BIAS gradients sum_rows dz come off the tensor cores too: the n1 operand tile ends in a constant
// chunk whose first column is 1 (times the tile's scale; column 128 of G2), and the input tile carries a (scaled) 1 in its
// first padding column (column in_dim of G1; W1' is zero there, so the forward pass does not see it).
constexpr uint32_t kColAcc = 0, kColF = 64, kColG2 = 128, kColG1 = 272, kColGh = 400;   // 64 | 64 | 144 | <= 128 | 32

// shared-memory map (byte offsets); Kp = in_dim rounded up to 32 or 64, AS = odd stride of the partial-logit rows
struct Map {
  uint32_t X, N1, BS, BS2, W1, W2, DL, Whf, b1f, b2f, b1s, b2s, bhf, dbh, xch, xlog, red, bars, rowp, raw, total;
  uint32_t x_piece, w1_piece, x_buf, n1_lo;
  int xlog_in_x;                 // 1: the head's partial logits live in the idle X buffer (see below), `xlog` unused
};
// The partial logits [4 column quarters][128 rows][AS] are written after the second layer and read before the
// weight-gradient group is issued; in the update kernels the OTHER X buffer is idle over exactly that span (the previous
// tile's fc1 weight gradient has completed, the next tile's rows are staged only after the block-wide barrier that
// follows the last read), so they borrow it whenever they fit.  The rollout kernels have no BS / BS2 / DL / dbh regions
// and keep a private one.
// raw_floats > 0: a landing buffer for the NEXT tile's input rows as one bulk (TMA) copy -- used when the rows of a tile
// are contiguous in global memory (no gather, row stride = in_dim), so that no thread issues row-divergent loads.
// vhead: scalar (value) head whose weight gradient accumulates in registers: no n2 / dlogits operand tiles.
// early: the NEXT tile's fc1 is issued at the second synchronisation of the tile (see the kernel): both X buffers are busy
// over the span the partial logits live, so they get a private region.
// rowp_floats > 0: a landing buffer for a tile's per-row loss scalars (see rowp_ok in the kernel), when it fits.
__host__ __device__ inline Map make_map(int Kp, int AS, bool update, int raw_floats = 0, bool vhead = false, bool early = false,
                                        int rowp_floats = 0) {
  Map m;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.x_piece = (uint32_t)(Kp / 8) * kRowChunk;
  m.w1_piece = (uint32_t)(Kp / 8) * kWChunk;
  const uint32_t ones = update ? 2 * kRowChunk : 0;  // the constant [1 0 .. 0] columns behind the n1 tile (bias gradient)
  m.x_buf = kNP * m.x_piece;
  m.X = take(2 * m.x_buf);                         // two buffers: this tile's rows and the next tile's
  m.N1 = take(kNP * 8 * kRowChunk + ones);          // [hi | lo | ones]: ONE B operand of the fc2 weight gradient (N = 144)
  m.n1_lo = 8 * kRowChunk;
  m.BS = take(update ? kNP * 8 * kRowChunk : 0);    // dz2, then dz1
  m.BS2 = take(update && !vhead ? kNP * 8 * kRowChunk : 0);   // n2 (operand of the head weight gradient)
  m.W1 = take(kNP * m.w1_piece);
  m.W2 = take(kNP * 8 * kWChunk);
  m.DL = take(update && !vhead ? kNP * 2 * kRowChunk : 0);
  m.Whf = take(kMaxA * 64 * 4);
  m.b1f = take(64 * 4);
  m.b2f = take(64 * 4);
  m.b1s = take(64 * 4);                             // biases times 2 log2(e): tanh straight from the accumulator
  m.b2s = take(64 * 4);
  m.bhf = take(2 * kMaxA * 4);                       // head biases (folded), then the row sums of the folded head weights
  m.dbh = take(update ? kBM * AS * 4 : 0);
  m.xch = take(2 * 4 * kBM * 8);
  m.xlog_in_x = (update && !early && (uint32_t)(4 * kBM * AS * 4) <= kNP * m.x_piece) ? 1 : 0;
  m.xlog = m.xlog_in_x ? 0u : take(4 * kBM * AS * 4);
  m.red = take(16 * 16 * 4 * 3);
  m.bars = take(64);
  m.rowp = take((uint32_t)rowp_floats * 4);          // actor [4][128], critic [3][128 * group]
  m.raw = take((uint32_t)raw_floats * 4);
  m.total = o;
  return m;
}

// ---- prologue: fold the LayerNorm affines into the following Linear (net.cuh), split into fp16 pairs, stage.
// The plain form, kept for the update kernels: their main loop's code generation is sensitive to what precedes it
// (the spread-out form below made the actor update 1.4 % slower), and 9 us of prologue are 0.5 % of their run time.
__device__ __forceinline__ void stage_weights_serial(unsigned char* sm, const Map& m, const MlpArgs& a, bool with_w1, bool update) {
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  const int D = L.in_dim, Kp = a.Kp, tid = threadIdx.x;
  for (int e = tid; with_w1 && e < 64 * Kp; e += kThreads) {
    const int j = e / Kp, k = e - j * Kp;
    float v = 0.0f;
    if (k < D) v = P[L.w1 + j * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
    put_pair_u(sm + m.W1, m.w1_piece, 64, j, k, v * kSW);
  }
  for (int e = tid; e < 64 * 64; e += kThreads) {
    const int j = e >> 6, k = e & 63;
    put_pair_u(sm + m.W2, 8 * kWChunk, 64, j, k, P[L.w2 + j * 64 + k] * P[L.g1 + k] * kSW);
  }
  float* Whf = reinterpret_cast<float*>(sm + m.Whf);
  for (int e = tid; e < kMaxA * 64; e += kThreads) {
    const int aidx = e >> 6, k = e & 63;
    float v = 0.0f;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    if (h < L.n_heads) v = P[L.wh[h] + (aidx - off) * 64 + k] * P[L.g2 + k];
    Whf[aidx * 64 + k] = v;
  }
  if (tid < 64) {
    const int j = tid;
    float s1 = P[L.b1 + j];
    if (L.use_fn) for (int k = 0; k < D; ++k) s1 = fmaf(P[L.w1 + j * D + k], P[L.fn_b + k], s1);
    reinterpret_cast<float*>(sm + m.b1f)[j] = s1;
    reinterpret_cast<float*>(sm + m.b1s)[j] = s1 * kTanhScale;
    float s2 = P[L.b2 + j];
    for (int k = 0; k < 64; ++k) s2 = fmaf(P[L.w2 + j * 64 + k], P[L.be1 + k], s2);
    reinterpret_cast<float*>(sm + m.b2f)[j] = s2;
    reinterpret_cast<float*>(sm + m.b2s)[j] = s2 * kTanhScale;
  } else if (tid < 64 + kMaxA) {
    const int aidx = tid - 64;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    float s = 0.0f, sw = 0.0f;
    if (h < L.n_heads) {
      s = P[L.bh[h] + (aidx - off)];
      for (int k = 0; k < 64; ++k) {
        const float w = P[L.wh[h] + (aidx - off) * 64 + k];
        s = fmaf(w, P[L.be2 + k], s);
        sw = fmaf(w, P[L.g2 + k], sw);
      }
    }
    reinterpret_cast<float*>(sm + m.bhf)[aidx] = s;
    reinterpret_cast<float*>(sm + m.bhf)[kMaxA + aidx] = sw;
  }
  float* dbh = reinterpret_cast<float*>(sm + m.dbh);
  for (int e = tid; update && e < kBM * (L.a_total | 1); e += kThreads) dbh[e] = 0.0f;
}

// ---- the same for the rollout kernels, whose launches are short (70 / 30 us) and many (51 per iteration).
// Every launch of every CTA pays for it, and all of it is L2 latency: the loops have compile-time trip counts so that a
// thread's loads go out together, and the three bias folds (64 + 64 + 16 dot products) are spread over all 16 warps
// -- two elements per lane and a shuffle reduction -- instead of 80 threads walking 64 dependent steps each.
template <int KP, bool SWF>
__device__ __forceinline__ void stage_weights_tc(unsigned char* sm, const Map& m, const MlpArgs& a, bool with_w1, bool update) {
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  const int D = L.in_dim, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (with_w1) {
#pragma unroll
    for (int it = 0; it < 64 * KP / kThreads; ++it) {
      const int e = tid + it * kThreads, j = e / KP, k = e % KP;
      float v = 0.0f;
      if (k < D) v = P[L.w1 + j * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
      put_pair_u(sm + m.W1, m.w1_piece, 64, j, k, v * kSW);
    }
  }
#pragma unroll
  for (int it = 0; it < 64 * 64 / kThreads; ++it) {
    const int e = tid + it * kThreads, j = e >> 6, k = e & 63;
    put_pair_u(sm + m.W2, 8 * kWChunk, 64, j, k, P[L.w2 + j * 64 + k] * P[L.g1 + k] * kSW);
  }
  float* Whf = reinterpret_cast<float*>(sm + m.Whf);
#pragma unroll
  for (int it = 0; it < kMaxA * 64 / kThreads; ++it) {
    const int e = tid + it * kThreads, aidx = e >> 6, k = e & 63;
    float v = 0.0f;
    int h = 0, off = 0;
    while (h < L.n_heads && aidx >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
    if (h < L.n_heads) v = P[L.wh[h] + (aidx - off) * 64 + k] * P[L.g2 + k];
    Whf[aidx * 64 + k] = v;
  }
  auto warp_sum = [](float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
  };
  // b1' = b1 + W1 beta_fn, b2' = b2 + W2 beta_1: warp w folds rows 4w .. 4w + 3
  float p1[4], p2[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int j = 4 * warp + i;
    p1[i] = 0.0f;
    if (L.use_fn) {                                    // (a wide first layer, mlp_wide.cu, leaves its bias to this kernel too)
      for (int k = lane; k < D; k += 32) p1[i] = fmaf(P[L.w1 + j * D + k], P[L.fn_b + k], p1[i]);
    }
    p2[i] = fmaf(P[L.w2 + j * 64 + lane + 32], P[L.be1 + lane + 32], P[L.w2 + j * 64 + lane] * P[L.be1 + lane]);
  }
  // head biases: warp w folds logit w
  float ph = 0.0f, pw = 0.0f;
  int hh = 0, hoff = 0;
  while (hh < L.n_heads && warp >= hoff + L.head_dim[hh]) { hoff += L.head_dim[hh]; ++hh; }
  if (hh < L.n_heads) {
    const float* wr = P + L.wh[hh] + (warp - hoff) * 64;
    ph = fmaf(wr[lane + 32], P[L.be2 + lane + 32], wr[lane] * P[L.be2 + lane]);
    if (SWF) pw = fmaf(wr[lane + 32], P[L.g2 + lane + 32], wr[lane] * P[L.g2 + lane]);   // only the fused head exchange reads it
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) { p1[i] = warp_sum(p1[i]); p2[i] = warp_sum(p2[i]); }
  ph = warp_sum(ph);
  if (SWF) pw = warp_sum(pw);
  if (lane < 4) {
    const int j = 4 * warp + lane;
    const float s1 = P[L.b1 + j] + (lane == 0 ? p1[0] : lane == 1 ? p1[1] : lane == 2 ? p1[2] : p1[3]);
    const float s2 = P[L.b2 + j] + (lane == 0 ? p2[0] : lane == 1 ? p2[1] : lane == 2 ? p2[2] : p2[3]);
    reinterpret_cast<float*>(sm + m.b1f)[j] = s1;
    reinterpret_cast<float*>(sm + m.b1s)[j] = s1 * kTanhScale;
    reinterpret_cast<float*>(sm + m.b2f)[j] = s2;
    reinterpret_cast<float*>(sm + m.b2s)[j] = s2 * kTanhScale;
  }
  if (lane == 0) {
    reinterpret_cast<float*>(sm + m.bhf)[warp] = (hh < L.n_heads) ? P[L.bh[hh] + (warp - hoff)] + ph : 0.0f;
    reinterpret_cast<float*>(sm + m.bhf)[kMaxA + warp] = pw;
  }
  float* dbh = reinterpret_cast<float*>(sm + m.dbh);
  for (int e = tid; update && e < kBM * (L.a_total | 1); e += kThreads) dbh[e] = 0.0f;
}

// ---- MMA issue (one elected thread) ---------------------------------------------------------------------
// Operand addresses and strides arrive in 16-BYTE UNITS (shared-memory base >> 4 plus compile-time offsets), so a
// descriptor costs one uniform add (umma::smem_desc16).
constexpr uint32_t kRowChunk16 = kRowChunk >> 4, kWChunk16 = kWChunk >> 4;
// forward-type GEMM: ACC[128][64] = A_hi . B_hi^T + A_hi . B_lo^T + A_lo . B_hi^T; A K-major 128-row tile
template <int KSTEPS>
__device__ __forceinline__ void issue_rowgemm(uint32_t acc, uint32_t a16, uint32_t a_piece16,
                                              uint32_t b16, uint32_t b_piece16, uint32_t b_lbo, uint32_t b_sbo,
                                              uint32_t b_kstep16, uint32_t idesc) {
#pragma unroll
  for (int ks = 0; ks < KSTEPS; ++ks) {
    const uint32_t ao = a16 + (uint32_t)ks * 2 * kRowChunk16, bo = b16 + (uint32_t)ks * b_kstep16;
    const uint64_t ah = umma::smem_desc16(ao, kRowChunk, 128), al = umma::smem_desc16(ao + a_piece16, kRowChunk, 128);
    const uint64_t bh = umma::smem_desc16(bo, b_lbo, b_sbo), bl = umma::smem_desc16(bo + b_piece16, b_lbo, b_sbo);
    umma::mma_f16(acc, ah, bh, idesc, ks > 0);
    umma::mma_f16(acc, ah, bl, idesc, true);
    umma::mma_f16(acc, al, bh, idesc, true);
  }
}
// weight-gradient GEMM: G[128 = (hi cols | lo cols) of A][N = (hi cols | lo cols [| ones]) of B] += A^T . B over the
// 128 tile rows; both MN-major.  One instruction per 16 rows yields all four products in four accumulator blocks
// (same scale: they are simply added when they leave the kernel).
__device__ __forceinline__ void issue_wgrad(uint32_t g, uint32_t a16, uint32_t b16, uint32_t idesc, bool accum) {
#pragma unroll
  for (int ks = 0; ks < kBM / 16; ++ks) {
    const uint64_t da = umma::smem_desc16(a16 + (uint32_t)ks * 16, 128, kRowChunk);
    const uint64_t db = umma::smem_desc16(b16 + (uint32_t)ks * 16, 128, kRowChunk);
    umma::mma_f16(g, da, db, idesc, accum || ks > 0);
  }
}

// this thread's 16 accumulator columns (still carrying the product of the two operand scales)
__device__ __forceinline__ void load_acc16(uint32_t lane_addr, int cq, float (&v)[16], uint32_t col) {
  umma::tmem_ld16(lane_addr + col + 16 * cq, v);
}

// activation -> LayerNorm (no affine) of this thread's 16 columns.  In: accumulator columns t[] = z / inv (inv = 1 over
// the product of the operand scales; 1 for rows that arrive unscaled).  Out: t[] = kSN n (what the operand tile and the
// backward pass hold), (mean, rstd) of the row.
// tanh takes its argument pre-scaled: 2 log2(e) (z + b) = fma(t, 2 log2(e) inv, bias_s) with bias_s = 2 log2(e) b, so the
// unscaling, the bias add and the scaling are one FFMA; the normalisation is one FFMA per element.
__device__ __forceinline__ void act_ln16(float (&t)[16], float inv, const float* __restrict__ bias, const float* __restrict__ bias_s,
                                         int relu, float2* xch, int& slot, int cq, int r, int q4, float& mean, float& rstd) {
  float s = 0.0f, q = 0.0f;
  const float ts = kTanhScale * inv;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float y = fmaf(t[i], ts, bias_s[i]);
    t[i] = relu ? fmaxf(fmaf(t[i], inv, bias[i]), 0.0f) : ((i & 1) ? tanh_scaled_fma(y) : tanh_scaled(y));
    s += t[i];
    q = fmaf(t[i], t[i], q);
  }
  const float2 tot = row_sum2(s, q, xch, slot, cq, r, q4);
  mean = tot.x * (1.0f / 64.0f);
  const float var = fmaxf(fmaf(-mean, mean, tot.y * (1.0f / 64.0f)), 0.0f);
  rstd = 1.0f / sqrtf(var + kLnEps);
  const float rs = rstd * kSN, nmr = -(mean * rstd) * kSN;
#pragma unroll
  for (int i = 0; i < 16; ++i) t[i] = fmaf(t[i], rs, nmr);
}

// dn (gradient w.r.t. the LayerNorm output, times 1 / dn_inv) -> out_s dz (gradient w.r.t. the pre-activation, times the
// scale of the operand tile it goes to), in place; n[] = kSN n as act_ln16 left it:
// dz = rstd (dn - mean(dn) - n mean(dn n)) act'(z), every power-of-two factor folded into the three row coefficients
__device__ __forceinline__ void ln_act_bwd16(float (&dn)[16], const float (&n)[16], float mean, float rstd, int relu, float dn_inv,
                                             float out_s, float2* xch, int& slot, int cq, int r, int q4) {
  float s1 = 0.0f, s2 = 0.0f;
#pragma unroll
  for (int i = 0; i < 16; ++i) { s1 += dn[i]; s2 = fmaf(dn[i], n[i], s2); }
  const float2 tot = row_sum2(s1, s2, xch, slot, cq, r, q4);
  const float c1 = rstd * (dn_inv * out_s);
  const float c0 = -(tot.x * (1.0f / 64.0f)) * c1, c2 = -(tot.y * (1.0f / 64.0f)) * (c1 * (1.0f / (kSN * kSN)));
  const float sd = (1.0f / rstd) * (1.0f / kSN), nz = -(mean * rstd) * kSN;
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    float da;
    if (relu) da = (n[i] != nz) ? 1.0f : 0.0f;            // t == 0  <=>  kSN n == fma(0, rstd kSN, nz), bit for bit
    else { const float t = fmaf(n[i], sd, mean); da = fmaf(-t, t, 1.0f); }
    dn[i] = fmaf(n[i], c2, fmaf(dn[i], c1, c0)) * da;
  }
}

// ---- input rows: gather, LayerNorm over in_dim (mlp.py:44-45,56-57), fp16 pairs -> X operand tile
// ONES: column in_dim (the first padding column; in_dim < KP) is set to 1 -- see the TMEM column map
template <int KP, bool ONES>
__device__ __forceinline__ void stage_x(const MlpArgs& a, long long tile, unsigned char* xs, uint32_t x_piece, float2* xch,
                                        int& slot, int cq, int r, int q4, const float* raw = nullptr) {
  constexpr int XC = KP / 4;                         // input columns per thread
  const int D = a.L.in_dim;
  const long long g = tile * kBM + r;
  const bool valid = g < a.rows;
  float xv[XC];
  if (raw) {                                          // the tile's rows already sit in shared memory
    const float* row = raw + r * D;
#pragma unroll
    for (int i = 0; i < XC; ++i) {
      const int c = XC * cq + i;
      xv[i] = (c < D) ? row[c] : 0.0f;
    }
  } else {
    long long srow = valid ? (a.index ? a.index[g] : g) : 0;
    if (a.x_div != 1) srow /= a.x_div;                 // (a 64-bit division: only the grouped critic rows pay for it)
    const float* __restrict__ row = a.x + srow * (long long)a.ld;
#pragma unroll
    for (int i = 0; i < XC; ++i) {
      const int c = XC * cq + i;
      xv[i] = (valid && c < D) ? __ldg(row + c) : 0.0f;
    }
  }
  if (a.L.use_fn) {
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < XC; ++i) s += xv[i];
    const float mean = row_sum2(s, 0.0f, xch, slot, cq, r, q4).x / (float)D;
    float q = 0.0f;
#pragma unroll
    for (int i = 0; i < XC; ++i) {
      xv[i] = (XC * cq + i < D) ? xv[i] - mean : 0.0f;
      q = fmaf(xv[i], xv[i], q);
    }
    const float rstd = 1.0f / sqrtf(row_sum2(q, 0.0f, xch, slot, cq, r, q4).x / (float)D + kLnEps);
#pragma unroll
    for (int i = 0; i < XC; ++i) xv[i] *= rstd * kSX;
  } else {
    // raw (un-normalised) inputs are the only operands whose magnitude the network does not bound: saturate them at
    // the fp16 range of the scaled hi halves so that an outlier becomes +-4094 instead of inf -> NaN in the MMA (the
    // fp32 FMA path has no such limit; MPE observations are O(1), see DESIGN.md section 4.1 "operand range")
#pragma unroll
    for (int i = 0; i < XC; ++i) xv[i] = fminf(fmaxf(xv[i], -kRawMax), kRawMax) * kSX;
  }
  if (ONES) {
#pragma unroll
    for (int i = 0; i < XC; ++i) if (XC * cq + i == D) xv[i] = kSX;
  }
#pragma unroll
  for (int h = 0; h < XC / 8; ++h) {
    uint4 hi, lo;
    split8u(xv + 8 * h, hi, lo);
    const uint32_t off = (uint32_t)(XC / 8 * cq + h) * kRowChunk + (uint32_t)r * 16;
    *reinterpret_cast<uint4*>(xs + off) = hi;
    *reinterpret_cast<uint4*>(xs + x_piece + off) = lo;
  }
}

// single categorical head, A <= AMAX logits in registers: PPO clipped surrogate + entropy (r_mappo.py:115-141,
// act.py:107-162).  Same arithmetic as actor_loss_row (mlp_rows.cuh) with exp() taken once per logit.
template <int AMAX>
__device__ __forceinline__ void actor_loss_regs(const mappo_loss_config& c, int A, const float (&l)[AMAX], int act, float old_lp,
                                                float adv, float active, float (&dl)[AMAX], float (&sums)[4]) {
  float m = l[0];
#pragma unroll
  for (int q = 1; q < AMAX; ++q) if (q < A) m = fmaxf(m, l[q]);
  float e[AMAX], s = 0.0f;
#pragma unroll
  for (int q = 0; q < AMAX; ++q) { e[q] = (q < A) ? __expf(l[q] - m) : 0.0f; s += e[q]; }
  const float inv = __frcp_rn(s), ls = m + __logf(s);
  float ent = 0.0f, lp_act = 0.0f;
#pragma unroll
  for (int q = 0; q < AMAX; ++q) {
    if (q < A) {
      const float lp = l[q] - ls;
      e[q] *= inv;                                       // p
      ent = fmaf(-e[q], lp, ent);
      if (q == act) lp_act = lp;
    }
  }
  const float lo = 1.0f - c.clip_param, hi = 1.0f + c.clip_param;
  const float ratio = __expf(lp_act - old_lp);
  const float s1 = ratio * adv, s2 = fminf(fmaxf(ratio, lo), hi) * adv;
  const bool inside = (ratio >= lo) && (ratio <= hi);
  const float gcoef = (inside || s1 < s2) ? s1 : 0.0f;   // torch.min / clamp gradient conventions, see actor_loss_row
#pragma unroll
  for (int q = 0; q < AMAX; ++q) {
    dl[q] = 0.0f;
    if (q < A) {
      const float lp = l[q] - ls;
      const float d_obj = -gcoef * ((q == act ? 1.0f : 0.0f) - e[q]);
      const float d_ent = c.entropy_coef * e[q] * (lp + ent);
      dl[q] = active * (d_obj + d_ent);
    }
  }
  sums[0] = -fminf(s1, s2) * active; sums[1] = ent * active; sums[2] = ratio; sums[3] = active;
}

// which instantiations run the EARLY schedule (see the kernel): update, fused first layer, and a shared-memory map
// that still fits with a private partial-logit region (32-wide input tile and at most 8 logits, or the critic)
__host__ __device__ constexpr bool early_tile(int mode, int kp, int amax, bool wide) {
  return (mode == 2 || mode == 3 || mode == 5) && !wide && (kp == 32 ? amax <= 8 : mode == 3);
}

// floats of the per-row loss-scalar landing buffer of one tile
__host__ __device__ constexpr int rowp_floats(int mode, int group) { return mode == 3 ? 3 * kBM * group : 4 * kBM; }

// ---- the kernel ------------------------------------------------------------------------------------------
// MODE: 0 rollout actor, 1 rollout critic, 2 update actor, 3 update critic; recurrent networks (mlp_gru.cu) use
// 4 = base features only (n2 rows written to a.feat) and 5 = base backward from dLoss/dn2 rows read from a.feat.
// KP: padded input width (32 | 64).
// AMAX: register width of the head (1 critic; 8 = single categorical head with <= 8 logits and no availability
// mask, losses evaluated in registers; 16 = generic heads through the row functions of mlp_rows.cuh).
// Per tile the update issues three groups of MMAs -- {fc2}, {fc2 dgrad | head wgrad, fc2 wgrad}, {fc1 of the NEXT
// tile | fc1 wgrad} -- so there are three CTA-wide synchronisations per tile.  Inside a group the MMAs whose result
// the threads wait for come first and commit to `bar`; the weight gradients behind them commit to `gbar`, which is
// only waited for where one of their operand buffers is about to be overwritten (they run in the background).  The
// next tile's input rows are staged into the second X buffer while the second group runs on the tensor cores.
// WIDE: the first layer lives outside (mlp_wide.cu): z1 rows come from a.z1 and dLoss/dz1 goes back in place.
// AFIX > 0: the head has exactly AFIX logits (compile-time loop bounds); ACT: 0 tanh, 1 ReLU, -1 read it from the layout.
template <int MODE, int KP, int AMAX, bool WIDE, int AFIX, int ACT>
__global__ void __launch_bounds__(kThreads, 1) mlp_tc_kernel(const MlpArgs a) {
  constexpr bool UPDATE = (MODE == 2 || MODE == 3 || MODE == 5);
  constexpr bool HEAD = (MODE < 4);
  constexpr bool VHEAD = (MODE == 3);                                  // critic update: head gradient in registers
  // EARLY (update kernels whose shared-memory map has room for a private partial-logit region): fc1 owns a second
  // accumulator (TMEM columns 64..127), the NEXT tile's input rows are staged while fc2 runs, and its fc1 is issued
  // together with the data-gradient GEMM at the tile's second synchronisation -- the tile boundary has no MMA wait left.  DEFER: operand stores only the weight-gradient GEMMs read (n2, dlogits) and the head-bias
  // sums wait until the data-gradient GEMM is running; the head weight gradient joins the last group of the tile.
  constexpr bool EARLY = early_tile(MODE, KP, AMAX, WIDE);
  constexpr bool DEFER = UPDATE && !WIDE;
  constexpr uint32_t kF = EARLY ? kColF : kColAcc;
  // 1 / (product of the operand scales) of the three row GEMMs, and of the weight-gradient accumulators
  constexpr float kInvFc1 = WIDE ? 1.0f : 1.0f / (kSX * kSW), kInvFc2 = 1.0f / (kSN * kSW), kInvDg = 1.0f / (kSD * kSW);
  constexpr float kInvG2 = 1.0f / (kSD * kSN), kInvG1 = 1.0f / (kSD * kSX), kInvGh = 1.0f / (kSN * kSD);
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  __shared__ double lred[4][kThreads / 32];
  __shared__ float vn_s[2];                                            // ValueNorm mean and 1 / sd (critic update)
  const NetLayout& L = a.L;
  const int A = AFIX > 0 ? AFIX : L.a_total, relu = ACT >= 0 ? ACT : L.use_relu, D = L.in_dim, AS = A | 1;
  const Map m = make_map(KP, AS, UPDATE, a.raw_ok ? kBM * D : 0, VHEAD || !HEAD, EARLY, a.rowp_ok ? rowp_floats(MODE, a.group) : 0);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + m.bars);           // completion of the last committed MMA group
  uint64_t* tbar = bar + 1;                                           // arrival of a bulk-copied input tile
  uint64_t* gbar = bar + 2;                                           // completion of the weight-gradient MMAs (background)
  uint64_t* fbar = bar + 3;                                           // EARLY: completion of fc1
  uint64_t* pbar = bar + 4;                                           // arrival of a tile's per-row loss scalars
  if (tid == 0) {
    umma::mbar_init(bar, 1); umma::mbar_init(tbar, 1); umma::mbar_init(gbar, 1); umma::mbar_init(fbar, 1);
    umma::mbar_init(pbar, 1);
    umma::fence_mbar_init();
  }
  // per-row scalars of the register loss path (action, old log-prob, advantage, active mask): with contiguous rows they
  // arrive as four 512-byte bulk copies per tile, requested a tile ahead by one thread -- no thread holds a global
  // load across the second layer, and nothing of their DRAM latency is exposed
  constexpr bool ROWP = (MODE == 2 && AMAX == 8) || MODE == 3;
  const float* rowp = reinterpret_cast<const float*>(sm + m.rowp);
  const bool rowp_active = ROWP && (MODE == 3 ? a.loss.use_value_active_masks : a.loss.use_policy_active_masks) && a.active;
  uint32_t pph = 0;
  auto rowp_tile = [&](long long t) { return ROWP && a.rowp_ok && t < a.ntiles && (t + 1) * kBM <= a.rows; };
  auto issue_rowp = [&](long long t) {                                // one thread
    if (!rowp_tile(t)) return;
    const uint32_t per_row = MODE == 3 ? (uint32_t)a.group : 1u;        // critic: `group` terms share a row's value
    const uint32_t dst = umma::smem_u32(sm + m.rowp), mb = umma::smem_u32(pbar), bytes = kBM * 4u * per_row;
    const uint32_t ncopy = MODE == 3 ? (rowp_active ? 3u : 2u) : (rowp_active ? 4u : 3u);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(bytes * ncopy) : "memory");
    auto copy = [&](uint32_t slot_, const float* src_) {
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   ::"r"(dst + slot_ * bytes), "l"(src_ + t * (long long)(kBM * per_row)), "r"(bytes), "r"(mb) : "memory");
    };
    if (MODE == 3) {
      copy(0u, a.returns); copy(1u, a.value_preds);
      if (rowp_active) copy(2u, a.active);
    } else {
      copy(0u, a.actions); copy(1u, a.old_logp); copy(2u, a.adv);
      if (rowp_active) copy(3u, a.active);
    }
  };
  const float* raw = reinterpret_cast<const float*>(sm + m.raw);
  const uint32_t raw_bytes = (uint32_t)(kBM * D) * 4u;
  uint32_t tph = 0;
  // tile t arrives by bulk copy iff it is a full tile, not this CTA's first, and the layout allows it
  auto bulk_tile = [&](long long t) { return a.raw_ok && !WIDE && t < a.ntiles && (t + 1) * kBM <= a.rows; };
  auto issue_bulk = [&](long long t) {                                // one thread
    if (!bulk_tile(t)) return;
    const uint32_t dst = umma::smem_u32(sm + m.raw), mb = umma::smem_u32(tbar);
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mb), "r"(raw_bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(a.x + t * (long long)(kBM * D)), "r"(raw_bytes), "r"(mb) : "memory");
  };
  auto stage_next = [&](long long t, unsigned char* xs, int& slot_ref) {
    if (bulk_tile(t)) {
      umma::mbar_wait(tbar, tph); tph ^= 1;
      stage_x<KP, UPDATE>(a, t, xs, m.x_piece, reinterpret_cast<float2*>(sm + m.xch), slot_ref, cq, r, q4, raw);
    } else {
      stage_x<KP, UPDATE>(a, t, xs, m.x_piece, reinterpret_cast<float2*>(sm + m.xch), slot_ref, cq, r, q4);
    }
  };
  if (MODE == 3 && tid == 32) {                                        // valuenorm.py:31-35,56-65 (ValueNorm.normalize)
    float mean = 0.0f, sd = 1.0f;
    if (a.loss.use_value_norm) {
      const float db = fmaxf(a.vn_stats[2], 1e-5f);
      mean = a.vn_stats[0] / db;
      sd = sqrtf(fmaxf(a.vn_stats[1] / db - mean * mean, 1e-2f));
    }
    vn_s[0] = mean; vn_s[1] = 1.0f / sd;
  }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, UPDATE ? kTmemCols : 128);
  if (UPDATE) stage_weights_serial(sm, m, a, !WIDE, UPDATE);
  else stage_weights_tc<KP, (HEAD && AMAX == 8)>(sm, m, a, !WIDE, UPDATE);
  if (UPDATE && tid < kBM) {                                            // the ones columns (written once)
    const uint4 one = make_uint4(0x00005400u, 0u, 0u, 0u), zero = make_uint4(0u, 0u, 0u, 0u);   // fp16(kSN) = 64
    auto put = [&](unsigned char* base) {
      *reinterpret_cast<uint4*>(base + tid * 16) = one;
      *reinterpret_cast<uint4*>(base + kRowChunk + tid * 16) = zero;
    };
    put(sm + m.N1 + kNP * 8 * kRowChunk);
  }
  int slot = 0;
  float2* xch = reinterpret_cast<float2*>(sm + m.xch);
  long long tile = blockIdx.x;
  if (!WIDE) stage_x<KP, UPDATE>(a, tile, sm + m.X, m.x_piece, xch, slot, cq, r, q4);
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  // MMA groups are issued by one elected lane of warps 0 / 1 / 2 (three different schedulers) in turn
  // (column quarter 3: the quarter-0 warps already carry the per-row extras -- loss sums, dlogits tile, head-bias sums)
  const bool issuer_fc2 = (warp == 12), issuer_bwd = (warp == 13), issuer_next = (warp == 14);
  if (issuer_next && umma::elect_one()) { issue_bulk(tile + gridDim.x); issue_rowp(tile); }
  const uint32_t tb = tmem_base_s;
  const uint32_t lane_addr = tb + ((uint32_t)(32 * q4) << 16);
  const uint32_t x_bytes = m.x_buf;
  const uint32_t b16 = umma::smem_u32(sm) >> 4;                         // operand addresses in 16-byte units
  const uint32_t sX = b16 + (m.X >> 4), sN1 = b16 + (m.N1 >> 4), sBS = b16 + (m.BS >> 4), sB2 = b16 + (m.BS2 >> 4);
  const uint32_t sW1 = b16 + (m.W1 >> 4), sW2 = b16 + (m.W2 >> 4), sDL = b16 + (m.DL >> 4);
  const uint32_t x_piece16 = m.x_piece >> 4, w1_piece16 = m.w1_piece >> 4, x_bytes16 = x_bytes >> 4, n1_lo16 = m.n1_lo >> 4;
  const float* Whf = reinterpret_cast<const float*>(sm + m.Whf);
  const float* b1f = reinterpret_cast<const float*>(sm + m.b1f) + 16 * cq;
  const float* b2f = reinterpret_cast<const float*>(sm + m.b2f) + 16 * cq;
  const float* b1s = reinterpret_cast<const float*>(sm + m.b1s) + 16 * cq;
  const float* b2s = reinterpret_cast<const float*>(sm + m.b2s) + 16 * cq;
  const float* bhf = reinterpret_cast<const float*>(sm + m.bhf);
  constexpr uint32_t kIdFwd = umma::idesc_f16(128, 64, 0, 0), kIdDgrad = umma::idesc_f16(128, 64, 0, 1);
  constexpr uint32_t kIdG2 = umma::idesc_f16(128, 128 + 16, 1, 1), kIdG1 = umma::idesc_f16(128, kNP * KP, 1, 1);
  constexpr uint32_t kIdGh = umma::idesc_f16(128, 32, 1, 1);
  if (!WIDE && issuer_next && umma::elect_one()) {
    issue_rowgemm<KP / 16>(tb + kF, sX, x_piece16, sW1, w1_piece16, kWChunk, 128, 2 * kWChunk16, kIdFwd);
    umma::mma_commit(EARLY ? fbar : bar);
  }
  uint32_t ph = 0, gph = 0, fph = 0;
  int buf = 0;
  bool first = true;
  float db1[WIDE ? 16 : 1], ghv[VHEAD ? 16 : 1];      // WIDE: the fc1 weight gradient (and its ones column) lives outside
  double lacc[4] = {0.0, 0.0, 0.0, 0.0};
  if (UPDATE) {
#pragma unroll
    for (int i = 0; i < (WIDE ? 16 : 1); ++i) db1[i] = 0.0f;
#pragma unroll
    for (int i = 0; i < (VHEAD ? 16 : 1); ++i) ghv[i] = 0.0f;
  }

  for (; tile < a.ntiles; tile += gridDim.x) {
    const long long g = tile * kBM + r;
    const bool valid = g < a.rows;
    const bool has_next = tile + gridDim.x < a.ntiles;
    const long long src = (valid && a.index) ? a.index[g] : (valid ? g : 0);
    // rows this tile will only ask for once it is under way (z1 of a wide first layer, dLoss/dn2 of a recurrent base)
    // are exposed DRAM latency with one tile per CTA: one thread asks the L2 for the NEXT tile's now
    if ((WIDE || MODE == 5) && has_next && tid == 96) {
      const long long g0 = (tile + gridDim.x) * kBM;
      const uint32_t bytes = (uint32_t)min((long long)kBM, a.rows - g0) * 256u;
      if (WIDE) umma::l2_prefetch(a.z1 + g0 * 64, bytes);
      if (MODE == 5) umma::l2_prefetch(a.feat + g0 * 64, bytes);
    }
    // per-row scalars of the register loss path: requested once fc2 is issued (below), consumed after the second layer
    float p_actf = 0.0f;                              // converted at its use: no wait on the load
    float p_old = 0.0f, p_adv = 0.0f, p_active = 1.0f;
    // ---- layer 1 epilogue
    float n1[16], mean1, rstd1;
    if (WIDE) {
      if (UPDATE && !first) { umma::mbar_wait(gbar, gph); gph ^= 1; }     // previous tile's G2 still reads N1 / BS
      const float4* zp = reinterpret_cast<const float4*>(a.z1 + g * 64 + 16 * cq);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float4 z = valid ? zp[i] : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        n1[4 * i] = z.x; n1[4 * i + 1] = z.y; n1[4 * i + 2] = z.z; n1[4 * i + 3] = z.w;
      }
    } else {
      if (EARLY) { umma::mbar_wait(fbar, fph); fph ^= 1; }
      else { umma::mbar_wait(bar, ph); ph ^= 1; }
      umma::tc_fence_after();
      load_acc16(lane_addr, cq, n1, kF);
    }
    act_ln16(n1, kInvFc1, b1f, b1s, relu, xch, slot, cq, r, q4, mean1, rstd1);
    store_cols16u(sm + m.N1, m.n1_lo, r, cq, n1);
    umma::fence_async_smem();
    umma::tc_fence_before();
    __syncthreads();
    if (issuer_fc2 && umma::elect_one()) {
      umma::tc_fence_after();
      issue_rowgemm<4>(tb + kColAcc, sN1, n1_lo16, sW2, 8 * kWChunk16, kWChunk, 128, 2 * kWChunk16, kIdFwd);
      umma::mma_commit(bar);
    }
    // rollout: the next tile's input rows are staged while fc2 runs, and its fc1 is issued as soon as every warp has
    // its fc2 accumulators in registers, so both GEMMs hide under epilogue work
    if (!UPDATE && !WIDE && has_next) stage_next(tile + gridDim.x, sm + m.X + (buf ^ 1) * x_bytes, slot);
    if (MODE == 2 && AMAX == 8 && valid && !rowp_tile(tile)) {
      p_actf = __ldg(a.actions + src);
      p_old = __ldg(a.old_logp + src);
      p_adv = __ldg(a.adv + src);
      if (a.loss.use_policy_active_masks && a.active) p_active = __ldg(a.active + src);
    }
    if (MODE == 3 && valid && cq == 0 && !rowp_tile(tile)) {              // value-loss inputs of this row's terms: warm the L1
      const long long o = src * a.group, e = o + a.group - 1;
      umma::l1_prefetch(a.returns + o); umma::l1_prefetch(a.returns + e);
      umma::l1_prefetch(a.value_preds + o); umma::l1_prefetch(a.value_preds + e);
      if (a.loss.use_value_active_masks && a.active) { umma::l1_prefetch(a.active + o); umma::l1_prefetch(a.active + e); }
    }
    if (EARLY) {                                                          // the same under fc2 of the update
      if (!first) { umma::mbar_wait(gbar, gph); gph ^= 1; }               // previous tile's G1 / Gh: frees the other X buffer, BS, n2, dlogits
      if (has_next) stage_next(tile + gridDim.x, sm + m.X + (buf ^ 1) * x_bytes, slot);
    }
    // ---- layer 2 epilogue + head
    float n2[16], mean2, rstd2;
    umma::mbar_wait(bar, ph); ph ^= 1;
    if (UPDATE && !WIDE && !EARLY && !first) { umma::mbar_wait(gbar, gph); gph ^= 1; }   // previous tile's G1: frees the other X buffer and BS
    umma::tc_fence_after();
    load_acc16(lane_addr, cq, n2, kColAcc);
    if (!UPDATE && !WIDE) {
      umma::fence_async_smem();
      umma::tc_fence_before();
      __syncthreads();
      if (has_next && issuer_next && umma::elect_one()) {
        issue_bulk(tile + 2 * (long long)gridDim.x);
        umma::tc_fence_after();
        issue_rowgemm<KP / 16>(tb + kColAcc, sX + (buf ^ 1) * x_bytes16, x_piece16, sW1, w1_piece16, kWChunk,
                               128, 2 * kWChunk16, kIdFwd);
        umma::mma_commit(bar);
      }
    }
    // FUSE (the register-loss head of at most 8 logits): the head's partial dot products are taken on the activations BEFORE the LayerNorm,
    // inside the MUFU-paced activation loop, and travel with the row's LayerNorm partial sums through ONE exchange:
    // logit = b' + rstd (sum_k t_k w'_k - mean sum_k w'_k).  Otherwise: LayerNorm exchange, then the logits' own.
    constexpr bool FUSE = HEAD && AMAX == 8;                 // (the single-logit critic head measured 6 % slower in the rollout)
    float* xlog = reinterpret_cast<float*>(sm + (m.xlog_in_x ? m.X + (buf ^ 1) * x_bytes : m.xlog));
    float l[AMAX];
    if (FUSE) {
      float s = 0.0f, q = 0.0f;
      const float ts = kTanhScale * kInvFc2;
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float y = fmaf(n2[i], ts, b2s[i]);
        n2[i] = relu ? fmaxf(fmaf(n2[i], kInvFc2, b2f[i]), 0.0f) : ((i & 1) ? tanh_scaled_fma(y) : tanh_scaled(y));
        s += n2[i];
        q = fmaf(n2[i], n2[i], q);
      }
      float* mine = xlog + (size_t)(cq * kBM + r) * AS;
#pragma unroll
      for (int qq = 0; qq < AMAX; ++qq) {
        if (qq < A) {
          const float4* w4 = reinterpret_cast<const float4*>(Whf + qq * 64 + 16 * cq);
          float d = 0.0f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 w = w4[i];
            d = fmaf(n2[4 * i], w.x, fmaf(n2[4 * i + 1], w.y, fmaf(n2[4 * i + 2], w.z, fmaf(n2[4 * i + 3], w.w, d))));
          }
          mine[qq] = d;
        }
      }
      const float2 tot = row_sum2(s, q, xch, slot, cq, r, q4);            // its barrier also publishes the partial dots
      mean2 = tot.x * (1.0f / 64.0f);
      const float var = fmaxf(fmaf(-mean2, mean2, tot.y * (1.0f / 64.0f)), 0.0f);
      rstd2 = 1.0f / sqrtf(var + kLnEps);
      const float rs = rstd2 * kSN, nmr = -(mean2 * rstd2) * kSN;
#pragma unroll
      for (int i = 0; i < 16; ++i) n2[i] = fmaf(n2[i], rs, nmr);
#pragma unroll
      for (int qq = 0; qq < AMAX; ++qq) {
        l[qq] = 0.0f;
        if (qq < A) {
          const float* p = xlog + (size_t)r * AS + qq;
          const float dsum = (p[0] + p[(size_t)kBM * AS]) + (p[(size_t)2 * kBM * AS] + p[(size_t)3 * kBM * AS]);
          l[qq] = fmaf(rstd2, fmaf(-mean2, bhf[kMaxA + qq], dsum), bhf[qq]);
        }
      }
    } else {
      act_ln16(n2, kInvFc2, b2f, b2s, relu, xch, slot, cq, r, q4, mean2, rstd2);
    }
    if (UPDATE && HEAD && !VHEAD && !DEFER) store_cols16u(sm + m.BS2, 8 * kRowChunk, r, cq, n2);
    if (MODE == 4 && valid) {
      float4* fo = reinterpret_cast<float4*>(a.feat + g * 64 + 16 * cq);
#pragma unroll
      for (int i = 0; i < 4; ++i)
        fo[i] = make_float4(n2[4 * i] * (1.0f / kSN), n2[4 * i + 1] * (1.0f / kSN), n2[4 * i + 2] * (1.0f / kSN), n2[4 * i + 3] * (1.0f / kSN));
    }
    if (HEAD && !FUSE) {
      float* mine = xlog + (size_t)(cq * kBM + r) * AS;
#pragma unroll
      for (int q = 0; q < AMAX; ++q) {
        if (q < A) {
          const float4* w4 = reinterpret_cast<const float4*>(Whf + q * 64 + 16 * cq);
          float s = 0.0f;
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            const float4 w = w4[i];
            s = fmaf(n2[4 * i], w.x, fmaf(n2[4 * i + 1], w.y, fmaf(n2[4 * i + 2], w.z, fmaf(n2[4 * i + 3], w.w, s))));
          }
          mine[q] = s * (1.0f / kSN);                       // n2[] holds kSN n2
        }
      }
      bar_named(1 + q4, 128);
#pragma unroll
      for (int q = 0; q < AMAX; ++q) {
        l[q] = 0.0f;
        if (q < A) {
          const float* p = xlog + (size_t)r * AS + q;
          l[q] = bhf[q] + ((p[0] + p[(size_t)kBM * AS]) + (p[(size_t)2 * kBM * AS] + p[(size_t)3 * kBM * AS]));
        }
      }
    }
    if (!HEAD) {
#pragma unroll
      for (int q = 0; q < AMAX; ++q) l[q] = 0.0f;
    }
    if (!UPDATE) {
      if (cq == 0 && valid) {
        if (MODE == 0) actor_rollout_row(a, g, l);
        if (MODE == 1) for (int mm = 0; mm < a.group; ++mm) a.values_out[g * a.group + mm] = l[0];
      }
      if (WIDE) continue;
      buf ^= 1;
      continue;
    }
    // ---- losses: every thread of a row evaluates them (no idle warps, no shared-memory round trip for dlogits);
    // the column-quarter-0 thread keeps the sums and writes the dlogits operand of the head weight gradient
    float dl[AMAX];
#pragma unroll
    for (int q = 0; q < AMAX; ++q) dl[q] = 0.0f;
    if (HEAD) {
      double mine[4] = {0.0, 0.0, 0.0, 0.0};
      if (MODE == 2 && AMAX == 8) {
        float sums[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        if (rowp_tile(tile)) {                                            // a full tile: every row is valid
          umma::mbar_wait(pbar, pph); pph ^= 1;
          p_actf = rowp[r]; p_old = rowp[kBM + r]; p_adv = rowp[2 * kBM + r];
          if (rowp_active) p_active = rowp[3 * kBM + r];
        }
        if (valid) actor_loss_regs<AMAX>(a.loss, A, l, (int)p_actf, p_old, p_adv, p_active, dl, sums);
        else {
#pragma unroll
          for (int q = 0; q < AMAX; ++q) dl[q] = 0.0f;
        }
        if (cq == 0) {
#pragma unroll
          for (int q = 0; q < 4; ++q) lacc[q] += (double)sums[q];
        }
      } else if (MODE == 2) {
        float lt[kMaxA], dt[kMaxA];
#pragma unroll
        for (int q = 0; q < kMaxA; ++q) { lt[q] = (q < AMAX) ? l[q < AMAX ? q : 0] : 0.0f; dt[q] = 0.0f; }
        if (valid) actor_loss_row(a, src, lt, dt, mine);
#pragma unroll
        for (int q = 0; q < AMAX; ++q) dl[q] = dt[q];
      } else {
        const bool act_on = a.loss.use_value_active_masks && a.active;
        if (rowp_tile(tile)) {                                            // a full tile, its loss inputs in shared memory
          umma::mbar_wait(pbar, pph); pph ^= 1;
          const int per = kBM * a.group;
          dl[0] = critic_loss_row_pre(a, rowp, rowp + per, act_on ? rowp + 2 * per : nullptr, (long long)r * a.group, l[0],
                                      vn_s[0], vn_s[1], cq == 0, mine);
        } else {
          dl[0] = valid ? critic_loss_row_pre(a, a.returns, a.value_preds, act_on ? a.active : nullptr, src * a.group, l[0],
                                              vn_s[0], vn_s[1], cq == 0, mine) : 0.0f;
        }
#pragma unroll
        for (int q = 1; q < AMAX; ++q) dl[q] = 0.0f;
      }
      if (cq == 0) {
#pragma unroll
        for (int q = 0; q < 4; ++q) lacc[q] += mine[q];
      }
    }
    // what only the head's weight / bias gradients need: the dlogits operand tile, the head-bias sums, (DEFER) n2
    auto head_operands = [&]() {
      if (HEAD && cq == 0) {
        float d16[16];
        float* dbh = reinterpret_cast<float*>(sm + m.dbh) + r * AS;
#pragma unroll
        for (int q = 0; q < 16; ++q) {
          d16[q] = (q < AMAX) ? dl[q < AMAX ? q : 0] : 0.0f;
          if (q < AMAX && q < A) dbh[q] += d16[q];
          d16[q] *= kSD;
        }
        if (!VHEAD) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            uint4 hi, lo;
            split8u(d16 + 8 * h, hi, lo);
            const uint32_t off = (uint32_t)h * kRowChunk + (uint32_t)r * 16;
            *reinterpret_cast<uint4*>(sm + m.DL + off) = hi;
            *reinterpret_cast<uint4*>(sm + m.DL + 2 * kRowChunk + off) = lo;
          }
        }
      }
      if (VHEAD) {
#pragma unroll
        for (int i = 0; i < 16; ++i) ghv[i] = fmaf(dl[0], n2[i], ghv[i]);    // G_v[k] += dv (kSN n2[k])
      }
      if (DEFER && HEAD && !VHEAD) store_cols16u(sm + m.BS2, 8 * kRowChunk, r, cq, n2);
    };
    if (!DEFER) head_operands();
    // ---- back through the head and layer 2
    float dz[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) dz[i] = 0.0f;
    if (MODE == 5 && valid) {                                             // dLoss/dn2 from the recurrent layer
      const float4* fi = reinterpret_cast<const float4*>(a.feat + g * 64 + 16 * cq);
#pragma unroll
      for (int i = 0; i < 4; ++i) { const float4 t = fi[i]; dz[4 * i] = t.x; dz[4 * i + 1] = t.y; dz[4 * i + 2] = t.z; dz[4 * i + 3] = t.w; }
    }
#pragma unroll
    for (int q = 0; q < AMAX; ++q) {
      if (HEAD && q < A) {
        const float4* w4 = reinterpret_cast<const float4*>(Whf + q * 64 + 16 * cq);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float4 w = w4[i];
          dz[4 * i] = fmaf(dl[q], w.x, dz[4 * i]); dz[4 * i + 1] = fmaf(dl[q], w.y, dz[4 * i + 1]);
          dz[4 * i + 2] = fmaf(dl[q], w.z, dz[4 * i + 2]); dz[4 * i + 3] = fmaf(dl[q], w.w, dz[4 * i + 3]);
        }
      }
    }
    ln_act_bwd16(dz, n2, mean2, rstd2, relu, 1.0f, kSD, xch, slot, cq, r, q4);
    store_cols16u(sm + m.BS, 8 * kRowChunk, r, cq, dz);
    umma::fence_async_smem();
    umma::tc_fence_before();
    __syncthreads();
    if (issuer_bwd && umma::elect_one()) {
      umma::tc_fence_after();
      issue_rowgemm<4>(tb + kColAcc, sBS, 8 * kRowChunk16, sW2, 8 * kWChunk16, 128, kWChunk, 16, kIdDgrad);
      umma::mma_commit(bar);                                                                    // dn1 = dz2 W2'
      if (HEAD && !VHEAD && !DEFER) issue_wgrad(tb + kColGh, sB2, sDL, kIdGh, !first);   // Gh += n2^T dlogits
      issue_wgrad(tb + kColG2, sBS, sN1, kIdG2, !first);               // G2 += dz2^T [n1 | 1]
      umma::mma_commit(gbar);                                          // (frees BS for dz1: ahead of the next tile's fc1)
    }
    if (issuer_next && has_next && umma::elect_one()) {              // (another lane quadrant shares the issue time)
      umma::tc_fence_after();
      if (ROWP) issue_rowp(tile + gridDim.x);                        // this tile's scalars were read before the barrier
      if (EARLY) {                                                                              // the next tile's fc1
        issue_bulk(tile + 2 * (long long)gridDim.x);
        issue_rowgemm<KP / 16>(tb + kF, sX + (buf ^ 1) * x_bytes16, x_piece16, sW1, w1_piece16, kWChunk,
                               128, 2 * kWChunk16, kIdFwd);
        umma::mma_commit(fbar);
      }
    }
    if (DEFER) head_operands();
    // ---- next tile's input rows, staged while the tensor cores work
    if (!WIDE && !EARLY && has_next) stage_next(tile + gridDim.x, sm + m.X + (buf ^ 1) * x_bytes, slot);
    // ---- layer 1 backward
    umma::mbar_wait(bar, ph); ph ^= 1;
    umma::tc_fence_after();
    load_acc16(lane_addr, cq, dz, kColAcc);
    ln_act_bwd16(dz, n1, mean1, rstd1, relu, kInvDg, WIDE ? 1.0f : kSD, xch, slot, cq, r, q4);
    if (WIDE) {
#pragma unroll
      for (int i = 0; i < 16; ++i) db1[WIDE ? i : 0] += dz[i];
    }
    if (WIDE) {                                                          // dLoss/dz1 leaves for wgrad1_wide_kernel
      if (valid) {
        float4* zp = reinterpret_cast<float4*>(a.z1 + g * 64 + 16 * cq);
#pragma unroll
        for (int i = 0; i < 4; ++i) zp[i] = make_float4(dz[4 * i], dz[4 * i + 1], dz[4 * i + 2], dz[4 * i + 3]);
      }
      first = false;
      continue;
    }
    umma::mbar_wait(gbar, gph); gph ^= 1;                                 // this tile's G2 has read dz2 out of BS
    store_cols16u(sm + m.BS, 8 * kRowChunk, r, cq, dz);
    umma::fence_async_smem();
    umma::tc_fence_before();
    __syncthreads();
    if (issuer_next && umma::elect_one()) {
      umma::tc_fence_after();
      if (!EARLY && has_next) {
        issue_bulk(tile + 2 * (long long)gridDim.x);
        issue_rowgemm<KP / 16>(tb + kColAcc, sX + (buf ^ 1) * x_bytes16, x_piece16, sW1, w1_piece16, kWChunk,
                               128, 2 * kWChunk16, kIdFwd);
        umma::mma_commit(bar);
      }
      if (HEAD && !VHEAD && DEFER) issue_wgrad(tb + kColGh, sB2, sDL, kIdGh, !first);   // Gh += n2^T dlogits
      issue_wgrad(tb + kColG1, sBS, sX + buf * x_bytes16, kIdG1, !first);   // G1 += dz1^T [x | 1]
      umma::mma_commit(gbar);
    }
    buf ^= 1;
    first = false;
  }
  // ---- epilogue: weight gradients TMEM -> two partial rows per CTA (the hi-half and the lo-half lanes of the stacked A operand), in the flat
  // parameter layout (G in the W slots, db in the b slots; LayerNorm-affine slots stay zero, optim.cu fills them)
  if (UPDATE) {
    if (!first) { umma::mbar_wait(gbar, gph); gph ^= 1; }               // the last tile's weight gradients
    umma::tc_fence_after();
    float* __restrict__ gp0 = a.grad_partials + (size_t)(2 * blockIdx.x) * L.total;
    if (HEAD) for (int e = tid; e < 2 * L.total; e += kThreads) gp0[e] = 0.0f;   // MODE 5: the caller zeroed the buffer
    float* red = reinterpret_cast<float*>(sm + m.red);                    // [2][16 warps][16]
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      if (WIDE) {
        float v1 = db1[WIDE ? i : 0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v1 += __shfl_xor_sync(0xffffffffu, v1, o);
        if (lane == 0) red[warp * 16 + i] = v1;
      }
      if (VHEAD) {
        float v3 = ghv[VHEAD ? i : 0];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v3 += __shfl_xor_sync(0xffffffffu, v3, o);
        if (lane == 0) red[512 + warp * 16 + i] = v3;
      }
    }
    __syncthreads();                                                      // zero fill + red visible
    float* __restrict__ gp = gp0 + (size_t)(q4 >> 1) * L.total;           // lanes 64..127: the lo halves of A, second partial row
    const int mrow = r & 63;
    {
      float u[16], w[16];
      umma::tmem_ld16(lane_addr + kColG2 + 16 * cq, u);
      umma::tmem_ld16(lane_addr + kColG2 + 64 + 16 * cq, w);
#pragma unroll
      for (int i = 0; i < 16; ++i) gp[L.w2 + mrow * 64 + 16 * cq + i] = (u[i] + w[i]) * kInvG2;
      if (!WIDE && cq < KP / 16) {
        umma::tmem_ld16(lane_addr + kColG1 + 16 * cq, u);
        umma::tmem_ld16(lane_addr + kColG1 + KP + 16 * cq, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int k = 16 * cq + i;
          if (k < D) gp[L.w1 + mrow * D + k] = (u[i] + w[i]) * kInvG1;
          if (k == D) gp[L.b1 + mrow] = u[i] * kInvG1;                  // the ones column of the input tile (holds kSX)
        }
      }
      if (cq == 0) {                                                      // bias gradients: the ones chunk (holds kSN)
        umma::tmem_ld16(lane_addr + kColG2 + 128, u);
        gp[L.b2 + mrow] = u[0] * kInvG2;
      }
      if (HEAD && !VHEAD && cq == 3) {                                    // warps 12..15: head gradient, lanes = hidden unit
        umma::tmem_ld16(lane_addr + kColGh, u);
        umma::tmem_ld16(lane_addr + kColGh + 16, w);
        int h = 0, off = 0;
#pragma unroll
        for (int q = 0; q < AMAX; ++q) {
          if (q < A) {
            while (q >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
            gp[L.wh[h] + (q - off) * 64 + mrow] = (u[q] + w[q]) * kInvGh;
          }
        }
      }
    }
    if (tid < 64) {
      const int c = tid >> 4, i = tid & 15;
      if (WIDE) {
        float s1 = 0.0f;
        for (int q = 0; q < 4; ++q) s1 += red[(4 * c + q) * 16 + i];
        gp0[L.b1 + tid] = s1;
      }
      if (VHEAD) {
        float s3 = 0.0f;
        for (int q = 0; q < 4; ++q) s3 += red[512 + (4 * c + q) * 16 + i];
        gp0[L.wh[0] + tid] = s3 * (1.0f / kSN);                          // accumulated against kSN n2
      }
    } else if (HEAD && tid < 64 + A) {
      const int q = tid - 64;
      const float* dbh = reinterpret_cast<const float*>(sm + m.dbh);
      float s = 0.0f;
      for (int rr = 0; rr < kBM; ++rr) s += dbh[rr * AS + q];
      int h = 0, off = 0;
      while (q >= off + L.head_dim[h]) { off += L.head_dim[h]; ++h; }
      gp0[L.bh[h] + (q - off)] = s;
    }
    // loss sums: block reduction in double
    for (int q = 0; q < 4; ++q) {
      double v = lacc[q];
      for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
      if (lane == 0) lred[q][warp] = v;
    }
    __syncthreads();
    if (HEAD && tid < 4) {
      double v = 0.0;
      for (int w = 0; w < kThreads / 32; ++w) v += lred[tid][w];
      a.loss_partials[(size_t)(2 * blockIdx.x) * 4 + tid] = v;
      a.loss_partials[(size_t)(2 * blockIdx.x + 1) * 4 + tid] = 0.0;
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, UPDATE ? kTmemCols : 128);
}

// ---- host side ---------------------------------------------------------------------------------------------
constexpr int kMaxDynSmem = 232448 - 2048;          // 227 KB per CTA minus this kernel's static shared memory

static int sm_count() {
  static int sms = 0;
  if (sms == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return 148;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  }
  return sms;
}

int grid_for(long long rows) {
  const long long ntiles = (rows + kBM - 1) / kBM;
  const int sms = sm_count();
  return (int)(ntiles < sms ? (ntiles < 1 ? 1 : ntiles) : sms);
}

bool wide_supports(const mappo_net_config* cfg);                                      // mlp_wide.cu
long long wide_workspace_bytes(long long rows);
int launch_wide_fc1(const MlpArgs& a, void* workspace, cudaStream_t stream, const char* who);
int launch_wide_wgrad1(const MlpArgs& a, void* workspace, cudaStream_t stream, const char* who);
int launch_wide_fc1_packed(const MlpArgs& a, void* workspace, const void* packed, cudaStream_t stream, const char* who);
int launch_wide_wgrad1_packed(const MlpArgs& a, void* workspace, const void* packed, cudaStream_t stream, const char* who);

bool supports(const mappo_net_config* cfg) {
  const bool wide = wide_supports(cfg);
  if (wide && cfg->is_critic) return true;
  if (!wide && !(cfg && cfg->hidden == 64 && cfg->layer_N == 1 && !cfg->recurrent && cfg->in_dim < kWideFrom)) return false;
  int A = 0;
  for (int h = 0; h < cfg->n_heads; ++h) A += cfg->head_dim[h];
  if (cfg->is_critic) A = 1;
  const int Kp = (wide || cfg->in_dim < 32) ? 32 : 64;
  // the maps of the instantiations launch() picks for this network: update (MODE 3 for a critic: no n2 / dlogits tiles)
  // with its early_tile() setting, and rollout
  const bool critic = cfg->is_critic != 0;
  const bool early = early_tile(critic ? 3 : 2, Kp, critic ? 1 : (A <= 8 ? 8 : 16), wide);
  return make_map(Kp, A | 1, true, 0, critic, early).total <= (uint32_t)kMaxDynSmem &&
         (critic || make_map(Kp, A | 1, true, 0, false, false).total <= (uint32_t)kMaxDynSmem) &&   // (availability masks: AMAX 16)
         make_map(Kp, A | 1, false).total <= (uint32_t)kMaxDynSmem;
}

template <int MODE, int KP, int AMAX, bool WIDE, int AFIX, int ACT>
static int launch_inst(MlpArgs& a, cudaStream_t stream, const char* who) {
  constexpr bool kUpd = (MODE == 2 || MODE == 3 || MODE == 5);
  constexpr bool kEarly = early_tile(MODE, KP, AMAX, WIDE);
  // bulk-copied input tiles: contiguous rows only, and only if the landing buffer still fits
  auto al16 = [](const void* p_) { return (reinterpret_cast<uintptr_t>(p_) & 15u) == 0; };
  const bool vh = (MODE == 3 || MODE == 5);
  auto fits = [&](int raw_f, int rowp_f) {
    return make_map(KP, a.L.a_total | 1, kUpd, raw_f, vh, kEarly, rowp_f).total <= (uint32_t)kMaxDynSmem;
  };
  // per-row loss scalars by bulk copy (a few KB), then bulk-copied input tiles (the landing buffer is the larger one):
  // contiguous rows only, and only if the buffers still fit
  a.rowp_ok = 0;
  if (MODE == 2) a.rowp_ok = (AMAX == 8 && a.index == nullptr && al16(a.actions) && al16(a.old_logp) && al16(a.adv) && al16(a.active)) ? 1 : 0;
  if (MODE == 3)
    a.rowp_ok = (a.index == nullptr && a.group >= 1 && a.group <= kRowpGroup && al16(a.returns) && al16(a.value_preds) && al16(a.active)) ? 1 : 0;
  const int rowp_f = rowp_floats(MODE, a.group);
  if (a.rowp_ok && !fits(0, rowp_f)) a.rowp_ok = 0;
  a.raw_ok = 0;
  if (!WIDE && a.index == nullptr && a.x_div == 1 && a.ld == a.L.in_dim && al16(a.x) && fits(kBM * a.L.in_dim, a.rowp_ok ? rowp_f : 0))
    a.raw_ok = 1;
  const Map m = make_map(KP, a.L.a_total | 1, kUpd, a.raw_ok ? kBM * a.L.in_dim : 0, vh, kEarly, a.rowp_ok ? rowp_f : 0);
  auto kern = mlp_tc_kernel<MODE, KP, AMAX, WIDE, AFIX, ACT>;
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kMaxDynSmem));
    attr_set = true;
  }
  MAPPO_REQUIRE(m.total <= kMaxDynSmem, "%s: shared-memory map too large (%u B)", who, m.total);
  kern<<<grid_for(a.rows), kThreads, m.total, stream>>>(a);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

// the common shapes get compile-time loop bounds and activation: tanh (the reference scripts' setting)
// and, for a single categorical head, exactly 5 logits (every moving MPE agent); everything else takes the generic build
template <int MODE, int KP, int AMAX, bool WIDE = false>
static int launch_one(MlpArgs& a, cudaStream_t stream, const char* who) {
  if constexpr (AMAX == 8) {
    if (!a.L.use_relu && a.L.a_total == 5) return launch_inst<MODE, KP, AMAX, WIDE, 5, 0>(a, stream, who);
  }
  if constexpr (AMAX == 1) {
    if (!a.L.use_relu) return launch_inst<MODE, KP, AMAX, WIDE, 0, 0>(a, stream, who);
  }
  return launch_inst<MODE, KP, AMAX, WIDE, 0, -1>(a, stream, who);
}

template <int MODE>
int launch(MlpArgs& a, const mappo_net_config* cfg, cudaStream_t stream, const char* who) {
  a.L = make_layout(*cfg);
  if (a.x_div < 1) a.x_div = 1;
  a.Kp = cfg->in_dim < 32 ? 32 : 64;               // strictly: the update kernels need one padding column
  a.ntiles = (a.rows + kBM - 1) / kBM;
  constexpr bool kCritic = (MODE == 1 || MODE == 3 || MODE >= 4);        // one register column for the head
  if constexpr (kCritic) {
    if (cfg->in_dim >= kWideFrom) {             // wide first layer: fc1 GEMM, fused 64-wide kernel, fc1 weight-gradient GEMM
      MAPPO_REQUIRE(a.workspace, "%s: in_dim %d needs a workspace of mappo_mlp_workspace_bytes()", who, cfg->in_dim);
      a.Kp = 32;
      a.z1 = static_cast<float*>(a.workspace);
      if (a.packed) { if (int rc = launch_wide_fc1_packed(a, a.workspace, a.packed, stream, who)) return rc; }
      else if (int rc = launch_wide_fc1(a, a.workspace, stream, who)) return rc;
      if (int rc = launch_one<MODE, 32, 1, true>(a, stream, who)) return rc;
      if (MODE == 3 && a.packed) return launch_wide_wgrad1_packed(a, a.workspace, a.packed, stream, who);
      if (MODE == 3 || MODE == 5) return launch_wide_wgrad1(a, a.workspace, stream, who);
      return 0;
    }
    if (a.Kp == 32) return launch_one<MODE, 32, 1>(a, stream, who);
    return launch_one<MODE, 64, 1>(a, stream, who);
  } else {
  const bool small_head = cfg->n_heads == 1 && a.L.a_total <= 8 && a.avail == nullptr;
  if (cfg->in_dim >= kWideFrom) {                 // wide observation (e.g. simple_spread with >= 12 agents): same three-kernel chain
    MAPPO_REQUIRE(a.workspace, "%s: in_dim %d needs a workspace of mappo_mlp_workspace_bytes()", who, cfg->in_dim);
    a.Kp = 32;
    a.z1 = static_cast<float*>(a.workspace);
    if (int rc = launch_wide_fc1(a, a.workspace, stream, who)) return rc;
    if (int rc = small_head ? launch_one<MODE, 32, 8, true>(a, stream, who) : launch_one<MODE, 32, 16, true>(a, stream, who)) return rc;
    if (MODE == 2) return launch_wide_wgrad1(a, a.workspace, stream, who);
    return 0;
  }
  if (small_head) {
    if (a.Kp == 32) return launch_one<MODE, 32, 8>(a, stream, who);
    return launch_one<MODE, 64, 8>(a, stream, who);
  }
  if (a.Kp == 32) return launch_one<MODE, 32, 16>(a, stream, who);
  return launch_one<MODE, 64, 16>(a, stream, who);
  }
}

template int launch<0>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);
template int launch<1>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);
template int launch<2>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);
template int launch<3>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);
template int launch<4>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);
template int launch<5>(MlpArgs&, const mappo_net_config*, cudaStream_t, const char*);

}  // namespace tc
}  // namespace mappo
