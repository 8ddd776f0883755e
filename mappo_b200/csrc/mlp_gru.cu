// Recurrent (GRU) networks: RNNLayer (onpolicy/algorithms/utils/rnn.py:7-77) between the MLP base and the head of
// R_Actor / R_Critic (r_actor_critic.py:43-107,146-165), as a chain of streaming kernels around the fused 64-wide
// kernel of mlp_tc.cu.  Positions p = s * nc + c: step s of chunk (sequence) c; rollout = one step, nc = rows.
//
//   mlp_tc_kernel<MODE 4>   n2[p]   = base(obs[src(p)])                      (features, all positions at once)
//   gru_fwd_kernel          h[p]    = GRU(n2[p], h[p - nc] * mask[p])          (sequential over s inside the CTA; tensor
//                                                                             cores for n2.Wih' and hm.Whh; stores the
//                                                                             gates and the masked previous hidden)
//   gru_head_kernel         y = LN(h[p]) -> head -> sample | value | loss;    dh[p] = dLoss/dh[p] through head and LN
//   gru_bwd_kernel          s = L-1 .. 0: gate derivatives, dn2[p] = dgi.Wih', carry = (dgh.Whh + z dh) * mask[p]
//   wgrad1_wide_kernel      dW_ih = dgi^T n2, dW_hh = dgh^T hm, bias sums     (mlp_wide.cu, generic form)
//   mlp_tc_kernel<MODE 5>   base backward from dn2[p]
//
// GRU (torch.nn.GRU, gate order r, z, n):  r = s(Wir x + bir + Whr h + bhr), z likewise,
// n = tanh(Win x + bin + r (Whn h + bhn)),  h' = (1 - z) n + z h;  x = LN-affine(n2) is folded into Wih' / bih'
// (net.cuh), the LayerNorm after the GRU into the head.  The reference's training branch (rnn.py:30-77) runs cuDNN
// over the mask-free segments of a chunk and multiplies the hidden state by the mask at segment starts: identical to
// multiplying by mask[p] at every step.
#include "mlp_tc_common.cuh"

namespace mappo {
namespace tc {

struct GruArgs {
  NetLayout L;
  const float* params;
  long long nc; int len;
  const int64_t* index;          // [len * nc] sample rows (step-major) or NULL (src = p)
  const float* n2;               // [len * nc][64]
  const float* h0;               // rnn states at sample rows [*][64]; step 0 reads row src(0, c)
  const float* masks;            // [*] at sample rows
  const float* mpos;             // update: masks gathered into position order ([len * nc]), so that no step waits for index[p] first
  float* hs;                     // [len * nc][64]  hidden after each step (rollout: the next rnn-state slot)
  float* hm;                     // [len * nc][64]  masked previous hidden, or NULL
  float* gates;                  // [len * nc][256] r | z | n | (Whn h + bhn) -> backward: dgi_n | P_r | P_z | dgh_n
  float* dh;                     // backward: [len * nc][64] dLoss/dh direct (in), dLoss/dn2 (out, in place)
  // which GRU layer this sweep runs (recurrent_N = 2 runs the sweeps once per layer): offsets of its four parameter
  // blocks, the LayerNorm affine folded into W_ih (layer 0: fc2's; -1 = none: layer 1 reads raw hidden states), and the
  // row stride of h0 (the stored rnn states are [*][recurrent_N][64])
  int wih, whh, bih, bhh, fold_g, fold_be, h0_ld;
};

constexpr uint32_t kGChunk = 192 * 16;                   // bytes of one 8-column chunk of a 192-row weight tile
constexpr uint32_t kGPiece = 8 * kGChunk;                // one fp16 piece of a [192][64] weight
constexpr uint32_t kTile = 2 * 8 * kRowChunk;            // a [128][64] operand tile, both pieces

constexpr uint32_t kStageFloats = 32 * 64;                 // one lane quarter's staging tile: 32 rows x 64 floats
struct GruMap { uint32_t Wih, Whh, T0, T1, T2, bias, bars, stage, total; };
__host__ __device__ inline GruMap make_gru_map(bool bwd) {
  GruMap m;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.Wih = take(2 * kGPiece);
  m.Whh = take(2 * kGPiece);
  m.T0 = take(kTile);
  m.T1 = take(kTile);
  m.T2 = take(bwd ? kTile : 0);
  m.bias = take(4 * 64 * 4);
  m.bars = take(64);
  m.stage = take(bwd ? 4 * kStageFloats * 4 : 2 * 4 * kStageFloats * 4);    // row staging: one (bwd) or two (fwd) tiles per lane quarter
  m.total = o;
  return m;
}

// 1 / (1 + exp(-x)) with ex2.approx and rcp.approx (each within 1 ulp of its result; the exact reciprocal's refinement
// steps were half of a forward step's gate arithmetic): two MUFU operations and two FMAs per element, like tanh_fast
__device__ __forceinline__ float sigmoid_f(float x) {
  float e, r;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(x * -1.4426950408889634f));
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(e + 1.0f));
  return r;
}

__device__ __forceinline__ void stage_gru_weights(unsigned char* sm, const GruMap& m, const GruArgs& a) {
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  for (int e = threadIdx.x; e < 192 * 64; e += kThreads) {
    const int j = e >> 6, k = e & 63;
    put_pair(sm + m.Wih, kGPiece, 192, j, k, P[a.wih + e] * (a.fold_g >= 0 ? P[a.fold_g + k] : 1.0f));
    put_pair(sm + m.Whh, kGPiece, 192, j, k, P[a.whh + e]);
  }
  float* b = reinterpret_cast<float*>(sm + m.bias);       // [0,128): r, z (input + hidden bias); [128,192): b_in; [192,256): b_hn
  if (threadIdx.x < 192) {
    const int j = threadIdx.x;
    float s = P[a.bih + j];
    if (a.fold_be >= 0) for (int k = 0; k < 64; ++k) s = fmaf(P[a.wih + j * 64 + k], P[a.fold_be + k], s);
    if (j < 128) b[j] = s + P[a.bhh + j];
    else { b[j] = s; b[j + 64] = P[a.bhh + j]; }
  }
}

__device__ __forceinline__ void load16(const float* __restrict__ p, float (&v)[16]) {
  const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int i = 0; i < 4; ++i) { const float4 t = __ldg(q + i); v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w; }
}
__device__ __forceinline__ void store16(float* __restrict__ p, const float (&v)[16]) {
  float4* q = reinterpret_cast<float4*>(p);
#pragma unroll
  for (int i = 0; i < 4; ++i) q[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
// ---- coalesced row stores ----------------------------------------------------------------------------------
// Thread (row, column quarter) owns 16 columns of its row (the TMEM lane mapping), so a direct st.global.v4 touches 32
// rows per warp instruction: 32 L1 wavefronts for 512 bytes.  Measured with clock64 (C4): those stores were 15 of the 20
// thousand cycles of a forward step.  The four warps that share a lane quarter (32 rows) therefore pass each array
// through a 32 x 64 staging tile -- 16-byte pieces XOR-swizzled by the row, conflict-free in both directions -- and
// write two whole rows per warp instruction.
__device__ __forceinline__ float4* stage_piece(float* st, int row, int piece) {
  return reinterpret_cast<float4*>(st) + row * 16 + (piece ^ (row & 7));
}
__device__ __forceinline__ void put_rows(float* st, int lane, int cq, const float (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) *stage_piece(st, lane, 4 * cq + i) = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}
// gdst: row 0 of the quarter's 32 rows, `stride` floats between rows; rows >= nvalid are padding of a ragged tile
__device__ __forceinline__ void flush_rows(float* st, float* __restrict__ gdst, int stride, int nvalid, int lane, int cq) {
  const int j = 32 * cq + lane, pc = j & 15, r0 = j >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = r0 + 8 * i;
    if (row < nvalid) *reinterpret_cast<float4*>(gdst + (long long)row * stride + 4 * pc) = *stage_piece(st, row, pc);
  }
}

// The same for loads: the quarter's 128 threads fetch whole rows (two per warp instruction), several arrays in flight,
// and each array then takes one trip through the staging tile to reach the threads that own its columns.
__device__ __forceinline__ void fetch_rows(const float* __restrict__ gsrc, int stride, int nvalid, int lane, int cq, float4 (&q)[4]) {
  const int j = 32 * cq + lane, pc = j & 15, r0 = j >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int row = r0 + 8 * i;
    q[i] = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (row < nvalid) q[i] = __ldg(reinterpret_cast<const float4*>(gsrc + (long long)row * stride + 4 * pc));
  }
}
__device__ __forceinline__ void deal_rows(float* st, int bar_id, int lane, int cq, const float4 (&q)[4], float (&v)[16]) {
  const int j = 32 * cq + lane, pc = j & 15, r0 = j >> 4;
#pragma unroll
  for (int i = 0; i < 4; ++i) *stage_piece(st, r0 + 8 * i, pc) = q[i];
  bar_named(bar_id, 128);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float4 t = *stage_piece(st, lane, 4 * cq + i);
    v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
  }
  bar_named(bar_id, 128);                                               // the tile is free again
}

// acc0 + 2^-12 acc1 of this thread's 16 columns
__device__ __forceinline__ void tmem_pair16(uint32_t a0, uint32_t a1, float (&v)[16]) {
  float w[16];
  umma::tmem_ld16(a0, v);
  umma::tmem_ld16(a1, w);
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = fmaf(w[i], kLoInv, v[i]);
}

// ---- forward sweep ---------------------------------------------------------------------------------------
// TMEM: D1 = n2.Wih'^T (+ hm.Whh^T on the r, z columns): acc0 at column 0, acc1 at 192 (192 columns each);
//       D2 = hm.Whn^T: acc0 at 384, acc1 at 448.
__global__ void __launch_bounds__(kThreads, 1) gru_fwd_kernel(const GruArgs a) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  const GruMap m = make_gru_map(false);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + m.bars);
  if (tid == 0) { umma::mbar_init(bar, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 512);
  stage_gru_weights(sm, m, a);
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const uint32_t la = tb + ((uint32_t)(32 * q4) << 16) + 16 * cq;
  const uint32_t sIh = umma::smem_u32(sm + m.Wih), sHh = umma::smem_u32(sm + m.Whh);
  const uint32_t sN2 = umma::smem_u32(sm + m.T0), sHm = umma::smem_u32(sm + m.T1);
  const float* bias = reinterpret_cast<const float*>(sm + m.bias) + 16 * cq;
  constexpr uint32_t id192 = umma::idesc_f16(128, 192, 0, 0), id128 = umma::idesc_f16(128, 128, 0, 0),
                     id64 = umma::idesc_f16(128, 64, 0, 0);
  uint32_t ph = 0;
  const long long ntiles = (a.nc + kBM - 1) / kBM;
  float* stage = reinterpret_cast<float*>(sm + m.stage) + (size_t)q4 * 2 * kStageFloats;   // this quarter's two staging tiles
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long c = tile * kBM + r;
    const bool valid = c < a.nc;
    const long long c0 = tile * kBM + 32 * q4;                           // first row of this lane quarter
    const int nvalid = (int)max(0LL, min(32LL, a.nc - c0));
    // row 0 of the quarter in a position-indexed array with `stride` floats per row
    auto store_rows_b = [&](float* arr, long long p0, int stride, const float (&v)[16]) {
      float* st = stage + kStageFloats;
      put_rows(st, lane, cq, v);
      bar_named(1 + q4, 128);
      flush_rows(st, arr + p0 * stride, stride, nvalid, lane, cq);
    };
    float h[16], x[16], mask;
    {
      const long long src = valid ? (a.index ? a.index[c] : c) : 0;
      if (valid) { load16(a.h0 + src * a.h0_ld + 16 * cq, h); load16(a.n2 + c * 64 + 16 * cq, x); mask = a.mpos ? __ldg(a.mpos + c) : __ldg(a.masks + src); }
      else {
#pragma unroll
        for (int i = 0; i < 16; ++i) h[i] = x[i] = 0.0f;
        mask = 0.0f;
      }
    }
    for (int s = 0; s < a.len; ++s) {
      const long long p = (long long)s * a.nc + c;
#pragma unroll
      for (int i = 0; i < 16; ++i) h[i] *= mask;                          // hm = h_prev * mask[p]
      store_cols16(sm + m.T0, r, cq, x);
      store_cols16(sm + m.T1, r, cq, h);
      const long long p0 = (long long)s * a.nc + c0;
      if (a.hm) store_rows_b(a.hm, p0, 64, h);
      umma::fence_async_smem();
      umma::tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        umma::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) {
          const uint32_t ao = (uint32_t)ks * 2 * kRowChunk, bo = (uint32_t)ks * 2 * kGChunk;
          const uint64_t xh = umma::smem_desc(sN2 + ao, kRowChunk, 128), xl = umma::smem_desc(sN2 + 8 * kRowChunk + ao, kRowChunk, 128);
          const uint64_t hh = umma::smem_desc(sHm + ao, kRowChunk, 128), hl = umma::smem_desc(sHm + 8 * kRowChunk + ao, kRowChunk, 128);
          const uint64_t ih = umma::smem_desc(sIh + bo, kGChunk, 128), il = umma::smem_desc(sIh + kGPiece + bo, kGChunk, 128);
          const uint64_t wh = umma::smem_desc(sHh + bo, kGChunk, 128), wl = umma::smem_desc(sHh + kGPiece + bo, kGChunk, 128);
          const uint64_t nh = umma::smem_desc(sHh + bo + 128 * 16, kGChunk, 128), nl = umma::smem_desc(sHh + kGPiece + bo + 128 * 16, kGChunk, 128);
          umma::mma_f16(tb, xh, ih, id192, ks > 0);
          umma::mma_f16(tb + 192, xh, il, id192, ks > 0);
          umma::mma_f16(tb + 192, xl, ih, id192, true);
          umma::mma_f16(tb, hh, wh, id128, true);
          umma::mma_f16(tb + 192, hh, wl, id128, true);
          umma::mma_f16(tb + 192, hl, wh, id128, true);
          umma::mma_f16(tb + 384, hh, nh, id64, ks > 0);
          umma::mma_f16(tb + 448, hh, nl, id64, ks > 0);
          umma::mma_f16(tb + 448, hl, nh, id64, true);
        }
        umma::mma_commit(bar);
      }
      // next step's inputs travel while the tensor cores work
      float xn[16], mask_n = 0.0f;
      if (s + 1 < a.len && valid) {
        const long long pn = p + a.nc;
        load16(a.n2 + pn * 64 + 16 * cq, xn);
        mask_n = a.mpos ? __ldg(a.mpos + pn) : __ldg(a.masks + (a.index ? a.index[pn] : pn));
      } else {
#pragma unroll
        for (int i = 0; i < 16; ++i) xn[i] = 0.0f;
      }
      umma::mbar_wait(bar, ph); ph ^= 1;
      umma::tc_fence_after();
      float g[16], t[16];
      // The gate rows leave one array behind the arithmetic: array k is put into one staging tile, array k + 1 is
      // computed (MUFU-bound), then the quarter's barrier and the flush of array k -- so the barrier skew and the
      // shared-memory round trip hide under the next gate's exponentials.  Tiles A and B alternate; a tile is rewritten
      // only after the barrier that follows its flush.  (hm used B at the top of the step; a CTA-wide barrier lies between.)
      float* gates = a.gates;
      float* stA = stage;
      float* stB = stage + kStageFloats;
      tmem_pair16(la, la + 192, g);                                      // r
#pragma unroll
      for (int i = 0; i < 16; ++i) g[i] = sigmoid_f(g[i] + bias[i]);
      if (gates) put_rows(stA, lane, cq, g);
      tmem_pair16(la + 384, la + 448, t);                                // Whn h + bhn
#pragma unroll
      for (int i = 0; i < 16; ++i) t[i] += bias[192 + i];
      if (gates) {
        bar_named(1 + q4, 128);
        flush_rows(stA, gates + p0 * 256, 256, nvalid, lane, cq);
        put_rows(stB, lane, cq, t);
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) t[i] *= g[i];
      tmem_pair16(la + 128, la + 192 + 128, g);                          // n
#pragma unroll
      for (int i = 0; i < 16; ++i) t[i] = tanh_fast(g[i] + bias[128 + i] + t[i]);
      if (gates) {
        bar_named(1 + q4, 128);
        flush_rows(stB, gates + 192 + p0 * 256, 256, nvalid, lane, cq);
        put_rows(stA, lane, cq, t);
      }
      tmem_pair16(la + 64, la + 192 + 64, g);                            // z
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        g[i] = sigmoid_f(g[i] + bias[64 + i]);
        h[i] = fmaf(g[i], h[i] - t[i], t[i]);                            // (1 - z) n + z hm
      }
      if (gates) {
        bar_named(1 + q4, 128);
        flush_rows(stA, gates + 128 + p0 * 256, 256, nvalid, lane, cq);
        put_rows(stB, lane, cq, g);
        bar_named(1 + q4, 128);
        flush_rows(stB, gates + 64 + p0 * 256, 256, nvalid, lane, cq);
      }
      put_rows(stA, lane, cq, h);
      bar_named(1 + q4, 128);
      flush_rows(stA, a.hs + p0 * 64, 64, nvalid, lane, cq);
#pragma unroll
      for (int i = 0; i < 16; ++i) x[i] = xn[i];
      mask = mask_n;
      umma::tc_fence_before();                                           // TMEM reads done before the next step's MMAs
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 512);
}

// ---- backward sweep ----------------------------------------------------------------------------------------
// Per step (descending): dt = dh[p] + carry;  dz~ = dt (hm - n) z (1 - z);  dn~ = dt (1 - z)(1 - n^2);
// dr~ = dn~ ghn r (1 - r);  dgi = (dr~, dz~, dn~), dgh = (dr~, dz~, dn~ r);  dn2 = dgi . Wih',
// carry = (dgh . Whh + dt z) * mask[p].  The gate slots of `gates` are overwritten with dgi_n | dr~ | dz~ | dgh_n.
// TMEM: dn2 acc0 / acc1 at columns 0 / 64, dhm at 128 / 192.
__global__ void __launch_bounds__(kThreads, 1) gru_bwd_kernel(const GruArgs a) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  const GruMap m = make_gru_map(true);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + m.bars);
  if (tid == 0) { umma::mbar_init(bar, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 256);
  stage_gru_weights(sm, m, a);
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const uint32_t la = tb + ((uint32_t)(32 * q4) << 16) + 16 * cq;
  const uint32_t sIh = umma::smem_u32(sm + m.Wih), sHh = umma::smem_u32(sm + m.Whh);
  const uint32_t sT0 = umma::smem_u32(sm + m.T0), sT1 = umma::smem_u32(sm + m.T1), sT2 = umma::smem_u32(sm + m.T2);
  constexpr uint32_t idD = umma::idesc_f16(128, 64, 0, 1);
  // acc += A (K-major [128][64] tile) . W[gate rows][:]   (W read MN-major: contraction over its 64 rows)
  auto dgrad = [&](uint32_t acc0, uint32_t acc1, uint32_t a_tile, uint32_t w_base, int gate, bool accum) {
#pragma unroll
    for (int ks = 0; ks < 4; ++ks) {
      const uint32_t ao = a_tile + (uint32_t)ks * 2 * kRowChunk, bo = w_base + (uint32_t)gate * 1024 + (uint32_t)ks * 256;
      const uint64_t ah = umma::smem_desc(ao, kRowChunk, 128), al = umma::smem_desc(ao + 8 * kRowChunk, kRowChunk, 128);
      const uint64_t bh = umma::smem_desc(bo, 128, kGChunk), bl = umma::smem_desc(bo + kGPiece, 128, kGChunk);
      umma::mma_f16(acc0, ah, bh, idD, accum || ks > 0);
      umma::mma_f16(acc1, ah, bl, idD, accum || ks > 0);
      umma::mma_f16(acc1, al, bh, idD, true);
    }
  };
  uint32_t ph = 0;
  const long long ntiles = (a.nc + kBM - 1) / kBM;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long c = tile * kBM + r;
    const bool valid = c < a.nc;
    const long long c0 = tile * kBM + 32 * q4;                           // first row of this lane quarter
    const int nvalid = (int)max(0LL, min(32LL, a.nc - c0));
    float* stage = reinterpret_cast<float*>(sm + m.stage) + (size_t)q4 * kStageFloats;
    // coalesced row store (see put_rows): one staging tile here, so a second barrier frees it for the next array.  A slot
    // is only overwritten after every thread of the quarter has CONSUMED what it loaded from it (the loads feed
    // arithmetic that precedes the put in program order), since the flush writes other threads' rows.
    auto store_rows = [&](float* arr, long long p0, int stride, const float (&v)[16]) {
      put_rows(stage, lane, cq, v);
      bar_named(1 + q4, 128);
      flush_rows(stage, arr + p0 * stride, stride, nvalid, lane, cq);
      bar_named(1 + q4, 128);
    };
    float carry[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) carry[i] = 0.0f;
    for (int s = a.len - 1; s >= 0; --s) {
      const long long p = (long long)s * a.nc + c;
      // The sweep is one dependent chain per CTA (loads -> gate derivatives -> MMAs -> carry), so nothing else hides
      // the DRAM latency of a step's 192 KB: one thread asks the L2 for the NEXT step's rows (contiguous per tile:
      // positions are step-major) while this step runs.
      if (tid == 32) {
        long long tl = tile;
        int sn = s - 1;
        if (sn < 0) { tl = tile + gridDim.x; sn = a.len - 1; }
        if (tl < ntiles) {
          const long long c0 = tl * kBM, p0 = (long long)sn * a.nc + c0;
          const uint32_t nr = (uint32_t)min((long long)kBM, a.nc - c0);
          umma::l2_prefetch(a.gates + p0 * 256, nr * 1024u);
          umma::l2_prefetch(a.dh + p0 * 64, nr * 256u);
          umma::l2_prefetch(a.hm + p0 * 64, nr * 256u);
        }
      }
      float dt[16], u[16], v[16], w[16], dd[16];
      float mask = 0.0f;
      const long long p0 = (long long)s * a.nc + c0;
      {
        float4 q0[4], q1[4], q2[4], q3[4];                               // four arrays in flight, whole rows
        fetch_rows(a.dh + p0 * 64, 64, nvalid, lane, cq, q0);
        fetch_rows(a.gates + 64 + p0 * 256, 256, nvalid, lane, cq, q1);  // z
        fetch_rows(a.gates + 128 + p0 * 256, 256, nvalid, lane, cq, q2); // n
        fetch_rows(a.hm + p0 * 64, 64, nvalid, lane, cq, q3);            // hm
        if (valid) mask = a.mpos ? __ldg(a.mpos + p) : __ldg(a.masks + (a.index ? a.index[p] : p));
        deal_rows(stage, 1 + q4, lane, cq, q0, dt);
        deal_rows(stage, 1 + q4, lane, cq, q1, u);
        deal_rows(stage, 1 + q4, lane, cq, q2, v);
        deal_rows(stage, 1 + q4, lane, cq, q3, w);
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        dt[i] += carry[i];
        dd[i] = dt[i] * u[i];                                           // dt z
        w[i] = dt[i] * (w[i] - v[i]) * u[i] * (1.0f - u[i]);            // dz~
        v[i] = dt[i] * (1.0f - u[i]) * fmaf(-v[i], v[i], 1.0f);         // dn~
      }
      store_cols16(sm + m.T1, r, cq, w);
      store_cols16(sm + m.T2, r, cq, v);
      {
        float4 q0[4], q1[4];
        fetch_rows(a.gates + p0 * 256, 256, nvalid, lane, cq, q0);       // r
        fetch_rows(a.gates + 192 + p0 * 256, 256, nvalid, lane, cq, q1); // ghn
        deal_rows(stage, 1 + q4, lane, cq, q0, u);
        deal_rows(stage, 1 + q4, lane, cq, q1, dt);
      }
      store_rows(a.gates + 128, p0, 256, w);                            // dz~ (the slot held n: consumed above)
#pragma unroll
      for (int i = 0; i < 16; ++i) w[i] = v[i] * dt[i] * u[i] * (1.0f - u[i]);   // dr~ (consumes r and ghn)
      store_rows(a.gates, p0, 256, v);                                  // dgi_n (the slot held r)
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] *= u[i];                        // dgh_n = dn~ r
      store_cols16(sm + m.T0, r, cq, w);
      store_rows(a.gates + 64, p0, 256, w);                             // dr~ (the slot held z)
      store_rows(a.gates + 192, p0, 256, v);                            // dgh_n (the slot held ghn)
      umma::fence_async_smem();
      umma::tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        umma::tc_fence_after();
        dgrad(tb, tb + 64, sT0, sIh, 0, false);
        dgrad(tb, tb + 64, sT1, sIh, 1, true);
        dgrad(tb, tb + 64, sT2, sIh, 2, true);
        dgrad(tb + 128, tb + 192, sT0, sHh, 0, false);
        dgrad(tb + 128, tb + 192, sT1, sHh, 1, true);
        umma::mma_commit(bar);
      }
      umma::mbar_wait(bar, ph); ph ^= 1;
      store_cols16(sm + m.T0, r, cq, v);                                // dgh_n
      umma::fence_async_smem();
      umma::tc_fence_before();
      __syncthreads();
      if (tid == 0) {
        umma::tc_fence_after();
        dgrad(tb + 128, tb + 192, sT0, sHh, 2, true);
        umma::mma_commit(bar);
      }
      umma::mbar_wait(bar, ph); ph ^= 1;
      umma::tc_fence_after();
      tmem_pair16(la, la + 64, u);                                      // dn2
      store_rows(a.dh, p0, 64, u);                                      // (dh[p] was consumed at the top of the step)
      tmem_pair16(la + 128, la + 192, u);                               // dgh . Whh
#pragma unroll
      for (int i = 0; i < 16; ++i) carry[i] = (u[i] + dd[i]) * mask;
      umma::tc_fence_before();
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 256);
}

}  // namespace tc
}  // namespace mappo

// ---- head on the hidden state (fp32 CUDA cores: the head is 64 x A, A <= 16) -------------------------------------
#include "mlp_fma.cuh"

namespace mappo {
namespace tc {

struct HeadSmem {
  static constexpr int BM = 128;
  static constexpr int DL = 20;                          // dlogits row stride: 16 logits + 4 (rows 16 B aligned)
  static constexpr int Whf = 0;                          // [16][64] Wh[a][k] gr[k]
  static constexpr int bhf = Whf + kMaxA * 64;           // [16]     bh + Wh ber
  static constexpr int TY = bhf + kMaxA;                 // [BM][kLD] y = LN(h) without the affine
  static constexpr int rstd = TY + BM * kLD;             // [BM]
  static constexpr int dlog = rstd + BM;                 // [BM][DL]
  static constexpr int total = dlog + BM * DL;
  static constexpr size_t bytes = sizeof(float) * (size_t)total;
};

// MODE 0 actor rollout (sample / evaluate), 1 critic rollout, 2 actor update, 3 critic update.
// Row g of this launch is position p = g; its sample row is index[g] (update) or g.  h: [rows][64] hidden states.
// Per tile of 128 rows: (A) LayerNorm, one warp per row; (B) logits with two threads per row (32 columns each, combined
// by one shuffle), the loss on the even thread; (C) head weight gradient G_h[a][k] += dlog[r][a] y[r][k] as a register
// tile -- thread (row group of 32, 4 logits, 4 columns) does 16 FMAs per two 16-byte shared-memory loads and keeps its
// 4 x 4 block in registers across all tiles; (D) dy = dlog . Wh' and (E) back through the LayerNorm.
template <int MODE>
__global__ void __launch_bounds__(256, 2) gru_head_kernel(const MlpArgs a, const float* __restrict__ h, float* __restrict__ dh) {
  constexpr bool UPDATE = (MODE >= 2);
  constexpr int BM = HeadSmem::BM, NT = 256, NTY = NT / 16, RPT = BM / NTY, DL = HeadSmem::DL;
  using S = HeadSmem;
  extern __shared__ __align__(16) float smf[];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4, warp = tid >> 5, lane = tid & 31;
  const NetLayout& L = a.L;
  const float* __restrict__ P = a.params;
  const int A = L.a_total;
  for (int e = tid; e < kMaxA * 64; e += NT) {
    const int aidx = e >> 6, k = e & 63;
    float v = 0.0f;
    int hh = 0, off = 0;
    while (hh < L.n_heads && aidx >= off + L.head_dim[hh]) { off += L.head_dim[hh]; ++hh; }
    if (hh < L.n_heads) v = P[L.wh[hh] + (aidx - off) * 64 + k] * P[L.gr + k];
    smf[S::Whf + e] = v;
  }
  if (tid < kMaxA) {
    int hh = 0, off = 0;
    while (hh < L.n_heads && tid >= off + L.head_dim[hh]) { off += L.head_dim[hh]; ++hh; }
    float s = 0.0f;
    if (hh < L.n_heads) {
      s = P[L.bh[hh] + (tid - off)];
      for (int k = 0; k < 64; ++k) s = fmaf(P[L.wh[hh] + (tid - off) * 64 + k], P[L.ber + k], s);
    }
    smf[S::bhf + tid] = s;
  }
  // phase C mapping: row group (32 rows), logit block (4 logits), column block (4 columns)
  const int c_k4 = tid & 15, c_ab = (tid >> 4) & 3, c_rg = tid >> 6;
  float gacc[4][4], sb[4];
  double lacc[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
  for (int j = 0; j < 4; ++j) { sb[j] = 0.0f; gacc[j][0] = gacc[j][1] = gacc[j][2] = gacc[j][3] = 0.0f; }
  const long long ntiles = (a.rows + BM - 1) / BM;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    __syncthreads();
    if (tid == 0 && tile + gridDim.x < ntiles) {                          // the next tile's rows: ask the L2 for them now
      const long long g0 = (tile + gridDim.x) * BM;
      umma::l2_prefetch(h + g0 * 64, (uint32_t)min((long long)BM, a.rows - g0) * 256u);
    }
    // (A) y = LN(h), one warp per row; a warp's rows are loaded eight at a time so that one DRAM round trip serves eight
    // rows instead of one
    constexpr int kRowsPerWarp = BM / (NT / 32), kBatch = 8;
    static_assert(kRowsPerWarp % kBatch == 0, "rows per warp");
#pragma unroll 1
    for (int rb = 0; rb < kRowsPerWarp; rb += kBatch) {
      float w0[kBatch], w1[kBatch];
#pragma unroll
      for (int i = 0; i < kBatch; ++i) {
        const long long g = tile * BM + warp + (NT / 32) * (rb + i);
        w0[i] = w1[i] = 0.0f;
        if (g < a.rows) { w0[i] = __ldg(h + g * 64 + lane); w1[i] = __ldg(h + g * 64 + lane + 32); }
      }
#pragma unroll
      for (int i = 0; i < kBatch; ++i) {
      const int r = warp + (NT / 32) * (rb + i);
      float v0 = w0[i], v1 = w1[i];
      float s = v0 + v1;
      for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      const float mean = s * (1.0f / 64.0f);
      v0 -= mean; v1 -= mean;
      float q = v0 * v0 + v1 * v1;
      for (int o = 16; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      const float rstd = 1.0f / sqrtf(q * (1.0f / 64.0f) + kLnEps);
      smf[S::TY + r * kLD + lane] = v0 * rstd;
      smf[S::TY + r * kLD + lane + 32] = v1 * rstd;
      if (lane == 0) smf[S::rstd + r] = rstd;
      }
    }
    __syncthreads();
    {                                                                     // (B) logits: two threads per row
      const int row = tid >> 1, half = tid & 1;
      const long long g = tile * BM + row;
      float l[kMaxA];
#pragma unroll
      for (int q = 0; q < kMaxA; ++q) l[q] = 0.0f;
      const float* yrow = smf + S::TY + row * kLD + 32 * half;
      const float* wbase = smf + S::Whf + 32 * half;
#pragma unroll 2
      for (int k4 = 0; k4 < 8; ++k4) {
        const float4 n = *reinterpret_cast<const float4*>(yrow + 4 * k4);
#pragma unroll
        for (int q = 0; q < kMaxA; ++q) {
          if (q < A) {
            const float4 w = *reinterpret_cast<const float4*>(wbase + q * 64 + 4 * k4);
            l[q] = fmaf(n.x, w.x, fmaf(n.y, w.y, fmaf(n.z, w.z, fmaf(n.w, w.w, l[q]))));
          }
        }
      }
#pragma unroll
      for (int q = 0; q < kMaxA; ++q) {
        const float other = __shfl_xor_sync(0xffffffffu, l[q], 1);
        l[q] = smf[S::bhf + q] + (half ? other + l[q] : l[q] + other);    // the same sum on both threads of the row
      }
      if (half == 0) {
        float* dl = smf + S::dlog + row * DL;
        if (g < a.rows) {
          const long long src = a.index ? a.index[g] : g;
          if (MODE == 0) actor_rollout_row(a, g, l);
          if (MODE == 1) a.values_out[g] = l[0];
          if (MODE == 2) actor_loss_row(a, src, l, dl, lacc);
          if (MODE == 3) {
            dl[0] = critic_loss_row(a, src, l[0], lacc);
            for (int q = 1; q < kMaxA; ++q) dl[q] = 0.0f;
          }
        } else if (UPDATE) {
          for (int q = 0; q < kMaxA; ++q) dl[q] = 0.0f;
        }
      }
    }
    if (!UPDATE) continue;
    __syncthreads();
    {                                                                     // (C) G_h += dlog^T y over this thread's 32 rows
      const float* yb = smf + S::TY + 4 * c_k4 + 32 * c_rg * kLD;
      const float* db = smf + S::dlog + 4 * c_ab + 32 * c_rg * DL;
#pragma unroll 4
      for (int rr = 0; rr < 32; ++rr) {
        const float4 y = *reinterpret_cast<const float4*>(yb + rr * kLD);
        const float4 d = *reinterpret_cast<const float4*>(db + rr * DL);
        const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          gacc[j][0] = fmaf(dv[j], y.x, gacc[j][0]); gacc[j][1] = fmaf(dv[j], y.y, gacc[j][1]);
          gacc[j][2] = fmaf(dv[j], y.z, gacc[j][2]); gacc[j][3] = fmaf(dv[j], y.w, gacc[j][3]);
          sb[j] += dv[j];
        }
      }
    }
    float acc[RPT][4];                                                    // (D) dy = dlog . Wh'
#pragma unroll
    for (int i = 0; i < RPT; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0.0f;
    for (int aa = 0; aa < A; ++aa) {
      const float4 w = *reinterpret_cast<const float4*>(smf + S::Whf + aa * 64 + 4 * tx);
#pragma unroll
      for (int i = 0; i < RPT; ++i) {
        const float d = smf[S::dlog + (ty + NTY * i) * DL + aa];
        acc[i][0] = fmaf(d, w.x, acc[i][0]); acc[i][1] = fmaf(d, w.y, acc[i][1]);
        acc[i][2] = fmaf(d, w.z, acc[i][2]); acc[i][3] = fmaf(d, w.w, acc[i][3]);
      }
    }
#pragma unroll
    for (int i = 0; i < RPT; ++i) {                                       // (E) through the LayerNorm: dh = rstd (dy - mean(dy) - y mean(dy y))
      const int r = ty + NTY * i;
      const long long g = tile * BM + r;
      const float4 y = *reinterpret_cast<const float4*>(smf + S::TY + r * kLD + 4 * tx);
      const float s1 = hw_sum((acc[i][0] + acc[i][1]) + (acc[i][2] + acc[i][3])) * (1.0f / 64.0f);
      const float s2 = hw_sum((acc[i][0] * y.x + acc[i][1] * y.y) + (acc[i][2] * y.z + acc[i][3] * y.w)) * (1.0f / 64.0f);
      const float rstd = smf[S::rstd + r];
      if (g < a.rows)
        *reinterpret_cast<float4*>(dh + g * 64 + 4 * tx) =
            make_float4(rstd * (acc[i][0] - s1 - y.x * s2), rstd * (acc[i][1] - s1 - y.y * s2),
                        rstd * (acc[i][2] - s1 - y.z * s2), rstd * (acc[i][3] - s1 - y.w * s2));
    }
  }
  if (!UPDATE) return;
  // head slots of partial row blockIdx.x (the buffer was zeroed by the caller): the four row groups' register blocks are
  // summed in group order through shared memory (the y tile and the dlogits tile are free now)
  __syncthreads();
  float* red = smf + S::TY;                                               // [4 row groups][16][64]
  float* redb = smf + S::dlog;                                            // [4 row groups][16]
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    *reinterpret_cast<float4*>(red + c_rg * (kMaxA * 64) + (4 * c_ab + j) * 64 + 4 * c_k4) =
        make_float4(gacc[j][0], gacc[j][1], gacc[j][2], gacc[j][3]);
    if (c_k4 == 0) redb[c_rg * kMaxA + 4 * c_ab + j] = sb[j];
  }
  __syncthreads();
  float* __restrict__ gp = a.grad_partials + (size_t)blockIdx.x * L.total;
  for (int o = tid; o < A * 64; o += NT) {
    const int aa = o >> 6, k = o & 63;
    int hh = 0, off = 0;
    while (aa >= off + L.head_dim[hh]) { off += L.head_dim[hh]; ++hh; }
    gp[L.wh[hh] + (aa - off) * 64 + k] = (red[o] + red[kMaxA * 64 + o]) + (red[2 * kMaxA * 64 + o] + red[3 * kMaxA * 64 + o]);
    if (k == 0) gp[L.bh[hh] + (aa - off)] = (redb[aa] + redb[kMaxA + aa]) + (redb[2 * kMaxA + aa] + redb[3 * kMaxA + aa]);
  }
  __shared__ double lred[4][8];
  for (int q = 0; q < 4; ++q) {
    double v = lacc[q];
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    if (lane == 0) lred[q][warp] = v;
  }
  __syncthreads();
  if (tid < 4) {
    double v = 0.0;
    for (int w = 0; w < NT / 32; ++w) v += lred[tid][w];
    a.loss_partials[(size_t)blockIdx.x * 4 + tid] = v;
  }
}

// ---- host side ---------------------------------------------------------------------------------------------
int grid_for(long long rows);                                                        // mlp_tc.cu
template <int MODE> int launch(MlpArgs& a, const mappo_net_config* cfg, cudaStream_t stream, const char* who);
}  // namespace tc
}  // namespace mappo

#include "mlp_wide.cuh"

namespace mappo {
namespace tc {

constexpr int kGruMaxSmem = 232448 - 1024;

static int check_rnn(const mappo_net_config* cfg, const char* who) {
  if (int rc = check_net(cfg, who)) return rc;
  MAPPO_REQUIRE(cfg->recurrent, "%s: not a recurrent network", who);
  if (cfg->in_dim >= kWideFrom && !wide_fits(cfg->in_dim))
    MAPPO_UNSUPPORTED("%s: in_dim %d is too wide for the staged first layer", who, cfg->in_dim);
  return 0;
}

static int set_attrs() {
  static bool done = false;
  if (done) return 0;
  MAPPO_CUDA(cudaFuncSetAttribute(gru_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGruMaxSmem));
  MAPPO_CUDA(cudaFuncSetAttribute(gru_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kGruMaxSmem));
  MAPPO_CUDA(cudaFuncSetAttribute(gru_head_kernel<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HeadSmem::bytes));
  MAPPO_CUDA(cudaFuncSetAttribute(gru_head_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HeadSmem::bytes));
  MAPPO_CUDA(cudaFuncSetAttribute(gru_head_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HeadSmem::bytes));
  MAPPO_CUDA(cudaFuncSetAttribute(gru_head_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)HeadSmem::bytes));
  done = true;
  return 0;
}

static int head_grid(long long rows) {
  const long long nt = (rows + HeadSmem::BM - 1) / HeadSmem::BM;
  const int g = 2 * grid_for(rows);                       // two CTAs per SM; <= the partial rows of an update
  return (int)(nt < g ? nt : g);
}

// floats of workspace per position: rollout keeps the base features only (+ per GRU layer of a stacked network a
// contiguous copy of its incoming and outgoing state); an update keeps n2, dh and per layer hs | hm | gates, + 1 for the
// gathered masks; + the wide first layer's scratch (z1 [64] and row statistics [2]) when the base reads > 64 columns
long long rnn_workspace_floats(long long positions, int update, int in_dim, int layers) {
  const int extra = layers > 1 ? (update ? (layers - 1) * 384 : layers * 128) : 0;
  return positions * ((update ? 513 : 64) + extra + (in_dim >= kWideFrom ? 66 : 0)) + (layers > 1 ? 64 : 0);   // + alignment slack
}

// the sweep arguments of GRU layer l (rnn.py:13: layer l > 0 reads the raw hidden states of the layer below)
static void set_layer(GruArgs& g, const NetLayout& L, int l) {
  if (l == 0) { g.wih = L.wih; g.whh = L.whh; g.bih = L.bih; g.bhh = L.bhh; g.fold_g = L.g2; g.fold_be = L.be2; }
  else { g.wih = L.wih1; g.whh = L.whh1; g.bih = L.bih1; g.bhh = L.bhh1; g.fold_g = -1; g.fold_be = -1; }
}

// mpos[p] = masks[index[p]]: the sweeps read their step's mask with the rest of the step's rows instead of behind a
// dependent index load
__global__ void gather_masks_kernel(const float* __restrict__ masks, const int64_t* __restrict__ index, float* __restrict__ mpos,
                                    long long n) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    mpos[i] = __ldg(masks + index[i]);
}

// One forward step for `rows` rows (rollout / evaluation): base -> GRU (recurrent_N layers) -> head.
// h_in / h_out: [rows][recurrent_N][64] (rnn.py:24-28: every layer's state is multiplied by the mask before the step).
template <int MODE>
int rnn_forward(MlpArgs& a, const mappo_net_config* cfg, const float* h_in, const float* masks, float* h_out,
                cudaStream_t stream, const char* who) {
  if (int rc = check_rnn(cfg, who)) return rc;
  MAPPO_REQUIRE(a.params && a.x && a.rows > 0 && a.ld >= cfg->in_dim && h_in && masks && h_out && a.workspace,
                "%s: bad arguments", who);
  if (int rc = set_attrs()) return rc;
  const int RN = cfg->recurrent;
  const long long rows = a.rows;
  float* ws = static_cast<float*>(a.workspace);
  float* n2 = ws;
  a.feat = n2;
  a.workspace = n2 + rows * 64;                          // the wide first layer's scratch, if the base needs one
  if (int rc = launch<4>(a, cfg, stream, who)) return rc;
  GruArgs g{};
  g.L = a.L; g.params = a.params; g.nc = rows; g.len = 1; g.index = nullptr; g.masks = masks; g.h0_ld = 64;
  const float* h_top = h_out;
  if (RN == 1) {
    set_layer(g, a.L, 0);
    g.n2 = n2; g.h0 = h_in; g.hs = h_out;
    gru_fwd_kernel<<<grid_for(rows), kThreads, make_gru_map(false).total, stream>>>(g);
    MAPPO_CHECK_LAUNCH(who);
  } else {
    // stacked layers: each sweep works on contiguous [rows][64] copies of its state slice
    float* lay = ws + ((rows * (64 + (cfg->in_dim >= kWideFrom ? 66 : 0)) + 63) & ~63LL);      // 256-byte aligned rows
    const float* x = n2;
    for (int l = 0; l < RN; ++l) {
      float* hin = lay + (size_t)(2 * l) * rows * 64;
      float* hs = lay + (size_t)(2 * l + 1) * rows * 64;
      MAPPO_CUDA(cudaMemcpy2DAsync(hin, 256, h_in + 64 * l, (size_t)RN * 256, 256, (size_t)rows, cudaMemcpyDeviceToDevice, stream));
      set_layer(g, a.L, l);
      g.n2 = x; g.h0 = hin; g.hs = hs;
      gru_fwd_kernel<<<grid_for(rows), kThreads, make_gru_map(false).total, stream>>>(g);
      MAPPO_CHECK_LAUNCH(who);
      MAPPO_CUDA(cudaMemcpy2DAsync(h_out + 64 * l, (size_t)RN * 256, hs, 256, 256, (size_t)rows, cudaMemcpyDeviceToDevice, stream));
      x = hs;
    }
    h_top = x;
  }
  gru_head_kernel<MODE><<<head_grid(rows), 256, HeadSmem::bytes, stream>>>(a, h_top, nullptr);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

// One PPO minibatch of n_chunks sequences of chunk_len steps: positions p = s * n_chunks + c, sample row index[p].
// rnn_states: [*][recurrent_N][64] at sample rows.
template <int MODE>
int rnn_update(MlpArgs& a, const mappo_net_config* cfg, long long n_chunks, int chunk_len, const float* rnn_states,
               const float* masks, int n_partials, cudaStream_t stream, const char* who) {
  if (int rc = check_rnn(cfg, who)) return rc;
  const long long B = n_chunks * chunk_len;
  MAPPO_REQUIRE(a.params && a.x && n_chunks > 0 && chunk_len > 0 && a.ld >= cfg->in_dim && rnn_states && masks &&
                a.index && a.workspace && a.grad_partials && a.loss_partials, "%s: bad arguments", who);
  MAPPO_REQUIRE(n_partials >= 2 * grid_for(B), "%s: %d partial rows, need %d", who, n_partials, 2 * grid_for(B));
  if (int rc = set_attrs()) return rc;
  const int RN = cfg->recurrent;
  a.rows = B;
  a.L = make_layout(*cfg);
  const size_t P = (size_t)a.L.total;
  MAPPO_CUDA(cudaMemsetAsync(a.grad_partials, 0, sizeof(float) * P * (size_t)n_partials, stream));
  MAPPO_CUDA(cudaMemsetAsync(a.loss_partials, 0, sizeof(double) * 4 * (size_t)n_partials, stream));
  float* ws = static_cast<float*>(a.workspace);
  const int wide = cfg->in_dim >= kWideFrom ? 66 : 0;
  float *n2 = ws, *dh = ws + B * 192;
  float *hs[kMaxRecurrentN], *hm[kMaxRecurrentN], *gates[kMaxRecurrentN];
  hs[0] = ws + B * 64; hm[0] = ws + B * 128; gates[0] = ws + B * 256;
  float* mpos = ws + B * (512 + wide);
  for (int l = 1; l < RN; ++l) {
    float* base = ws + ((B * (513 + wide) + 63) & ~63LL) + (size_t)(l - 1) * B * 384;     // 256-byte aligned rows
    hs[l] = base; hm[l] = base + B * 64; gates[l] = base + B * 128;
  }
  a.workspace = ws + B * 512;                            // the wide first layer's scratch, if the base needs one
  gather_masks_kernel<<<(int)((B + 255) / 256 < 1184 ? (B + 255) / 256 : 1184), 256, 0, stream>>>(masks, a.index, mpos, B);
  MAPPO_CHECK_LAUNCH(who);
  a.feat = n2;
  if (int rc = launch<4>(a, cfg, stream, who)) return rc;
  GruArgs g{};
  g.L = a.L; g.params = a.params; g.nc = n_chunks; g.len = chunk_len; g.index = a.index;
  g.masks = masks; g.mpos = mpos; g.dh = dh; g.h0_ld = 64 * RN;
  for (int l = 0; l < RN; ++l) {                          // forward sweeps, bottom up
    set_layer(g, a.L, l);
    g.n2 = l == 0 ? n2 : hs[l - 1]; g.h0 = rnn_states + 64 * l; g.hs = hs[l]; g.hm = hm[l]; g.gates = gates[l];
    gru_fwd_kernel<<<grid_for(n_chunks), kThreads, make_gru_map(false).total, stream>>>(g);
    MAPPO_CHECK_LAUNCH(who);
  }
  gru_head_kernel<MODE><<<head_grid(B), 256, HeadSmem::bytes, stream>>>(a, hs[RN - 1], dh);
  MAPPO_CHECK_LAUNCH(who);
  for (int l = RN - 1; l >= 0; --l) {                     // backward sweeps, top down: dh becomes dLoss/d(input of layer l)
    set_layer(g, a.L, l);
    g.n2 = l == 0 ? n2 : hs[l - 1]; g.h0 = rnn_states + 64 * l; g.hs = hs[l]; g.hm = hm[l]; g.gates = gates[l];
    gru_bwd_kernel<<<grid_for(n_chunks), kThreads, make_gru_map(true).total, stream>>>(g);
    MAPPO_CHECK_LAUNCH(who);
    WideArgs w{};
    w.rows = B; w.index = nullptr; w.x_div = 1; w.vec = 1; w.ld = 256; w.D = 192; w.dz_ld = 64; w.norm = 0;
    w.total = a.L.total; w.transposed = 1; w.grad_partials = a.grad_partials;
    // ONE pass over the 256-column gradient rows `dgi_n | dr~ | dz~ | dgh_n` yields both weight gradients: dW_ih from the
    // columns [0, 192) against n2 (stored rotated: n, r, z -> r, z, n), dW_hh from [64, 256) against hm
    w.D = 256; w.win = 192;
    w.x = gates[l]; w.dz1 = const_cast<float*>(g.n2); w.rot = 128; w.out_w = g.wih; w.out_b = g.bih;
    w.dz2 = hm[l]; w.off2 = 64; w.rot2 = 0; w.out_w2 = g.whh; w.out_b2 = g.bhh;
    if (int rc = launch_wgrad_generic(w, stream, who)) return rc;
  }
  a.feat = dh;
  return launch<5>(a, cfg, stream, who);
}

template int rnn_forward<0>(MlpArgs&, const mappo_net_config*, const float*, const float*, float*, cudaStream_t, const char*);
template int rnn_forward<1>(MlpArgs&, const mappo_net_config*, const float*, const float*, float*, cudaStream_t, const char*);
template int rnn_update<2>(MlpArgs&, const mappo_net_config*, long long, int, const float*, const float*, int, cudaStream_t, const char*);
template int rnn_update<3>(MlpArgs&, const mappo_net_config*, long long, int, const float*, const float*, int, cudaStream_t, const char*);

}  // namespace tc
}  // namespace mappo

using namespace mappo;

extern "C" long long mappo_rnn_workspace_bytes(const mappo_net_config* cfg, long long positions, int update) {
  if (!cfg || positions < 0) return -1;
  return cfg->recurrent ? 4 * tc::rnn_workspace_floats(positions, update, cfg->in_dim, cfg->recurrent) : 0;
}

extern "C" int mappo_actor_forward_rnn(const mappo_net_config* cfg, const float* params, const float* obs, int obs_ld,
                                       long long rows, const float* rnn_states_in, const float* masks,
                                       const float* available_actions, int deterministic, uint64_t seed, uint64_t step,
                                       long long row_offset, const float* given_actions, float* rnn_states_out,
                                       float* actions_out, float* logp_out, float* entropy_out,
                                       const int64_t* step_offset, void* workspace, void* stream) {
  MAPPO_REQUIRE(logp_out && (actions_out || given_actions), "actor_forward_rnn: null outputs");
  MAPPO_REQUIRE(!cfg || !cfg->is_critic, "actor_forward_rnn: critic config");
  MlpArgs a{};
  a.params = params; a.x = obs; a.ld = obs_ld; a.rows = rows; a.avail = available_actions;
  a.deterministic = deterministic; a.k0 = (uint32_t)(seed & 0xFFFFFFFFull); a.step = (uint32_t)step;
  a.step_dev = step_offset;
  a.row_offset = row_offset; a.actions_out = actions_out; a.logp_out = logp_out;
  a.given_actions = given_actions; a.entropy_out = entropy_out; a.workspace = workspace;
  return tc::rnn_forward<0>(a, cfg, rnn_states_in, masks, rnn_states_out, (cudaStream_t)stream, "actor_forward_rnn");
}

extern "C" int mappo_critic_forward_rnn(const mappo_net_config* cfg, const float* params, const float* share_obs,
                                        int obs_ld, long long rows, int share_rows_div, const float* rnn_states_in,
                                        const float* masks, float* rnn_states_out, float* values_out, void* workspace,
                                        void* stream) {
  MAPPO_REQUIRE(values_out && share_rows_div >= 1, "critic_forward_rnn: bad outputs");
  MAPPO_REQUIRE(!cfg || cfg->is_critic, "critic_forward_rnn: actor config");
  MlpArgs a{};
  a.params = params; a.x = share_obs; a.ld = obs_ld; a.rows = rows; a.group = 1; a.values_out = values_out;
  a.x_div = share_rows_div; a.workspace = workspace;
  return tc::rnn_forward<1>(a, cfg, rnn_states_in, masks, rnn_states_out, (cudaStream_t)stream, "critic_forward_rnn");
}

extern "C" int mappo_actor_update_rnn(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                                      const float* obs, int obs_ld, long long n_chunks, int chunk_len,
                                      const int64_t* index, const float* rnn_states, const float* masks,
                                      const float* actions, const float* old_logp, const float* adv,
                                      const float* active_masks, const float* available_actions, float* grad_partials,
                                      int n_partials, double* loss_partials, void* workspace, void* stream) {
  MAPPO_REQUIRE(loss && actions && old_logp && adv, "actor_update_rnn: null pointer");
  MAPPO_REQUIRE(!cfg || !cfg->is_critic, "actor_update_rnn: critic config");
  MlpArgs a{};
  a.params = params; a.x = obs; a.ld = obs_ld; a.index = index; a.x_div = 1; a.actions = actions;
  a.old_logp = old_logp; a.adv = adv; a.active = active_masks; a.avail = available_actions; a.loss = *loss;
  a.grad_partials = grad_partials; a.loss_partials = loss_partials; a.workspace = workspace;
  return tc::rnn_update<2>(a, cfg, n_chunks, chunk_len, rnn_states, masks, n_partials, (cudaStream_t)stream,
                           "actor_update_rnn");
}

extern "C" int mappo_critic_update_rnn(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                                       const float* share_obs, int obs_ld, long long n_chunks, int chunk_len,
                                       const int64_t* index, int share_rows_div, const float* rnn_states,
                                       const float* masks, const float* value_preds, const float* returns,
                                       const float* active_masks, const float* vn_stats, float* grad_partials,
                                       int n_partials, double* loss_partials, void* workspace, void* stream) {
  MAPPO_REQUIRE(loss && value_preds && returns && share_rows_div >= 1, "critic_update_rnn: null pointer");
  MAPPO_REQUIRE(!cfg || cfg->is_critic, "critic_update_rnn: actor config");
  MAPPO_REQUIRE(!loss->use_value_norm || vn_stats, "critic_update_rnn: use_value_norm without statistics");
  MlpArgs a{};
  a.params = params; a.x = share_obs; a.ld = obs_ld; a.index = index; a.x_div = share_rows_div; a.group = 1;
  a.value_preds = value_preds; a.returns = returns; a.active = active_masks; a.vn_stats = vn_stats; a.loss = *loss;
  a.grad_partials = grad_partials; a.loss_partials = loss_partials; a.workspace = workspace;
  return tc::rnn_update<3>(a, cfg, n_chunks, chunk_len, rnn_states, masks, n_partials, (cudaStream_t)stream,
                           "critic_update_rnn");
}
