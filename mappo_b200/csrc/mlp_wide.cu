// First layer of a WIDE-input network (in_dim > 64: the centralised critic of many-agent scenarios, share_obs =
// M * obs_dim = 460 for simple_spread with 10 agents) as two streaming tensor-core GEMMs around the fused 64-wide kernel
// of mlp_tc.cu:
//
//   fc1_wide_kernel     z1[rows][64]  = LayerNorm_feat(x)[rows][D] . W1'^T          (forward; also writes the row
//                                                                                     statistics mean, rstd)
//   mlp_tc_kernel<WIDE> reads z1 where it would read its own fc1 accumulator, and writes dz1 = dLoss/dz1 in place
//   wgrad1_wide_kernel  G1[64][D]     = sum_rows dz1[row][:]^T x^[row][:]           (weight gradient, split over rows and
//                                                                                     column groups; partial rows of the
//                                                                                     CTAs go to the W1 slots of the
//                                                                                     main kernel's gradient partials)
//
// Both kernels are bound by reading x once from HBM (D * 4 bytes per row); the products use the fp16-pair scheme of
// mlp_tc.cu (hi*hi + 2^-12 (hi*lo + lo*hi), fp32 accumulation in TMEM).  x rows are normalised without the affine,
// which is folded into W1' (net.cuh), so G1 / db1 are gradients with respect to the folded weights as everywhere else.
//
// Reference: mlp.py:44-59 (feature_norm, fc1), r_actor_critic.py:146-165, r_mappo.py:52-89,151-156.
#include "mlp_tc_common.cuh"
#include "mlp_wide.cuh"

namespace mappo {
namespace tc {

// 8 consecutive floats of a row, columns [c, c + 8), zero beyond D
__device__ __forceinline__ void load8(const float* __restrict__ row, int c, int D, bool ok, int vec, float (&v)[8]) {
  if (ok && vec && c + 8 <= D) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(row + c));
    const float4 b = __ldg(reinterpret_cast<const float4*>(row + c + 4));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (ok && c + i < D) ? __ldg(row + c + i) : 0.0f;
  }
}

// ---- forward ---------------------------------------------------------------------------------------------
// Shared memory: W1' as the K-major B operand (both fp16 pieces, K padded to a multiple of 16), two X buffers of
// one 64-column K-chunk each (128 rows, both pieces), the row-exchange buffers, two mbarriers (one per X buffer).
// nhalf = 2 | 4: a CTA keeps 32 | 16 of the 64 output columns of W1' (in_dim up to ~1,250 | ~2,500; x is then read once
// per part).
struct FwdMap { uint32_t W, w_piece, w_chunk, X, xch, bars, total; int K16; };
__host__ __device__ inline FwdMap make_fwd_map(int D, int nhalf = 1) {
  FwdMap m;
  m.K16 = (D + 15) / 16;
  m.w_chunk = kWChunk / (uint32_t)nhalf;                               // bytes of one 8-column chunk of the staged W1' rows
  m.w_piece = (uint32_t)(2 * m.K16) * m.w_chunk;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.W = take(2 * m.w_piece);
  m.X = take(2 * 2 * 8 * kRowChunk);
  m.xch = take(2 * 4 * kBM * 8);
  m.bars = take(64);
  m.total = o;
  return m;
}

// One pass over x: the staged operand is d = x - x[row][0] (exactly representable differences keep the fp16 pairs
// accurate); the LayerNorm is applied algebraically when the accumulators leave TMEM,
//   LN(x) . W'^T = rstd (d . W'^T - (mean - x0) colsum(W')),
// with the row sums of d and d^2 gathered while staging.  The next 64-column chunk travels in registers while the
// current one is converted and multiplied.
__global__ void __launch_bounds__(kThreads, 1) fc1_wide_kernel(const WideArgs a) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  __shared__ float cs[64];                                              // colsum_k W'[j][k]
  const NetLayout& L = a.L;
  const int D = L.in_dim;
  const FwdMap m = make_fwd_map(D, a.nhalf);
  const int JN = 64 / a.nhalf, j0 = JN * blockIdx.y, ncol = JN / 4;   // this CTA's output columns; columns per thread
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + m.bars);          // bars[b]: MMAs that read X buffer b
  if (tid == 0) { umma::mbar_init(bars, 1); umma::mbar_init(bars + 1, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 128);
  {
    const float* __restrict__ P = a.params;
    const int Kp = 16 * m.K16;
    for (int e = tid; e < JN * Kp; e += kThreads) {
      const int j = e / Kp, k = e - j * Kp;
      float v = 0.0f;
      if (k < D) v = P[L.w1 + (j0 + j) * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
      put_pair(sm + m.W, m.w_piece, JN, j, k, v);
    }
    for (int j = warp; j < JN; j += kThreads / 32) {                   // one warp per output row, fixed lane order
      float t = 0.0f;
      for (int k = lane; k < D; k += 32) t += P[L.w1 + (j0 + j) * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      if (lane == 0) cs[j] = t;
    }
  }
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const uint32_t lane_addr = tb + ((uint32_t)(32 * q4) << 16);
  const uint32_t sX = umma::smem_u32(sm + m.X), sW = umma::smem_u32(sm + m.W);
  constexpr uint32_t kXBuf = 2 * 8 * kRowChunk;
  const uint32_t kId = umma::idesc_f16(128, JN, 0, 0);
  float2* xch = reinterpret_cast<float2*>(sm + m.xch);
  int slot = 0;
  const int nkc = (m.K16 + 3) / 4;                                     // 64-column K-chunks per tile
  const long long ntiles = (a.rows + kBM - 1) / kBM;
  uint32_t n = 0;                                                      // MMA groups committed so far

  struct RowInfo { const float* row; long long g; float x0; bool valid; };
  auto row_info = [&](long long tile) {
    RowInfo ri;
    ri.g = tile * kBM + r;
    ri.valid = ri.g < a.rows;
    const long long srow = ri.valid ? ((a.index ? a.index[ri.g] : ri.g) / a.x_div) : 0;
    ri.row = a.x + srow * (long long)a.ld;
    ri.x0 = (ri.valid && L.use_fn) ? __ldg(ri.row) : 0.0f;
    return ri;
  };
  auto load_chunk = [&](const RowInfo& ri, int kc, float (&v)[16]) {
    const int c0 = 64 * kc + 16 * cq;
    if (16 * cq < 16 * min(4, m.K16 - 4 * kc)) {
      load8(ri.row, c0, D, ri.valid, a.vec, *reinterpret_cast<float(*)[8]>(v));
      load8(ri.row, c0 + 8, D, ri.valid, a.vec, *reinterpret_cast<float(*)[8]>(v + 8));
    }
  };
  auto wait_commit = [&](uint32_t c) { umma::mbar_wait(bars + (c & 1), (c >> 1) & 1); };
  auto drain = [&](long long g, float ms, float rstd) {                 // accumulators -> z1 row g
    wait_commit(n - 1);
    umma::tc_fence_after();
    float u[16], w[16];                                                 // the first ncol of them are this thread's
    umma::tmem_ld16(lane_addr + ncol * cq, u);
    umma::tmem_ld16(lane_addr + 64 + ncol * cq, w);
    if (g < a.rows) {
      float4* out = reinterpret_cast<float4*>(a.z1 + g * 64 + j0 + ncol * cq);
      const float* c16 = cs + ncol * cq;
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (4 * i < ncol) out[i] = make_float4(rstd * (fmaf(w[4 * i], kLoInv, u[4 * i]) - ms * c16[4 * i]),
                             rstd * (fmaf(w[4 * i + 1], kLoInv, u[4 * i + 1]) - ms * c16[4 * i + 1]),
                             rstd * (fmaf(w[4 * i + 2], kLoInv, u[4 * i + 2]) - ms * c16[4 * i + 2]),
                             rstd * (fmaf(w[4 * i + 3], kLoInv, u[4 * i + 3]) - ms * c16[4 * i + 3]));
    }
    umma::tc_fence_before();
    __syncthreads();
  };

  long long tile = blockIdx.x;
  long long prev_g = -1;
  float prev_ms = 0.0f, prev_rstd = 1.0f;
  if (tile < ntiles) {
    RowInfo cur = row_info(tile);
    float pre[16];
    load_chunk(cur, 0, pre);
    while (tile < ntiles) {
      float s = 0.0f, q = 0.0f;
      RowInfo nxt = cur;
      const long long ntile = tile + gridDim.x;
      for (int kc = 0; kc < nkc; ++kc) {
        float v[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = pre[i];
        if (kc + 1 < nkc) load_chunk(cur, kc + 1, pre);                 // next chunk of this tile ...
        else if (ntile < ntiles) { nxt = row_info(ntile); load_chunk(nxt, 0, pre); }   // ... or the first of the next
        const uint32_t b = n & 1;
        const int ksteps = min(4, m.K16 - 4 * kc);
        if (n >= 2) wait_commit(n - 2);                                // X buffer b is free again
        if (cq < ksteps) {
          const int c0 = 64 * kc + 16 * cq;
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            v[i] = (cur.valid && c0 + i < D) ? v[i] - cur.x0 : 0.0f;
            s += v[i];
            q = fmaf(v[i], v[i], q);
          }
          store_cols16(sm + m.X + b * kXBuf, r, cq, v);
        }
        umma::fence_async_smem();
        umma::tc_fence_before();
        __syncthreads();
        if (kc == 0 && prev_g >= 0) drain(prev_g, prev_ms, prev_rstd);  // previous tile leaves TMEM before fc1 restarts
        if (tid == 0) {
          umma::tc_fence_after();
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t ao = sX + b * kXBuf + (uint32_t)ks * 2 * kRowChunk;
            const uint32_t bo = sW + (uint32_t)(4 * kc + ks) * 2 * m.w_chunk;
            const uint64_t ah = umma::smem_desc(ao, kRowChunk, 128), al = umma::smem_desc(ao + 8 * kRowChunk, kRowChunk, 128);
            const uint64_t bh = umma::smem_desc(bo, m.w_chunk, 128), bl = umma::smem_desc(bo + m.w_piece, m.w_chunk, 128);
            const bool acc = kc > 0 || ks > 0;
            umma::mma_f16(tb, ah, bh, kId, acc);
            umma::mma_f16(tb + 64, ah, bl, kId, acc);
            umma::mma_f16(tb + 64, al, bh, kId, true);
          }
          umma::mma_commit(bars + b);
        }
        ++n;
      }
      // row statistics (mlp.py:44-45: LayerNorm over in_dim, eps 1e-5) from the sums of d and d^2
      float ms = 0.0f, rstd = 1.0f;
      if (L.use_fn) {
        const float2 tot = row_sum2(s, q, xch, slot, cq, r, q4);
        ms = tot.x / (float)D;
        const float var = fmaxf(tot.y / (float)D - ms * ms, 0.0f);
        rstd = 1.0f / sqrtf(var + kLnEps);
      }
      if (cq == 0 && blockIdx.y == 0 && cur.valid && a.xstats) { a.xstats[2 * cur.g] = cur.x0 + ms; a.xstats[2 * cur.g + 1] = rstd; }
      prev_g = cur.g; prev_ms = ms; prev_rstd = rstd;
      cur = nxt;
      tile = ntile;
    }
    drain(prev_g, prev_ms, prev_rstd);
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 128);
}

// ---- packed inputs -------------------------------------------------------------------------------------------
// With num_mini_batch == 1 every PPO epoch reads the same share_obs rows, and their LayerNorm does not depend on the
// weights (the affine is folded into W1').  wide_pack_kernel therefore converts them ONCE per iteration into the
// operand image the tensor cores read -- per 128-row tile: [piece hi | lo][8-column chunk][row][8 fp16] of
// x^ = LN(x), zero rows / columns as padding -- and fc1_packed_kernel streams those tiles with bulk copies
// (cp.async.bulk + mbarrier transaction counts) through a 3-stage ring straight into tcgen05.mma: no thread touches an
// input element.  Warp roles: warp 0 one producer lane, warp 1 one MMA-issuing lane, warps 2..5 drain the two
// alternating TMEM accumulators to z1.
__host__ __device__ inline long long packed_tile_bytes(int D) { return 2ll * (2 * ((D + 15) / 16)) * kRowChunk; }

__global__ void __launch_bounds__(kThreads, 1) wide_pack_kernel(const WideArgs a, unsigned char* __restrict__ packed) {
  __shared__ float2 xch_s[2 * 4 * kBM];
  const NetLayout& L = a.L;
  const int D = L.in_dim, K16 = (D + 15) / 16, C8 = 2 * K16, nkc = (K16 + 3) / 4;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
  const long long ntiles = (a.rows + kBM - 1) / kBM;
  const long long tile_bytes = packed_tile_bytes(D);
  int slot = 0;
  for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
    const long long g = tile * kBM + r;
    const bool valid = g < a.rows;
    const float* __restrict__ row = a.x + (valid ? g : 0) * (long long)a.ld;
    float ms = 0.0f, rstd = 1.0f, x0 = 0.0f;
    if (L.use_fn) {                                                     // mlp.py:44-45: LayerNorm over in_dim, eps 1e-5
      x0 = valid ? __ldg(row) : 0.0f;
      float s = 0.0f, q = 0.0f;
      for (int kc = 0; kc < nkc; ++kc) {
        const int c0 = 64 * kc + 16 * cq;
        if (c0 < 16 * K16) {
          float v[16];
          load8(row, c0, D, valid, a.vec, *reinterpret_cast<float(*)[8]>(v));
          load8(row, c0 + 8, D, valid, a.vec, *reinterpret_cast<float(*)[8]>(v + 8));
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            const float d = (valid && c0 + i < D) ? v[i] - x0 : 0.0f;
            s += d;
            q = fmaf(d, d, q);
          }
        }
      }
      const float2 tot = row_sum2(s, q, xch_s, slot, cq, r, q4);
      ms = tot.x / (float)D;
      rstd = 1.0f / sqrtf(fmaxf(tot.y / (float)D - ms * ms, 0.0f) + kLnEps);
    }
    if (cq == 0 && valid && a.xstats) { a.xstats[2 * g] = x0 + ms; a.xstats[2 * g + 1] = rstd; }
    unsigned char* __restrict__ out = packed + tile * tile_bytes;
    for (int kc = 0; kc < nkc; ++kc) {
      const int c0 = 64 * kc + 16 * cq;
      if (c0 >= 16 * K16) continue;
      float v[16];
      load8(row, c0, D, valid, a.vec, *reinterpret_cast<float(*)[8]>(v));
      load8(row, c0 + 8, D, valid, a.vec, *reinterpret_cast<float(*)[8]>(v + 8));
#pragma unroll
      for (int i = 0; i < 16; ++i) v[i] = (valid && c0 + i < D) ? ((v[i] - x0) - ms) * rstd : 0.0f;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint4 hi, lo;
        split8(v + 8 * h, hi, lo);
        const long long off = (long long)(8 * kc + 2 * cq + h) * kRowChunk + (long long)r * 16;
        *reinterpret_cast<uint4*>(out + off) = hi;                      // a warp writes 512 contiguous bytes
        *reinterpret_cast<uint4*>(out + (long long)C8 * kRowChunk + off) = lo;
      }
    }
  }
}

constexpr int kPkThreads = 192, kPkStages = 3;
constexpr uint32_t kPkStage = 2 * 8 * kRowChunk;                        // one 64-column K-chunk, both pieces: 32 KB
struct PkMap { uint32_t W, w_piece, X, bars, total; int K16; };
__host__ __device__ inline PkMap make_pk_map(int D) {
  PkMap m;
  m.K16 = (D + 15) / 16;
  m.w_piece = (uint32_t)(2 * m.K16) * kWChunk;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.W = take(2 * m.w_piece);
  m.X = take(kPkStages * kPkStage);
  m.bars = take(128);
  m.total = o;
  return m;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(umma::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(umma::smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(dst), "l"(src), "r"(bytes), "r"(umma::smem_u32(bar)) : "memory");
}

__global__ void __launch_bounds__(kPkThreads, 1) fc1_packed_kernel(const WideArgs a, const unsigned char* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  const NetLayout& L = a.L;
  const int D = L.in_dim;
  const PkMap m = make_pk_map(D);
  const int tid = threadIdx.x, warp = tid >> 5;
  uint64_t* full = reinterpret_cast<uint64_t*>(sm + m.bars);          // [3] stage landed (transaction bytes)
  uint64_t* empty = full + kPkStages;                                  // [3] stage read by its MMAs (tcgen05.commit)
  uint64_t* acc_full = empty + kPkStages;                              // [2] tile accumulated (tcgen05.commit)
  uint64_t* acc_empty = acc_full + 2;                                  // [2] accumulator drained (4 epilogue warps)
  if (tid == 0) {
    for (int i = 0; i < kPkStages; ++i) { umma::mbar_init(full + i, 1); umma::mbar_init(empty + i, 1); }
    for (int i = 0; i < 2; ++i) { umma::mbar_init(acc_full + i, 1); umma::mbar_init(acc_empty + i, 4); }
    umma::fence_mbar_init();
  }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 256);
  {
    const float* __restrict__ P = a.params;
    const int Kp = 16 * m.K16;
    for (int e = tid; e < 64 * Kp; e += kPkThreads) {
      const int j = e / Kp, k = e - j * Kp;
      float v = 0.0f;
      if (k < D) v = P[L.w1 + j * D + k] * (L.use_fn ? P[L.fn_g + k] : 1.0f);
      put_pair(sm + m.W, m.w_piece, 64, j, k, v);
    }
  }
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const int nkc = (m.K16 + 3) / 4, C8 = 2 * m.K16;
  const long long ntiles = (a.rows + kBM - 1) / kBM;
  const long long tile_bytes = packed_tile_bytes(D);
  if (warp == 0) {
    if (umma::elect_one()) {                                           // ---- producer
      const uint32_t sX = umma::smem_u32(sm + m.X);
      uint32_t it = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const unsigned char* src = packed + tile * tile_bytes;
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const uint32_t s = it % kPkStages, ph = (it / kPkStages) & 1;
          umma::mbar_wait(empty + s, ph ^ 1);
          const uint32_t bytes = (uint32_t)min(8, C8 - 8 * kc) * kRowChunk;
          mbar_expect_tx(full + s, 2 * bytes);
          bulk_g2s(sX + s * kPkStage, src + (long long)(8 * kc) * kRowChunk, bytes, full + s);
          bulk_g2s(sX + s * kPkStage + 8 * kRowChunk, src + (long long)(C8 + 8 * kc) * kRowChunk, bytes, full + s);
        }
      }
    }
  } else if (warp == 1) {
    if (umma::elect_one()) {                                           // ---- MMA issue
      const uint32_t sX = umma::smem_u32(sm + m.X), sW = umma::smem_u32(sm + m.W);
      constexpr uint32_t kId = umma::idesc_f16(128, 64, 0, 0);
      uint32_t it = 0, t = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
        const uint32_t b = t & 1, acc = tb + 128 * b;
        umma::mbar_wait(acc_empty + b, ((t >> 1) & 1) ^ 1);
        umma::tc_fence_after();
        for (int kc = 0; kc < nkc; ++kc, ++it) {
          const uint32_t s = it % kPkStages, ph = (it / kPkStages) & 1;
          umma::mbar_wait(full + s, ph);
          umma::tc_fence_after();
          const int ksteps = min(4, m.K16 - 4 * kc);
          for (int ks = 0; ks < ksteps; ++ks) {
            const uint32_t ao = sX + s * kPkStage + (uint32_t)ks * 2 * kRowChunk;
            const uint32_t bo = sW + (uint32_t)(4 * kc + ks) * 2 * kWChunk;
            const uint64_t ah = umma::smem_desc(ao, kRowChunk, 128), al = umma::smem_desc(ao + 8 * kRowChunk, kRowChunk, 128);
            const uint64_t bh = umma::smem_desc(bo, kWChunk, 128), bl = umma::smem_desc(bo + m.w_piece, kWChunk, 128);
            const bool accum = kc > 0 || ks > 0;
            umma::mma_f16(acc, ah, bh, kId, accum);
            umma::mma_f16(acc + 64, ah, bl, kId, accum);
            umma::mma_f16(acc + 64, al, bh, kId, true);
          }
          umma::mma_commit(empty + s);
        }
        umma::mma_commit(acc_full + b);
      }
    }
  } else {                                                             // ---- epilogue: warps 2..5, TMEM lanes 32 (warp % 4)
    const int q4 = warp & 3, lane = tid & 31, r = 32 * q4 + lane;
    const uint32_t lane_addr = tb + ((uint32_t)(32 * q4) << 16);
    uint32_t t = 0;
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++t) {
      const uint32_t b = t & 1;
      umma::mbar_wait(acc_full + b, (t >> 1) & 1);
      umma::tc_fence_after();
      const long long g = tile * kBM + r;
#pragma unroll
      for (int cq = 0; cq < 4; ++cq) {
        float u[16], w[16];
        umma::tmem_ld16x2(lane_addr + 128 * b + 16 * cq, lane_addr + 128 * b + 64 + 16 * cq, u, w);
        if (g < a.rows) {
          float4* out = reinterpret_cast<float4*>(a.z1 + g * 64 + 16 * cq);
#pragma unroll
          for (int i = 0; i < 4; ++i)
            out[i] = make_float4(fmaf(w[4 * i], kLoInv, u[4 * i]), fmaf(w[4 * i + 1], kLoInv, u[4 * i + 1]),
                                 fmaf(w[4 * i + 2], kLoInv, u[4 * i + 2]), fmaf(w[4 * i + 3], kLoInv, u[4 * i + 3]));
        }
      }
      umma::tc_fence_before();
      __syncwarp();
      if (umma::elect_one()) mbar_arrive(acc_empty + b);
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 256);
}

// ---- weight gradient ---------------------------------------------------------------------------------------
// CTA (column group cg, row split sp) accumulates G[128 = dz1 hi | dz1 lo][colw] over its 64-row half tiles in ONE TMEM
// block (x^ hi and x^ lo are equally scaled and go into the same columns); DUAL: a second dz operand (dz2) accumulates
// against the same x^ tiles into the block at column 256.  Operands are MN-major (the contraction runs over the rows): A =
// dz1 [64 rows][128], B = x^ [64 rows][colw] per piece.  One operand buffer; the next half tile's values are
// prefetched into registers while the tensor cores read the current one.
constexpr int kHR = 64;                                    // rows per half tile
constexpr uint32_t kHChunk = kHR * 16;                     // bytes of one 8-column chunk of a 64-row operand tile
constexpr int kMaxColw = 256;
constexpr int kXChunksPerThread = 4;                       // ceil(32 chunks / 8 column subsets)

constexpr int kDzStride = 68;                              // floats per staged dz row (64 + 4: conflict-free float4 reads)
struct WgMap { uint32_t A, A2, B, b_piece, stage, stage_dz, stage_dz2, xs_stride, bars, total; };
__host__ __device__ inline WgMap make_wg_map(int colw, bool dual = false) {
  WgMap m;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.A = take(16 * kHChunk);
  m.A2 = take(dual ? 16 * kHChunk : 0);
  m.b_piece = (uint32_t)(colw / 8) * kHChunk;
  m.B = take(2 * m.b_piece);
  m.xs_stride = (uint32_t)colw + 4;                        // floats per staged x row
  m.stage = take(kHR * m.xs_stride * 4);                   // landing tiles of the asynchronous, row-coalesced loads: x ...
  m.stage_dz = take(kHR * kDzStride * 4);                  // ... and dz
  m.stage_dz2 = take(dual ? kHR * kDzStride * 4 : 0);
  m.bars = take(64);
  m.total = o;
  return m;
}

// 16-byte asynchronous global -> shared copy of `bytes` (0..16) valid bytes, zero-filling the rest
__device__ __forceinline__ void cp_async16(void* dst, const void* src, int bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(umma::smem_u32(dst)), "l"(src), "r"(bytes) : "memory");
}

template <bool COLSUM, bool DUAL>
__global__ void __launch_bounds__(kThreads, 1) wgrad1_wide_kernel(const WideArgs a) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  __shared__ float csred[2][8][kXChunksPerThread * 8];
  const int D = a.D, colw = a.colw;
  const WgMap m = make_wg_map(colw, DUAL);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cg = blockIdx.x % a.ncg, sp = blockIdx.x / a.ncg;
  const int col0 = cg * colw;
  const int hr = 32 * (warp & 1) + lane;                  // row of the half tile this thread stages
  const int cs = warp >> 1;                               // column subset: chunks cs, cs + 8, ...
  const int nchx = colw / 8;
  uint64_t* bar = reinterpret_cast<uint64_t*>(sm + m.bars);
  if (tid == 0) { umma::mbar_init(bar, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 512);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const uint32_t sA = umma::smem_u32(sm + m.A), sA2 = umma::smem_u32(sm + m.A2), sB = umma::smem_u32(sm + m.B);
  const uint32_t idesc = umma::idesc_f16(128, colw, 1, 1);
  const long long nht = (a.rows + kHR - 1) / kHR;

  float xr[kXChunksPerThread][8], dr[8], dr2[8], mean = 0.0f, rstd = 1.0f;
  bool valid = false;
  const float sx = a.norm ? kSN : kSD;                    // x is LayerNorm output (C5) or a gradient array (GRU chain)
  // 16-byte aligned rows: the next half tile is copied asynchronously into shared-memory landing tiles, consecutive
  // threads fetching consecutive 16-byte units of a row (coalesced: a warp touches 4 lines, not 32), the moment the
  // current tile has been read out; each thread then picks up its (row, chunk) values from the landing tile.
  float mean_n = 0.0f, rstd_n = 1.0f;
  bool valid_n = false;
  float* xs = reinterpret_cast<float*>(sm + m.stage);
  float* dzs = reinterpret_cast<float*>(sm + m.stage_dz);
  float* dzs2 = reinterpret_cast<float*>(sm + m.stage_dz2);
  const int xs_stride = (int)m.xs_stride, upr = colw / 4;     // 16-byte units per staged x row
  (void)upr;
  auto issue_async = [&](long long ht) {
    {
      const long long g = ht * kHR + hr;
      valid_n = g < a.rows;
      if (a.norm && valid_n) { mean_n = __ldg(a.xstats + 2 * g); rstd_n = __ldg(a.xstats + 2 * g + 1); }
      else { mean_n = 0.0f; rstd_n = 1.0f; }
    }
    // warp w fetches rows w, w + 16, ...: the row quantities are warp-uniform, lanes walk the 16-byte units of the row
    for (int rr = warp; rr < kHR; rr += kThreads / 32) {
      const long long g = ht * kHR + rr;
      const bool ok = g < a.rows;
      const long long srow = ok ? ((a.index ? a.index[g] : g) / a.x_div) : 0;
      const float* __restrict__ src = a.x + srow * (long long)a.ld + col0;
      float* dst = xs + rr * xs_stride;
      const int lim = ok ? D - col0 : 0;                                // valid floats of this row from col0 on
      for (int q = lane; q < upr; q += 32) {
        const int bytes = max(0, min(4, lim - 4 * q)) * 4;
        cp_async16(dst + 4 * q, bytes ? src + 4 * q : a.x, bytes);
      }
    }
    for (int rr = 2 * warp + (lane >> 4); rr < kHR; rr += 2 * (kThreads / 32)) {
      const long long g = ht * kHR + rr;
      const bool ok = g < a.rows;
      const int q = lane & 15;
      cp_async16(dzs + rr * kDzStride + 4 * q, a.dz1 + (ok ? g : 0) * a.dz_ld + 4 * q, ok ? 16 : 0);
      if (DUAL) cp_async16(dzs2 + rr * kDzStride + 4 * q, a.dz2 + (ok ? g : 0) * a.dz_ld + 4 * q, ok ? 16 : 0);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto take_async = [&]() {
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();                                                    // everybody's copies have landed
#pragma unroll
    for (int i = 0; i < kXChunksPerThread; ++i) {
      if (cs + 8 * i < nchx) {
        const float4* p = reinterpret_cast<const float4*>(xs + hr * xs_stride + 8 * (cs + 8 * i));
        const float4 u0 = p[0], u1 = p[1];
        xr[i][0] = u0.x; xr[i][1] = u0.y; xr[i][2] = u0.z; xr[i][3] = u0.w;
        xr[i][4] = u1.x; xr[i][5] = u1.y; xr[i][6] = u1.z; xr[i][7] = u1.w;
      }
    }
    const float4* p = reinterpret_cast<const float4*>(dzs + hr * kDzStride + 8 * cs);
    const float4 d0 = p[0], d1 = p[1];
    dr[0] = d0.x; dr[1] = d0.y; dr[2] = d0.z; dr[3] = d0.w; dr[4] = d1.x; dr[5] = d1.y; dr[6] = d1.z; dr[7] = d1.w;
    if (DUAL) {
      const float4* p2 = reinterpret_cast<const float4*>(dzs2 + hr * kDzStride + 8 * cs);
      const float4 e0 = p2[0], e1 = p2[1];
      dr2[0] = e0.x; dr2[1] = e0.y; dr2[2] = e0.z; dr2[3] = e0.w; dr2[4] = e1.x; dr2[5] = e1.y; dr2[6] = e1.z; dr2[7] = e1.w;
    }
    valid = valid_n; mean = mean_n; rstd = rstd_n;
    __syncthreads();                                                    // landing tiles free for the next half tile
  };
  const bool use_async = a.vec && (a.dz_ld & 3) == 0;
  auto prefetch = [&](long long ht) {
    const long long g = ht * kHR + hr;
    valid = g < a.rows;
    const long long srow = valid ? ((a.index ? a.index[g] : g) / a.x_div) : 0;
    const float* __restrict__ row = a.x + srow * (long long)a.ld;
    if (a.norm && valid) { mean = __ldg(a.xstats + 2 * g); rstd = __ldg(a.xstats + 2 * g + 1); }
    else { mean = 0.0f; rstd = 1.0f; }
#pragma unroll
    for (int i = 0; i < kXChunksPerThread; ++i) {
      const int ch = cs + 8 * i;
      if (ch < nchx) load8(row, col0 + 8 * ch, D, valid, a.vec, xr[i]);
    }
    if (valid) {
      const float4 p = __ldg(reinterpret_cast<const float4*>(a.dz1 + g * a.dz_ld + 8 * cs));
      const float4 q = __ldg(reinterpret_cast<const float4*>(a.dz1 + g * a.dz_ld + 8 * cs + 4));
      dr[0] = p.x; dr[1] = p.y; dr[2] = p.z; dr[3] = p.w; dr[4] = q.x; dr[5] = q.y; dr[6] = q.z; dr[7] = q.w;
      if (DUAL) {
        const float4 p2 = __ldg(reinterpret_cast<const float4*>(a.dz2 + g * a.dz_ld + 8 * cs));
        const float4 q2 = __ldg(reinterpret_cast<const float4*>(a.dz2 + g * a.dz_ld + 8 * cs + 4));
        dr2[0] = p2.x; dr2[1] = p2.y; dr2[2] = p2.z; dr2[3] = p2.w; dr2[4] = q2.x; dr2[5] = q2.y; dr2[6] = q2.z; dr2[7] = q2.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) { dr[i] = 0.0f; dr2[i] = 0.0f; }
    }
  };

  float csum[kXChunksPerThread][8];
#pragma unroll
  for (int i = 0; i < kXChunksPerThread; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) csum[i][e] = 0.0f;
  uint32_t n = 0;
  long long ht = sp;
  if (ht < nht) { if (use_async) issue_async(ht); else prefetch(ht); }
  for (; ht < nht; ht += a.nsplit) {
    if (use_async) {
      take_async();
      if (ht + a.nsplit < nht) issue_async(ht + a.nsplit);
    }
    if (n > 0) umma::mbar_wait(bar, (n - 1) & 1);                       // the operand buffer is free again
    {
      // uniformly scaled operand pairs (mlp_tc_common.cuh): both halves of an operand carry one power-of-two scale, folded
      // into the normalisation here and taken out of the accumulators (and the column sums) when they leave the kernel
      uint4 hi, lo;
      const float rs = rstd * sx;
#pragma unroll
      for (int i = 0; i < kXChunksPerThread; ++i) {
        const int ch = cs + 8 * i;
        if (ch < nchx) {
          float v[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            v[e] = (valid && col0 + 8 * ch + e < D) ? (xr[i][e] - mean) * rs : 0.0f;
            if (COLSUM) csum[i][e] += v[e];
          }
          split8u(v, hi, lo);
          const uint32_t off = (uint32_t)ch * kHChunk + (uint32_t)hr * 16;
          *reinterpret_cast<uint4*>(sm + m.B + off) = hi;
          *reinterpret_cast<uint4*>(sm + m.B + m.b_piece + off) = lo;
        }
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) dr[e] *= kSD;
      split8u(dr, hi, lo);
      const uint32_t off = (uint32_t)cs * kHChunk + (uint32_t)hr * 16;
      *reinterpret_cast<uint4*>(sm + m.A + off) = hi;
      *reinterpret_cast<uint4*>(sm + m.A + 8 * kHChunk + off) = lo;
      if (DUAL) {
#pragma unroll
        for (int e = 0; e < 8; ++e) dr2[e] *= kSD;
        split8u(dr2, hi, lo);
        *reinterpret_cast<uint4*>(sm + m.A2 + off) = hi;
        *reinterpret_cast<uint4*>(sm + m.A2 + 8 * kHChunk + off) = lo;
      }
    }
    umma::fence_async_smem();
    umma::tc_fence_before();
    __syncthreads();
    if (tid == 0) {
      umma::tc_fence_after();
#pragma unroll
      for (int ks = 0; ks < kHR / 16; ++ks) {
        const uint64_t da = umma::smem_desc(sA + (uint32_t)ks * 256, 128, kHChunk);
        const uint64_t dh = umma::smem_desc(sB + (uint32_t)ks * 256, 128, kHChunk);
        const uint64_t dl = umma::smem_desc(sB + m.b_piece + (uint32_t)ks * 256, 128, kHChunk);
        umma::mma_f16(tb, da, dh, idesc, n > 0 || ks > 0);
        umma::mma_f16(tb, da, dl, idesc, true);
        if (DUAL) {
          const uint64_t da2 = umma::smem_desc(sA2 + (uint32_t)ks * 256, 128, kHChunk);
          umma::mma_f16(tb + 256, da2, dh, idesc, n > 0 || ks > 0);
          umma::mma_f16(tb + 256, da2, dl, idesc, true);
        }
      }
      umma::mma_commit(bar);
    }
    ++n;
    if (!use_async && ht + a.nsplit < nht) prefetch(ht + a.nsplit);
  }
  // ---- G -> the W1 slots of partial rows 2 sp (lanes 0..63: dz1 hi) and 2 sp + 1 (lanes 64..127: dz1 lo, scaled)
  if (n > 0) {
    umma::mbar_wait(bar, (n - 1) & 1);
    umma::tc_fence_after();
    const int q4 = warp & 3, cq = warp >> 2, r = 32 * q4 + lane;
    const uint32_t lane_addr = tb + ((uint32_t)(32 * q4) << 16);
    float* __restrict__ gp = a.grad_partials + (size_t)(2 * sp + (q4 >> 1)) * a.total;
    const int j = r & 63;
    const float ginv = 1.0f / (sx * kSD);
    const int win = DUAL ? a.win : D;
    for (int blk = cq; blk < colw / 16; blk += 4) {
      float u[16], w[16];
      umma::tmem_ld16(lane_addr + 16 * blk, u);
      if (DUAL) umma::tmem_ld16(lane_addr + 256 + 16 * blk, w);
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const int k = col0 + 16 * blk + i;
        if (k < win) {
          const float val = u[i] * ginv;
          if (a.transposed) gp[a.out_w + (size_t)((k + a.rot) % win) * 64 + j] = val;
          else gp[a.out_w + (size_t)j * D + k] = val;
        }
        if (DUAL && k >= a.off2 && k < a.off2 + win)
          gp[a.out_w2 + (size_t)((k - a.off2 + a.rot2) % win) * 64 + j] = w[i] * ginv;
      }
    }
  }
  if (COLSUM && a.out_b >= 0) {                            // column sums of x over this CTA's rows -> partial row 2 sp
#pragma unroll
    for (int i = 0; i < kXChunksPerThread; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        float v = csum[i][e];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) csred[warp & 1][cs][i * 8 + e] = v;
      }
    __syncthreads();
    if (tid < 8 * kXChunksPerThread * 8) {
      const int c8 = tid / (kXChunksPerThread * 8), ie = tid % (kXChunksPerThread * 8);
      const int ch = c8 + 8 * (ie >> 3), k = col0 + 8 * ch + (ie & 7);
      if (ch < nchx && k < D) {
        const float v = (csred[0][c8][ie] + csred[1][c8][ie]) * (1.0f / sx);
        const int win = DUAL ? a.win : D;
        if (k < win) a.grad_partials[(size_t)(2 * sp) * a.total + a.out_b + (k + a.rot) % win] = v;
        if (DUAL && k >= a.off2 && k < a.off2 + win)
          a.grad_partials[(size_t)(2 * sp) * a.total + a.out_b2 + (k - a.off2 + a.rot2) % win] = v;
      }
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 512);
}

// ---- weight gradient from packed inputs ------------------------------------------------------------------------
// G1[64][D] = sum_rows dz1^T x^ with x^ read from the packed tiles: the bytes of a tile that served as the K-major A
// operand of the forward are, unchanged, the MN-major B operand here (umma.cuh).  CTA (column group of up to 16
// 8-column chunks, row split) walks its 128-row tiles through a 2-stage ring: warp 0's elected lane fetches the group's
// hi and lo chunk ranges with two bulk copies per tile, warps 2..5 turn the tile's dz1 rows into the stacked fp16-pair A
// operand ([hi columns | lo columns] = M 128), warp 1's elected lane issues 8 K-steps x 2 MMAs into two TMEM blocks.
constexpr int kPgChunks = 16;                                           // 8-column chunks per column group (N <= 128)
constexpr uint32_t kPgB = 2 * kPgChunks * kRowChunk;                    // one B stage: both pieces, 64 KB
constexpr uint32_t kPgA = 16 * kRowChunk;                               // one A stage: dz1 hi | lo, 32 KB
struct PgMap { uint32_t A, B, bars, total; };
__host__ __device__ inline PgMap make_pg_map() {
  PgMap m;
  uint32_t o = 0;
  auto take = [&](uint32_t bytes) { uint32_t r = o; o += (bytes + 127u) & ~127u; return r; };
  m.A = take(2 * kPgA);
  m.B = take(2 * kPgB);
  m.bars = take(128);
  m.total = o;
  return m;
}

__global__ void __launch_bounds__(kPkThreads, 1) wgrad1_packed_kernel(const WideArgs a, const unsigned char* __restrict__ packed) {
  extern __shared__ __align__(1024) unsigned char sm[];
  __shared__ uint32_t tmem_base_s;
  const int D = a.D, K16 = (D + 15) / 16, C8 = 2 * K16;
  const PgMap m = make_pg_map();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int cg = blockIdx.x % a.ncg, sp = blockIdx.x / a.ncg;
  const int ch0 = cg * kPgChunks, nch = min(kPgChunks, C8 - ch0), N = 8 * nch;   // this group's chunks; N is a multiple of 16
  uint64_t* b_full = reinterpret_cast<uint64_t*>(sm + m.bars);        // [2] B stage landed
  uint64_t* a_full = b_full + 2;                                       // [2] A stage converted (4 warps)
  uint64_t* st_empty = a_full + 2;                                     // [2] stage read by its MMAs
  uint64_t* done = st_empty + 2;                                       // all MMAs of this CTA
  if (tid == 0) {
    for (int i = 0; i < 2; ++i) { umma::mbar_init(b_full + i, 1); umma::mbar_init(a_full + i, 4); umma::mbar_init(st_empty + i, 1); }
    umma::mbar_init(done, 1);
    umma::fence_mbar_init();
  }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, 256);
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tb = tmem_base_s;
  const long long ntiles = (a.rows + kBM - 1) / kBM;
  const long long tile_bytes = packed_tile_bytes(D);
  uint32_t my_tiles = 0;
  for (long long t = sp; t < ntiles; t += a.nsplit) ++my_tiles;
  if (warp == 0) {
    if (umma::elect_one()) {                                           // ---- producer: x^ chunks [ch0, ch0 + nch), hi and lo
      const uint32_t sB = umma::smem_u32(sm + m.B);
      const uint32_t bytes = (uint32_t)nch * kRowChunk;
      uint32_t it = 0;
      for (long long t = sp; t < ntiles; t += a.nsplit, ++it) {
        const uint32_t s = it & 1, ph = (it >> 1) & 1;
        umma::mbar_wait(st_empty + s, ph ^ 1);
        const unsigned char* src = packed + t * tile_bytes + (long long)ch0 * kRowChunk;
        mbar_expect_tx(b_full + s, 2 * bytes);
        bulk_g2s(sB + s * kPgB, src, bytes, b_full + s);
        bulk_g2s(sB + s * kPgB + kPgChunks * kRowChunk, src + (long long)C8 * kRowChunk, bytes, b_full + s);
      }
    }
  } else if (warp == 1) {
    if (umma::elect_one()) {                                           // ---- MMA issue
      const uint32_t sA = umma::smem_u32(sm + m.A), sB = umma::smem_u32(sm + m.B);
      const uint32_t idesc = umma::idesc_f16(128, N, 1, 1);
      for (uint32_t it = 0; it < my_tiles; ++it) {
        const uint32_t s = it & 1, ph = (it >> 1) & 1;
        umma::mbar_wait(a_full + s, ph);
        umma::mbar_wait(b_full + s, ph);
        umma::tc_fence_after();
#pragma unroll
        for (int ks = 0; ks < kBM / 16; ++ks) {
          const uint64_t da = umma::smem_desc(sA + s * kPgA + (uint32_t)ks * 256, 128, kRowChunk);
          const uint64_t dh = umma::smem_desc(sB + s * kPgB + (uint32_t)ks * 256, 128, kRowChunk);
          const uint64_t dl = umma::smem_desc(sB + s * kPgB + kPgChunks * kRowChunk + (uint32_t)ks * 256, 128, kRowChunk);
          umma::mma_f16(tb, da, dh, idesc, it > 0 || ks > 0);
          umma::mma_f16(tb + 128, da, dl, idesc, it > 0 || ks > 0);
        }
        umma::mma_commit(st_empty + s);
      }
      umma::mma_commit(done);
    }
  } else {                                                             // ---- warps 2..5: dz1 rows -> A operand, then the epilogue
    const int r = 32 * (warp - 2) + lane;                              // row of the tile this thread converts
    uint32_t it = 0;
    for (long long t = sp; t < ntiles; t += a.nsplit, ++it) {
      const uint32_t s = it & 1, ph = (it >> 1) & 1;
      const long long g = t * kBM + r;
      float v[64];
      if (g < a.rows) {
        const float4* src = reinterpret_cast<const float4*>(a.dz1 + g * a.dz_ld);
#pragma unroll
        for (int i = 0; i < 16; ++i) { const float4 q = __ldg(src + i); v[4 * i] = q.x; v[4 * i + 1] = q.y; v[4 * i + 2] = q.z; v[4 * i + 3] = q.w; }
      } else {
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = 0.0f;
      }
      umma::mbar_wait(st_empty + s, ph ^ 1);
#pragma unroll
      for (int cq = 0; cq < 4; ++cq) store_cols16(sm + m.A + s * kPgA, r, cq, v + 16 * cq);
      umma::fence_async_smem();
      __syncwarp();
      if (umma::elect_one()) mbar_arrive(a_full + s);
    }
    // G -> the W1 slots of partial rows 2 sp (lanes 0..63: dz1 hi) and 2 sp + 1 (lanes 64..127: dz1 lo, scaled)
    if (my_tiles > 0) {
      umma::mbar_wait(done, 0);
      umma::tc_fence_after();
      const int q4 = warp & 3, rr = 32 * q4 + lane;
      const uint32_t lane_addr = tb + ((uint32_t)(32 * q4) << 16);
      float* __restrict__ gp = a.grad_partials + (size_t)(2 * sp + (q4 >> 1)) * a.total + a.out_w;
      const int j = rr & 63;
      const float lane_scale = (q4 >> 1) ? kLoInv : 1.0f;
      for (int blk = 0; blk < N / 16; ++blk) {
        float u[16], w[16];
        umma::tmem_ld16x2(lane_addr + 16 * blk, lane_addr + 128 + 16 * blk, u, w);
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          const int k = 8 * ch0 + 16 * blk + i;
          if (k < D) gp[(size_t)j * D + k] = fmaf(w[i], kLoInv, u[i]) * lane_scale;
        }
      }
    }
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tb, 256);
}

// ---- host side ---------------------------------------------------------------------------------------------
constexpr int kWideMaxSmem = 232448 - 1024;

int grid_for(long long rows);                                             // mlp_tc.cu

bool wide_supports(const mappo_net_config* cfg) {
  return cfg && cfg->hidden == 64 && cfg->layer_N == 1 && !cfg->recurrent && cfg->in_dim >= kWideFrom &&
         make_fwd_map(cfg->in_dim, 4).total <= (uint32_t)kWideMaxSmem;
}

bool wide_fits(int in_dim) { return make_fwd_map(in_dim, 4).total <= (uint32_t)kWideMaxSmem; }

// bytes of the caller-provided scratch: z1 / dz1 [rows][64] (the gradient overwrites the activations in place) and the
// row statistics [rows][2]
long long wide_workspace_bytes(long long rows) { return rows * 64 * 4 + rows * 2 * 4; }

static void fill(WideArgs& w, const MlpArgs& a, void* workspace) {
  w.L = a.L; w.params = a.params; w.x = a.x; w.ld = a.ld; w.rows = a.rows; w.index = a.index;
  w.x_div = a.x_div < 1 ? 1 : a.x_div;
  w.vec = ((reinterpret_cast<uintptr_t>(a.x) & 15u) == 0 && (a.ld & 3) == 0) ? 1 : 0;
  w.z1 = static_cast<float*>(workspace);
  w.xstats = w.z1 + a.rows * 64;
  w.dz1 = w.z1;
  w.D = a.L.in_dim; w.dz_ld = 64; w.norm = a.L.use_fn; w.total = a.L.total; w.out_w = a.L.w1;
  w.transposed = 0; w.rot = 0; w.out_b = -1;
}

int launch_wide_fc1(const MlpArgs& a, void* workspace, cudaStream_t stream, const char* who) {
  WideArgs w{};
  fill(w, a, workspace);
  w.nhalf = 1;
  while (w.nhalf < 4 && make_fwd_map(a.L.in_dim, w.nhalf).total > (uint32_t)kWideMaxSmem) w.nhalf *= 2;
  const FwdMap m = make_fwd_map(a.L.in_dim, w.nhalf);
  MAPPO_REQUIRE(m.total <= (uint32_t)kWideMaxSmem, "%s: in_dim %d too wide for the staged first layer", who, a.L.in_dim);
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(fc1_wide_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideMaxSmem));
    attr_set = true;
  }
  fc1_wide_kernel<<<dim3(grid_for(a.rows), w.nhalf), kThreads, m.total, stream>>>(w);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

// packed inputs (see wide_pack_kernel): bytes of the tile image for `rows` rows, the packing pass, the streamed forward
long long wide_packed_bytes(int in_dim, long long rows) { return ((rows + kBM - 1) / kBM) * packed_tile_bytes(in_dim); }

bool wide_pack_ok(const MlpArgs& a) {
  return a.index == nullptr && a.x_div <= 1 && make_pk_map(a.L.in_dim).total <= (uint32_t)kWideMaxSmem;
}

int launch_wide_pack(const MlpArgs& a, void* workspace, void* packed, cudaStream_t stream, const char* who) {
  MAPPO_REQUIRE(wide_pack_ok(a) && packed && (reinterpret_cast<uintptr_t>(packed) & 15u) == 0,
                "%s: inputs cannot be packed (row gather, in_dim %d too wide, or unaligned buffer)", who, a.L.in_dim);
  WideArgs w{};
  fill(w, a, workspace);
  wide_pack_kernel<<<grid_for(a.rows), kThreads, 0, stream>>>(w, static_cast<unsigned char*>(packed));
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

int launch_wide_fc1_packed(const MlpArgs& a, void* workspace, const void* packed, cudaStream_t stream, const char* who) {
  WideArgs w{};
  fill(w, a, workspace);
  const PkMap m = make_pk_map(a.L.in_dim);
  MAPPO_REQUIRE(wide_pack_ok(a), "%s: packed inputs do not apply to this call", who);
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(fc1_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kWideMaxSmem));
    attr_set = true;
  }
  fc1_packed_kernel<<<grid_for(a.rows), kPkThreads, m.total, stream>>>(w, static_cast<const unsigned char*>(packed));
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

int launch_wide_wgrad1_packed(const MlpArgs& a, void* workspace, const void* packed, cudaStream_t stream, const char* who) {
  WideArgs w{};
  fill(w, a, workspace);
  w.grad_partials = a.grad_partials;
  MAPPO_REQUIRE(wide_pack_ok(a) && (w.dz_ld & 3) == 0, "%s: packed inputs do not apply to this call", who);
  const int C8 = 2 * ((w.D + 15) / 16);
  w.ncg = (C8 + kPgChunks - 1) / kPgChunks;
  const long long ntiles = (w.rows + kBM - 1) / kBM;
  long long ns = grid_for(w.rows) / w.ncg;                               // <= the main kernel's CTAs (2 partial rows each)
  if (ns < 1) ns = 1;
  if (ns > ntiles) ns = ntiles;
  w.nsplit = (int)ns;
  const PgMap m = make_pg_map();
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(wgrad1_packed_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)m.total));
    attr_set = true;
  }
  wgrad1_packed_kernel<<<w.ncg * w.nsplit, kPkThreads, m.total, stream>>>(w, static_cast<const unsigned char*>(packed));
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

// G[j][k] = sum_rows dz[row][j] x[row][k] into the partial rows, split over rows (nsplit <= grid_for(rows)) and
// column groups of at most 240 columns
int launch_wgrad_generic(WideArgs& w, cudaStream_t stream, const char* who) {
  const int D = w.D;
  w.ncg = (D + kMaxColw - 1) / kMaxColw;
  w.colw = (((D + w.ncg - 1) / w.ncg) + 15) & ~15;
  const long long nht = (w.rows + kHR - 1) / kHR;
  long long ns = grid_for(w.rows) / w.ncg;                               // <= the main kernel's CTAs (2 partial rows each)
  if (ns < 1) ns = 1;
  if (ns > nht) ns = nht;
  w.nsplit = (int)ns;
  const bool dual = w.dz2 != nullptr;
  MAPPO_REQUIRE(!dual || (w.ncg == 1 && w.transposed && w.out_b >= 0 && w.win > 0 && w.off2 + w.win <= D),
                "%s: the dual weight gradient needs one column group (D <= %d), transposed outputs and both bias slots", who, kMaxColw);
  const WgMap m = make_wg_map(w.colw, dual);
  static bool attr_set = false;
  if (!attr_set) {
    MAPPO_CUDA(cudaFuncSetAttribute(wgrad1_wide_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    MAPPO_CUDA(cudaFuncSetAttribute(wgrad1_wide_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    MAPPO_CUDA(cudaFuncSetAttribute(wgrad1_wide_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 216 * 1024));
    attr_set = true;
  }
  MAPPO_REQUIRE(m.total <= 216 * 1024, "%s: weight-gradient shared-memory map too large (%u B)", who, m.total);
  if (dual) wgrad1_wide_kernel<true, true><<<w.ncg * w.nsplit, kThreads, m.total, stream>>>(w);
  else if (w.out_b >= 0) wgrad1_wide_kernel<true, false><<<w.ncg * w.nsplit, kThreads, m.total, stream>>>(w);
  else wgrad1_wide_kernel<false, false><<<w.ncg * w.nsplit, kThreads, m.total, stream>>>(w);
  MAPPO_CHECK_LAUNCH(who);
  return 0;
}

int launch_wide_wgrad1(const MlpArgs& a, void* workspace, cudaStream_t stream, const char* who) {
  WideArgs w{};
  fill(w, a, workspace);
  w.grad_partials = a.grad_partials;
  return launch_wgrad_generic(w, stream, who);
}

}  // namespace tc
}  // namespace mappo
