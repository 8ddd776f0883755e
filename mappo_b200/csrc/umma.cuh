// tcgen05 / TMEM / mbarrier primitives for sm_100a (inline PTX), and the shared-memory operand layout the
// fused policy kernels use.
//
// Operand layout ("chunk-major", no swizzle).  A [R rows][C cols] fp32/tf32 tile is stored as
//     byte offset(r, c) = ((c / 4) * R + r) * 16 + (c % 4) * 4
// i.e. 16-byte chunks of 4 consecutive columns, all rows of a chunk contiguous.  One 8-row x 16-byte block is a
// tcgen05 "core matrix" (128 contiguous bytes), so the SAME bytes serve as
//   * a K-major operand  (MN index = row, K index = col): 8-row groups 128 B apart (SBO), the next 4 columns
//     R*16 B further (LBO); one K = 8 instruction step advances the start address by 2*R*16 B;
//   * an MN-major operand (K index = row, MN index = col): the next 4 MN elements R*16 B further (SBO), the next
//     8 K rows 128 B further (LBO / start-address advance per K = 8 step).
// A thread that owns row r writes its 4-column chunks with st.shared.v4: the 32 lanes of a warp write 512
// contiguous bytes, which is conflict-free.  Layouts were verified on a B200 by tests/test_gpu_umma.py.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace mappo {
namespace umma {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// ---- mbarrier ----------------------------------------------------------------------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {}
}

// ---- proxy / tcgen05 fences --------------------------------------------------------------------------
// generic-proxy shared-memory writes -> visible to the async proxy (tcgen05.mma operand reads)
// ask the L2 for `bytes` (a multiple of 16) of global memory ahead of the loads that will want them: no registers,
// no shared memory, one instruction from one thread
__device__ __forceinline__ void l2_prefetch(const void* gptr, uint32_t bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(gptr), "r"(bytes) : "memory");
}
// one line into the L1 ahead of a load whose address is known early but whose value is wanted late (no register held)
__device__ __forceinline__ void l1_prefetch(const void* gptr) { asm volatile("prefetch.global.L1 [%0];" ::"l"(gptr) : "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// ---- TMEM allocation (one full warp calls alloc / dealloc) ---------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}

// ---- descriptors ---------------------------------------------------------------------------------------
// shared-memory matrix descriptor, no swizzle (cute::UMMA::SmemDescriptor: start[0,14) LBO[16,30) SBO[32,46)
// version[46,48) = 1, layout_type[61,64) = 0), all byte quantities >> 4
// (no field masks: shared-memory addresses are < 256 KB and the strides < 256 KB, so every field fits its 14 bits;
// without the masks the compiler folds a descriptor into one or two uniform-datapath instructions)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) + ((lbo_bytes >> 4) << 16)) | ((uint64_t)((sbo_bytes >> 4) | (1u << 14)) << 32);
}
// same, from an address that is already in 16-byte units (base >> 4 computed once, offsets added as constants)
__device__ __forceinline__ uint64_t smem_desc16(uint32_t addr16, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)(addr16 + ((lbo_bytes >> 4) << 16)) | ((uint64_t)((sbo_bytes >> 4) | (1u << 14)) << 32);
}
// one elected lane of a CONVERGED warp (call under a warp-uniform condition).  Issuing tcgen05.mma / commit / bulk
// copies under `warp == w && elect_one()` instead of `tid == 0` lets the compiler emit them straight on the uniform
// datapath; under a per-thread condition it wraps every one of them in an ELECT / BRA.U.ANY serialisation loop.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
}
// instruction descriptor for kind::tf32, fp32 accumulate (cute::UMMA::InstrDescriptor)
__host__ __device__ constexpr uint32_t idesc_tf32(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// instruction descriptor for kind::f16 with bf16 operands, fp32 accumulate
__host__ __device__ constexpr uint32_t idesc_bf16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) |
         ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
// kind::f16 with fp16 operands
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N, int a_mn_major, int b_mn_major) {
  return (1u << 4) | ((uint32_t)a_mn_major << 15) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
      :
      : "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "r"(0u)
      : "memory");
}

__device__ __forceinline__ void mma_f16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  mma_bf16(d_tmem, a_desc, b_desc, idesc, accumulate);   // same instruction; the operand format lives in idesc
}

// D[tmem] (+)= A[smem] * B[smem]; issued by ONE thread
__device__ __forceinline__ void mma_tf32(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  const uint32_t acc = accumulate ? 1u : 0u;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, {%5, %5, %5, %5}, p;\n\t}"
      :
      : "r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "r"(0u)
      : "memory");
}
// all MMAs issued so far by this thread -> one arrival on `bar` when they have completed
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// ---- TMEM -> registers: this warp's 32 lanes (lane field of taddr = 32 * (warp % 4)), 16 consecutive columns
// asynchronous form: the registers are valid only after tmem_ld_wait()
__device__ __forceinline__ void tmem_ld16_async(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  tmem_ld16_async(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 16; ++i) v[i] = __uint_as_float(r[i]);
}
// two 16-column blocks in flight under one wait
__device__ __forceinline__ void tmem_ld16x2(uint32_t taddr0, uint32_t taddr1, float (&v)[16], float (&w)[16]) {
  uint32_t r0[16], r1[16];
  tmem_ld16_async(taddr0, r0);
  tmem_ld16_async(taddr1, r1);
  tmem_ld_wait();
#pragma unroll
  for (int i = 0; i < 16; ++i) { v[i] = __uint_as_float(r0[i]); w[i] = __uint_as_float(r1[i]); }
}

// ---- operand tiles ---------------------------------------------------------------------------------------
__host__ __device__ constexpr uint32_t tile_off(int R, int r, int c) { return (uint32_t)(((c >> 2) * R + r) * 16 + (c & 3) * 4); }
// 16-bit elements: 8 columns per 16-byte chunk
__host__ __device__ constexpr uint32_t tile_off16(int R, int r, int c) { return (uint32_t)(((c >> 3) * R + r) * 16 + (c & 7) * 2); }

// fp32 -> (hi, lo) with hi = tf32(x) (round to nearest, ties away: cvt.rna) and lo = tf32(x - hi)
__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
  uint32_t h;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
  hi = __uint_as_float(h);
  uint32_t l;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(x - hi));
  lo = __uint_as_float(l);
}

}  // namespace umma
}  // namespace mappo
