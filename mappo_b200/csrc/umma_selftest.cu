// Hardware self-test of the tcgen05 operand layouts in umma.cuh: one CTA stages two fp32 tiles in the chunk-major
// layout, issues K/8 tf32 MMAs with caller-given descriptor strides, and writes the TMEM accumulator back.
// tests/test_gpu_umma.py checks the three operand combinations the policy kernels use against numpy.
#include "common.cuh"
#include "umma.cuh"
#include <cuda_bf16.h>

namespace mappo {

struct UmmaProbe {
  const float* A; int a_rows, a_cols;
  const float* B; int b_rows, b_cols;
  int M, N, K, a_mn, b_mn;
  uint32_t a_lbo, a_sbo, a_kstep, b_lbo, b_sbo, b_kstep;
  float* D;
  int tmem_cols;
  int bf16;
};

__global__ void __launch_bounds__(128, 1) umma_probe_kernel(const UmmaProbe p) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  unsigned char* sa = smem;
  const int esz = p.bf16 ? 2 : 4;
  unsigned char* sb = smem + (size_t)p.a_rows * p.a_cols * esz;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int e = tid; e < p.a_rows * p.a_cols; e += 128) {
    const int r = e / p.a_cols, c = e - r * p.a_cols;
    if (p.bf16) *reinterpret_cast<__nv_bfloat16*>(sa + umma::tile_off16(p.a_rows, r, c)) = __float2bfloat16(p.A[e]);
    else *reinterpret_cast<float*>(sa + umma::tile_off(p.a_rows, r, c)) = p.A[e];
  }
  for (int e = tid; e < p.b_rows * p.b_cols; e += 128) {
    const int r = e / p.b_cols, c = e - r * p.b_cols;
    if (p.bf16) *reinterpret_cast<__nv_bfloat16*>(sb + umma::tile_off16(p.b_rows, r, c)) = __float2bfloat16(p.B[e]);
    else *reinterpret_cast<float*>(sb + umma::tile_off(p.b_rows, r, c)) = p.B[e];
  }
  if (tid == 0) { umma::mbar_init(&bar, 1); umma::fence_mbar_init(); }
  if (warp == 0) umma::tmem_alloc(&tmem_base_s, (uint32_t)p.tmem_cols);
  umma::fence_async_smem();
  umma::tc_fence_before();
  __syncthreads();
  umma::tc_fence_after();
  const uint32_t tbase = tmem_base_s;
  if (tid == 0) {
    const uint32_t idesc = p.bf16 ? umma::idesc_bf16(p.M, p.N, p.a_mn, p.b_mn) : umma::idesc_tf32(p.M, p.N, p.a_mn, p.b_mn);
    const int kk = p.bf16 ? 16 : 8;
    const uint32_t a0 = umma::smem_u32(sa), b0 = umma::smem_u32(sb);
    for (int k = 0; k < p.K / kk; ++k) {
      const uint64_t da = umma::smem_desc(a0 + k * p.a_kstep, p.a_lbo, p.a_sbo);
      const uint64_t db = umma::smem_desc(b0 + k * p.b_kstep, p.b_lbo, p.b_sbo);
      if (p.bf16) umma::mma_bf16(tbase, da, db, idesc, k > 0);
      else umma::mma_tf32(tbase, da, db, idesc, k > 0);
    }
    umma::mma_commit(&bar);
  }
  umma::mbar_wait(&bar, 0);
  umma::tc_fence_after();
  for (int c0 = 0; c0 < p.N; c0 += 16) {
    float v[16];
    umma::tmem_ld16(tbase + ((uint32_t)(32 * warp) << 16) + (uint32_t)c0, v);
    for (int i = 0; i < 16; ++i) p.D[(size_t)(32 * warp + lane) * p.N + c0 + i] = v[i];
  }
  umma::tc_fence_before();
  __syncthreads();
  if (warp == 0) umma::tmem_dealloc(tbase, (uint32_t)p.tmem_cols);
}

}  // namespace mappo

using namespace mappo;

extern "C" int mappo_selftest_umma(const float* A, int a_rows, int a_cols, const float* B, int b_rows, int b_cols,
                                   int M, int N, int K, int a_mn_major, int b_mn_major, uint32_t a_lbo,
                                   uint32_t a_sbo, uint32_t a_kstep, uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep,
                                   float* D, int use_bf16, void* stream) {
  MAPPO_REQUIRE(A && B && D, "selftest_umma: null pointer");
  MAPPO_REQUIRE(M == 128 && N >= 16 && N <= 256 && (N % 16) == 0 && K >= 8 && (K % (use_bf16 ? 16 : 8)) == 0, "selftest_umma: bad shape");
  MAPPO_REQUIRE((a_rows % 8) == 0 && (a_cols % 8) == 0 && (b_rows % 8) == 0 && (b_cols % 8) == 0, "selftest_umma: tile shape");
  UmmaProbe p{};
  p.A = A; p.a_rows = a_rows; p.a_cols = a_cols; p.B = B; p.b_rows = b_rows; p.b_cols = b_cols;
  p.M = M; p.N = N; p.K = K; p.a_mn = a_mn_major; p.b_mn = b_mn_major;
  p.a_lbo = a_lbo; p.a_sbo = a_sbo; p.a_kstep = a_kstep; p.b_lbo = b_lbo; p.b_sbo = b_sbo; p.b_kstep = b_kstep;
  p.D = D; p.bf16 = use_bf16;
  int cols = 32;
  while (cols < N) cols <<= 1;
  p.tmem_cols = cols;
  const size_t smem = ((size_t)a_rows * a_cols + (size_t)b_rows * b_cols) * (use_bf16 ? 2 : 4);
  MAPPO_REQUIRE(smem <= 200 * 1024, "selftest_umma: tiles too large");
  MAPPO_CUDA(cudaFuncSetAttribute(umma_probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  umma_probe_kernel<<<1, 128, smem, (cudaStream_t)stream>>>(p);
  MAPPO_CHECK_LAUNCH("umma_probe_kernel");
  return 0;
}
