// Minibatch row gather: dst[i][:] = src[(index[i] / row_div) * row_stride + :width].
//   the NumPy fancy-index gathers of feed_forward_generator / recurrent_generator / naive_recurrent_generator
//   onpolicy/utils/shared_buffer.py:262-282,440-455,330-345 and separated_buffer.py:193-213
// The training path fuses this gather into the update kernels (index argument); this standalone kernel backs the
// generator API.  One warp per row, float4 when the row is 16-byte aligned.
#include "common.cuh"

namespace mappo {

__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ src, long long row_stride, int row_div, int width,
                   const int64_t* __restrict__ index, long long n, float* __restrict__ dst) {
  const int lane = threadIdx.x & 31;
  const long long warp = (blockIdx.x * (long long)blockDim.x + threadIdx.x) >> 5;
  const long long nwarps = ((long long)gridDim.x * blockDim.x) >> 5;
  for (long long i = warp; i < n; i += nwarps) {
    const long long r = index[i] / row_div;
    const float* __restrict__ s = src + r * row_stride;
    float* __restrict__ d = dst + i * (long long)width;
    const bool vec = ((width & 3) == 0) && ((row_stride & 3) == 0) &&
                     ((reinterpret_cast<uintptr_t>(src) & 15) == 0) && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0);
    if (vec) {
      for (int c = lane; c < (width >> 2); c += 32)
        reinterpret_cast<float4*>(d)[c] = reinterpret_cast<const float4*>(s)[c];
    } else {
      for (int c = lane; c < width; c += 32) d[c] = s[c];
    }
  }
}

}  // namespace mappo

using namespace mappo;

extern "C" int mappo_gather_rows(const float* src, long long row_stride, int row_div, int width,
                                 const int64_t* index, long long n, float* dst, void* stream) {
  MAPPO_REQUIRE(src && index && dst && width > 0 && n > 0 && row_div >= 1 && row_stride >= 0, "gather_rows: bad arguments");
  long long blocks = (n * 32 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gather_rows_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, row_stride, row_div, width, index, n, dst);
  MAPPO_CHECK_LAUNCH("gather_rows_kernel");
  return 0;
}
