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

// one-hot `actions_env` rows (the reference's np.eye(n)[actions], shared/mpe_runner.py:110-121; MultiDiscrete: the
// heads' one-hots concatenated) -> float action ids, one thread per (row, head): index of the first maximum
struct OneHotArgs { int n_heads; int dim[MAPPO_MAX_HEADS]; int off[MAPPO_MAX_HEADS]; int width; };

template <typename T>
__global__ void __launch_bounds__(256)
onehot_to_ids_kernel(const T* __restrict__ onehot, long long rows, OneHotArgs a, float* __restrict__ ids, int ids_ld) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows * a.n_heads) return;
  const long long r = i / a.n_heads;
  const int h = (int)(i - r * a.n_heads);
  const T* __restrict__ p = onehot + r * a.width + a.off[h];
  int best = 0;
  T bv = p[0];
  for (int q = 1; q < a.dim[h]; ++q) { const T v = p[q]; if (v > bv) { bv = v; best = q; } }
  ids[r * ids_ld + h] = (float)best;
}

// the forward direction on the device: float ids [rows][ids_ld] -> uint8 one-hot [rows][width], one thread per output byte
__global__ void __launch_bounds__(256)
ids_to_onehot_kernel(const float* __restrict__ ids, long long rows, int ids_ld, OneHotArgs a, uint8_t* __restrict__ out,
                     int32_t* __restrict__ bad) {
  const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (i >= rows * a.width) return;
  const long long r = i / a.width;
  const int c = (int)(i - r * a.width);
  int h = 0;
  while (h + 1 < a.n_heads && c >= a.off[h + 1]) ++h;
  const int id = (int)ids[r * ids_ld + h];
  out[i] = (id == c - a.off[h]) ? 1 : 0;
  if (bad && c == a.off[h] && (id < 0 || id >= a.dim[h])) atomicExch(bad, 1);     // np.eye(n)[id] would raise IndexError
}

}  // namespace mappo

using namespace mappo;

extern "C" int mappo_ids_to_onehot_u8(const float* ids, long long rows, int ids_ld, int n_heads, const int* head_dim_host,
                                      uint8_t* onehot, int32_t* bad_flag, void* stream) {
  MAPPO_REQUIRE(ids && onehot && head_dim_host && rows > 0 && n_heads >= 1 && n_heads <= MAPPO_MAX_HEADS && ids_ld >= n_heads,
                "ids_to_onehot_u8: bad arguments");
  OneHotArgs a{};
  a.n_heads = n_heads;
  int w = 0;
  for (int h = 0; h < n_heads; ++h) {
    MAPPO_REQUIRE(head_dim_host[h] >= 1, "ids_to_onehot_u8: head_dim");
    a.dim[h] = head_dim_host[h]; a.off[h] = w; w += head_dim_host[h];
  }
  a.width = w;
  ids_to_onehot_kernel<<<ceil_div(rows * w, 256), 256, 0, (cudaStream_t)stream>>>(ids, rows, ids_ld, a, onehot, bad_flag);
  MAPPO_CHECK_LAUNCH("ids_to_onehot_kernel");
  return 0;
}

extern "C" int mappo_onehot_to_ids(const void* onehot, int dtype, long long rows, int n_heads, const int* head_dim_host,
                                   float* ids, int ids_ld, void* stream) {
  MAPPO_REQUIRE(onehot && ids && head_dim_host && rows > 0 && n_heads >= 1 && n_heads <= MAPPO_MAX_HEADS && ids_ld >= n_heads,
                "onehot_to_ids: bad arguments");
  OneHotArgs a{};
  a.n_heads = n_heads;
  int w = 0;
  for (int h = 0; h < n_heads; ++h) {
    MAPPO_REQUIRE(head_dim_host[h] >= 1, "onehot_to_ids: head_dim");
    a.dim[h] = head_dim_host[h]; a.off[h] = w; w += head_dim_host[h];
  }
  a.width = w;
  const int grid = ceil_div(rows * n_heads, 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0) onehot_to_ids_kernel<float><<<grid, 256, 0, st>>>(static_cast<const float*>(onehot), rows, a, ids, ids_ld);
  else if (dtype == 1) onehot_to_ids_kernel<double><<<grid, 256, 0, st>>>(static_cast<const double*>(onehot), rows, a, ids, ids_ld);
  else if (dtype == 2) onehot_to_ids_kernel<uint8_t><<<grid, 256, 0, st>>>(static_cast<const uint8_t*>(onehot), rows, a, ids, ids_ld);
  else MAPPO_REQUIRE(false, "onehot_to_ids: dtype %d (0 = float32, 1 = float64, 2 = uint8)", dtype);
  MAPPO_CHECK_LAUNCH("onehot_to_ids_kernel");
  return 0;
}

extern "C" int mappo_gather_rows(const float* src, long long row_stride, int row_div, int width,
                                 const int64_t* index, long long n, float* dst, void* stream) {
  MAPPO_REQUIRE(src && index && dst && width > 0 && n > 0 && row_div >= 1 && row_stride >= 0, "gather_rows: bad arguments");
  long long blocks = (n * 32 + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  gather_rows_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, row_stride, row_div, width, index, n, dst);
  MAPPO_CHECK_LAUNCH("gather_rows_kernel");
  return 0;
}

// ---- several device -> device copies in one launch (SharedReplayBuffer.insert) ------------------------------------
namespace mappo {
// every block walks every segment with a grid stride: 16-byte pieces where both pointers allow it, 4-byte words otherwise
__global__ void __launch_bounds__(256) copy_segments_kernel(const mappo_copy_list l) {
  const long long t0 = (long long)blockIdx.x * blockDim.x + threadIdx.x, nt = (long long)gridDim.x * blockDim.x;
  for (int s = 0; s < l.n; ++s) {
    const char* __restrict__ src = static_cast<const char*>(l.src[s]);
    char* __restrict__ dst = static_cast<char*>(l.dst[s]);
    const long long bytes = l.bytes[s];
    const bool wide = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15u) == 0;
    if (wide) {
      const long long n16 = bytes >> 4;
      for (long long i = t0; i < n16; i += nt) reinterpret_cast<uint4*>(dst)[i] = reinterpret_cast<const uint4*>(src)[i];
      for (long long i = (n16 << 2) + t0; i < (bytes >> 2); i += nt)
        reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
    } else {
      for (long long i = t0; i < (bytes >> 2); i += nt)
        reinterpret_cast<uint32_t*>(dst)[i] = reinterpret_cast<const uint32_t*>(src)[i];
    }
  }
}
}  // namespace mappo

extern "C" int mappo_copy_segments(const mappo_copy_list* list, void* stream) {
  MAPPO_REQUIRE(list && list->n >= 1 && list->n <= 8, "copy_segments: 1..8 segments");
  long long total = 0;
  for (int s = 0; s < list->n; ++s) {
    MAPPO_REQUIRE(list->src[s] && list->dst[s] && list->bytes[s] > 0 && (list->bytes[s] & 3) == 0 &&
                  ((reinterpret_cast<uintptr_t>(list->src[s]) | reinterpret_cast<uintptr_t>(list->dst[s])) & 3u) == 0,
                  "copy_segments: segment %d needs 4-byte aligned pointers and a size that is a multiple of 4", s);
    total += list->bytes[s];
  }
  long long blocks = (total / 16 + 255) / 256;
  if (blocks < 1) blocks = 1;
  if (blocks > 148 * 8) blocks = 148 * 8;
  mappo::copy_segments_kernel<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(*list);
  MAPPO_CHECK_LAUNCH("copy_segments_kernel");
  return 0;
}
