// Returns / GAE / ValueNorm / advantage statistics.  float32, reference operation order, compiled with
// -fmad=false so every product and sum rounds separately exactly as numpy / torch-CPU do.
//   SharedReplayBuffer.compute_returns     onpolicy/utils/shared_buffer.py:168-224
//   ValueNorm                               onpolicy/utils/valuenorm.py:31-78
//   advantage prologue of R_MAPPO.train     onpolicy/algorithms/r_mappo/r_mappo.py:174-182
#include "common.cuh"

namespace mappo {

constexpr int kGaeThreads = 256;
constexpr int kRedThreads = 256;

// ValueNorm.running_mean_var (valuenorm.py:31-35): returns mean and sqrt(var), all float32 ops.
__device__ __forceinline__ void vn_mean_std(const float* st, float& mean, float& sd) {
  const float db = fmaxf(st[2], 1e-5f);                 // debiasing_term.clamp(min=epsilon)
  mean = st[0] / db;
  const float mean_sq = st[1] / db;
  const float var = fmaxf(mean_sq - mean * mean, 1e-2f);
  sd = sqrtf(var);
}

// block-wide sum of three doubles; result valid in thread 0
__device__ __forceinline__ void block_sum3(double& a, double& b, double& c) {
  __shared__ double sh[3][32];
  for (int o = 16; o > 0; o >>= 1) {
    a += __shfl_down_sync(0xffffffffu, a, o);
    b += __shfl_down_sync(0xffffffffu, b, o);
    c += __shfl_down_sync(0xffffffffu, c, o);
  }
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) { sh[0][w] = a; sh[1][w] = b; sh[2][w] = c; }
  __syncthreads();
  if (w == 0) {
    const int nw = (blockDim.x + 31) >> 5;
    a = l < nw ? sh[0][l] : 0.0; b = l < nw ? sh[1][l] : 0.0; c = l < nw ? sh[2][l] : 0.0;
    for (int o = 16; o > 0; o >>= 1) {
      a += __shfl_down_sync(0xffffffffu, a, o);
      b += __shfl_down_sync(0xffffffffu, b, o);
      c += __shfl_down_sync(0xffffffffu, c, o);
    }
  }
}

// One thread per trajectory r; backward scan over t.  Loads at fixed t are coalesced across r.
__global__ void __launch_bounds__(kGaeThreads)
gae_kernel(const mappo_gae_config cfg, const float* __restrict__ rewards, float* __restrict__ value_preds,
           float* __restrict__ returns, const float* __restrict__ masks, const float* __restrict__ bad_masks,
           const float* __restrict__ active_masks, const float* __restrict__ next_value,
           const float* __restrict__ vn_stats, float* __restrict__ adv_raw, double* __restrict__ partials) {
  const int T = cfg.T, R = cfg.R;
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  double s1 = 0.0, s2 = 0.0, cnt = 0.0;
  if (r < R) {
    const bool norm = cfg.use_value_norm != 0;
    float mean = 0.f, sd = 1.f;
    if (norm) vn_mean_std(vn_stats, mean, sd);
    const float g = cfg.gamma, gl = cfg.gamma_lambda;
    const size_t S = (size_t)R;
    if (cfg.use_gae) {
      float v_next = next_value ? next_value[r] : value_preds[(size_t)T * S + r];
      if (next_value) value_preds[(size_t)T * S + r] = v_next;            // shared_buffer.py:175,207
      float dn_next = norm ? (v_next * sd + mean) : v_next;                // denormalize: x*sqrt(var)+mean
      float gae = 0.f;
      for (int t = T - 1; t >= 0; --t) {
        const float v = value_preds[(size_t)t * S + r];
        const float dn = norm ? (v * sd + mean) : v;
        const float m1 = masks[(size_t)(t + 1) * S + r];
        const float rew = rewards[(size_t)t * S + r];
        // delta = r + gamma * V(t+1) * mask(t+1) - V(t)      (:209-211 / :217-218)
        const float delta = (rew + (g * dn_next) * m1) - dn;
        if (cfg.use_proper_time_limits) {
          // gae = delta + gamma*lambda * gae * mask ; gae *= bad_mask   (:181-183 / :188-189)
          gae = delta + (gl * gae) * m1;
          gae = gae * (bad_masks ? bad_masks[(size_t)(t + 1) * S + r] : 1.0f);
        } else {
          gae = delta + (gl * m1) * gae;                                   // (:212 / :219)
        }
        const float ret = gae + dn;
        returns[(size_t)t * S + r] = ret;
        const float adv = ret - dn;                                        // r_mappo.py:174-177
        adv_raw[(size_t)t * S + r] = adv;
        const bool active = active_masks ? (active_masks[(size_t)t * S + r] != 0.0f) : true;
        if (active) { s1 += (double)adv; s2 += (double)adv * (double)adv; cnt += 1.0; }
        dn_next = dn;
      }
    } else {
      float ret_next = next_value ? next_value[r] : returns[(size_t)T * S + r];
      returns[(size_t)T * S + r] = ret_next;                              // :192 / :222
      for (int t = T - 1; t >= 0; --t) {
        const float m1 = masks[(size_t)(t + 1) * S + r];
        const float rew = rewards[(size_t)t * S + r];
        const float v = value_preds[(size_t)t * S + r];
        const float dn = norm ? (v * sd + mean) : v;
        float ret = ((ret_next * g) * m1) + rew;
        if (cfg.use_proper_time_limits) {
          const float bm = bad_masks ? bad_masks[(size_t)(t + 1) * S + r] : 1.0f;
          ret = ret * bm + (1.0f - bm) * dn;                               // :194-203
        }
        returns[(size_t)t * S + r] = ret;
        const float adv = ret - dn;
        adv_raw[(size_t)t * S + r] = adv;
        const bool active = active_masks ? (active_masks[(size_t)t * S + r] != 0.0f) : true;
        if (active) { s1 += (double)adv; s2 += (double)adv * (double)adv; cnt += 1.0; }
        ret_next = ret;
      }
    }
  }
  block_sum3(s1, s2, cnt);
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * 3 + 0] = s1;
    partials[(size_t)blockIdx.x * 3 + 1] = s2;
    partials[(size_t)blockIdx.x * 3 + 2] = cnt;
  }
}

// fixed-order reduction of per-block partials -> stats3 (single block)
__global__ void __launch_bounds__(kRedThreads)
stats_reduce_kernel(const double* __restrict__ partials, int n, double* __restrict__ stats3) {
  double a = 0.0, b = 0.0, c = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    a += partials[(size_t)i * 3 + 0]; b += partials[(size_t)i * 3 + 1]; c += partials[(size_t)i * 3 + 2];
  }
  block_sum3(a, b, c);
  if (threadIdx.x == 0) { stats3[0] = a; stats3[1] = b; stats3[2] = c; }
}

__global__ void __launch_bounds__(256)
adv_normalize_kernel(const float* __restrict__ raw, float* __restrict__ out, long long count,
                     const double* __restrict__ stats3) {
  const double n = stats3[2];
  const double mu = stats3[0] / n;
  double var = stats3[1] / n - mu * mu;
  var = var > 0.0 ? var : 0.0;
  const float mean = (float)mu;                       // np.nanmean -> float32
  const float sd = (float)sqrt(var);                  // np.nanstd (population) -> float32
  const float denom = sd + 1e-5f;                     // r_mappo.py:182
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = (raw[i] - mean) / denom;
}

__global__ void __launch_bounds__(kRedThreads)
moments_kernel(const float* __restrict__ x, long long count, double* __restrict__ partials) {
  double a = 0.0, b = 0.0, c = 0.0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = x[i];
    a += (double)v; b += (double)(v * v);              // (input_vector ** 2) is rounded to f32 first
    c += 1.0;
  }
  block_sum3(a, b, c);
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * 3 + 0] = a; partials[(size_t)blockIdx.x * 3 + 1] = b;
    partials[(size_t)blockIdx.x * 3 + 2] = c;
  }
}

// r_mappo.py:174-177 from the buffer contents, with the moments over active entries
__global__ void __launch_bounds__(kRedThreads)
advantages_kernel(const float* __restrict__ returns, const float* __restrict__ value_preds,
                  const float* __restrict__ active, const float* __restrict__ vn_stats, long long count,
                  float* __restrict__ adv_raw, double* __restrict__ partials) {
  float mean = 0.f, sd = 1.f;
  if (vn_stats) vn_mean_std(vn_stats, mean, sd);
  double a = 0.0, b = 0.0, c = 0.0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x) {
    const float v = value_preds[i];
    const float dn = vn_stats ? (v * sd + mean) : v;
    const float adv = returns[i] - dn;
    adv_raw[i] = adv;
    if (!active || active[i] != 0.0f) { a += (double)adv; b += (double)adv * (double)adv; c += 1.0; }
  }
  block_sum3(a, b, c);
  if (threadIdx.x == 0) {
    partials[(size_t)blockIdx.x * 3 + 0] = a; partials[(size_t)blockIdx.x * 3 + 1] = b;
    partials[(size_t)blockIdx.x * 3 + 2] = c;
  }
}

// ValueNorm.update (valuenorm.py:38-54), per_element_update=False: weight = beta
__global__ void valuenorm_apply_kernel(float* st, const double* __restrict__ stats3, float w, float omw) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    const float bm = (float)(stats3[0] / stats3[2]);
    const float bsq = (float)(stats3[1] / stats3[2]);
    st[0] = st[0] * w + bm * omw;
    st[1] = st[1] * w + bsq * omw;
    st[2] = st[2] * w + omw;                            // 1.0 * (1.0 - weight)
  }
}

__global__ void __launch_bounds__(256)
valuenorm_transform_kernel(const float* __restrict__ st, const float* __restrict__ x, float* __restrict__ out,
                           long long count, int denorm) {
  float mean, sd;
  vn_mean_std(st, mean, sd);
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < count;
       i += (long long)gridDim.x * blockDim.x)
    out[i] = denorm ? (x[i] * sd + mean) : ((x[i] - mean) / sd);
}

// popart.py:52-70 (the two assignments at the end of PopArt.update), float32 like torch
__global__ void popart_rescale_kernel(const float* __restrict__ old_st, const float* __restrict__ new_st,
                                      float* __restrict__ w, float* __restrict__ b, int n_in, int n_out) {
  float om, os, nm, ns;
  vn_mean_std(old_st, om, os);
  vn_mean_std(new_st, nm, ns);
  for (int e = threadIdx.x; e < n_in * n_out; e += blockDim.x) w[e] = w[e] * os / ns;
  for (int j = threadIdx.x; j < n_out; j += blockDim.x) b[j] = (os * b[j] + om - nm) / ns;
}

static inline int grid_for(long long count, int threads, int cap) {
  long long g = (count + threads - 1) / threads;
  if (g < 1) g = 1;
  return (int)(g > cap ? cap : g);
}

}  // namespace mappo

using namespace mappo;

extern "C" int mappo_gae_num_partials(int R) { return ceil_div(R, kGaeThreads); }
extern "C" int mappo_moments_num_partials(long long count) { return grid_for(count, kRedThreads * 8, 148 * 8); }

extern "C" int mappo_gae_returns(const mappo_gae_config* cfg, const float* rewards, float* value_preds,
                                 float* returns, const float* masks, const float* bad_masks,
                                 const float* active_masks, const float* next_value, const float* vn_stats,
                                 float* advantages_raw, double* adv_partials, void* stream) {
  MAPPO_REQUIRE(cfg && rewards && value_preds && returns && masks && advantages_raw && adv_partials,
                "gae_returns: null pointer");
  MAPPO_REQUIRE(cfg->T > 0 && cfg->R > 0, "gae_returns: bad sizes T=%d R=%d", cfg->T, cfg->R);
  MAPPO_REQUIRE(!cfg->use_value_norm || vn_stats, "gae_returns: use_value_norm without statistics");
  gae_kernel<<<ceil_div(cfg->R, kGaeThreads), kGaeThreads, 0, (cudaStream_t)stream>>>(
      *cfg, rewards, value_preds, returns, masks, bad_masks, active_masks, next_value, vn_stats, advantages_raw,
      adv_partials);
  MAPPO_CHECK_LAUNCH("gae_kernel");
  return 0;
}

extern "C" int mappo_adv_stats_reduce(const double* partials, int num_partials, double* stats3, void* stream) {
  MAPPO_REQUIRE(partials && stats3 && num_partials > 0, "adv_stats_reduce: bad arguments");
  stats_reduce_kernel<<<1, kRedThreads, 0, (cudaStream_t)stream>>>(partials, num_partials, stats3);
  MAPPO_CHECK_LAUNCH("stats_reduce_kernel");
  return 0;
}

extern "C" int mappo_adv_normalize(const float* advantages_raw, float* advantages, long long count,
                                   const double* stats3, void* stream) {
  MAPPO_REQUIRE(advantages_raw && advantages && stats3 && count > 0, "adv_normalize: bad arguments");
  adv_normalize_kernel<<<grid_for(count, 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(advantages_raw, advantages,
                                                                                         count, stats3);
  MAPPO_CHECK_LAUNCH("adv_normalize_kernel");
  return 0;
}

extern "C" int mappo_moments(const float* x, long long count, double* partials, void* stream) {
  MAPPO_REQUIRE(x && partials && count > 0, "moments: bad arguments");
  moments_kernel<<<mappo_moments_num_partials(count), kRedThreads, 0, (cudaStream_t)stream>>>(x, count, partials);
  MAPPO_CHECK_LAUNCH("moments_kernel");
  return 0;
}

// ---- allocation-free variant of the batch-moments pair (no counterpart in the reference; see the header) -------------
// The per-block partial sums that mappo_moments hands back through caller scratch live here instead: a file-scope
// device array, present once on every GPU that loads the library until the process exits (28,416 bytes).
constexpr int kStaticMaxPartials = 148 * 8;                       // the launch cap of moments_kernel (grid_for above)
static __device__ __align__(256) double g_static_partials[kStaticMaxPartials * 3];

// In host code `g_static_partials` is only a stand-in for the symbol: its device address is looked up once per
// device, with that device current, and it is that pointer the kernels receive.
static int static_partials_ptr(double** out) {
  constexpr int kMaxDev = 64;
  static double* cache[kMaxDev] = {nullptr};
  int dev = 0;
  MAPPO_CUDA(cudaGetDevice(&dev));
  MAPPO_REQUIRE(dev >= 0 && dev < kMaxDev, "moments_static: device ordinal %d out of range", dev);
  if (cache[dev] == nullptr) {
    double* p = nullptr;
    MAPPO_CUDA(cudaGetSymbolAddress((void**)&p, g_static_partials));
    MAPPO_REQUIRE(p != nullptr, "moments_static: cudaGetSymbolAddress returned a null pointer");
    cache[dev] = p;
  }
  *out = cache[dev];
  return 0;
}

extern "C" long long mappo_moments_static_max_count(void) { return MAPPO_MOMENTS_STATIC_MAX_COUNT; }

extern "C" int mappo_moments_static(const float* x, long long count, double* stats3, void* stream) {
  MAPPO_REQUIRE(x && stats3 && count > 0, "moments_static: bad arguments");
  MAPPO_REQUIRE(count <= MAPPO_MOMENTS_STATIC_MAX_COUNT,
                "moments_static: count %lld exceeds the maximum of %lld elements this allocation-free entry point "
                "accepts (use mappo_moments with caller scratch)", count, (long long)MAPPO_MOMENTS_STATIC_MAX_COUNT);
  const int nparts = mappo_moments_num_partials(count);
  MAPPO_REQUIRE(nparts <= kStaticMaxPartials, "moments_static: %d partial rows exceed the static scratch (%d)", nparts,
                kStaticMaxPartials);
  double* scratch = nullptr;
  if (int rc = static_partials_ptr(&scratch)) return rc;
  // nothing has to start at zero: every launched block overwrites its own row, and the reduction reads exactly the
  // rows this launch wrote (the array keeps the previous call's contents beyond them)
  moments_kernel<<<nparts, kRedThreads, 0, (cudaStream_t)stream>>>(x, count, scratch);
  MAPPO_CHECK_LAUNCH("moments_kernel (static scratch)");
  stats_reduce_kernel<<<1, kRedThreads, 0, (cudaStream_t)stream>>>(scratch, nparts, stats3);
  MAPPO_CHECK_LAUNCH("stats_reduce_kernel (static scratch)");
  return 0;
}

extern "C" int mappo_advantages(const float* returns, const float* value_preds, const float* active_masks,
                                const float* vn_stats, long long count, float* advantages_raw, double* partials,
                                void* stream) {
  MAPPO_REQUIRE(returns && value_preds && advantages_raw && partials && count > 0, "advantages: bad arguments");
  advantages_kernel<<<mappo_moments_num_partials(count), kRedThreads, 0, (cudaStream_t)stream>>>(
      returns, value_preds, active_masks, vn_stats, count, advantages_raw, partials);
  MAPPO_CHECK_LAUNCH("advantages_kernel");
  return 0;
}

extern "C" int mappo_popart_rescale(const float* old_stats, const float* new_stats, float* head_w, float* head_b, int n_in,
                                    int n_out, void* stream) {
  MAPPO_REQUIRE(old_stats && new_stats && head_w && head_b && n_in > 0 && n_out > 0, "popart_rescale: bad arguments");
  popart_rescale_kernel<<<1, 64, 0, (cudaStream_t)stream>>>(old_stats, new_stats, head_w, head_b, n_in, n_out);
  MAPPO_CHECK_LAUNCH("popart_rescale_kernel");
  return 0;
}

extern "C" int mappo_valuenorm_apply(float* vn_stats, const double* stats3, double beta, void* stream) {
  MAPPO_REQUIRE(vn_stats && stats3, "valuenorm_apply: null pointer");
  valuenorm_apply_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(vn_stats, stats3, (float)beta, (float)(1.0 - beta));
  MAPPO_CHECK_LAUNCH("valuenorm_apply_kernel");
  return 0;
}

extern "C" int mappo_valuenorm_transform(const float* vn_stats, const float* x, float* out, long long count,
                                         int denormalize, void* stream) {
  MAPPO_REQUIRE(vn_stats && x && out && count > 0, "valuenorm_transform: bad arguments");
  valuenorm_transform_kernel<<<grid_for(count, 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(vn_stats, x, out, count,
                                                                                               denormalize);
  MAPPO_CHECK_LAUNCH("valuenorm_transform_kernel");
  return 0;
}
