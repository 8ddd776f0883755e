// Gradient reduction, LayerNorm-affine chain rule, global-norm clipping and Adam.
//   nn.utils.clip_grad_norm_(params, max_grad_norm)     onpolicy/algorithms/r_mappo/r_mappo.py:143-146,157-160
//   torch.optim.Adam(lr, eps=opti_eps, weight_decay)    onpolicy/algorithms/r_mappo/algorithm/rMAPPOPolicy.py:31-37
//   train_info accumulation                              r_mappo.py:187-219
#include "net.cuh"

namespace mappo {

// column sums over the per-CTA partial rows, fixed order (deterministic): 32 row slices per column are summed
// independently (rows s, s + 32, ...: ~10 loads in flight per thread at 296 partial rows) and combined in slice order
constexpr int kRedSlices = 32, kRedCols = 32;
template <typename GT>
__global__ void __launch_bounds__(kRedSlices * kRedCols)
grad_reduce_kernel(const float* __restrict__ gp, const double* __restrict__ lp, int blocks, long long P,
                   GT* __restrict__ gsum, double* __restrict__ lsum) {
  __shared__ float part[kRedSlices][kRedCols];
  const int c = threadIdx.x % kRedCols, sl = threadIdx.x / kRedCols;
  const long long e = blockIdx.x * (long long)kRedCols + c;
  float s = 0.0f;
  if (e < P)
    for (int b = sl; b < blocks; b += kRedSlices) s += gp[(size_t)b * P + e];
  part[sl][c] = s;
  __syncthreads();
  if (sl == 0 && e < P) {
    float t = part[0][c];
#pragma unroll
    for (int q = 1; q < kRedSlices; ++q) t += part[q][c];
    gsum[e] = (GT)t;
  }
  if (blockIdx.x == 0 && threadIdx.x < 4) {
    double t = 0.0;
    for (int b = 0; b < blocks; ++b) t += lp[(size_t)b * 4 + threadIdx.x];
    lsum[threadIdx.x] = t;
  }
}

constexpr int kOptThreads = 1024;

// one Linear's slots: dW[j][k] = g[k] G[j][k] + be[k] db[j];  dg[k] = sum_j W[j][k] G[j][k];  dbe[k] = sum_j W[j][k] db[j]
template <typename GT>
__device__ __forceinline__ void unfold_linear(const float* __restrict__ P, const GT* __restrict__ G, float* __restrict__ out,
                                              int w_off, int b_off, int rows, int cols, int g_off, int be_off,
                                              bool accumulate_affine) {
  const int tid = threadIdx.x;
  for (int e = tid; e < rows * cols; e += kOptThreads) {
    const int j = e / cols, k = e - j * cols;
    out[w_off + e] = (g_off >= 0) ? fmaf(P[g_off + k], (float)G[w_off + e], P[be_off + k] * (float)G[b_off + j])
                                  : (float)G[w_off + e];
  }
  for (int j = tid; j < rows; j += kOptThreads) out[b_off + j] = (float)G[b_off + j];
  if (g_off >= 0) {
    // 64 columns at a time, 16 row slices per column (rows s, s + 16, ...), combined in slice order
    __shared__ float pdg[16][64], pdbe[16][64];
    const int kk = tid & 63, sl = tid >> 6;
    for (int k0 = 0; k0 < cols; k0 += 64) {
      const int k = k0 + kk;
      float dg = 0.0f, dbe = 0.0f;
      if (k < cols) {
        for (int j = sl; j < rows; j += 16) {
          const float w = P[w_off + j * cols + k];
          dg = fmaf(w, (float)G[w_off + j * cols + k], dg);
          dbe = fmaf(w, (float)G[b_off + j], dbe);
        }
      }
      pdg[sl][kk] = dg; pdbe[sl][kk] = dbe;
      __syncthreads();
      if (sl == 0 && k < cols) {
        float a = pdg[0][kk], b = pdbe[0][kk];
#pragma unroll
        for (int q = 1; q < 16; ++q) { a += pdg[q][kk]; b += pdbe[q][kk]; }
        if (accumulate_affine) { out[g_off + k] += a; out[be_off + k] += b; }
        else { out[g_off + k] = a; out[be_off + k] = b; }
      }
      __syncthreads();
    }
  }
}

template <typename GT>
__global__ void __launch_bounds__(kOptThreads)
adam_step_kernel(const NetLayout L, const mappo_adam_config ad, float* __restrict__ params, float* __restrict__ m,
                 float* __restrict__ v, const GT* __restrict__ gsum, const double* __restrict__ lsum,
                 int denom_slot, long long rows_total, int apply_update, float* __restrict__ g, float* info,
                 long long* step_dev) {
  const int tid = threadIdx.x;
  const int D = L.in_dim, H = L.H;
  // 1. chain rule through the folded affines (net.cuh)
  unfold_linear(params, gsum, g, L.w1, L.b1, H, D, L.use_fn ? L.fn_g : -1, L.fn_b, false);
  unfold_linear(params, gsum, g, L.w2, L.b2, H, H, L.g1, L.be1, false);
  for (int l = 1; l < L.layer_N; ++l)      // fc2[l] follows the LayerNorm of fc2[l - 1] (mlp.py:26-28)
    unfold_linear(params, gsum, g, layer_w(L, l), layer_b(L, l), H, H, layer_g(L, l - 1), layer_be(L, l - 1), false);
  // what follows the last hidden LayerNorm: the heads, or (recurrent) the GRU input weights; the GRU's own LayerNorm
  // then precedes the heads, and W_hh sees the raw hidden state (no affine)
  int hg = layer_g(L, L.layer_N - 1), hbe = layer_be(L, L.layer_N - 1);
  if (L.recurrent) {
    unfold_linear(params, gsum, g, L.wih, L.bih, 3 * H, H, L.g2, L.be2, false);
    unfold_linear(params, gsum, g, L.whh, L.bhh, 3 * H, H, -1, -1, false);
    if (L.rnn_layers > 1) {             // layer 1 reads the raw hidden state of layer 0: no affine in front of either matrix
      unfold_linear(params, gsum, g, L.wih1, L.bih1, 3 * H, H, -1, -1, false);
      unfold_linear(params, gsum, g, L.whh1, L.bhh1, 3 * H, H, -1, -1, false);
    }
    hg = L.gr; hbe = L.ber;
  }
  for (int h = 0; h < L.n_heads; ++h) {
    unfold_linear(params, gsum, g, L.wh[h], L.bh[h], L.head_dim[h], H, hg, hbe, h > 0);
    __syncthreads();                                    // heads accumulate into the same affine slots
  }
  __syncthreads();
  // a head outside the optimiser (PopArt on cuda): its G has done its work in the chain rule above; take it out of the
  // norm, of grad_out and of the update
  auto in_head = [&](int e) {
    for (int h = 0; h < L.n_heads; ++h)
      if ((e >= L.wh[h] && e < L.wh[h] + L.head_dim[h] * H) || (e >= L.bh[h] && e < L.bh[h] + L.head_dim[h])) return true;
    return false;
  };
  if (ad.frozen_head) {
    for (int e = tid; e < L.total; e += kOptThreads) if (in_head(e)) g[e] = 0.0f;
    __syncthreads();
  }
  // 2. loss denominator, global L2 norm
  const double denom = lsum[denom_slot];
  const float inv = (float)(1.0 / denom);
  double sq = 0.0;
  for (int e = tid; e < L.total; e += kOptThreads) {
    const float x = g[e] * inv;
    g[e] = x;
    sq += (double)x * (double)x;
  }
  __shared__ double red[32];
  for (int o = 16; o > 0; o >>= 1) sq += __shfl_down_sync(0xffffffffu, sq, o);
  if ((tid & 31) == 0) red[tid >> 5] = sq;
  __syncthreads();
  if (tid < 32) {
    double t = red[tid];
    for (int o = 16; o > 0; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
    if (tid == 0) red[0] = t;
  }
  __syncthreads();
  const float norm = (float)sqrt(red[0]);
  float clip = 1.0f;
  if (ad.use_max_grad_norm) clip = fminf(ad.max_grad_norm / (norm + 1e-6f), 1.0f);   // clip_grad_norm_: clamp(coef, max=1)
  // 3. Adam (torch.optim.Adam, no amsgrad)
  if (apply_update) {
    float bc1 = ad.bias_correction1, bc2s = ad.bias_correction2_sqrt;
    if (step_dev) {                                     // device-resident step count (CUDA-graph replays): t = *step_dev + 1
      const double t = (double)(*step_dev + 1);
      bc1 = (float)(1.0 - pow(ad.beta1_exact, t));
      bc2s = (float)sqrt(1.0 - pow(ad.beta2_exact, t));
    }
    const float step_size = ad.lr / bc1;
    for (int e = tid; e < L.total; e += kOptThreads) {
      if (ad.frozen_head && in_head(e)) continue;
      float grad = g[e] * clip;
      const float p = params[e];
      if (ad.weight_decay != 0.0f) grad = fmaf(ad.weight_decay, p, grad);
      const float mm = m[e] + (grad - m[e]) * ad.one_minus_beta1;                   // exp_avg.lerp_(grad, 1 - beta1)
      const float vv = v[e] * ad.beta2 + ad.one_minus_beta2 * grad * grad;          // mul_(beta2).addcmul_(g, g, 1 - beta2)
      m[e] = mm; v[e] = vv;
      const float den = sqrtf(vv) / bc2s + ad.eps;
      params[e] = p - step_size * (mm / den);
    }
  }
  if (step_dev && apply_update) {
    __syncthreads();                                    // every thread has read the count
    if (tid == 0) *step_dev += 1;
  }
  // 4. train_info accumulation: {loss, grad_norm, entropy, ratio}
  if (info && tid == 0) {
    info[0] += (float)(lsum[0] / denom);
    info[1] += norm;
    if (!L.is_critic) {
      info[2] += (float)(lsum[1] / denom);
      info[3] += (float)(lsum[2] / ((double)rows_total * (double)L.n_heads));
    }
  }
}

}  // namespace mappo

using namespace mappo;

template <typename GT>
static int grad_reduce_any(const float* grad_partials, const double* loss_partials, int blocks, long long P, GT* grad_sum,
                           double* loss_sum, void* stream) {
  MAPPO_REQUIRE(grad_partials && loss_partials && grad_sum && loss_sum && blocks > 0 && P > 0, "grad_reduce: bad arguments");
  grad_reduce_kernel<GT><<<ceil_div(P, kRedCols), kRedSlices * kRedCols, 0, (cudaStream_t)stream>>>(
      grad_partials, loss_partials, blocks, P, grad_sum, loss_sum);
  MAPPO_CHECK_LAUNCH("grad_reduce_kernel");
  return 0;
}

extern "C" int mappo_grad_reduce(const float* grad_partials, const double* loss_partials, int blocks, long long P,
                                 float* grad_sum, double* loss_sum, void* stream) {
  return grad_reduce_any<float>(grad_partials, loss_partials, blocks, P, grad_sum, loss_sum, stream);
}

extern "C" int mappo_grad_reduce_f64(const float* grad_partials, const double* loss_partials, int blocks, long long P,
                                     double* grad_sum, double* loss_sum, void* stream) {
  return grad_reduce_any<double>(grad_partials, loss_partials, blocks, P, grad_sum, loss_sum, stream);
}

template <typename GT>
static int adam_step_any(const mappo_net_config* cfg, const mappo_adam_config* adam, float* params, float* exp_avg,
                         float* exp_avg_sq, const GT* grad_sum, const double* loss_sum, int denominator_slot,
                         long long rows_total, int apply_update, float* grad_out, float* info_out, long long* step_dev,
                         void* stream) {
  if (int rc = check_net(cfg, "adam_step")) return rc;
  MAPPO_REQUIRE(adam && params && grad_sum && loss_sum && grad_out, "adam_step: null pointer (grad_out is required)");
  MAPPO_REQUIRE(!apply_update || (exp_avg && exp_avg_sq), "adam_step: null optimiser state");
  MAPPO_REQUIRE(denominator_slot >= 0 && denominator_slot < 4 && rows_total > 0, "adam_step: bad denominator");
  const NetLayout L = make_layout(*cfg);
  adam_step_kernel<GT><<<1, kOptThreads, 0, (cudaStream_t)stream>>>(L, *adam, params, exp_avg, exp_avg_sq, grad_sum,
                                                                   loss_sum, denominator_slot, rows_total, apply_update,
                                                                   grad_out, info_out, step_dev);
  MAPPO_CHECK_LAUNCH("adam_step_kernel");
  return 0;
}

extern "C" int mappo_adam_step(const mappo_net_config* cfg, const mappo_adam_config* adam, float* params,
                               float* exp_avg, float* exp_avg_sq, const float* grad_sum, const double* loss_sum,
                               int denominator_slot, long long rows_total, int apply_update, float* grad_out,
                               float* info_out, long long* step_dev, void* stream) {
  return adam_step_any<float>(cfg, adam, params, exp_avg, exp_avg_sq, grad_sum, loss_sum, denominator_slot, rows_total,
                              apply_update, grad_out, info_out, step_dev, stream);
}

extern "C" int mappo_adam_step_f64(const mappo_net_config* cfg, const mappo_adam_config* adam, float* params,
                                   float* exp_avg, float* exp_avg_sq, const double* grad_sum, const double* loss_sum,
                                   int denominator_slot, long long rows_total, int apply_update, float* grad_out,
                                   float* info_out, long long* step_dev, void* stream) {
  return adam_step_any<double>(cfg, adam, params, exp_avg, exp_avg_sq, grad_sum, loss_sum, denominator_slot, rows_total,
                               apply_update, grad_out, info_out, step_dev, stream);
}
