// Row-level pieces shared by the two fused policy kernels (mlp.cu: fp32 FMA tiles; mlp_tc.cu: tcgen05 tensor-core
// tiles): the argument block, the categorical head (sample / mode / log-prob / entropy), the PPO clipped-surrogate
// loss and the clipped value loss with their derivatives.
// Reference: act.py:41-162, distributions.py:14-68, r_mappo.py:52-136, utils/util.py:23-29.
#pragma once
#include "net.cuh"
#include "philox.cuh"

namespace mappo {

constexpr int kLD = 68;            // tile row stride in floats: 64 + 4 (rows 16 B aligned, banks rotate by 4)
constexpr int kDL = 17;            // dlogits row stride
constexpr int kMaxA = 16;          // max total action logits
constexpr uint32_t kStreamSample = 0x53414D50u;   // 'SAMP' (oracle/policy.py STREAM_SAMPLE)
constexpr float kLnEps = 1e-5f;

struct MlpArgs {
  NetLayout L;
  const float* params;
  const float* x; int ld; long long rows; const int64_t* index; int x_div;
  int Kp;                          // in_dim rounded up to a multiple of 4
  long long ntiles;
  // rollout, actor
  const float* avail; int deterministic; uint32_t k0; uint32_t step; long long row_offset;
  const int64_t* step_dev;         // device-resident offset added to `step` (CUDA-graph replays), or NULL
  float* actions_out; float* logp_out; const float* given_actions; float* entropy_out;
  // rollout, critic
  float* values_out; int group;
  // update
  const float* actions; const float* old_logp; const float* adv; const float* active;
  const float* value_preds; const float* returns; const float* vn_stats;
  mappo_loss_config loss;
  float* grad_partials; double* loss_partials;
  // wide first layer (mlp_wide.cu): caller scratch; z1 = its [rows][64] part (activations in, gradient out)
  void* workspace; float* z1;
  const void* packed;              // mlp_wide.cu: the input rows as packed fp16-pair operand tiles (wide_pack_kernel), or NULL
  // recurrent networks (mlp_gru.cu): n2 rows out (base features) / dLoss/dn2 rows in (base backward), [rows][64]
  float* feat;
  int raw_ok;                      // mlp_tc.cu: input tiles are contiguous and arrive by bulk copy
  int rowp_ok;                     // mlp_tc.cu: so do the per-row loss scalars (actions, old log-probs, advantages, active masks)
};

// log-softmax over l[0..n): returns logsumexp, writes logp in place
__device__ __forceinline__ void log_softmax(float* l, int n) {
  float m = l[0];
  for (int q = 1; q < n; ++q) m = fmaxf(m, l[q]);
  float s = 0.0f;
  for (int q = 0; q < n; ++q) s += expf(l[q] - m);
  const float lse = m + logf(s);
  for (int q = 0; q < n; ++q) l[q] -= lse;
}

// Rollout: sample / mode + log-prob for each head (act.py:41-81, distributions.py:14-28)
__device__ __forceinline__ void actor_rollout_row(const MlpArgs& a, long long g, float* l) {
  const NetLayout& L = a.L;
  int off = 0;
  float ent_sum = 0.0f;
  for (int h = 0; h < L.n_heads; ++h) {
    const int n = L.head_dim[h];
    float* lh = l + off;
    if (a.avail && h == 0)
      for (int q = 0; q < n; ++q) if (a.avail[g * n + q] == 0.0f) lh[q] = -1e10f;     // distributions.py:66-67
    log_softmax(lh, n);
    int pick = 0;
    if (a.given_actions) {
      pick = (int)a.given_actions[g * L.n_heads + h];
      float ent = 0.0f;
      for (int q = 0; q < n; ++q) ent -= expf(lh[q]) * lh[q];
      ent_sum += ent;
    } else if (a.deterministic) {
      float best = lh[0];
      for (int q = 1; q < n; ++q) if (lh[q] > best) { best = lh[q]; pick = q; }
    } else {
      const uint32_t stp = a.step + (a.step_dev ? (uint32_t)__ldg(a.step_dev) : 0u);
      const u32x4 rnd = philox4x32_10((uint32_t)(a.row_offset + g), stp, (uint32_t)h, 0u, a.k0, kStreamSample);
      const float u = u32_to_unit_f32(rnd.x);
      float cum = 0.0f;
      pick = n - 1;
      for (int q = 0; q < n; ++q) { cum += expf(lh[q]); if (u < cum) { pick = q; break; } }
    }
    if (a.actions_out) a.actions_out[g * L.n_heads + h] = (float)pick;
    a.logp_out[g * L.n_heads + h] = lh[pick];
    off += n;
  }
  if (a.entropy_out) a.entropy_out[g] = ent_sum / (float)L.n_heads;
}

// Update, actor: PPO clipped surrogate + entropy for one row; writes dlogits (unnormalised) and loss sums.
// r_mappo.py:115-141, act.py:107-162.
__device__ __forceinline__ void actor_loss_row(const MlpArgs& a, long long src, float* l, float* __restrict__ dl,
                                               double* acc4) {
  const NetLayout& L = a.L;
  const float active = (a.loss.use_policy_active_masks && a.active) ? a.active[src] : 1.0f;
  const float adv = a.adv[src];
  const float lo = 1.0f - a.loss.clip_param, hi = 1.0f + a.loss.clip_param;
  const float ecoef = a.loss.entropy_coef / (float)L.n_heads;
  int off = 0;
  float obj_sum = 0.0f, ent_sum = 0.0f, ratio_sum = 0.0f;
  for (int h = 0; h < L.n_heads; ++h) {
    const int n = L.head_dim[h];
    float* lh = l + off;
    if (a.avail && h == 0)
      for (int q = 0; q < n; ++q) if (a.avail[src * n + q] == 0.0f) lh[q] = -1e10f;
    log_softmax(lh, n);
    const int act = (int)a.actions[src * L.n_heads + h];
    const float ratio = expf(lh[act] - a.old_logp[src * L.n_heads + h]);
    const float s1 = ratio * adv, s2 = fminf(fmaxf(ratio, lo), hi) * adv;
    obj_sum += fminf(s1, s2);
    ratio_sum += ratio;
    float ent = 0.0f;
    for (int q = 0; q < n; ++q) ent -= expf(lh[q]) * lh[q];
    ent_sum += ent;
    // d(-obj)/dlogp_act: torch.min ties split evenly and clamp passes the gradient on its closed range
    const bool inside = (ratio >= lo) && (ratio <= hi);
    const float gcoef = (inside || s1 < s2) ? ratio * adv : 0.0f;
    for (int q = 0; q < n; ++q) {
      const float p = expf(lh[q]);
      const float d_obj = -gcoef * ((q == act ? 1.0f : 0.0f) - p);
      const float d_ent = ecoef * p * (lh[q] + ent);        // d(-c H)/dl = c p (logp + H)
      dl[off + q] = active * (d_obj + d_ent);
    }
    off += n;
  }
  for (int q = off; q < kMaxA; ++q) dl[q] = 0.0f;
  acc4[0] += (double)(-obj_sum * active);
  acc4[1] += (double)(ent_sum / (float)L.n_heads * active);
  acc4[2] += (double)ratio_sum;
  acc4[3] += (double)active;
}

__device__ __forceinline__ float huber(float e, float d) { return fabsf(e) <= d ? 0.5f * e * e : d * (fabsf(e) - 0.5f * d); }
__device__ __forceinline__ float huber_d(float e, float d) { return fabsf(e) <= d ? e : (e > 0.0f ? d : -d); }

// Update, critic: value loss of the `group` terms that share this row's value (r_mappo.py:52-89).
__device__ __forceinline__ float critic_loss_row(const MlpArgs& a, long long src, float v, double* acc4) {
  const mappo_loss_config& c = a.loss;
  float mean = 0.0f, sd = 1.0f;
  if (c.use_value_norm) {                                   // ValueNorm.normalize, valuenorm.py:31-35,56-65
    const float db = fmaxf(a.vn_stats[2], 1e-5f);
    mean = a.vn_stats[0] / db;
    sd = sqrtf(fmaxf(a.vn_stats[1] / db - mean * mean, 1e-2f));
  }
  float dv = 0.0f;
  for (int m = 0; m < a.group; ++m) {
    const long long o = src * a.group + m;
    const float active = (c.use_value_active_masks && a.active) ? a.active[o] : 1.0f;
    const float vp = a.value_preds[o];
    const float target = c.use_value_norm ? (a.returns[o] - mean) / sd : a.returns[o];
    const float diff = v - vp;
    const float vclip = vp + fminf(fmaxf(diff, -c.clip_param), c.clip_param);
    const float e_c = target - vclip, e_o = target - v;
    const float l_c = c.use_huber_loss ? huber(e_c, c.huber_delta) : 0.5f * e_c * e_c;
    const float l_o = c.use_huber_loss ? huber(e_o, c.huber_delta) : 0.5f * e_o * e_o;
    const float g_o = -(c.use_huber_loss ? huber_d(e_o, c.huber_delta) : e_o);
    const bool in_clip = (diff >= -c.clip_param) && (diff <= c.clip_param);
    const float g_c = in_clip ? -(c.use_huber_loss ? huber_d(e_c, c.huber_delta) : e_c) : 0.0f;
    float loss, g;
    if (!c.use_clipped_value_loss) { loss = l_o; g = g_o; }
    else if (l_o > l_c) { loss = l_o; g = g_o; }
    else if (l_o < l_c) { loss = l_c; g = g_c; }
    else { loss = l_o; g = 0.5f * (g_o + g_c); }            // torch.max ties split evenly
    dv += active * g * c.value_loss_coef;
    acc4[0] += (double)(loss * active);
    acc4[1] += (double)active;
  }
  return dv;
}

// The same with the ValueNorm statistics taken once per kernel (mean, 1 / sd handed in: the target is
// (return - mean) * inv_sd, one rounding away from the division above) and the loss sums only where they are kept
// (`sums`: warp-uniform in the tensor-core kernels, where four threads share a row but one of them owns the sums).
// ret / vpred / act: the three per-term arrays (global, or a tile's copy in shared memory), o0: this row's first term.
__device__ __forceinline__ float critic_loss_row_pre(const MlpArgs& a, const float* ret, const float* vpred, const float* act,
                                                     long long o0, float v, float mean, float inv_sd, bool sums, double* acc4) {
  const mappo_loss_config& c = a.loss;
  float dv = 0.0f;
  for (int m = 0; m < a.group; ++m) {
    const long long o = o0 + m;
    const float active = act ? act[o] : 1.0f;
    const float vp = vpred[o];
    const float target = c.use_value_norm ? (ret[o] - mean) * inv_sd : ret[o];
    const float diff = v - vp;
    const float vclip = vp + fminf(fmaxf(diff, -c.clip_param), c.clip_param);
    const float e_c = target - vclip, e_o = target - v;
    const float l_c = c.use_huber_loss ? huber(e_c, c.huber_delta) : 0.5f * e_c * e_c;
    const float l_o = c.use_huber_loss ? huber(e_o, c.huber_delta) : 0.5f * e_o * e_o;
    const float g_o = -(c.use_huber_loss ? huber_d(e_o, c.huber_delta) : e_o);
    const bool in_clip = (diff >= -c.clip_param) && (diff <= c.clip_param);
    const float g_c = in_clip ? -(c.use_huber_loss ? huber_d(e_c, c.huber_delta) : e_c) : 0.0f;
    float loss, g;
    if (!c.use_clipped_value_loss) { loss = l_o; g = g_o; }
    else if (l_o > l_c) { loss = l_o; g = g_o; }
    else if (l_o < l_c) { loss = l_c; g = g_c; }
    else { loss = l_o; g = 0.5f * (g_o + g_c); }            // torch.max ties split evenly
    dv += active * g * c.value_loss_coef;
    if (sums) { acc4[0] += (double)(loss * active); acc4[1] += (double)active; }
  }
  return dv;
}

}  // namespace mappo
