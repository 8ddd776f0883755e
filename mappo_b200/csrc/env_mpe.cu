// MPE env step / reset kernels: one thread per (env, agent), float64 physics like the reference.
// Compiled with -fmad=false: the reference evaluates every product and sum with its own rounding
// (numpy f64), and the oracle (oracle/mpe.py) is matched to ~1 ulp of libm.
//
// Reference functions restated here (all fused into one launch per env step):
//   _set_action                   onpolicy/envs/mpe/environment.py:194-256
//   apply_action_force            onpolicy/envs/mpe/core.py:229-238
//   apply_environment_force       core.py:241-263  + get_entity_collision_force core.py:290-321
//   integrate_state               core.py:265-278
//   update_agent_state            core.py:280-287
//   reward / observation          scenarios/simple_spread.py:72-103, simple_speaker_listener.py:69-98,
//                                 simple_reference.py:62-97; collaborative sum environment.py:141-143
//   _get_done + auto-reset        environment.py:179-185, envs/env_wrappers.py:145-154
//   reset_world                   simple_spread.py:32-45, simple_speaker_listener.py:38-63, simple_reference.py:34-60
// plus simple_adversary.py:9-137, simple_push.py:8-104, simple_tag.py:7-144: per-agent size /
// accel / max_speed (core.py:229-238,265-278, environment.py:243-246), agent-landmark contacts (core.py:290-321),
// individual (non-collaborative) rewards.
#include "common.cuh"
#include "philox.cuh"

namespace mappo {

constexpr double kDt = 0.1, kDamp = 0.25, kContactForce = 1e2, kContactMargin = 1e-3, kSensitivity = 5.0;
constexpr uint32_t kStreamReset = 0x52455345u;  // must match oracle/mpe.py STREAM_RESET

__constant__ double c_colors_sl[9] = {0.65, 0.15, 0.15, 0.15, 0.65, 0.15, 0.15, 0.15, 0.65};
__constant__ double c_colors_ref[9] = {0.75, 0.25, 0.25, 0.25, 0.75, 0.25, 0.25, 0.25, 0.75};

struct ScnInfo {
  int dim_c, n_goals, collide;
  int shared_goal;          // one goal landmark, the same for every agent (agent.goal_a = goal)
  int collaborative;        // world.collaborative: every agent receives the sum (environment.py:141-143)
  int landmark_collide;     // landmarks are obstacles (simple_tag.py:32-34)
  double agent_size, landmark_scale, landmark_size;
};

__host__ __device__ inline ScnInfo scn_info(int scn) {
  ScnInfo s;
  s.shared_goal = 0; s.collaborative = 1; s.landmark_collide = 0; s.landmark_size = 0.05;
  if (scn == MAPPO_SCN_SIMPLE_SPREAD) { s.dim_c = 2; s.n_goals = 0; s.collide = 1; s.agent_size = 0.15; s.landmark_scale = 0.8; }
  else if (scn == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) { s.dim_c = 3; s.n_goals = 1; s.collide = 0; s.agent_size = 0.075; s.landmark_scale = 1.0; }
  else if (scn == MAPPO_SCN_SIMPLE_REFERENCE) { s.dim_c = 10; s.n_goals = 2; s.collide = 0; s.agent_size = 0.05; s.landmark_scale = 0.8; }
  else if (scn == MAPPO_SCN_SIMPLE_ADVERSARY) {
    s.dim_c = 2; s.n_goals = 1; s.collide = 0; s.agent_size = 0.15; s.landmark_scale = 1.0; s.shared_goal = 1; s.collaborative = 0;
  } else if (scn == MAPPO_SCN_SIMPLE_PUSH) {
    s.dim_c = 2; s.n_goals = 1; s.collide = 1; s.agent_size = 0.05; s.landmark_scale = 0.8; s.shared_goal = 1; s.collaborative = 0;
  } else if (scn == MAPPO_SCN_SIMPLE_WORLD_COMM) {  // simple_world_comm.py:7-53: tag physics + food, forests, a speaking leader
    s.dim_c = 4; s.n_goals = 0; s.collide = 1; s.agent_size = 0.045; s.landmark_scale = 0.8; s.collaborative = 0;
    s.landmark_collide = 1; s.landmark_size = 0.2;
  } else if (scn == MAPPO_SCN_SIMPLE_CRYPTO) {  // simple_crypto.py:21-47: nobody moves, everybody speaks; goal + key landmark
    s.dim_c = 4; s.n_goals = 2; s.collide = 0; s.agent_size = 0.05; s.landmark_scale = 1.0; s.collaborative = 0;
  } else {  // simple_tag: sizes / accel / max_speed depend on the agent's side, see agent_* below
    s.dim_c = 2; s.n_goals = 0; s.collide = 1; s.agent_size = 0.05; s.landmark_scale = 0.8; s.collaborative = 0;
    s.landmark_collide = 1; s.landmark_size = 0.2;
  }
  return s;
}
// adversaries come first (agent i is one iff i < nadv); simple_adversary / simple_push have exactly one
__host__ __device__ inline int n_adversaries(int scn, int cfg_nadv) {
  if (scn == MAPPO_SCN_SIMPLE_ADVERSARY || scn == MAPPO_SCN_SIMPLE_PUSH || scn == MAPPO_SCN_SIMPLE_CRYPTO) return 1;
  if (scn == MAPPO_SCN_SIMPLE_TAG) return cfg_nadv > 0 ? cfg_nadv : 3;              // simple_tag.py:12 default
  if (scn == MAPPO_SCN_SIMPLE_WORLD_COMM) return cfg_nadv > 0 ? cfg_nadv : 4;       // simple_world_comm.py:12 default
  return 0;
}
// landmarks behind the `num_landmarks` obstacles the flag counts: simple_world_comm appends 2 food and 2 forests
// (simple_world_comm.py:27-50); they are observed like landmarks but do not collide
__host__ __device__ inline int n_extra_landmarks(int scn) { return scn == MAPPO_SCN_SIMPLE_WORLD_COMM ? 4 : 0; }
constexpr double kFoodSize = 0.03, kForestSize = 0.3;
__host__ __device__ inline double agent_size_of(int scn, int i, int nadv) {
  if (scn == MAPPO_SCN_SIMPLE_TAG) return i < nadv ? 0.075 : 0.05;                   // simple_tag.py:22
  if (scn == MAPPO_SCN_SIMPLE_WORLD_COMM) return i < nadv ? 0.075 : 0.045;           // simple_world_comm.py:23
  return scn_info(scn).agent_size;
}
__host__ __device__ inline double agent_accel_of(int scn, int i, int nadv) {         // 0: accel None
  return (scn == MAPPO_SCN_SIMPLE_TAG || scn == MAPPO_SCN_SIMPLE_WORLD_COMM) ? (i < nadv ? 3.0 : 4.0) : 0.0;   // simple_tag.py:23
}
__host__ __device__ inline double agent_max_speed_of(int scn, int i, int nadv) {     // 0: max_speed None
  return (scn == MAPPO_SCN_SIMPLE_TAG || scn == MAPPO_SCN_SIMPLE_WORLD_COMM) ? (i < nadv ? 1.0 : 1.3) : 0.0;   // simple_tag.py:24
}

__host__ __device__ inline int obs_dim_of(int scn, int M, int L, int agent, int nadv = 0) {
  nadv = n_adversaries(scn, nadv);
  if (scn == MAPPO_SCN_SIMPLE_SPREAD) return 4 + 2 * L + 4 * (M - 1);
  if (scn == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) return agent == 0 ? 3 : 2 + 2 * L + 3;
  if (scn == MAPPO_SCN_SIMPLE_REFERENCE) return 2 + 2 * L + 3 + 10;
  if (scn == MAPPO_SCN_SIMPLE_ADVERSARY) return (agent < nadv ? 0 : 2) + 2 * L + 2 * (M - 1);
  if (scn == MAPPO_SCN_SIMPLE_PUSH) return agent < nadv ? 2 + 2 * L + 2 * (M - 1) : 2 + 2 + 3 + 2 * L + 3 * L + 2 * (M - 1);
  if (scn == MAPPO_SCN_SIMPLE_CRYPTO) return agent == 0 ? 4 : 8;                      // simple_crypto.py:160-171
  if (scn == MAPPO_SCN_SIMPLE_WORLD_COMM) {                                           // simple_world_comm.py:281-288
    const int Lt = L + 4, good = M - nadv;
    return agent < nadv ? 4 + 2 * Lt + 2 * (M - 1) + 2 * good + 2 + 4 : 4 + 2 * Lt + 2 * (M - 1) + 2 + 2 * (good - 1);
  }
  const int good_others = (M - nadv) - (agent < nadv ? 0 : 1);
  return 4 + 2 * L + 2 * (M - 1) + 2 * good_others;
}
__host__ __device__ inline int obs_off_of(int scn, int M, int L, int agent, int nadv = 0) {
  int o = 0;
  for (int i = 0; i < agent; ++i) o += obs_dim_of(scn, M, L, i, nadv);
  return o;
}
__host__ __device__ inline bool is_movable(int scn, int agent) {
  if (scn == MAPPO_SCN_SIMPLE_CRYPTO) return false;                                   // simple_crypto.py:33
  return !(scn == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER && agent == 0);
}
__host__ __device__ inline bool is_silent(int scn, int agent) {
  if (scn == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) return agent == 1;
  if (scn == MAPPO_SCN_SIMPLE_WORLD_COMM) return agent != 0;                          // only the leader speaks (:20)
  return !(scn == MAPPO_SCN_SIMPLE_REFERENCE || scn == MAPPO_SCN_SIMPLE_CRYPTO);      // every other scenario: silent = True
}

// np.logaddexp(0, y)   (numpy npy_logaddexp with x = 0)
__device__ __forceinline__ double logaddexp0(double y) {
  if (y == 0.0) return 0.6931471805599453094;  // x == y: x + LOGE2
  const double tmp = 0.0 - y;
  if (tmp > 0) return 0.0 + log1p(exp(-tmp));
  if (tmp <= 0) return y + log1p(exp(tmp));
  return tmp;  // NaN
}

// np.sum over M per-agent rewards with numpy's pairwise unrolling (see oracle/mpe.py np_sum_small)
__device__ __forceinline__ double np_sum_small(const double* v, int stride, int n) {
  if (n < 8) {
    double r = v[0];
    for (int i = 1; i < n; ++i) r = r + v[i * stride];
    return r;
  }
  double r[8];
  for (int j = 0; j < 8; ++j) r[j] = v[j * stride];
  int i = 8;
  for (; i < n - (n % 8); i += 8)
    for (int j = 0; j < 8; ++j) r[j] = r[j] + v[(i + j) * stride];
  double res = ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
  for (; i < n; ++i) res = res + v[i * stride];
  return res;
}

// Shared-memory carve-up for one block of EPB envs (all agent-major so a warp has consecutive envs).
struct EnvSmem {
  double* pos;    // [2][M][EPB]   agent positions (old, then new, then reset)
  double* vel;    // [2][M][EPB]   agent velocities after the step / reset (simple_tag observes the others')
  double* lm;     // [2][L][EPB]
  double* aux;    // [L*M + M][EPB]  landmark distances, then individual rewards
  int* comm;      // [M][EPB]
  int* goal;      // [M][EPB]
  float* tile;    // [EPB][W + pad]  obs rows staged for a coalesced store
};

__device__ __forceinline__ EnvSmem carve(unsigned char* base, int M, int L, int EPB, int tile_ld) {
  EnvSmem s;
  s.pos = reinterpret_cast<double*>(base);
  s.vel = s.pos + 2 * M * EPB;
  s.lm = s.vel + 2 * M * EPB;
  s.aux = s.lm + 2 * L * EPB;
  const int aux_rows = L * M + M;
  s.comm = reinterpret_cast<int*>(s.aux + aux_rows * EPB);
  s.goal = s.comm + M * EPB;
  s.tile = reinterpret_cast<float*>(s.goal + M * EPB);
  (void)tile_ld;
  return s;
}

static inline int tile_ld_of(int W) { return (W % 2 == 0) ? W + 1 : W; }   // odd row stride: conflict-free columns
static size_t env_smem_bytes(int M, int L, int EPB, int W) {
  const int aux_rows = L * M + M;
  return sizeof(double) * (size_t)EPB * (4 * M + 2 * L + aux_rows) + sizeof(int) * (size_t)EPB * 2 * M +
         sizeof(float) * (size_t)EPB * tile_ld_of(W) + sizeof(int) * (size_t)EPB;
}

// Write agent i's observation for env slot e into the staged tile (f64 -> f32 round-to-nearest, as
// the buffer's float32 assignment does, utils/shared_buffer.py:96-97).
template <int SCN>
__device__ __forceinline__ void write_obs(const EnvSmem& s, int M, int L, int EPB, int ld, int e, int i,
                                          double vx, double vy, int nadv) {
  float* o = s.tile + (size_t)e * ld + obs_off_of(SCN, M, L - n_extra_landmarks(SCN), i, nadv);
  const double px = s.pos[(0 * M + i) * EPB + e], py = s.pos[(1 * M + i) * EPB + e];
  int k = 0;
  if (SCN == MAPPO_SCN_SIMPLE_SPREAD) {
    o[k++] = (float)vx; o[k++] = (float)vy; o[k++] = (float)px; o[k++] = (float)py;
    for (int l = 0; l < L; ++l) {
      o[k++] = (float)(s.lm[(0 * L + l) * EPB + e] - px);
      o[k++] = (float)(s.lm[(1 * L + l) * EPB + e] - py);
    }
    for (int j = 0; j < M; ++j) {
      if (j == i) continue;
      o[k++] = (float)(s.pos[(0 * M + j) * EPB + e] - px);
      o[k++] = (float)(s.pos[(1 * M + j) * EPB + e] - py);
    }
    for (int j = 0; j < 2 * (M - 1); ++j) o[k++] = 0.0f;  // other.state.c of silent agents
  } else if (SCN == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) {
    if (i == 0) {
      const int g = s.goal[0 * EPB + e];
      for (int c = 0; c < 3; ++c) o[k++] = (float)c_colors_sl[g * 3 + c];
    } else {
      o[k++] = (float)vx; o[k++] = (float)vy;
      for (int l = 0; l < L; ++l) {
        o[k++] = (float)(s.lm[(0 * L + l) * EPB + e] - px);
        o[k++] = (float)(s.lm[(1 * L + l) * EPB + e] - py);
      }
      const int cid = s.comm[0 * EPB + e];
      for (int c = 0; c < 3; ++c) o[k++] = (c == cid) ? 1.0f : 0.0f;
    }
  } else if (SCN == MAPPO_SCN_SIMPLE_REFERENCE) {
    o[k++] = (float)vx; o[k++] = (float)vy;
    for (int l = 0; l < L; ++l) {
      o[k++] = (float)(s.lm[(0 * L + l) * EPB + e] - px);
      o[k++] = (float)(s.lm[(1 * L + l) * EPB + e] - py);
    }
    const int g = s.goal[i * EPB + e];
    for (int c = 0; c < 3; ++c) o[k++] = (float)c_colors_ref[g * 3 + c];
    const int cid = s.comm[(1 - i) * EPB + e];
    for (int c = 0; c < 10; ++c) o[k++] = (c == cid) ? 1.0f : 0.0f;
  } else if (SCN == MAPPO_SCN_SIMPLE_WORLD_COMM) {
    // simple_world_comm.py:214-288.  The last two landmarks are the forests: an agent is hidden in one when it overlaps it
    // (is_collision); another agent is visible when both share a forest, when neither is in any forest, or to the leader.
    const int f1 = L - 2, f2 = L - 1;
    auto in_f = [&](int ag, int f) {
      const double dx = s.pos[(0 * M + ag) * EPB + e] - s.lm[(0 * L + f) * EPB + e];
      const double dy = s.pos[(1 * M + ag) * EPB + e] - s.lm[(1 * L + f) * EPB + e];
      return sqrt(dx * dx + dy * dy) < agent_size_of(SCN, ag, nadv) + kForestSize;
    };
    const bool inf1 = in_f(i, f1), inf2 = in_f(i, f2), adv = i < nadv;
    o[k++] = (float)vx; o[k++] = (float)vy; o[k++] = (float)px; o[k++] = (float)py;
    for (int l = 0; l < L; ++l) {
      o[k++] = (float)(s.lm[(0 * L + l) * EPB + e] - px);
      o[k++] = (float)(s.lm[(1 * L + l) * EPB + e] - py);
    }
    unsigned vis = 0;
    for (int j = 0; j < M; ++j) {
      if (j == i) continue;
      const bool o1 = in_f(j, f1), o2 = in_f(j, f2);
      const bool v = (inf1 && o1) || (inf2 && o2) || (!inf1 && !o1 && !inf2 && !o2) || i == 0;
      if (v) vis |= 1u << j;
      o[k++] = v ? (float)(s.pos[(0 * M + j) * EPB + e] - px) : 0.0f;
      o[k++] = v ? (float)(s.pos[(1 * M + j) * EPB + e] - py) : 0.0f;
    }
    if (!adv) { o[k++] = inf1 ? 1.0f : -1.0f; o[k++] = inf2 ? 1.0f : -1.0f; }          // good agents: in_forest first
    for (int j = nadv; j < M; ++j) {                                                   // velocities of the other GOOD agents
      if (j == i) continue;
      const bool v = (vis >> j) & 1u;
      o[k++] = v ? (float)s.vel[(0 * M + j) * EPB + e] : 0.0f;
      o[k++] = v ? (float)s.vel[(1 * M + j) * EPB + e] : 0.0f;
    }
    if (adv) {
      o[k++] = inf1 ? 1.0f : -1.0f; o[k++] = inf2 ? 1.0f : -1.0f;
      const int said = s.comm[0 * EPB + e];                                            // the leader's word
      for (int c = 0; c < 4; ++c) o[k++] = (c == said) ? 1.0f : 0.0f;
    }
  } else if (SCN == MAPPO_SCN_SIMPLE_CRYPTO) {
    // simple_crypto.py:127-171: landmark l's colour is the unit vector e_l of the 4 colour channels (:54-58);
    // goal slot 0 = the goal landmark, slot 1 = the landmark whose colour is the speaker's key
    const int g = s.goal[0 * EPB + e], key = s.goal[1 * EPB + e], said = s.comm[2 * EPB + e];
    if (i == 2) for (int c = 0; c < 4; ++c) o[k++] = (c == g) ? 1.0f : 0.0f;           // speaker: goal colour, key
    if (i >= 1) for (int c = 0; c < 4; ++c) o[k++] = (c == key) ? 1.0f : 0.0f;         // listener: key, utterance
    if (i <= 1) for (int c = 0; c < 4; ++c) o[k++] = (c == said) ? 1.0f : 0.0f;        // adversary: utterance only
  } else {
    // simple_adversary.py:116-137, simple_push.py:81-104, simple_tag.py:125-144
    const bool adv = i < nadv;
    const int g = s.goal[i * EPB + e];
    if (SCN != MAPPO_SCN_SIMPLE_ADVERSARY) { o[k++] = (float)vx; o[k++] = (float)vy; }
    if (SCN == MAPPO_SCN_SIMPLE_TAG) { o[k++] = (float)px; o[k++] = (float)py; }
    if (SCN != MAPPO_SCN_SIMPLE_TAG && !adv) {
      o[k++] = (float)(s.lm[(0 * L + g) * EPB + e] - px);
      o[k++] = (float)(s.lm[(1 * L + g) * EPB + e] - py);
      if (SCN == MAPPO_SCN_SIMPLE_PUSH)                                  // agent.color: 0.25, goal channel + 0.5
        for (int c = 0; c < 3; ++c) o[k++] = (float)(c == g + 1 ? 0.25 + 0.5 : 0.25);
    }
    for (int l = 0; l < L; ++l) {
      o[k++] = (float)(s.lm[(0 * L + l) * EPB + e] - px);
      o[k++] = (float)(s.lm[(1 * L + l) * EPB + e] - py);
    }
    if (SCN == MAPPO_SCN_SIMPLE_PUSH && !adv)
      for (int l = 0; l < L; ++l)                                        // landmark.color: 0.1, channel l + 1 + 0.8
        for (int c = 0; c < 3; ++c) o[k++] = (float)(c == l + 1 ? 0.1 + 0.8 : 0.1);
    for (int j = 0; j < M; ++j) {
      if (j == i) continue;
      o[k++] = (float)(s.pos[(0 * M + j) * EPB + e] - px);
      o[k++] = (float)(s.pos[(1 * M + j) * EPB + e] - py);
    }
    if (SCN == MAPPO_SCN_SIMPLE_TAG)
      for (int j = nadv; j < M; ++j) {                                   // velocities of the other GOOD agents
        if (j == i) continue;
        o[k++] = (float)s.vel[(0 * M + j) * EPB + e];
        o[k++] = (float)s.vel[(1 * M + j) * EPB + e];
      }
  }
}

// Coalesced copy of the staged obs tile to obs_out[n0 .. n0+cnt) (rows are contiguous in global memory).
__device__ __forceinline__ void store_tile(const float* tile, int ld, int W, int cnt, float* __restrict__ dst) {
  const int total = cnt * W;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int r = idx / W, c = idx - r * W;
    dst[idx] = tile[(size_t)r * ld + c];
  }
}

// Philox reset draws for env n: draw order [goal ids][agent x,y ...][landmark x,y ...]
template <int SCN>
__device__ __forceinline__ void draw_agent(uint32_t env_lo, uint32_t ep, uint32_t k0, int i, double& x, double& y) {
  const int base = scn_info(SCN).n_goals + 2 * i;
  x = u32_to_sym(philox_draw(env_lo, ep, base, k0, kStreamReset));
  y = u32_to_sym(philox_draw(env_lo, ep, base + 1, k0, kStreamReset));
}
template <int SCN>
__device__ __forceinline__ void draw_landmark(uint32_t env_lo, uint32_t ep, uint32_t k0, int M, int l, double& x, double& y) {
  const int base = scn_info(SCN).n_goals + 2 * M + 2 * l;
  x = u32_to_sym(philox_draw(env_lo, ep, base, k0, kStreamReset));
  y = u32_to_sym(philox_draw(env_lo, ep, base + 1, k0, kStreamReset));
  if (scn_info(SCN).landmark_scale != 1.0) { x = x * 0.8; y = y * 0.8; }
}
__device__ __forceinline__ int draw_goal(uint32_t env_lo, uint32_t ep, uint32_t k0, int g, int L) {
  const uint32_t d = philox_draw(env_lo, ep, g, k0, kStreamReset);
  return (int)(((uint64_t)d * (uint64_t)L) >> 32);
}

struct StepArgs {
  int N, M, L, T, EPB, W, ld, nadv;
  int Lc;               // colliding landmarks: the first Lc of the L (simple_world_comm: food and forests follow)
  uint32_t k0;
  long long env_id_offset;
  mappo_env_state st;
  mappo_env_state rs;   // reset source (valid if has_rs)
  int has_rs;
  const float* actions;
  int act_w;
  float* obs_out;
  float* rewards_out;
  float* masks_out;
  uint8_t* dones_out;
  float* indiv_out;
  int mode;             // 0 = step, 1 = reset (draw), 2 = reset from the state already in st
};

template <int SCN>
__global__ void __launch_bounds__(512) env_kernel(const StepArgs a) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int N = a.N, M = a.M, L = a.L, EPB = a.EPB, ld = a.ld;
  EnvSmem s = carve(smem_raw, M, L, EPB, ld);
  double* ind = s.aux + (size_t)L * M * EPB;                       // [M][EPB] individual rewards
  int* doneflag = reinterpret_cast<int*>(s.tile + (size_t)EPB * ld);  // [EPB]
  const ScnInfo si = scn_info(SCN);
  const int tid = threadIdx.x;
  const int i = tid / EPB, e = tid - i * EPB;
  const int n0 = blockIdx.x * EPB;
  const int n = n0 + e;
  const bool valid = (n < N);
  const int cnt = (N - n0 < EPB) ? (N - n0) : EPB;
  const bool stepping = (a.mode == 0);

  // ---- phase 1: load state into shared memory
  double px = 0, py = 0, vx = 0, vy = 0;
  int step = 0;
  if (valid) {
    px = a.st.p_pos[(size_t)(0 * M + i) * N + n]; py = a.st.p_pos[(size_t)(1 * M + i) * N + n];
    vx = a.st.p_vel[(size_t)(0 * M + i) * N + n]; vy = a.st.p_vel[(size_t)(1 * M + i) * N + n];
    s.pos[(0 * M + i) * EPB + e] = px; s.pos[(1 * M + i) * EPB + e] = py;
    s.vel[(0 * M + i) * EPB + e] = vx; s.vel[(1 * M + i) * EPB + e] = vy;
    s.goal[i * EPB + e] = a.st.goal[(size_t)i * N + n];
    s.comm[i * EPB + e] = a.st.comm[(size_t)i * N + n];
    step = a.st.step_count[n] + (stepping ? 1 : 0);                   // environment.py:118
    if (i == 0) doneflag[e] = stepping ? (step >= a.T) : (a.mode == 1);
  }
  for (int idx = tid; idx < EPB * L; idx += blockDim.x) {
    const int l = idx / EPB, e2 = idx - l * EPB;
    if (n0 + e2 < N) {
      s.lm[(0 * L + l) * EPB + e2] = a.st.lm_pos[(size_t)(0 * L + l) * N + n0 + e2];
      s.lm[(1 * L + l) * EPB + e2] = a.st.lm_pos[(size_t)(1 * L + l) * N + n0 + e2];
    }
  }
  __syncthreads();

  double total = 0.0, indiv = 0.0;
  int new_comm = -1;
  if (stepping) {
    // ---- phase 2: action -> force -> integrate (old positions of the other agents from smem)
    const bool movable = is_movable(SCN, i), silent = is_silent(SCN, i);
    if (valid) {
      const float* arow = a.actions + ((size_t)n * M + i) * a.act_w;
      double fx = 0.0, fy = 0.0;
      int head = 0;
      if (movable) {
        const int act = (int)arow[head++];
        double ux = 0.0, uy = 0.0;
        ux += (act == 1 ? 1.0 : 0.0) - (act == 2 ? 1.0 : 0.0);          // environment.py:224-225
        uy += (act == 3 ? 1.0 : 0.0) - (act == 4 ? 1.0 : 0.0);
        const double accel = agent_accel_of(SCN, i, a.nadv);
        if (accel != 0.0) {
          ux *= accel; uy *= accel;                                      // sensitivity = agent.accel
          fx = (1.0 * accel) * ux + 0.0; fy = (1.0 * accel) * uy + 0.0;  // (mass * accel) * u + noise
        } else {
          ux *= kSensitivity; uy *= kSensitivity;                        // environment.py:243-246
          fx = 1.0 * ux + 0.0; fy = 1.0 * uy + 0.0;                      // core.py:237-238 (mass 1, no noise)
        }
      }
      if (!silent) new_comm = (int)arow[head];
      if (si.collide) {
        const double my_size = agent_size_of(SCN, i, a.nadv);
        for (int j = 0; j < M; ++j) {                                    // core.py:241-263, pair order = j ascending
          if (j == i) continue;
          const double dist_min = (i < j) ? my_size + agent_size_of(SCN, j, a.nadv) : agent_size_of(SCN, j, a.nadv) + my_size;
          const double qx = s.pos[(0 * M + j) * EPB + e], qy = s.pos[(1 * M + j) * EPB + e];
          // the reference evaluates the pair once as (a,b) with a<b: delta = p_a - p_b
          const double dx = (i < j) ? (px - qx) : (qx - px), dy = (i < j) ? (py - qy) : (qy - py);
          const double dist = sqrt(dx * dx + dy * dy);
          const double pen = logaddexp0(-(dist - dist_min) / kContactMargin) * kContactMargin;
          double gx = kContactForce * dx / dist * pen, gy = kContactForce * dy / dist * pen;
          if (i > j) { gx = -gx; gy = -gy; }                             // force_b = -(1/ratio) * force
          fx = gx + fx; fy = gy + fy;
        }
        if (si.landmark_collide)
          for (int l = 0; l < a.Lc; ++l) {                               // (agent, landmark) pairs follow the agent pairs
            const double dist_min = my_size + si.landmark_size;
            const double dx = px - s.lm[(0 * L + l) * EPB + e], dy = py - s.lm[(1 * L + l) * EPB + e];
            const double dist = sqrt(dx * dx + dy * dy);
            const double pen = logaddexp0(-(dist - dist_min) / kContactMargin) * kContactMargin;
            const double gx = kContactForce * dx / dist * pen, gy = kContactForce * dy / dist * pen;
            fx = gx + fx; fy = gy + fy;                                  // core.py:318-320: force_a = +force
          }
      }
      if (movable) {                                                     // core.py:265-278
        vx = vx * (1 - kDamp); vy = vy * (1 - kDamp);
        vx += (fx / 1.0) * kDt; vy += (fy / 1.0) * kDt;
        const double vmax = agent_max_speed_of(SCN, i, a.nadv);
        if (vmax != 0.0) {                                               // core.py:272-277
          const double speed = sqrt(vx * vx + vy * vy);
          if (speed > vmax) { vx = vx / speed * vmax; vy = vy / speed * vmax; }
        }
        px += vx * kDt; py += vy * kDt;
      }
    }
    __syncthreads();                                                     // all reads of old positions done
    if (valid) {
      s.pos[(0 * M + i) * EPB + e] = px; s.pos[(1 * M + i) * EPB + e] = py;
      s.vel[(0 * M + i) * EPB + e] = vx; s.vel[(1 * M + i) * EPB + e] = vy;
      s.comm[i * EPB + e] = new_comm;                                    // core.py:280-287
    }
    __syncthreads();

    // ---- phase 3: rewards from the new state
    if (SCN == MAPPO_SCN_SIMPLE_SPREAD) {
      if (valid)
        for (int l = 0; l < L; ++l) {
          const double dx = px - s.lm[(0 * L + l) * EPB + e], dy = py - s.lm[(1 * L + l) * EPB + e];
          s.aux[(size_t)(l * M + i) * EPB + e] = sqrt(dx * dx + dy * dy);
        }
      __syncthreads();
      if (valid) {
        double rew = 0.0;
        for (int l = 0; l < L; ++l) {                                    // simple_spread.py:75-78
          double m = s.aux[(size_t)(l * M + 0) * EPB + e];
          for (int k = 1; k < M; ++k) { const double d = s.aux[(size_t)(l * M + k) * EPB + e]; m = d < m ? d : m; }
          rew -= m;
        }
        const double dist_min = si.agent_size + si.agent_size;
        for (int k = 0; k < M; ++k) {                                    // simple_spread.py:80-83 (includes k == i)
          const double dx = s.pos[(0 * M + k) * EPB + e] - px, dy = s.pos[(1 * M + k) * EPB + e] - py;
          if (sqrt(dx * dx + dy * dy) < dist_min) rew -= 1;
        }
        indiv = rew;
      }
    } else if (SCN == MAPPO_SCN_SIMPLE_ADVERSARY || SCN == MAPPO_SCN_SIMPLE_PUSH) {
      const int g = valid ? s.goal[0 * EPB + e] : 0;
      const double gx = s.lm[(0 * L + g) * EPB + e], gy = s.lm[(1 * L + g) * EPB + e];
      if (valid) {                                                       // every agent's distance to the goal landmark
        const double dx = px - gx, dy = py - gy;
        s.aux[(size_t)i * EPB + e] = sqrt(dx * dx + dy * dy);
      }
      __syncthreads();
      if (valid) {
        double good_min = s.aux[(size_t)a.nadv * EPB + e];
        for (int k = a.nadv + 1; k < M; ++k) { const double d = s.aux[(size_t)k * EPB + e]; good_min = d < good_min ? d : good_min; }
        if (SCN == MAPPO_SCN_SIMPLE_ADVERSARY) {                         // simple_adversary.py:69-113
          if (i < a.nadv) { const double dx = px - gx, dy = py - gy; indiv = -(dx * dx + dy * dy); }
          else {
            double adv_rew = 0.0;
            for (int k = 0; k < a.nadv; ++k) adv_rew = adv_rew + s.aux[(size_t)k * EPB + e];
            indiv = -good_min + adv_rew;
          }
        } else {                                                         // simple_push.py:62-78
          if (i < a.nadv) { const double dx = gx - px, dy = gy - py; indiv = good_min - sqrt(dx * dx + dy * dy); }
          else indiv = -s.aux[(size_t)i * EPB + e];
        }
      }
    } else if (SCN == MAPPO_SCN_SIMPLE_TAG) {                            // simple_tag.py:73-123
      if (valid) {
        double rew = 0.0;
        if (i < a.nadv) {
          for (int ag = a.nadv; ag < M; ++ag)
            for (int adv = 0; adv < a.nadv; ++adv) {
              const double dx = s.pos[(0 * M + ag) * EPB + e] - s.pos[(0 * M + adv) * EPB + e];
              const double dy = s.pos[(1 * M + ag) * EPB + e] - s.pos[(1 * M + adv) * EPB + e];
              if (sqrt(dx * dx + dy * dy) < agent_size_of(SCN, ag, a.nadv) + agent_size_of(SCN, adv, a.nadv)) rew += 10;
            }
        } else {
          for (int adv = 0; adv < a.nadv; ++adv) {
            const double dx = s.pos[(0 * M + adv) * EPB + e] - px, dy = s.pos[(1 * M + adv) * EPB + e] - py;
            if (sqrt(dx * dx + dy * dy) < agent_size_of(SCN, adv, a.nadv) + agent_size_of(SCN, i, a.nadv)) rew -= 10;
          }
          for (int p = 0; p < 2; ++p) {                                  // bound(), simple_tag.py:95-104
            const double x = fabs(p == 0 ? px : py);
            double b;
            if (x < 0.9) b = 0.0;
            else if (x < 1.0) b = (x - 0.9) * 10;
            else { const double ex = exp(2 * x - 2); b = ex < 10.0 ? ex : 10.0; }
            rew -= b;
          }
        }
        indiv = rew;
      }
    } else if (SCN == MAPPO_SCN_SIMPLE_WORLD_COMM) {                     // simple_world_comm.py:135-190
      if (valid) {
        auto dist = [&](double ax, double ay, double bx, double by) {
          const double dx = ax - bx, dy = ay - by;
          return sqrt(dx * dx + dy * dy);
        };
        auto apos = [&](int ag, int c) { return s.pos[(c * M + ag) * EPB + e]; };
        double rew;
        if (i < a.nadv) {
          double near = dist(apos(a.nadv, 0), apos(a.nadv, 1), px, py);
          for (int g = a.nadv + 1; g < M; ++g) { const double d = dist(apos(g, 0), apos(g, 1), px, py); near = d < near ? d : near; }
          rew = 0.0 - 0.1 * near;
          for (int ag = a.nadv; ag < M; ++ag)
            for (int adv = 0; adv < a.nadv; ++adv)
              if (dist(apos(ag, 0), apos(ag, 1), apos(adv, 0), apos(adv, 1)) <
                  agent_size_of(SCN, ag, a.nadv) + agent_size_of(SCN, adv, a.nadv)) rew += 5;
        } else {
          rew = 0.0;
          for (int adv = 0; adv < a.nadv; ++adv)
            if (dist(apos(adv, 0), apos(adv, 1), px, py) < agent_size_of(SCN, adv, a.nadv) + agent_size_of(SCN, i, a.nadv)) rew -= 5;
          for (int p = 0; p < 2; ++p) {                                  // 2 * bound(|x|)
            const double x = fabs(p == 0 ? px : py);
            double b;
            if (x < 0.9) b = 0.0;
            else if (x < 1.0) b = (x - 0.9) * 10;
            else { const double ex = exp(2 * x - 2); b = ex < 10.0 ? ex : 10.0; }
            rew -= 2 * b;
          }
          const int food0 = L - 4;
          double fmin_ = 0.0;
          for (int f = 0; f < 2; ++f) {
            const double fx = s.lm[(0 * L + food0 + f) * EPB + e], fy = s.lm[(1 * L + food0 + f) * EPB + e];
            if (dist(px, py, fx, fy) < agent_size_of(SCN, i, a.nadv) + kFoodSize) rew += 2;
            const double d = dist(fx, fy, px, py);
            fmin_ = (f == 0 || d < fmin_) ? d : fmin_;
          }
          rew += 0.05 * fmin_;
        }
        indiv = rew;
      }
    } else if (SCN == MAPPO_SCN_SIMPLE_CRYPTO) {                         // simple_crypto.py:90-125
      if (valid) {
        // sum((c - goal_colour)^2) for one-hot c: 0 if the word is the goal's channel else 2; 0 while nothing was said
        const int g = s.goal[0 * EPB + e];
        const int c_adv = s.comm[0 * EPB + e], c_good = s.comm[1 * EPB + e];
        const double adv_err = (c_adv < 0 || c_adv == g) ? 0.0 : 2.0;
        const double good_err = (c_good < 0 || c_good == g) ? 0.0 : 2.0;
        indiv = (i == 0) ? 0.0 - adv_err : adv_err + (0.0 - good_err);
      }
    } else if (valid) {
      const int who = (SCN == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) ? 1 : 1 - i;   // goal_a
      const int g = s.goal[((SCN == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) ? 0 : i) * EPB + e];
      const double dx = s.pos[(0 * M + who) * EPB + e] - s.lm[(0 * L + g) * EPB + e];
      const double dy = s.pos[(1 * M + who) * EPB + e] - s.lm[(1 * L + g) * EPB + e];
      indiv = -(dx * dx + dy * dy);
    }
    if (valid) ind[i * EPB + e] = indiv;
    __syncthreads();
    if (valid) total = si.collaborative ? np_sum_small(ind + e, EPB, M) : indiv;   // environment.py:141-143
  }

  // ---- phase 4: reset finished envs (env_wrappers.py:145-154) / all envs (mode 1)
  const bool done = valid && doneflag[e];
  if (done) {
    const uint32_t env_lo = (uint32_t)(a.env_id_offset + n);
    const uint32_t ep = (uint32_t)a.st.episode[n];
    if (a.has_rs) {
      px = a.rs.p_pos[(size_t)(0 * M + i) * N + n]; py = a.rs.p_pos[(size_t)(1 * M + i) * N + n];
      s.goal[i * EPB + e] = a.rs.goal[(size_t)i * N + n];
    } else {
      draw_agent<SCN>(env_lo, ep, a.k0, i, px, py);
      if (si.shared_goal) s.goal[i * EPB + e] = draw_goal(env_lo, ep, a.k0, 0, L);
      else s.goal[i * EPB + e] = (i < si.n_goals) ? draw_goal(env_lo, ep, a.k0, i, L) : -1;
    }
    vx = 0.0; vy = 0.0;
    s.pos[(0 * M + i) * EPB + e] = px; s.pos[(1 * M + i) * EPB + e] = py;
    s.vel[(0 * M + i) * EPB + e] = 0.0; s.vel[(1 * M + i) * EPB + e] = 0.0;
    s.comm[i * EPB + e] = -1;
  }
  for (int idx = tid; idx < EPB * L; idx += blockDim.x) {
    const int l = idx / EPB, e2 = idx - l * EPB, n2 = n0 + e2;
    if (n2 < N && doneflag[e2]) {
      double x, y;
      if (a.has_rs) { x = a.rs.lm_pos[(size_t)(0 * L + l) * N + n2]; y = a.rs.lm_pos[(size_t)(1 * L + l) * N + n2]; }
      else draw_landmark<SCN>((uint32_t)(a.env_id_offset + n2), (uint32_t)a.st.episode[n2], a.k0, M, l, x, y);
      s.lm[(0 * L + l) * EPB + e2] = x; s.lm[(1 * L + l) * EPB + e2] = y;
      a.st.lm_pos[(size_t)(0 * L + l) * N + n2] = x; a.st.lm_pos[(size_t)(1 * L + l) * N + n2] = y;
    }
  }
  __syncthreads();

  // ---- phase 5: observations -> staged tile -> coalesced store
  if (valid) write_obs<SCN>(s, M, L, EPB, ld, e, i, vx, vy, a.nadv);
  __syncthreads();
  store_tile(s.tile, ld, a.W, cnt, a.obs_out + (size_t)n0 * a.W);

  // ---- phase 6: state and per-agent outputs
  if (valid) {
    a.st.p_pos[(size_t)(0 * M + i) * N + n] = px; a.st.p_pos[(size_t)(1 * M + i) * N + n] = py;
    a.st.p_vel[(size_t)(0 * M + i) * N + n] = vx; a.st.p_vel[(size_t)(1 * M + i) * N + n] = vy;
    a.st.comm[(size_t)i * N + n] = s.comm[i * EPB + e];
    if (done) a.st.goal[(size_t)i * N + n] = s.goal[i * EPB + e];
    if (i == 0) {
      a.st.step_count[n] = (done || !stepping) ? 0 : step;
      if (done) a.st.episode[n] = a.st.episode[n] + 1;
    }
    if (stepping) {
      const size_t o = (size_t)n * M + i;
      a.rewards_out[o] = (float)total;
      if (a.masks_out) a.masks_out[o] = done ? 0.0f : 1.0f;              // shared/mpe_runner.py:130-131
      if (a.dones_out) a.dones_out[o] = done ? 1 : 0;
      if (a.indiv_out) a.indiv_out[o] = (float)indiv;
    }
  }
}

static int check_cfg(const mappo_env_config* cfg, const mappo_env_state* st) {
  MAPPO_REQUIRE(cfg && st, "env: null config/state");
  MAPPO_REQUIRE(cfg->scenario >= 0 && cfg->scenario <= MAPPO_SCN_SIMPLE_WORLD_COMM, "env: unknown scenario %d", cfg->scenario);
  MAPPO_REQUIRE(cfg->num_envs > 0 && cfg->num_agents > 0 && cfg->num_landmarks > 0 && cfg->episode_length > 0,
                "env: non-positive size");
  MAPPO_REQUIRE(cfg->num_agents <= 16 && cfg->num_landmarks + n_extra_landmarks(cfg->scenario) <= 16,
                "env: at most 16 agents and 16 landmarks");
  if (cfg->scenario == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER || cfg->scenario == MAPPO_SCN_SIMPLE_REFERENCE)
    MAPPO_REQUIRE(cfg->num_agents == 2 && cfg->num_landmarks == 3,
                  "env: only 2 agents is supported (3 landmarks), check the config");  // scenario asserts
  if (cfg->scenario == MAPPO_SCN_SIMPLE_ADVERSARY)
    MAPPO_REQUIRE(cfg->num_agents >= 2 && cfg->num_landmarks == cfg->num_agents - 1,
                  "env: simple_adversary has num_agents - 1 landmarks");               // simple_adversary.py:17
  if (cfg->scenario == MAPPO_SCN_SIMPLE_PUSH)
    MAPPO_REQUIRE(cfg->num_agents >= 2 && cfg->num_landmarks <= 2, "env: simple_push has at most 2 landmarks");  // simple_push.py:42
  if (cfg->scenario == MAPPO_SCN_SIMPLE_CRYPTO)
    MAPPO_REQUIRE(cfg->num_agents == 3 && cfg->num_landmarks <= 4,
                  "env: simple_crypto has 3 agents and at most 4 landmarks (dim_c colours)");  // simple_crypto.py:27-31,54-58
  if (cfg->scenario == MAPPO_SCN_SIMPLE_TAG || cfg->scenario == MAPPO_SCN_SIMPLE_WORLD_COMM) {
    const int nadv = n_adversaries(cfg->scenario, cfg->num_adversaries);
    MAPPO_REQUIRE(nadv > 0 && nadv < cfg->num_agents, "env: simple_tag / simple_world_comm need 0 < num_adversaries (%d) < num_agents", nadv);
  }
  MAPPO_REQUIRE(st->p_pos && st->p_vel && st->lm_pos && st->comm && st->goal && st->step_count && st->episode,
                "env: null state array");
  return 0;
}

static int launch_env(const mappo_env_config* cfg, StepArgs& a, cudaStream_t stream) {
  const int M = cfg->num_agents, L = cfg->num_landmarks + n_extra_landmarks(cfg->scenario);
  a.N = cfg->num_envs; a.M = M; a.L = L; a.Lc = cfg->num_landmarks; a.T = cfg->episode_length;
  a.nadv = n_adversaries(cfg->scenario, cfg->num_adversaries);
  a.W = obs_off_of(cfg->scenario, M, cfg->num_landmarks, M, a.nadv);
  a.ld = tile_ld_of(a.W);
  a.k0 = (uint32_t)(cfg->seed & 0xFFFFFFFFull);
  a.env_id_offset = cfg->env_id_offset;
  int EPB = 128;
  while (EPB > 32 && (EPB * M > 512 || env_smem_bytes(M, L, EPB, a.W) > 96 * 1024)) EPB >>= 1;
  MAPPO_REQUIRE(EPB * M <= 512, "env: block too large");
  a.EPB = EPB;
  const size_t smem = env_smem_bytes(M, L, EPB, a.W);
  MAPPO_REQUIRE(smem <= 227 * 1024, "env: obs row too wide for shared-memory staging (%zu B)", smem);
  const int grid = ceil_div(a.N, EPB);
#define LAUNCH_SCN(S)                                                                                   \
  do {                                                                                                  \
    if (smem > 48 * 1024)                                                                               \
      MAPPO_CUDA(cudaFuncSetAttribute(env_kernel<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
    env_kernel<S><<<grid, EPB * M, smem, stream>>>(a);                                                  \
  } while (0)
  if (cfg->scenario == MAPPO_SCN_SIMPLE_SPREAD) LAUNCH_SCN(MAPPO_SCN_SIMPLE_SPREAD);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_SPEAKER_LISTENER) LAUNCH_SCN(MAPPO_SCN_SIMPLE_SPEAKER_LISTENER);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_REFERENCE) LAUNCH_SCN(MAPPO_SCN_SIMPLE_REFERENCE);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_ADVERSARY) LAUNCH_SCN(MAPPO_SCN_SIMPLE_ADVERSARY);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_PUSH) LAUNCH_SCN(MAPPO_SCN_SIMPLE_PUSH);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_CRYPTO) LAUNCH_SCN(MAPPO_SCN_SIMPLE_CRYPTO);
  else if (cfg->scenario == MAPPO_SCN_SIMPLE_WORLD_COMM) LAUNCH_SCN(MAPPO_SCN_SIMPLE_WORLD_COMM);
  else LAUNCH_SCN(MAPPO_SCN_SIMPLE_TAG);
#undef LAUNCH_SCN
  MAPPO_CHECK_LAUNCH("env_kernel");
  return 0;
}

}  // namespace mappo

using namespace mappo;

extern "C" int mappo_env_obs_dim(int scenario, int M, int L, int agent) { return obs_dim_of(scenario, M, L, agent); }
extern "C" int mappo_env_obs_width(int scenario, int M, int L) { return obs_off_of(scenario, M, L, M); }
extern "C" int mappo_env_obs_dim_cfg(const mappo_env_config* c, int agent) {
  return c ? obs_dim_of(c->scenario, c->num_agents, c->num_landmarks, agent, c->num_adversaries) : -1;
}
extern "C" int mappo_env_num_landmarks(const mappo_env_config* c) {
  return c ? c->num_landmarks + n_extra_landmarks(c->scenario) : -1;
}
extern "C" int mappo_env_obs_width_cfg(const mappo_env_config* c) {
  return c ? obs_off_of(c->scenario, c->num_agents, c->num_landmarks, c->num_agents, c->num_adversaries) : -1;
}

extern "C" int mappo_env_reset(const mappo_env_config* cfg, const mappo_env_state* st, float* obs_out,
                               int use_given_state, void* stream) {
  if (int rc = check_cfg(cfg, st)) return rc;
  MAPPO_REQUIRE(obs_out, "env_reset: null obs_out");
  StepArgs a{};
  a.st = *st; a.has_rs = 0; a.obs_out = obs_out; a.mode = use_given_state ? 2 : 1;
  return launch_env(cfg, a, (cudaStream_t)stream);
}

extern "C" int mappo_env_step(const mappo_env_config* cfg, const mappo_env_state* st, const float* actions,
                              int act_width, float* obs_out, float* rewards_out, float* masks_out,
                              uint8_t* dones_out, float* indiv_out, const mappo_env_state* reset_src,
                              void* stream) {
  if (int rc = check_cfg(cfg, st)) return rc;
  MAPPO_REQUIRE(actions && obs_out && rewards_out, "env_step: null actions/obs_out/rewards_out");
  const int need = (cfg->scenario == MAPPO_SCN_SIMPLE_REFERENCE || cfg->scenario == MAPPO_SCN_SIMPLE_WORLD_COMM) ? 2 : 1;
  MAPPO_REQUIRE(act_width >= need, "env_step: act_width %d < %d heads", act_width, need);
  StepArgs a{};
  a.st = *st; a.actions = actions; a.act_w = act_width; a.obs_out = obs_out; a.rewards_out = rewards_out;
  a.masks_out = masks_out; a.dones_out = dones_out; a.indiv_out = indiv_out; a.mode = 0;
  if (reset_src) {
    MAPPO_REQUIRE(reset_src->p_pos && reset_src->lm_pos && reset_src->goal, "env_step: incomplete reset_src");
    a.rs = *reset_src; a.has_rs = 1;
  }
  return launch_env(cfg, a, (cudaStream_t)stream);
}
