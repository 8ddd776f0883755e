/* mappo_b200 -- C ABI of the B200-native MAPPO hot path (vectorised MPE rollout + GAE + PPO update).
 *
 * The reference (Chen001117/mappo, package `onpolicy`) is pure Python: it has no FFI on this path
 * (SURVEY.md §8b).  The drop-in boundary is therefore the Python classes in `mappo_b200/` that
 * mirror `onpolicy/`'s names and signatures; THIS header is the thin native layer underneath them
 * (north_star: "all calling hand-written sm_100a CUDA kernels through a thin C-ABI extension").
 * Each entry point cites the reference function(s) it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - extern "C", plain pointers and sizes only.  Every data pointer is a DEVICE pointer unless its
 *     name ends in `_host`.  The library never allocates, frees or keeps user memory: the caller
 *     (PyTorch in the Python mirror) owns every buffer and keeps it alive until the stream work ends.
 *   - Every call is stream-ordered on `stream` (a cudaStream_t passed as void*; NULL = legacy default
 *     stream), launches asynchronously and never synchronises the device.
 *   - Return value: 0 on success, a negative mappo_status on error; mappo_last_error() gives the text
 *     (thread-local).  The reference signals errors with Python assert/NotImplementedError
 *     (utils/shared_buffer.py:237, utils/util.py:37); the Python mirror turns a non-zero status into
 *     the same exception types.
 *   - There is no CPU fallback anywhere behind this ABI.
 */
#ifndef MAPPO_B200_H
#define MAPPO_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MAPPO_B200_ABI_VERSION 1

typedef enum {
  MAPPO_OK = 0,
  MAPPO_ERR_INVALID_ARG = -1,   /* bad size / null pointer / unsupported combination */
  MAPPO_ERR_UNSUPPORTED = -2,   /* valid in the reference, not built here (maps to NotImplementedError) */
  MAPPO_ERR_CUDA = -3           /* a CUDA runtime call or launch failed */
} mappo_status;

int mappo_abi_version(void);
/* compute capability the kernels were compiled for (100 = sm_100a) */
int mappo_sm_arch(void);
const char* mappo_last_error(void);
/* number of kernel launches issued through this library since load (bench.py's gpu_launches) */
uint64_t mappo_launch_count(void);

/* ------------------------------------------------------------------------------------------------
 * 1. MPE environment: device-resident batched vec env with in-kernel auto-reset.
 *    Replaces, per env step and for all envs at once:
 *      MultiAgentEnv.step/_set_action/_get_done/reset   onpolicy/envs/mpe/environment.py:117-163,179-256
 *      World.step and the four physics functions         onpolicy/envs/mpe/core.py:207-321
 *      Scenario.reset_world/reward/observation           onpolicy/envs/mpe/scenarios/simple_spread.py:32-103,
 *                                                        simple_speaker_listener.py:38-98, simple_reference.py:34-97
 *      SubprocVecEnv.step_async/step_wait/reset + worker auto-reset   onpolicy/envs/env_wrappers.py:140-157,257-272
 *      the runner glue done->mask                        onpolicy/runner/shared/mpe_runner.py:125-131
 * ---------------------------------------------------------------------------------------------- */
typedef enum {
  MAPPO_SCN_SIMPLE_SPREAD = 0,
  MAPPO_SCN_SIMPLE_SPEAKER_LISTENER = 1,
  MAPPO_SCN_SIMPLE_REFERENCE = 2
} mappo_scenario;

typedef struct {
  int32_t scenario;        /* mappo_scenario */
  int32_t num_envs;        /* N: envs in this shard */
  int32_t num_agents;      /* M */
  int32_t num_landmarks;   /* L */
  int32_t episode_length;  /* world_length: done when step count reaches it (environment.py:179-185) */
  int32_t reserved;
  uint64_t seed;           /* Philox key word 0 for reset draws */
  int64_t env_id_offset;   /* global id of local env 0: reset streams do not depend on the sharding */
} mappo_env_config;

/* World state, structure-of-arrays, float64 like the reference (core.py integrates in f64).
 * Component-major then entity-major then env: element (c, i, n) at [(c*M + i)*N + n]. */
typedef struct {
  double* p_pos;        /* [2][M][N] agent positions */
  double* p_vel;        /* [2][M][N] agent velocities */
  double* lm_pos;       /* [2][L][N] landmark positions */
  int32_t* comm;        /* [M][N] id of the one-hot utterance state.c of agent i, -1 = all zeros */
  int32_t* goal;        /* [M][N] landmark id of agent i's goal_b, -1 = none */
  int32_t* step_count;  /* [N] MultiAgentEnv.current_step */
  int32_t* episode;     /* [N] resets performed so far (Philox counter word 1) */
} mappo_env_state;

/* obs row width per env: sum over agents of that agent's obs dim (= share_obs dim for MPE) */
int mappo_env_obs_width(int scenario, int num_agents, int num_landmarks);
/* obs dim of one agent */
int mappo_env_obs_dim(int scenario, int num_agents, int num_landmarks, int agent);

/* envs.reset(): draw every env's initial state (Philox, counter = (env id, episode, block, 0)),
 * zero step counts, and write the observations.  obs_out: float32 [N][obs_width], agent-major inside
 * a row, i.e. exactly the reference's obs[N,M,D] (and one share_obs row) for homogeneous agents.
 * If `use_given_state` != 0 no state is drawn: obs are computed from the state already in `st`
 * (parity tests inject the reference's own reset states this way). */
int mappo_env_reset(const mappo_env_config* cfg, const mappo_env_state* st, float* obs_out,
                    int use_given_state, void* stream);

/* envs.step(actions): one fused kernel over envs x agents.
 *   actions      float32 [N][M][act_width] action ids exactly as SharedReplayBuffer.actions stores them
 *                (head 0 = physical move for movable agents, else the utterance; head 1 = utterance)
 *   obs_out      float32 [N][obs_width]; for envs that finish their episode this is the RESET obs
 *   rewards_out  float32 [N][M]  shared (summed) reward, the terminal step's on reset
 *   masks_out    float32 [N][M]  1 - done      (may be NULL)
 *   dones_out    uint8   [N][M]                (may be NULL)
 *   indiv_out    float32 [N][M]  info['individual_reward'] (may be NULL)
 *   reset_src    NULL -> finished envs re-draw their state from Philox.  Otherwise a state whose arrays
 *                (p_pos, lm_pos, goal) hold the reset state to install for finished envs (parity tests). */
int mappo_env_step(const mappo_env_config* cfg, const mappo_env_state* st, const float* actions,
                   int act_width, float* obs_out, float* rewards_out, float* masks_out,
                   uint8_t* dones_out, float* indiv_out, const mappo_env_state* reset_src, void* stream);


/* ------------------------------------------------------------------------------------------------
 * 2. Returns / GAE / ValueNorm / advantages.
 *    Replaces SharedReplayBuffer.compute_returns (all four branches)   onpolicy/utils/shared_buffer.py:168-224
 *             ValueNorm.running_mean_var/update/normalize/denormalize   onpolicy/utils/valuenorm.py:31-78
 *             the advantage prologue of R_MAPPO.train                   onpolicy/algorithms/r_mappo/r_mappo.py:174-182
 *    All arithmetic is float32 with the reference's operation order and separately rounded ops
 *    (the file is compiled without FMA contraction), so returns match numpy bit for bit.
 *    Arrays are time-major [T or T+1][R] with R = N*M (shared buffer) or N (separated buffer) trajectories.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int32_t T;                       /* episode_length */
  int32_t R;                       /* trajectories */
  float gamma;                     /* float32(args.gamma) */
  float gamma_lambda;              /* float32(args.gamma * args.gae_lambda), product taken in double */
  int32_t use_gae;
  int32_t use_proper_time_limits;
  int32_t use_value_norm;          /* ValueNorm or PopArt statistics in `vn_stats` */
  int32_t reserved;
} mappo_gae_config;

/* vn_stats: float32[3] = {running_mean, running_mean_sq, debiasing_term} (valuenorm.py:19-21), or NULL.
 * next_value: float32 [R] written to value_preds[T] first (shared_buffer.py:207), or NULL.
 * bad_masks: [T+1][R] or NULL (= ones).  active_masks: [T+1][R] or NULL (= ones).
 * Outputs: returns [T+1][R]; advantages_raw [T][R] = returns[:-1] - denorm(value_preds[:-1]);
 * adv_partials: double [mappo_gae_num_partials(R)][3] per-block (sum, sum of squares, count) over active
 * entries, reduced by mappo_adv_stats_reduce. */
int mappo_gae_num_partials(int R);
int mappo_gae_returns(const mappo_gae_config* cfg, const float* rewards, float* value_preds, float* returns,
                      const float* masks, const float* bad_masks, const float* active_masks,
                      const float* next_value, const float* vn_stats, float* advantages_raw,
                      double* adv_partials, void* stream);
/* Sum the per-block partials in a fixed order into stats3 = double {sum, sumsq, count}.  In a multi-GPU
 * job the caller all-reduces stats3 (3 doubles) between this call and mappo_adv_normalize. */
int mappo_adv_stats_reduce(const double* partials, int num_partials, double* stats3, void* stream);
/* advantages = (raw - mean) / (std + 1e-5) with population std (np.nanmean / np.nanstd, r_mappo.py:180-182). */
int mappo_adv_normalize(const float* advantages_raw, float* advantages, long long count, const double* stats3,
                        void* stream);

/* ValueNorm.update(batch) (valuenorm.py:38-54), split so that the batch moments can be all-reduced:
 *   moments: per-block partial {sum x, sum x^2, n} of `x[count]` (optionally only where weights != 0)
 *   -> mappo_adv_stats_reduce -> [all-reduce] -> mappo_valuenorm_apply: EMA update of vn_stats (float32[3]). */
int mappo_moments_num_partials(long long count);
int mappo_moments(const float* x, long long count, double* partials, void* stream);
int mappo_valuenorm_apply(float* vn_stats, const double* stats3, double beta, void* stream);
/* out = (x - mean) / sqrt(var)  or  out = x * sqrt(var) + mean  (valuenorm.py:56-78), elementwise float32 */
int mappo_valuenorm_transform(const float* vn_stats, const float* x, float* out, long long count, int denormalize,
                              void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAPPO_B200_H */
