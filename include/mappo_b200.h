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

#define MAPPO_B200_ABI_VERSION 6

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
  MAPPO_SCN_SIMPLE_REFERENCE = 2,
  /* scenarios/simple_adversary.py:9-137, simple_push.py:8-104, simple_tag.py:7-144
   * (individual rewards; simple_tag adds per-agent size / accel / max_speed and obstacle landmarks) */
  MAPPO_SCN_SIMPLE_ADVERSARY = 3,
  MAPPO_SCN_SIMPLE_PUSH = 4,
  MAPPO_SCN_SIMPLE_TAG = 5,
  /* scenarios/simple_crypto.py:21-171 (= simple_crypto_display.py without its prints): three immobile agents
   * (adversary, listener, speaker) exchanging 4-colour words; goal slot 0 = goal landmark, slot 1 = key landmark */
  MAPPO_SCN_SIMPLE_CRYPTO = 6,
  /* scenarios/simple_world_comm.py:7-288: simple_tag's physics (per-side size / accel / max_speed, obstacle landmarks)
   * plus 2 food and 2 forest landmarks stored behind the `num_landmarks` obstacles, forest-limited visibility, and a
   * leader (agent 0) with a movement and a 4-word communication head */
  MAPPO_SCN_SIMPLE_WORLD_COMM = 7
} mappo_scenario;

typedef struct {
  int32_t scenario;        /* mappo_scenario */
  int32_t num_envs;        /* N: envs in this shard */
  int32_t num_agents;      /* M */
  int32_t num_landmarks;   /* L */
  int32_t episode_length;  /* world_length: done when step count reaches it (environment.py:179-185) */
  int32_t num_adversaries; /* simple_tag: args.num_adversaries (0 = the scenario's default 3); the agents listed first.
                            * Occupies the word that was `reserved` (= 0) in ABI v2. */
  uint64_t seed;           /* Philox key word 0 for reset draws */
  int64_t env_id_offset;   /* global id of local env 0: reset streams do not depend on the sharding */
} mappo_env_config;

/* World state, structure-of-arrays, float64 like the reference (core.py integrates in f64).
 * Component-major then entity-major then env: element (c, i, n) at [(c*M + i)*N + n]. */
typedef struct {
  double* p_pos;        /* [2][M][N] agent positions */
  double* p_vel;        /* [2][M][N] agent velocities */
  double* lm_pos;       /* [2][L][N] landmark positions; L = mappo_env_num_landmarks(cfg) (= num_landmarks except for
                         * simple_world_comm: obstacles, then 2 food, then 2 forests) */
  int32_t* comm;        /* [M][N] id of the one-hot utterance state.c of agent i, -1 = all zeros */
  int32_t* goal;        /* [M][N] landmark id of agent i's goal_b, -1 = none */
  int32_t* step_count;  /* [N] MultiAgentEnv.current_step */
  int32_t* episode;     /* [N] resets performed so far (Philox counter word 1) */
} mappo_env_state;

/* obs row width per env: sum over agents of that agent's obs dim (= share_obs dim for MPE) */
int mappo_env_obs_width(int scenario, int num_agents, int num_landmarks);
/* obs dim of one agent */
int mappo_env_obs_dim(int scenario, int num_agents, int num_landmarks, int agent);
/* the same from a config (needed where the width depends on num_adversaries: simple_tag) */
int mappo_env_obs_dim_cfg(const mappo_env_config* cfg, int agent);
int mappo_env_obs_width_cfg(const mappo_env_config* cfg);
/* landmarks held in the state arrays: cfg->num_landmarks plus the scenario's own (simple_world_comm: + 4) */
int mappo_env_num_landmarks(const mappo_env_config* cfg);

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

/* envs.render('rgb_array'): headless rasteriser.
 *    Replaces MultiAgentEnv.render                      onpolicy/envs/mpe/environment.py:263-412
 *             Viewer.render / FilledPolygon / make_circle onpolicy/envs/mpe/rendering.py:44-112,223-250
 *    (pyglet window + fixed-function OpenGL in the reference: unavailable on this box; this is a CUDA
 *    restatement of that raster -- white clear, one 30-gon per entity blended SRC_ALPHA / ONE_MINUS_SRC_ALPHA,
 *    half-intensity outline of width `line_width`, communication circles last -- into a device buffer.) */
typedef struct {
  int32_t width, height;   /* rendering.Viewer(700, 700), environment.py:297; width % 4 == 0 */
  int32_t num_images;      /* envs to draw */
  int32_t num_entities;    /* E = M + L: world.entities = agents then landmarks (core.py:142-144) */
  int32_t res;             /* vertices per circle: 30 (rendering.make_circle) */
  int32_t reserved;
  double cam_range;        /* 2 (environment.py:8); shared viewer centred on the origin (environment.py:383-388) */
  double line_width;       /* glLineWidth(2.0), rendering.py:59 */
} mappo_render_config;
/*   env_ids      int32 [num_images] local env index of each image (NULL: image i = env i)
 *   entity_size  float32 [E]                 entity.size
 *   entity_rgba  float32 [num_images][E][4]  entity.color + alpha (agents 0.5, landmarks 1; environment.py:315,338)
 *   comm_dim     int32 [E] or NULL           world.dim_c for non-silent agents (communication circles), else 0
 *   unit_circle  float64 [res][2]            (cos, sin)(2 pi k / res), computed by the caller (math.cos / math.sin)
 *   rgb_out      uint8 [num_images][height][width][3], row 0 = top (rendering.py:103 flips the GL buffer) */
int mappo_env_render(const mappo_env_config* cfg, const mappo_env_state* st, const mappo_render_config* rc,
                     const int32_t* env_ids, const float* entity_size, const float* entity_rgba,
                     const int32_t* comm_dim, const double* unit_circle, uint8_t* rgb_out, void* stream);


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

/* advantages_raw = returns[:-1] - denorm(value_preds[:-1]) (r_mappo.py:174-177) recomputed from the buffer,
 * plus the per-block moments over active entries (as mappo_gae_returns produces them).  count = T*R elements;
 * partials: double [mappo_moments_num_partials(count)][3]. */
int mappo_advantages(const float* returns, const float* value_preds, const float* active_masks,
                     const float* vn_stats, long long count, float* advantages_raw, double* partials, void* stream);

/* ValueNorm.update(batch) (valuenorm.py:38-54), split so that the batch moments can be all-reduced:
 *   moments: per-block partial {sum x, sum x^2, n} of `x[count]` (optionally only where weights != 0)
 *   -> mappo_adv_stats_reduce -> [all-reduce] -> mappo_valuenorm_apply: EMA update of vn_stats (float32[3]). */
int mappo_moments_num_partials(long long count);
/* PopArt.update's rescale of the value head (popart.py:52-70, as it runs on cuda): with (mean, var) the debiased
 * statistics before (`old_stats`) and after (`new_stats`, both float32[3] like vn_stats) the EMA step,
 *   weight <- weight * sqrt(old_var) / sqrt(new_var),   bias <- (sqrt(old_var) * bias + old_mean - new_mean) / sqrt(new_var)
 * on head_w [n_out][n_in] and head_b [n_out] (slices of the critic's flat parameter buffer).  (ABI v6.) */
int mappo_popart_rescale(const float* old_stats, const float* new_stats, float* head_w, float* head_b, int n_in,
                         int n_out, void* stream);
int mappo_moments(const float* x, long long count, double* partials, void* stream);

/* Allocation-free variant of mappo_moments + mappo_adv_stats_reduce: stats3 (double[3] on the device) receives
 * {sum x, sum x^2 (x^2 rounded to float32 first), count} of x[count], the same numbers the pair above produces.
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot allocate
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The per-block partial sums
 * live in a file-scope `static __device__` array inside the library (148 * 8 rows of 3 doubles = 28,416 bytes, held by
 * every GPU that loads the library until the process exits); the call performs no cudaMalloc / cudaFree and calls no
 * allocating routine.  Limit: count <= MAPPO_MOMENTS_STATIC_MAX_COUNT elements, larger counts return
 * MAPPO_ERR_INVALID_ARG with a message (use mappo_moments with caller scratch).  All calls on one GPU share the
 * array: calls that overlap on the GPU (different streams, or different host threads / graphs running concurrently)
 * RACE -- serialise them on one stream.  The array keeps the previous call's contents; nothing in it needs zeroing
 * (every block overwrites its row and exactly the rows written are read back). */
#define MAPPO_MOMENTS_STATIC_MAX_COUNT (1LL << 28)
long long mappo_moments_static_max_count(void);
int mappo_moments_static(const float* x, long long count, double* stats3, void* stream);
int mappo_valuenorm_apply(float* vn_stats, const double* stats3, double beta, void* stream);
/* out = (x - mean) / sqrt(var)  or  out = x * sqrt(var) + mean  (valuenorm.py:56-78), elementwise float32 */
int mappo_valuenorm_transform(const float* vn_stats, const float* x, float* out, long long count, int denormalize,
                              void* stream);

/* ------------------------------------------------------------------------------------------------
 * 3. Actor / critic networks: fused forward (rollout) and fused forward + loss + backward (update).
 *    Replaces R_Actor.forward / evaluate_actions, R_Critic.forward   onpolicy/algorithms/r_mappo/algorithm/r_actor_critic.py:43-107,146-165
 *             MLPBase / MLPLayer (LayerNorm -> Linear -> act -> LayerNorm ...)   onpolicy/algorithms/utils/mlp.py:6-59
 *             ACTLayer + FixedCategorical (sample / mode / log_probs / entropy)  onpolicy/algorithms/utils/act.py:41-162,
 *                                                                                  onpolicy/algorithms/utils/distributions.py:14-68
 *             R_MAPPO.ppo_update losses and cal_value_loss                         onpolicy/algorithms/r_mappo/r_mappo.py:52-136
 *             huber_loss / mse_loss                                                onpolicy/utils/util.py:23-29
 *    Parameters of one network live in ONE flat float32 buffer in this order (the reference's
 *    nn.Module.parameters() order, minus the never-executed `fc_h` template, mlp.py:20-27):
 *      [feature_norm.weight[in], feature_norm.bias[in]]            if use_feature_norm
 *      fc1: W[H][in], b[H], ln_w[H], ln_b[H]
 *      fc2[l], l < layer_N: W[H][H], b[H], ln_w[H], ln_b[H]
 *      [rnn: W_ih[3H][H], W_hh[3H][H], b_ih[3H], b_hh[3H], norm_w[H], norm_b[H]]   if recurrent
 *      heads h < n_heads: W[head_dim[h]][H], b[head_dim[h]]        (critic: one head of dim 1 = v_out)
 *    Gradients use the same layout.
 * ---------------------------------------------------------------------------------------------- */
#define MAPPO_MAX_HEADS 4
typedef struct {
  int32_t in_dim;              /* obs dim D (actor) or share-obs dim S (critic) */
  int32_t hidden;              /* H; this build: 64 */
  int32_t layer_N;             /* this build: 1 */
  int32_t use_relu;            /* 0 = Tanh, 1 = ReLU (mlp.py:11) */
  int32_t use_feature_norm;    /* LayerNorm on the input (mlp.py:44-45) */
  int32_t recurrent;           /* 0 = MLP, 1 = one GRU layer + LayerNorm (rnn.py) */
  int32_t n_heads;             /* categorical heads (Discrete: 1, MultiDiscrete: 2); critic: 1 */
  int32_t head_dim[MAPPO_MAX_HEADS];
  int32_t is_critic;
} mappo_net_config;

long long mappo_net_param_count(const mappo_net_config* cfg);

/* Actor forward for the rollout (R_Actor.forward): rows = N*M observations.
 *   obs [rows][obs_ld] float32 (row stride obs_ld >= in_dim: rows may be slices of a wider share_obs row)
 *   available_actions [rows][head_dim[0]] or NULL; deterministic != 0 -> mode() else sample()
 *   sampling: inverse CDF on the softmax with u = Philox(counter = (row_offset + row, step, head, 0), key = (seed, 'SAMP'))
 *   actions_out, logp_out: float32 [rows][n_heads]  (act.py:65-81: per-head, concatenated)
 * Evaluate mode (R_Actor.evaluate_actions without the backward, act.py:107-162): if `given_actions`
 * (float32 [rows][n_heads]) is not NULL nothing is sampled; logp_out receives log_prob(given action) and
 * entropy_out (float32 [rows], may be NULL) the mean over heads of each row's entropy; actions_out may be NULL.
 * step_offset (int64 on the device, may be NULL) is added to `step` inside the kernel: a captured CUDA graph of the
 * rollout keeps its sampling counter in HBM and advances it between launches.  (ABI v3.)
 * workspace: mappo_mlp_workspace_bytes(cfg, rows) bytes of scratch for a wide observation (in_dim > 64, e.g.
 * simple_spread with 12 or more agents), else NULL.  (ABI v4.) */
int mappo_actor_forward(const mappo_net_config* cfg, const float* params, const float* obs, int obs_ld,
                        long long rows, const float* available_actions, int deterministic, uint64_t seed,
                        uint64_t step, long long row_offset, const float* given_actions, float* actions_out,
                        float* logp_out, float* entropy_out, const int64_t* step_offset, void* workspace,
                        void* stream);

/* Critic forward (R_Critic.forward / get_values): `rows` unique share-obs rows, each value written `group`
 * times: values_out float32 [rows][group] (group = M when the M agents of an env share one share_obs row). */
int mappo_critic_forward(const mappo_net_config* cfg, const float* params, const float* share_obs, int obs_ld,
                         long long rows, int group, float* values_out, void* workspace, void* stream);

/* Scratch bytes an actor or critic call over `rows` rows needs in `workspace` (device memory, 16-byte aligned): 0 for
 * in_dim <= 64 (pass NULL); for a wide input (in_dim > 64, e.g. the centralised critic's 460 = 10 agents x 46, or an
 * actor's 70 columns with 16 agents) the first layer runs as separate streaming GEMMs and its [rows][64] activations /
 * gradients plus [rows][2] row statistics live there. */
long long mappo_mlp_workspace_bytes(const mappo_net_config* cfg, long long rows);

typedef struct {
  float clip_param;            /* args.clip_param */
  float entropy_coef;          /* args.entropy_coef */
  float value_loss_coef;       /* args.value_loss_coef */
  float huber_delta;           /* args.huber_delta */
  int32_t use_huber_loss;
  int32_t use_clipped_value_loss;
  int32_t use_value_norm;      /* normalize(return) with vn_stats (ValueNorm or PopArt statistics) */
  int32_t use_policy_active_masks;
  int32_t use_value_active_masks;
  int32_t reserved;
} mappo_loss_config;

/* number of partial rows the update kernels write for `rows` rows of network `cfg` on the current device */
int mappo_update_num_blocks(const mappo_net_config* cfg, long long rows);

/* Which kernels serve network `cfg` (host-side decision, no device call; no counterpart in the reference):
 *   0  fp32 tile kernels (mlp.cu: layer_N > 1, hidden != 64, shared-memory map of the fused kernel too large, or
 *      MAPPO_MLP_IMPL=fma), 1  fused tcgen05 kernels (mlp_tc.cu), 2  the same behind a wide first layer (mlp_wide.cu),
 *   3  recurrent chain (mlp_gru.cu over the fused base); -1 for a configuration check_net() refuses. */
int mappo_mlp_kernel_path(const mappo_net_config* cfg);

/* Actor: forward + clipped-surrogate/entropy loss + backward over one minibatch (r_mappo.py:115-141).
 *   index: int64 [rows] row gather into the flat buffers (the minibatch permutation,
 *          shared_buffer.py:246-247), or NULL for the identity (num_mini_batch == 1: the loss is a mean,
 *          so the order only changes the summation order).
 *   actions, old_logp [*][n_heads]; adv [*]; active_masks [*] or NULL.
 * Outputs, one row per partial b < mappo_update_num_blocks(cfg, rows), UNNORMALISED sums (the 1/sum(active) factor,
 * the LayerNorm-affine chain rule and the global sum over blocks and ranks are applied by
 * mappo_grad_reduce / mappo_adam_step):
 *   grad_partials float32 [blocks][P];
 *   loss_partials double [blocks][4] = {sum -min(surr1,surr2)*active, sum entropy*active (mean over heads),
 *                                       sum ratio (all heads), sum active}
 * workspace: as for mappo_actor_forward.  (ABI v4.) */
int mappo_actor_update(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                       const float* obs, int obs_ld, long long rows, const int64_t* index, const float* actions,
                       const float* old_logp, const float* adv, const float* active_masks,
                       const float* available_actions, float* grad_partials, double* loss_partials, void* workspace,
                       void* stream);

/* Critic: forward + (clipped, huber/mse, value-normalised) value loss + backward (r_mappo.py:52-89,151-156).
 *   `rows` unique share-obs rows; each carries `group` loss terms: value_preds, returns, active_masks
 *   [rows][group] (group = M for the shared buffer with num_mini_batch == 1, else 1 with `index`).
 *   share_rows_div: with `index`, sample i reads share_obs row index[i] / share_rows_div (= M when the
 *   storage keeps ONE share_obs row per env and the reference's M replicated rows are a view of it), while
 *   its value_preds / returns / active_masks entries are read at index[i].  1 otherwise.
 *   vn_stats float32[3] AFTER this minibatch's ValueNorm.update (r_mappo.py:65-67) or NULL.
 *   loss_partials double [blocks][4] = {sum value_loss*active, sum active, 0, 0}
 *   packed_inputs: NULL, or the buffer mappo_critic_pack_inputs filled from these same share_obs rows (wide first
 *   layer, no row gather): the forward pass then streams the packed tiles instead of converting the rows again.
 *   (ABI v3.) */
int mappo_critic_update(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                        const float* share_obs, int obs_ld, long long rows, int group, const int64_t* index,
                        int share_rows_div,
                        const float* value_preds, const float* returns, const float* active_masks,
                        const float* vn_stats, float* grad_partials, double* loss_partials, void* workspace,
                        const void* packed_inputs, void* stream);

/* Wide first layer (in_dim > 64), num_mini_batch == 1: every PPO epoch reads the same share_obs rows and their
 * LayerNorm does not depend on the weights, so they can be converted ONCE per iteration into the fp16-pair operand
 * tiles the tensor cores read (per 128 rows: [hi | lo][8-column chunk][row][8 fp16] of LN(x)).
 *   mappo_critic_packed_bytes: bytes of that buffer for `rows` rows (0: this network does not pack)
 *   mappo_critic_pack_inputs:  fill it (and the row statistics inside `workspace`, which must be the workspace later
 *                              passed to mappo_critic_update) */
long long mappo_critic_packed_bytes(const mappo_net_config* cfg, long long rows);
int mappo_critic_pack_inputs(const mappo_net_config* cfg, const float* share_obs, int obs_ld, long long rows,
                             void* packed, void* workspace, void* stream);

/* ---- recurrent (GRU) networks: cfg->recurrent = 1 ---------------------------------------------------------
 *    Replaces RNNLayer.forward (rollout branch and the chunked training branch)   onpolicy/algorithms/utils/rnn.py:24-77
 *             inside R_Actor / R_Critic                                            r_actor_critic.py:43-107,146-165
 *             and the recurrent ppo_update (evaluate_actions on chunks)            r_mappo.py:91-164
 * rnn_states: float32 [rows][H] (recurrent_N = 1); masks: float32 [rows]; the hidden state is multiplied by the mask
 * before the step (rnn.py:26,60).  `workspace`: mappo_rnn_workspace_bytes(cfg, positions, update) bytes, 16-byte
 * aligned; positions = rows (forward) or n_chunks * chunk_len (update). */
long long mappo_rnn_workspace_bytes(const mappo_net_config* cfg, long long positions, int update);

int mappo_actor_forward_rnn(const mappo_net_config* cfg, const float* params, const float* obs, int obs_ld,
                            long long rows, const float* rnn_states_in, const float* masks,
                            const float* available_actions, int deterministic, uint64_t seed, uint64_t step,
                            long long row_offset, const float* given_actions, float* rnn_states_out,
                            float* actions_out, float* logp_out, float* entropy_out, const int64_t* step_offset,
                            void* workspace, void* stream);     /* step_offset: as mappo_actor_forward (ABI v5) */

/* share_obs row of sample i is share_obs[i / share_rows_div] (several agents' samples may share one row). */
int mappo_critic_forward_rnn(const mappo_net_config* cfg, const float* params, const float* share_obs, int obs_ld,
                             long long rows, int share_rows_div, const float* rnn_states_in, const float* masks,
                             float* rnn_states_out, float* values_out, void* workspace, void* stream);

/* One recurrent PPO minibatch: n_chunks sequences of chunk_len steps (recurrent_generator, shared_buffer.py:385-494;
 * naive_recurrent_generator :288-383 with chunk_len = episode_length).  index int64 [chunk_len][n_chunks]: the
 * flat sample row of step s of chunk c; obs / actions / old_logp / adv / masks / active_masks / rnn_states are the
 * flat per-sample arrays of the buffer and are read at those rows; the initial hidden state of chunk c is
 * rnn_states[index[0][c]].  Back-propagation through time covers the whole chunk.  grad_partials
 * float32 [n_partials][P] with n_partials = mappo_update_num_blocks(cfg, n_chunks * chunk_len) (zeroed here). */
int mappo_actor_update_rnn(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                           const float* obs, int obs_ld, long long n_chunks, int chunk_len, const int64_t* index,
                           const float* rnn_states, const float* masks, const float* actions, const float* old_logp,
                           const float* adv, const float* active_masks, const float* available_actions,
                           float* grad_partials, int n_partials, double* loss_partials, void* workspace, void* stream);

int mappo_critic_update_rnn(const mappo_net_config* cfg, const mappo_loss_config* loss, const float* params,
                            const float* share_obs, int obs_ld, long long n_chunks, int chunk_len,
                            const int64_t* index, int share_rows_div, const float* rnn_states, const float* masks,
                            const float* value_preds, const float* returns, const float* active_masks,
                            const float* vn_stats, float* grad_partials, int n_partials, double* loss_partials,
                            void* workspace, void* stream);

/* Sum the per-block partials in a fixed order: grad_sum float32 [P], loss_sum double [4].  In a multi-GPU job
 * the caller all-reduces both (sum) before mappo_adam_step. */
int mappo_grad_reduce(const float* grad_partials, const double* loss_partials, int blocks, long long P,
                      float* grad_sum, double* loss_sum, void* stream);

/* The same with grad_sum in float64: a data-parallel job keeps [actor gradient sums | critic gradient sums | the two
 * double[4] loss sums] in ONE float64 buffer, so that a ppo_update needs ONE all-reduce instead of one
 * per dtype; mappo_adam_step_f64 reads it.  (ABI v6.) */
int mappo_grad_reduce_f64(const float* grad_partials, const double* loss_partials, int blocks, long long P,
                          double* grad_sum, double* loss_sum, void* stream);

typedef struct {
  float lr, beta1, beta2, eps, weight_decay;
  float max_grad_norm;         /* args.max_grad_norm */
  int32_t use_max_grad_norm;
  float bias_correction1;      /* 1 - beta1^t  (t = this step, computed on the host in double) */
  float bias_correction2_sqrt; /* sqrt(1 - beta2^t) */
  float one_minus_beta1;       /* float(1.0 - beta1) formed in double on the host: the scalar torch hands to lerp_ */
  float one_minus_beta2;       /* float(1.0 - beta2), the `value` of addcmul_ (1 - 0.999f in fp32 is off by 1.3e-5 relative) */
  double beta1_exact;          /* the betas in double: only read when the step count lives on the device (below) */
  double beta2_exact;
  int32_t frozen_head;         /* != 0: the head's weight / bias slots are not optimiser parameters -- they are left out of
                                * the global norm and of the update (their gradient still feeds the LayerNorm-affine chain
                                * rule in front of the head).  --use_popart on cuda: PopArt's weight / bias are plain
                                * tensors, absent from critic.parameters() (popart.py:21-27, rMAPPOPolicy.py:35-37).  ABI v6 */
  int32_t reserved2;
} mappo_adam_config;

/* From the reduced sums: finish the chain rule through the folded LayerNorm affines, divide by the loss
 * denominator (loss_sum[denominator_slot], or `rows_total` if use_mean != 0), clip the global L2 norm
 * (nn.utils.clip_grad_norm_, r_mappo.py:143-146) and apply torch.optim.Adam's update in place
 * (rMAPPOPolicy.py:31-37).  grad_out (float32 [P], REQUIRED: it is also the kernel's scratch) receives the final (unclipped) gradient;
 * info_out (float32 [4], may be NULL) is ACCUMULATED with {loss, grad_norm, aux (entropy | 0), ratio mean | 0}.
 * step_dev (int64 on the device, may be NULL): optimiser step count kept in HBM so that a captured CUDA graph of the
 * whole iteration replays correctly -- when given, the bias corrections are computed in the kernel from
 * t = *step_dev + 1 (the host-computed fields of `adam` are ignored) and *step_dev is incremented when
 * apply_update != 0.  (ABI v3.) */
int mappo_adam_step(const mappo_net_config* cfg, const mappo_adam_config* adam, float* params, float* exp_avg,
                    float* exp_avg_sq, const float* grad_sum, const double* loss_sum, int denominator_slot,
                    long long rows_total, int apply_update, float* grad_out, float* info_out, long long* step_dev,
                    void* stream);

int mappo_adam_step_f64(const mappo_net_config* cfg, const mappo_adam_config* adam, float* params, float* exp_avg,
                        float* exp_avg_sq, const double* grad_sum, const double* loss_sum, int denominator_slot,
                        long long rows_total, int apply_update, float* grad_out, float* info_out, long long* step_dev,
                        void* stream);

/* ------------------------------------------------------------------------------------------------
 * 4. Minibatch gather (the generator API; the update kernels take `index` and gather in place).
 *    Replaces the fancy-index gathers of SharedReplayBuffer.feed_forward_generator / recurrent_generator /
 *    naive_recurrent_generator   onpolicy/utils/shared_buffer.py:262-282,330-345,440-455
 *    dst[i][0:width] = src[(index[i] / row_div) * row_stride + 0:width],  i < n;  index is the int64 permutation
 *    (torch.randperm on the CPU generator, shared_buffer.py:246,301,397) copied to the device.
 * ---------------------------------------------------------------------------------------------- */
int mappo_gather_rows(const float* src, long long row_stride, int row_div, int width, const int64_t* index,
                      long long n, float* dst, void* stream);

/* One-hot `actions_env` rows -> float action ids on the device: replaces the inverse of the runner's
 * np.eye(n)[actions] (onpolicy/runner/shared/mpe_runner.py:110-121; MultiDiscrete: the heads' one-hots concatenated,
 * :111-117) that MultiAgentEnv._set_action undoes per agent (environment.py:194-256).  onehot: [rows][sum(head_dim)]
 * of float32 (dtype 0), float64 (1) or uint8 (2) on the DEVICE; head_dim_host: n_heads ints on the HOST;
 * ids: [rows][ids_ld] float32, column h = index of the first maximum of head h.  (ABI v6.) */
/* The forward direction on the DEVICE: float32 ids [rows][ids_ld] (column h = action of head h) -> uint8 one-hot
 * [rows][sum(head_dim)].  Used when the ids handed to the runner's one-hot helper are a host copy the library made of a
 * device tensor (host_mirror pass-through): the one-hot is built next to the ids and crosses PCIe once, as bytes. */
int mappo_ids_to_onehot_u8(const float* ids, long long rows, int ids_ld, int n_heads, const int* head_dim_host,
                           uint8_t* onehot, int32_t* bad_flag /* device, set to 1 by an id outside its head, or NULL */,
                           void* stream);
/* HOST helper (no device work): the runner's one-hot `actions_env = np.eye(n)[actions]`
 * (onpolicy/runner/shared/mpe_runner.py:110-121) as uint8 [rows][sum(head_dim)] from int64 ids [rows][n_heads], both on
 * the host; one pass in C instead of a NumPy fancy-index scatter.  (ABI v6.) */
int mappo_host_onehot_u8(const int64_t* ids_host, long long rows, int n_heads, const int* head_dim_host,
                         uint8_t* out_host);
/* the same from float32 ids (what the policy returns: R_Actor stores actions as floats, act.py:41-81), sparing the
 * runner an int64 conversion pass */
int mappo_host_onehot_u8_f32(const float* ids_host, long long rows, int n_heads, const int* head_dim_host,
                             uint8_t* out_host);

/* Up to 8 device -> device copies in ONE launch: SharedReplayBuffer.insert (onpolicy/utils/shared_buffer.py:121-147)
 * stores six arrays per step; as separate copies they cost six launches of host time each step of a host-driven
 * rollout.  Pointers 4-byte aligned, sizes multiples of 4 bytes; segments must not overlap.  (No reference counterpart:
 * the reference's insert is NumPy assignment.) */
typedef struct mappo_copy_list {
  int32_t n;                   /* segments in use, 1..8 */
  int32_t reserved;
  const void* src[8];
  void* dst[8];
  long long bytes[8];
} mappo_copy_list;
int mappo_copy_segments(const mappo_copy_list* list, void* stream);
int mappo_onehot_to_ids(const void* onehot, int dtype, long long rows, int n_heads, const int* head_dim_host,
                        float* ids, int ids_ld, void* stream);

/* ------------------------------------------------------------------------------------------------
 * 5. Hardware self-test of the tcgen05 (5th-gen tensor core) operand layouts the policy kernels rely on: stages
 *    tiles A [a_rows][a_cols] and B [b_rows][b_cols] (fp32, tf32-representable values) in the library's
 *    shared-memory layout, issues `tcgen05.mma` instructions (M = 128; kind::tf32 with K = 8 per instruction, or kind::f16 on bf16 with
 *    K = 16 when use_bf16 != 0) with the given descriptor
 *    strides and writes the fp32 accumulator D [128][N] back.  No reference counterpart (test infrastructure).
 * ---------------------------------------------------------------------------------------------- */
int mappo_selftest_umma(const float* A, int a_rows, int a_cols, const float* B, int b_rows, int b_cols, int M, int N,
                        int K, int a_mn_major, int b_mn_major, uint32_t a_lbo, uint32_t a_sbo, uint32_t a_kstep,
                        uint32_t b_lbo, uint32_t b_sbo, uint32_t b_kstep, float* D, int use_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* MAPPO_B200_H */
