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

#ifdef __cplusplus
}
#endif
#endif /* MAPPO_B200_H */
