/*
 * bb_rollout.h — C ABI of the fused rollout path (libbball.so), SURVEY.md section 8(f) rows n1 + n2.
 *
 * The callers of the step path in the reference are plain torch: scripts/ppo.py `rollout()`
 * (ppo.py:60-141) runs Agent.forward (scripts/agent.py:134-154: observation normaliser ->
 * Linear 128->H, LayerNorm, ReLU -> Linear H->H, LayerNorm, ReLU -> actor head (19 logits in
 * buckets [2,8,3,2,2,2], scripts/env.py:102) + critic head -> six categorical samples,
 * scripts/action.py:30-33) as ~15 small kernels per step, copies observations / rewards /
 * dones three times (env.py:167-169), and synchronises with the host every step
 * (ppo_stats.py:26 `dones.any()`).  Here one kernel does the whole forward + sampling for
 * every environment, reads the observation rows straight from the step engine's column,
 * writes the sampled action into the engine's Action column and fills the rollout buffer
 * (scripts/buffers.py:4-21); a whole T-step rollout is ONE kernel launch when its grid is resident at once (each block
 * takes its worlds through policy rows and steps in turn; otherwise 2 kernels per step), with no host
 * synchronisation, and compute_advantages' GAE scan (ppo.py:144-166) is one kernel.
 *
 * Plain pointers and sizes only.  All pointers are DEVICE pointers unless stated.
 */
#ifndef BB_ROLLOUT_H
#define BB_ROLLOUT_H

#include <stdint.h>
#include "bb_engine.h"

#ifdef __cplusplus
extern "C" {
#endif

#define BB_OBS_DIM 128        /* Observations width, types.hpp:164-167 */
#define BB_NUM_LOGITS 19      /* sum of the action buckets [2,8,3,2,2,2], env.py:102 */
#define BB_NUM_ACTION_DIMS 6  /* Action fields, types.hpp Action / mgr.cpp:332 */
#define BB_POLICY_MAX_HIDDEN 32 /* ppo.py:256 builds Agent(num_channels=32, num_layers=2) */

/* Parameters of scripts/agent.py Agent (num_layers = 2), torch layouts (nn.Linear.weight is [out, in],
 * row-major), float32.  The caller owns the memory (typically the torch parameters themselves). */
typedef struct bb_policy {
    int32_t hidden;            /* num_channels; 1..32 */
    const float *obs_mean;     /* [128]  obs_norm.mean as float32                     agent.py:29 */
    const float *obs_inv_std;  /* [128]  rsqrt(obs_norm.var as float32 + 1e-5)         agent.py:30,36 */
    float obs_clamp;           /* 5.0 (agent.py:130); 0 = no clamp                     agent.py:37-38 */
    const float *w1, *b1;      /* backbone.0  Linear(128, H)      [H,128], [H]         agent.py:115 */
    const float *ln1_g, *ln1_b;/* backbone.1  LayerNorm(H)        [H], [H]  eps 1e-5   agent.py:116 */
    const float *w2, *b2;      /* backbone.3  Linear(H, H)        [H,H], [H]           agent.py:120 */
    const float *ln2_g, *ln2_b;/* backbone.4  LayerNorm(H)                              agent.py:121 */
    const float *wa, *ba;      /* actor       Linear(H, 19)       [19,H], [19]         agent.py:126 */
    const float *wc, *bc;      /* critic      Linear(H, 1)        [1,H], [1]           agent.py:127 */
} bb_policy;

/* One Agent.forward over n observation rows (agent.py:134-154).
 *   obs            f32, row r at obs + r * obs_stride (obs_stride = 256 for the view observations[:, agent_idx])
 *   stochastic     1: sample each bucket by inverse CDF; Threefry2x32(key (seed, row_offset + r), counter (draw, c)),
 *                     c = 0..2, gives the 32-bit words of buckets 2c and 2c+1, u = (word >> 8) * 2^-24
 *                     (action.py:30-33 draws with torch.multinomial; the torch RNG stream is not reproducible
 *                     outside torch, the distribution is the same);
 *                  0: argmax per bucket (action.py:21-23)
 *   actions_out    i32, 6 per row at actions_out + r * action_stride (may be NULL); pointing it at
 *                  Action column + agent_idx * 6 with stride 12 is `actions[:, agent_idx] = a` (env.py:147)
 *   logp_out       f32 [n]  sum over the six buckets of log_prob(action)   (may be NULL)
 *   value_out      f32 [n]  critic output                                    (may be NULL)
 *   logits_out     f32 [n,19] raw actor output                               (may be NULL; tests) */
int bb_policy_forward(const bb_policy *pol, const float *obs, int64_t obs_stride, int64_t n, int stochastic,
                      uint32_t seed, uint32_t draw, int64_t row_offset, int32_t *actions_out, int64_t action_stride,
                      float *logp_out, float *value_out, float *logits_out, void *cuda_stream);

/* scripts/buffers.py RolloutBuffer: [T, N, ...] float32, step-major. */
typedef struct bb_rollout_buffers {
    int32_t horizon;    /* T = num_rollout_steps (ppo.py:43: 32) */
    float *obs;         /* [T,N,128] */
    float *actions;     /* [T,N,6]   float, like the reference buffer */
    float *values;      /* [T,N]     critic output (normalised scale) */
    float *log_probs;   /* [T,N] */
    float *rewards;     /* [T,N] */
    float *not_dones;   /* [T,N]     1 - done */
    float *next_value;  /* [N]       agent.evaluate(obs after the last step), ppo.py:136-137 */
    /* episode statistics (scripts/ppo_stats.py:21-36), persistent across rollouts */
    float *cur_reward;  /* [N] running episode return */
    float *ep_length;   /* [N] running episode length */
    double *ep_sums;    /* [3] sums over finished episodes: return, length, count */
} bb_rollout_buffers;

/* Test hook for the tcgen05 building blocks of bb_policy_forward's tensor-core path (csrc/bb_umma.cuh):
 * D[128 x 32] = A[128 x K] . B[32 x K]^T as 3xTF32 through tcgen05.mma (kind::tf32, M = 128, N = 32, accumulator in tensor
 * memory, read back with tcgen05.ld).  A, B, D: device pointers, f32, row-major; K a multiple of 8, 8..128. */
int bb_umma_selftest(const float *a_dev, const float *b_dev, float *d_dev, int k, void *cuda_stream);

/* ppo.py rollout() (ppo.py:60-141) on the engine's worlds for the trainee `agent_idx`:
 * env.reset() (env.py:178-185: Reset flags = 1, zero trainee action, one step, flags = 0), then T x
 * { Agent.forward on observations[:, agent_idx]; actions[:, agent_idx] = a; [frozen opponent:
 * actions[:, 1 - agent_idx] = frozen.forward(observations[:, 1 - agent_idx]), env.py:126-143];
 * Manager::step; record obs / action / value / log-prob / reward / not-done; episode statistics },
 * then next_value = critic(observations after the last step).  Everything is enqueued on
 * `cuda_stream` (one launch of the looping kernel when the grid is resident at once -- up to ~9400 envs on a B200 --
 * otherwise 2 kernels per step, 3 with a frozen opponent; BB_ROLLOUT_LOOP=0 forces the chain) without host synchronisation.
 * `frozen` may be NULL (the other agent then keeps the action hardCodeDefenseSystem wrote).
 * `draw_base`: the sampling streams use draws draw_base .. draw_base + T - 1 (trainee) and
 * 0x80000000 + the same (frozen opponent). */
int bb_rollout(bb_engine *e, const bb_policy *trainee, const bb_policy *frozen, const bb_rollout_buffers *buf,
               int32_t agent_idx, uint32_t seed, uint32_t draw_base, int64_t world_offset, void *cuda_stream);

/* compute_advantages (ppo.py:144-166) up to `returns = advantages + values`:
 * values are first un-normalised with the value normaliser (agent.py:31-35: mean + sqrt(var + 1e-5) *
 * clamp(v, +-5)), then the reverse GAE scan per environment.
 *   value_mean / value_var   f64 [1], RunningMeanStd buffers of value_norm (agent.py:24-25)
 *   values_unnorm_out        f32 [T,N] (the `values` temporary of ppo.py:149; feeds value_norm.update)
 *   advantages_out, returns_out  f32 [T,N] (un-normalised; ppo.py:168-176 stay in torch) */
int bb_gae(const bb_rollout_buffers *buf, int64_t n, double gamma, double gae_lambda, const double *value_mean,
           const double *value_var, float value_clamp, float *values_unnorm_out, float *advantages_out,
           float *returns_out, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* BB_ROLLOUT_H */
