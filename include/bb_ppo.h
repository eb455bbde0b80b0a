/*
 * bb_ppo.h — C ABI of the fused PPO update (libbball.so): the caller on the far side of the
 * rollout path, scripts/ppo.py `update_policy` (ppo.py:180-216).
 *
 * The reference runs each of its 16 minibatch steps per iteration (4 epochs x 4 minibatches of
 * 65 536 rows at the ppo.py defaults) as ~200 small torch kernels: Agent.get_stats
 * (agent.py:171-178) forward, the clipped-PPO loss (ppo.py:192-208), autograd backward,
 * clip_grad_norm_ and Adam (ppo.py:210-214).  After the rollout fusion (bb_rollout.h) that
 * update is 96 % of a PPO iteration.  Here a minibatch step is three kernels:
 *
 *   bb_ppo_grad   forward + loss + backward of the 5 972-parameter policy over the minibatch rows
 *                 (gathered through the permutation), per-block partial gradients, then a
 *                 fixed-order reduction -> one flat gradient.  Deterministic (no atomics).
 *   (optional)    one flat all-reduce of that gradient across ranks (torch.distributed / NCCL).
 *   bb_ppo_adam   global-norm clip (nn.utils.clip_grad_norm_) + torch.optim.Adam step on the flat
 *                 parameter vector.
 *
 * Plain pointers and sizes only; all pointers are DEVICE pointers.
 */
#ifndef BB_PPO_H
#define BB_PPO_H

#include <stdint.h>
#include "bb_rollout.h"

#ifdef __cplusplus
extern "C" {
#endif

/* Flat parameter / gradient layout for Agent(num_channels = 32, num_layers = 2): the order of
 * `agent.parameters()` (agent.py:111-127), each tensor row-major as torch stores it. */
#define BB_PPO_HIDDEN 32
#define BB_PPO_OFF_W1 0        /* backbone.0.weight [32,128] */
#define BB_PPO_OFF_B1 4096     /* backbone.0.bias   [32]     */
#define BB_PPO_OFF_G1 4128     /* backbone.1.weight [32]  (LayerNorm gamma) */
#define BB_PPO_OFF_E1 4160     /* backbone.1.bias   [32]  (LayerNorm beta)  */
#define BB_PPO_OFF_W2 4192     /* backbone.3.weight [32,32]  */
#define BB_PPO_OFF_B2 5216     /* backbone.3.bias   [32]     */
#define BB_PPO_OFF_G2 5248     /* backbone.4.weight [32]     */
#define BB_PPO_OFF_E2 5280     /* backbone.4.bias   [32]     */
#define BB_PPO_OFF_WA 5312     /* actor.weight      [19,32]  */
#define BB_PPO_OFF_BA 5920     /* actor.bias        [19]     */
#define BB_PPO_OFF_WC 5939     /* critic.weight     [1,32]   */
#define BB_PPO_OFF_BC 5971     /* critic.bias       [1]      */
#define BB_PPO_NUM_PARAMS 5972
#define BB_PPO_NUM_STATS 4     /* sums over the minibatch rows: pg loss, value loss (x0.5 max), entropy, clipped-ratio count */

/* One minibatch of scripts/buffers.py RolloutBuffer.get_minibatch (flattened [T*N, ...] float32). */
typedef struct bb_ppo_batch {
    const float *obs;         /* [R,128]  raw observations (normalised inside, agent.py:27-39) */
    const float *actions;     /* [R,6]    float, as the reference buffer stores them */
    const float *log_probs;   /* [R]      behaviour log-prob (ppo.py:189 `lp`) */
    const float *values;      /* [R]      behaviour value   (ppo.py:189 `v`)  */
    const float *advantages;  /* [R]      normalised advantages (ppo.py:173-174) */
    const float *returns;     /* [R]      normalised returns    (ppo.py:176) */
    const int64_t *indices;   /* [m]      rows of this minibatch (torch.randperm slice, ppo.py:186-188); NULL = indices32 or 0..m-1 */
    int64_t num_rows;         /* m */
    float clip_coef;          /* ppo.py:48  0.2  */
    float ent_coef;           /* ppo.py:49  0.01 */
    float vf_coef;            /* ppo.py:50  1.0  */
    const int32_t *indices32; /* [m]      the same as 32-bit indices (a slice of bb_randperm's output); used when indices is NULL */
} bb_ppo_batch;

/* Bytes of scratch `bb_ppo_grad` needs on the current device (per-block partial gradients). */
int64_t bb_ppo_workspace_bytes(void);

/* loss = pg_loss + vf_coef * c_loss + ent_coef * entropy_loss (ppo.py:192-208) over the minibatch,
 * means over its m rows; grad_out[BB_PPO_NUM_PARAMS] = d loss / d parameters in the flat layout
 * above (what loss.backward() leaves in .grad); stats_out[BB_PPO_NUM_STATS] (may be NULL) = sums
 * over rows of: max(surr1, surr2); 0.5 * max(vf, vf_clipped); entropy; [ratio outside the clip]. */
int bb_ppo_grad(const bb_policy *pol, const bb_ppo_batch *batch, void *workspace, float *grad_out, float *stats_out,
                void *cuda_stream);

/* nn.utils.clip_grad_norm_(params, max_grad_norm) followed by torch.optim.Adam.step() (no weight
 * decay, no amsgrad) on flat vectors of n floats.  grad is first multiplied by grad_scale (1 / world
 * size after a summing all-reduce).  step_dev: int32 step counter on the device, incremented here.
 * grad_norm_out (may be NULL): the total norm before clipping. */
int bb_ppo_adam(float *params, const float *grad, float *exp_avg, float *exp_avg_sq, int32_t *step_dev, int64_t n,
                float grad_scale, float max_grad_norm, float lr, float beta1, float beta2, float eps,
                float *grad_norm_out, void *cuda_stream);

/* ---- compute_advantages after the GAE scan (ppo.py:167-176) ----
 * BB_PPO_MOMENT_COLS columns: the 128 observation columns, then values (un-normalised), returns, advantages. */
#define BB_PPO_MOMENT_COLS 131

/* Bytes of scratch bb_ppo_moments needs on the current device. */
int64_t bb_ppo_moments_workspace_bytes(void);

/* One pass over this rank's rollout rows: sums_out[c] = sum over rows of column c, sums_out[131 + c] = sum of
 * squares, float64, block-order reduction (bitwise reproducible).  Sums of several ranks add (one all-reduce). */
int bb_ppo_moments(const float *obs, const float *values_unnorm, const float *returns, const float *advantages, int64_t rows,
                   void *workspace, double *sums_out, void *cuda_stream);

/* bb_ppo_moments without a workspace argument, for callers that cannot allocate (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * The per-block partial sums go to a static device array inside the library
 * (BB_PPO_MOMENTS_STATIC_MAX_BLOCKS x 2 x 131 doubles = 1 073 152 bytes, held by every GPU that loads libbball.so until
 * the process exits); the call itself performs no cudaMalloc / cudaFree and calls no allocating library routine.
 * Same kernels, same grid and therefore the same bits in sums_out as bb_ppo_moments.
 * Limits: rows <= BB_PPO_MOMENTS_STATIC_MAX_ROWS and a device whose grid (2 blocks per SM) fits
 * BB_PPO_MOMENTS_STATIC_MAX_BLOCKS, i.e. up to 256 SMs; beyond either the call fails with a message in bb_last_error().
 * ALL calls on a device share the one array (it keeps the previous call's partials; nothing needs clearing because
 * every block overwrites its whole partial and only those are read): calls that overlap on the GPU -- different
 * streams, or different threads / processes' graphs replayed concurrently on one context -- race.  Serialise them on
 * one stream, or use bb_ppo_moments with a workspace per stream. */
#define BB_PPO_MOMENTS_STATIC_MAX_BLOCKS 512
#define BB_PPO_MOMENTS_STATIC_MAX_ROWS (1 << 24)
int bb_ppo_moments_static(const float *obs, const float *values_unnorm, const float *returns, const float *advantages,
                          int64_t rows, double *sums_out, void *cuda_stream);

/* RunningMeanStd state (agent.py:24-26): float64 mean[dim], var[dim], count[1] on the device, updated in place. */
typedef struct bb_running_norm { double *mean, *var, *count; } bb_running_norm;

/* ppo.py:167-176 given GLOBAL sums over `global_rows` rows (all ranks): obs_norm.update(obs); value_norm.update(values);
 * value_norm.update(returns) (agent.py:41-51, unbiased batch variance); advantages = (adv - mean) / (std + 1e-8);
 * values, returns = value_norm(.) with the UPDATED statistics, clamped to +-value_clamp (agent.py:36-39), on this rank's
 * `rows` elements in place.  obs_mean_f32 / obs_inv_std_f32 [128]: the float32 vectors bb_policy reads (agent.py:28-30,36).
 * scratch: 4 floats on the device. */
int bb_ppo_finish_advantages(const double *sums, double global_rows, const bb_running_norm *obs_norm,
                             const bb_running_norm *value_norm, float *obs_mean_f32, float *obs_inv_std_f32, float value_clamp,
                             float *values_unnorm, float *returns, float *advantages, int64_t rows, float *scratch,
                             void *cuda_stream);

/* torch.randperm(n) of ppo.py:186 as two kernels: element i gets the 32-bit key Threefry2x32(key (seed, stream),
 * counter (i, 0)).x0 and (key, i) pairs are radix-sorted (cub::DeviceRadixSort, stable).  A uniformly random
 * permutation up to ties between equal keys (n^2 / 2^33 expected pairs: 8 at n = 262 144), which keep index order.
 * perm_out: int32 [n].  107 us -> ~25 us per epoch at n = 262 144. */
int64_t bb_randperm_workspace_bytes(int64_t n);
int bb_randperm(int64_t n, uint32_t seed, uint32_t stream, void *workspace, int64_t workspace_bytes, int32_t *perm_out,
                void *cuda_stream);

/* ---- gradient all-reduce fused into the optimiser step, over NVLink peer memory (multi-GPU PPO) ----
 * The reference is single-GPU; with one process per GPU the 24 KB flat gradient has to be summed across ranks before
 * every Adam step (16 times per PPO iteration).  NCCL does that in ~30 us of latency per call; here the collective is
 * part of the optimiser kernel: every rank stores its gradient, as 8-byte words { value, sequence number }, straight into
 * a slot of every peer's exchange buffer (device memory mapped through CUDA IPC; remote stores pipeline over NVLink), then
 * polls its own local copy of the slots until every word carries the call's sequence number, and adds the ranks' values
 * in rank order (every rank computes the same bits) before the clip + Adam arithmetic of bb_ppo_adam.  One kernel, one
 * one-way NVLink trip, no fence, no NCCL call.
 *
 *   bb_p2p_create        allocates this rank's exchange buffer, returns an opaque context and writes the 64-byte CUDA IPC
 *                        handle of the buffer to handle_out (exchange the handles with any host-side all-gather)
 *   bb_p2p_connect       opens the peers' buffers (handles: world_size x 64 bytes in rank order)
 *   bb_ppo_adam_p2p      bb_ppo_adam with grad := (1 / world_size) * sum over ranks of grad.  Every rank must call it the
 *                        same number of times.  A peer that does not show up within timeout_ms makes the kernel give up
 *                        WITHOUT applying the step and sets *error_flag_dev (int32, device) to 1. */
typedef struct bb_p2p bb_p2p;
bb_p2p *bb_p2p_create(int32_t rank, int32_t world_size, int64_t num_floats, unsigned char handle_out[64]);
int bb_p2p_connect(bb_p2p *ctx, const unsigned char *handles);
void bb_p2p_destroy(bb_p2p *ctx);
int bb_ppo_adam_p2p(bb_p2p *ctx, float *params, const float *grad, float *exp_avg, float *exp_avg_sq, int32_t *step_dev, int64_t n,
                    float max_grad_norm, float lr, float beta1, float beta2, float eps, float *grad_norm_out,
                    int32_t *error_flag_dev, int32_t timeout_ms, void *cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* BB_PPO_H */
