/*
 * bb_rollout.cuh — fused policy forward / rollout bookkeeping / GAE kernels (include/bb_rollout.h).
 *
 * Included at the end of bb_engine.cu (it enqueues the step kernel through launch_step()).
 * Reference semantics: scripts/agent.py:108-154 (Agent), scripts/action.py (bucketed categoricals),
 * scripts/ppo.py:60-166 (rollout, compute_advantages), scripts/buffers.py, scripts/ppo_stats.py:21-36.
 *
 * bb_policy_kernel: one WARP per observation row, lane j = hidden channel j (the reference's
 * num_channels is 32 = one warp).  Lane j keeps row j of W1 (128 floats) and of W2 (32) in
 * registers for the whole kernel, so layer 1 is 128 FMAs per row against a broadcast read of the
 * normalised observation from shared memory; LayerNorm statistics are warp-shuffle reductions;
 * the two heads read a transposed [32][20] copy of (actor | critic) from shared memory; lanes 0..5
 * each own one action bucket for the softmax / inverse-CDF sampling.  fp32 FFMA throughout: the
 * contraction is 5.7 kFLOP per row — at 8192 rows a few microseconds on the CUDA cores — and
 * agreement with the torch fp32 reference to ~1e-6 matters more here than tensor-core peak.
 */
#pragma once
#include "../../include/bb_rollout.h"

namespace bb {

constexpr int kPolThreads = 128;
constexpr int kPolWarps = kPolThreads / 32;
constexpr int kHeads = BB_NUM_LOGITS + 1; /* 19 logits + value */

struct PolicyIo {
    const float *obs;
    int64_t obsStride, n, rowOffset;
    int stochastic, evaluateOnly;
    uint32_t seed, draw;
    int32_t *actionsOut;
    int64_t actionStride;
    float *logpOut, *valueOut, *logitsOut;
    float *bufObs, *bufActions; /* rollout buffer rows of this step */
    /* bookkeeping of the PREVIOUS step (its reward / done are in the engine columns now) */
    const float *rew, *done;
    int64_t rdStride;
    float *bufRewPrev, *bufNotDonePrev;
    float *curReward, *epLength;
    double *epSums;
    int32_t *resetCol; /* both Reset flags of every world are cleared (end of env.reset(), env.py:182) */
    float *nextValue;
};

__device__ __constant__ int kBucketOff[BB_NUM_ACTION_DIMS] = {0, 2, 10, 13, 15, 17}; /* env.py:102 */
__device__ __constant__ int kBucketCnt[BB_NUM_ACTION_DIMS] = {2, 8, 3, 2, 2, 2};

BBD float warp_sum(float v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

/* nn.LayerNorm(H) + ReLU on the value each lane < H holds (biased variance, eps 1e-5) */
BBD float layer_norm_relu(float v, bool active, int H, float g, float b)
{
    const float inv = 1.f / (float)H;
    const float mean = warp_sum(active ? v : 0.f) * inv;
    const float d = active ? v - mean : 0.f;
    const float var = warp_sum(d * d) * inv;
    const float y = d * rsqrtf(var + 1e-5f) * g + b;
    return active ? fmaxf(y, 0.f) : 0.f;
}

__global__ void __launch_bounds__(kPolThreads, 2) bb_policy_kernel(const bb_policy P, const PolicyIo io)
{
    __shared__ float sMean[BB_OBS_DIM], sInv[BB_OBS_DIM];
    __shared__ float sHeadT[BB_POLICY_MAX_HIDDEN][kHeads]; /* [k][j]: j < 19 actor rows, j == 19 critic */
    __shared__ float sHeadB[kHeads];
    __shared__ __align__(16) float sX[kPolWarps][BB_OBS_DIM];
    __shared__ float sH[kPolWarps][BB_POLICY_MAX_HIDDEN];
    __shared__ float sL[kPolWarps][kHeads];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int H = P.hidden;
    const bool active = lane < H;

    for (int i = threadIdx.x; i < BB_OBS_DIM; i += kPolThreads) { sMean[i] = P.obs_mean[i]; sInv[i] = P.obs_inv_std[i]; }
    for (int i = threadIdx.x; i < BB_POLICY_MAX_HIDDEN * kHeads; i += kPolThreads) {
        const int k = i / kHeads, j = i % kHeads;
        float w = 0.f;
        if (k < H) w = (j < BB_NUM_LOGITS) ? P.wa[j * H + k] : P.wc[k];
        sHeadT[k][j] = w;
    }
    if (threadIdx.x < kHeads) sHeadB[threadIdx.x] = threadIdx.x < BB_NUM_LOGITS ? P.ba[threadIdx.x] : P.bc[0];

    /* this lane's rows of W1 / W2 and its per-channel vectors, register-resident */
    float w1[BB_OBS_DIM], w2[BB_POLICY_MAX_HIDDEN];
#pragma unroll
    for (int k = 0; k < BB_OBS_DIM; k += 4) {
        float4 v = active ? *reinterpret_cast<const float4 *>(P.w1 + (int64_t)lane * BB_OBS_DIM + k) : make_float4(0.f, 0.f, 0.f, 0.f);
        w1[k] = v.x; w1[k + 1] = v.y; w1[k + 2] = v.z; w1[k + 3] = v.w;
    }
#pragma unroll
    for (int k = 0; k < BB_POLICY_MAX_HIDDEN; k++) w2[k] = (active && k < H) ? P.w2[lane * H + k] : 0.f;
    const float b1 = active ? P.b1[lane] : 0.f, g1 = active ? P.ln1_g[lane] : 0.f, e1 = active ? P.ln1_b[lane] : 0.f;
    const float b2 = active ? P.b2[lane] : 0.f, g2 = active ? P.ln2_g[lane] : 0.f, e2 = active ? P.ln2_b[lane] : 0.f;
    __syncthreads();

    const int64_t warpsTotal = (int64_t)gridDim.x * kPolWarps;
    for (int64_t r = (int64_t)blockIdx.x * kPolWarps + warp; r < io.n; r += warpsTotal) {
        /* ---- observation row: raw copy into the rollout buffer, normalise (agent.py:27-39) ---- */
        const float *orow = io.obs + r * io.obsStride;
#pragma unroll
        for (int i = 0; i < 4; i++) {
            const int k = lane + 32 * i;
            const float x = orow[k];
            if (io.bufObs) io.bufObs[r * BB_OBS_DIM + k] = x;
            float xn = (x - sMean[k]) * sInv[k];
            if (P.obs_clamp > 0.f) xn = fminf(fmaxf(xn, -P.obs_clamp), P.obs_clamp);
            sX[warp][k] = xn;
        }
        __syncwarp();
        /* ---- backbone.0-2: Linear(128,H) + LayerNorm + ReLU ---- */
        float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
        for (int k = 0; k < BB_OBS_DIM; k += 4) {
            const float4 x = *reinterpret_cast<const float4 *>(&sX[warp][k]);
            a0 = fmaf(w1[k], x.x, a0); a1 = fmaf(w1[k + 1], x.y, a1);
            a2 = fmaf(w1[k + 2], x.z, a2); a3 = fmaf(w1[k + 3], x.w, a3);
        }
        float h = layer_norm_relu(((a0 + a1) + (a2 + a3)) + b1, active, H, g1, e1);
        sH[warp][lane] = h;
        __syncwarp();
        /* ---- backbone.3-5: Linear(H,H) + LayerNorm + ReLU ---- */
        a0 = a1 = 0.f;
#pragma unroll
        for (int k = 0; k < BB_POLICY_MAX_HIDDEN; k += 2) {
            a0 = fmaf(w2[k], sH[warp][k], a0);
            a1 = fmaf(w2[k + 1], sH[warp][k + 1], a1);
        }
        __syncwarp();
        h = layer_norm_relu((a0 + a1) + b2, active, H, g2, e2);
        sH[warp][lane] = h;
        __syncwarp();
        /* ---- heads: lanes 0..18 actor logits, lane 19 critic ---- */
        float out = 0.f;
        if (lane < kHeads) {
            a0 = a1 = 0.f;
#pragma unroll
            for (int k = 0; k < BB_POLICY_MAX_HIDDEN; k += 2) {
                a0 = fmaf(sHeadT[k][lane], sH[warp][k], a0);
                a1 = fmaf(sHeadT[k + 1][lane], sH[warp][k + 1], a1);
            }
            out = (a0 + a1) + sHeadB[lane];
            sL[warp][lane] = out;
            if (io.logitsOut && lane < BB_NUM_LOGITS) io.logitsOut[r * BB_NUM_LOGITS + lane] = out;
        }
        __syncwarp();
        if (lane == BB_NUM_LOGITS) {
            if (io.evaluateOnly) { if (io.nextValue) io.nextValue[r] = out; }
            else if (io.valueOut) io.valueOut[r] = out;
        }
        /* ---- bucketed categoricals (action.py:5-33): lane b owns bucket b ---- */
        float logp = 0.f;
        if (!io.evaluateOnly && lane < BB_NUM_ACTION_DIMS) {
            const int off = kBucketOff[lane], cnt = kBucketCnt[lane];
            const float *l = &sL[warp][off];
            float m = l[0];
            for (int i = 1; i < cnt; i++) m = fmaxf(m, l[i]);
            float s = 0.f;
            for (int i = 0; i < cnt; i++) s += expf(l[i] - m);
            const float lse = m + logf(s);
            int a = 0;
            if (io.stochastic) {
                uint32_t x0, x1;
                threefry2x32(io.seed, (uint32_t)(io.rowOffset + r), io.draw, (uint32_t)lane, x0, x1);
                const float u = (float)((x0 ^ x1) >> 8) * 0x1p-24f;
                float c = 0.f;
                a = cnt - 1;
                for (int i = 0; i < cnt; i++) {
                    c += expf(l[i] - lse);
                    if (u < c) { a = i; break; }
                }
            } else {
                for (int i = 1; i < cnt; i++) if (l[i] > l[a]) a = i;
            }
            logp = l[a] - lse;
            if (io.actionsOut) io.actionsOut[r * io.actionStride + lane] = a;
            if (io.bufActions) io.bufActions[r * BB_NUM_ACTION_DIMS + lane] = (float)a;
        }
        if (!io.evaluateOnly) {
            /* log_probs.sum(dim=-1), agent.py:152: fixed order bucket 0..5 */
            float tot = 0.f;
#pragma unroll
            for (int b = 0; b < BB_NUM_ACTION_DIMS; b++) tot += __shfl_sync(0xffffffffu, logp, b);
            if (lane == 0 && io.logpOut) io.logpOut[r] = tot;
        }
        /* ---- bookkeeping of the previous step (ppo.py:84-85,128-134; ppo_stats.py:21-36) ---- */
        if (lane == 0 && io.bufRewPrev) {
            const float rw = io.rew[r * io.rdStride], dn = io.done[r * io.rdStride];
            io.bufRewPrev[r] = rw;
            io.bufNotDonePrev[r] = 1.0f - dn;
            if (io.curReward) {
                const float cr = io.curReward[r] + rw, el = io.epLength[r] + 1.f;
                if (dn != 0.f) {
                    atomicAdd(&io.epSums[0], (double)cr);
                    atomicAdd(&io.epSums[1], (double)el);
                    atomicAdd(&io.epSums[2], 1.0);
                }
                io.curReward[r] = cr * (1.f - dn);
                io.epLength[r] = el * (1.f - dn);
            }
        }
        if (lane < 2 && io.resetCol) io.resetCol[r * 2 + lane] = 0;
        __syncwarp();
    }
}

/* env.reset() prologue (env.py:179-180): Reset flags = 1, trainee action = 0 */
__global__ void bb_reset_prologue_kernel(int32_t *resetCol, int32_t *actionCol, int64_t n, int agentIdx)
{
    const int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= n) return;
    resetCol[w * 2] = 1;
    resetCol[w * 2 + 1] = 1;
    int2 *a = reinterpret_cast<int2 *>(actionCol + (w * 2 + agentIdx) * 6);
    a[0] = make_int2(0, 0); a[1] = make_int2(0, 0); a[2] = make_int2(0, 0);
}

/* compute_advantages (ppo.py:144-166): thread per environment, reverse scan over the horizon */
__global__ void bb_gae_kernel(const float *values, const float *nextValue, const float *rewards, const float *notDones, int T,
                              int64_t n, float gamma, float gammaLam, const double *vMean, const double *vVar, float clampV,
                              float *valuesUn, float *adv, float *ret)
{
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n) return;
    /* RunningMeanStd.forward(unnorm=True), agent.py:27-35 */
    const float mean = (float)vMean[0];
    const float sd = sqrtf((float)vVar[0] + 1e-5f);
    auto unnorm = [&](float v) {
        if (clampV > 0.f) v = fminf(fmaxf(v, -clampV), clampV);
        return mean + sd * v;
    };
    float nextV = unnorm(nextValue[e]);
    float nextNonTerminal = notDones[(int64_t)(T - 1) * n + e]; /* t == T-1 uses its own not_done (ppo.py:156-158) */
    float last = 0.f;
    for (int t = T - 1; t >= 0; t--) {
        const int64_t i = (int64_t)t * n + e;
        const float v = unnorm(values[i]);
        const float delta = rewards[i] + gamma * nextV * nextNonTerminal - v;
        last = delta + gammaLam * nextNonTerminal * last; /* python: gamma * gae_lambda is one double scalar */
        adv[i] = last;
        ret[i] = last + v;
        if (valuesUn) valuesUn[i] = v;
        nextV = v;
        nextNonTerminal = notDones[i]; /* for step t-1: next_non_terminal = not_dones[t] (ppo.py:160) */
    }
}

} // namespace bb
