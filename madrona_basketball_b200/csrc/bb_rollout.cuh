/*
 * bb_rollout.cuh — fused policy forward / rollout bookkeeping / GAE kernels (include/bb_rollout.h).
 *
 * Included at the end of bb_engine.cu (it enqueues the step kernel through launch_step()).
 * Reference semantics: scripts/agent.py:108-154 (Agent), scripts/action.py (bucketed categoricals),
 * scripts/ppo.py:60-166 (rollout, compute_advantages), scripts/buffers.py, scripts/ppo_stats.py:21-36.
 *
 * bb_policy_kernel: four threads per observation row against broadcast 16-byte reads of the weights in
 * shared memory (see the kernel).  The contraction is 5.7 kFLOP per row; agreement with the torch fp32
 * reference to ~1e-6 matters more here than tensor-core peak, so the three products run as 3xTF32
 * (mma.sync, operands split hi + lo, fp32 accumulate: fp32-level accuracy; quad_mma_tf32x3 below), with
 * the FFMA form kept under BB_POL_MMA=0.
 */
#pragma once
#include "../../include/bb_rollout.h"

namespace bb {

#ifndef BB_POL_THREADS
#define BB_POL_THREADS 256
#endif
constexpr int kPolThreads = BB_POL_THREADS; /* multiple of 32, >= 224 (the staging code addresses threads 0..211) */
constexpr int kHeads = BB_NUM_LOGITS + 1; /* 19 logits + value */

struct PolicyIo {
    const float *obs;
    int64_t obsStride, n, rowOffset;
    int stochastic, evaluateOnly;
    int earlyTrigger;         /* 1: griddepcontrol.launch_dependents once this grid runs (the following step kernel sits resident) */
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
    /* policy_rows8<true> (bb_rollout_loop_kernel): the tile's observation image in shared memory, the tile's first world and
     * which of a world's two rows (agent index) to read */
    const float4 *obsImg;
    int64_t obsImgWorld0;
    int obsImgRow;
};

/* ------------------------------------------------------------------------------------------------
 * Four threads ("quad") per observation row; thread q of the quad owns hidden channels q, q+4, .., q+28
 * (and head outputs q, q+4, .., q+16).  The weights stay in torch's [out][in] layout in shared memory
 * (rows padded by 4 floats): a thread reads its channel's weights for four consecutive inputs as one
 * 16-byte load - the same address for every row of the block (broadcast), bank-disjoint across the
 * four q - so each shared load feeds four FMAs and nothing is transposed.  LayerNorm statistics are
 * two xor-shuffles inside the quad; the hidden vector passes through a private shared row; the six
 * softmax buckets are spread over the quad.  At 8192 rows that is 1024 warps of ~3 k instructions
 * (history: warp-per-row 40 us, thread-per-row 29 us at 8192 rows - too few, too long warps).
 * ------------------------------------------------------------------------------------------------ */
constexpr int kH = BB_POLICY_MAX_HIDDEN;
constexpr int kCh = kH / 4;                 /* channels per thread */
constexpr int kRowsPerBlock = kPolThreads / 4;
constexpr int kW1Stride = BB_OBS_DIM + 4;   /* 132: rows q, q+1, .. land 4 banks apart */
constexpr int kW2Stride = kH + 4;           /* 36 */
constexpr int kHeadPerThread = kHeads / 4;  /* 5 */

template <int kRows>
struct __align__(16) PolicySmemT {
    float w1[kH][kW1Stride];
    float w2[kH][kW2Stride];
    float wh[kHeads][kW2Stride];            /* rows 0..18 actor, row 19 critic */
    float hid[kRows][kW2Stride];            /* a row's hidden vector (private to its quad) */
    float out[kRows][kHeads + 4];           /* a row's 19 logits + value */
    float mean[BB_OBS_DIM], inv[BB_OBS_DIM];
    float b1[kH], g1[kH], e1[kH], b2[kH], g2[kH], e2[kH], bh[kHeads];
};
using PolicySmem = PolicySmemT<kRowsPerBlock>;

BBD float quad_sum(float v)
{
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    return v;
}

/* nn.LayerNorm(H) + ReLU over the quad's 32 channels (thread q holds channels 4c + q); biased variance, eps 1e-5 */
BBD void layer_norm_relu(float (&v)[kCh], int q, int H, const float *g, const float *b)
{
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kCh; c++) s += v[c]; /* channels >= H hold 0 */
    const float inv = 1.f / (float)H;
    const float mean = quad_sum(s) * inv;
    s = 0.f;
#pragma unroll
    for (int c = 0; c < kCh; c++) {
        v[c] = (4 * c + q < H) ? v[c] - mean : 0.f;
        s = fmaf(v[c], v[c], s);
    }
    const float rstd = rsqrtf(quad_sum(s) * inv + 1e-5f);
#pragma unroll
    for (int c = 0; c < kCh; c++) v[c] = (4 * c + q < H) ? fmaxf(fmaf(v[c] * rstd, g[4 * c + q], b[4 * c + q]), 0.f) : 0.f;
}

/* ---- row-phase GEMVs on the tensor cores, fp32-accurate (3xTF32) ----
 * ncu (profiles/r01g_ppo_grad_*) showed the FFMA version bound by shared-memory bandwidth (73 % of the LSU wavefront
 * peak): a quad thread loaded 9 float4 per 32 FFMA, and broadcast does not multiply the 128 B/clk a warp can receive.
 * mma.sync.m16n8k8 shares fragments across the warp instead: the warp's 8 rows are rows 0..7 of the A tile (rows 8..15
 * are zero registers, never loaded), thread (g = lane >> 2, t = lane & 3) loads x[row g][k0 + t], x[row g][k0 + t + 4]
 * and, per 8-channel tile j, w[ch][k0 + t], w[ch][k0 + t + 4] with ch = 8j + 4 (g & 1) + (g >> 1).  That column
 * permutation makes the accumulator pair the thread receives (tile columns 2t, 2t + 1) the channels 4 (2j) + t and
 * 4 (2j + 1) + t — exactly the quad mapping (channel 4c + q) of the surrounding code, so nothing else moves.
 * Every operand is split x = hi + lo with hi = tf32(x), lo = tf32(x - hi) and the product is lo*hi + hi*lo + hi*hi,
 * accumulated in fp32: ~2^-21 relative per product, inside the parity bar against torch fp32 autograd. */
#ifndef BB_POL_MMA
#define BB_POL_MMA 1   /* bb_policy_kernel: the three products on the tensor cores (0: FFMA) */
#endif
#ifndef BB_PPO_MMA
#define BB_PPO_MMA 63   /* bit 0: layer 1, 1: layer 2, 2: heads, 3: backward heads, 4: backward layer 2, 5: weight gradients (tile phase) */
#endif
#ifndef BB_PPO_TRUNC_SPLIT
#define BB_PPO_TRUNC_SPLIT 1
#endif
BBD void tf32_split(float x, uint32_t &hi, uint32_t &lo)
{
#if BB_PPO_TRUNC_SPLIT
    hi = __float_as_uint(x) & 0xffffe000u;                 /* the tensor core reads only the top 19 bits anyway */
    lo = __float_as_uint(x - __uint_as_float(hi));         /* exact; its own low bits are dropped by the hardware (2^-21 |x|) */
#else
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(hi) : "f"(x));
    const float r = x - __uint_as_float(hi);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(lo) : "f"(r));
#endif
}
BBD void mma_tf32_half(float &c0, float &c1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1)
{
    float d2, d3;
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%10,%10};"
                 : "+f"(c0), "+f"(c1), "=f"(d2), "=f"(d3)
                 : "r"(a0), "r"(0u), "r"(a2), "r"(0u), "r"(b0), "r"(b1), "f"(0.f));
}
/* one k-step (8 inputs) of out[2j + e] += sum_k W[(8j + 4e + q) * WSTR + k] * x[k]: xa = x[k0 + t], xb = x[k0 + t + 4],
 * wk = W + (4 (g & 1) + (g >> 1)) * WSTR + t + k0 */
template <int WSTR, int NT>
BBD void quad_mma_step(float xa, float xb, const float *wk, float (&out)[2 * NT])
{
    uint32_t a0h, a0l, a2h, a2l;
    tf32_split(xa, a0h, a0l);
    tf32_split(xb, a2h, a2l);
#pragma unroll
    for (int j = 0; j < NT; j++) {
        const float *w = wk + 8 * j * WSTR;
        uint32_t b0h, b0l, b1h, b1l;
        tf32_split(w[0], b0h, b0l);
        tf32_split(w[4], b1h, b1l);
        mma_tf32_half(out[2 * j], out[2 * j + 1], a0l, a2l, b0h, b1h);
        mma_tf32_half(out[2 * j], out[2 * j + 1], a0h, a2h, b0l, b1l);
        mma_tf32_half(out[2 * j], out[2 * j + 1], a0h, a2h, b0h, b1h);
    }
}
/* out[2j + e] += sum_k W[(8j + 4e + q) * WSTR + k] * xrow[k], k < K (multiple of 8), j < NT; all 32 lanes must call it */
template <int K, int WSTR, int NT>
BBD void quad_mma_tf32x3(const float *xrow, const float *W, int lane, float (&out)[2 * NT])
{
    const int g = lane >> 2, t = lane & 3;
    const float *wbase = W + (4 * (g & 1) + (g >> 1)) * WSTR + t;
#pragma unroll 4
    for (int k0 = 0; k0 < K; k0 += 8) quad_mma_step<WSTR, NT>(xrow[k0 + t], xrow[k0 + t + 4], wbase + k0, out);
}

/* the buckets [2,8,3,2,2,2] (env.py:102) spread over the quad: q0 {0,2}, q1 {1}, q2 {3,4}, q3 {5} */
__device__ __constant__ int kBucketOff[BB_NUM_ACTION_DIMS] = {0, 2, 10, 13, 15, 17};
__device__ __constant__ int kBucketCnt[BB_NUM_ACTION_DIMS] = {2, 8, 3, 2, 2, 2};
__device__ __constant__ int kQuadBuckets[4][2] = {{0, 2}, {1, -1}, {3, 4}, {5, -1}};

/* One bucketed categorical (action.py:20-33) over the cnt <= kMax logits l[0..cnt): returns log_prob(a) = l[a] - logsumexp and
 * the action a -- stochastic: inverse CDF, the smallest a with u * s < e_0 + .. + e_a, u = (word >> 8) * 2^-24; otherwise argmax.
 * The exponentials are computed once and kept in registers (the loops are unrolled to kMax and predicated on cnt); sums run in
 * index order, so the bits are those of the plain loops. */
template <int kMax>
BBD float sample_bucket(const float *l, int cnt, bool stochastic, uint32_t word, int &aOut)
{
    float m = l[0];
#pragma unroll
    for (int i = 1; i < kMax; i++) if (i < cnt) m = fmaxf(m, l[i]);
    float e[kMax];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < kMax; i++) {
        e[i] = 0.f;
        if (i < cnt) { e[i] = expf(l[i] - m); s += e[i]; }
    }
    const float lse = m + logf(s);
    int a = 0;
    if (stochastic) {
        const float us = ((float)(word >> 8) * 0x1p-24f) * s;
        float c = 0.f;
        bool found = false;
        a = cnt - 1;
#pragma unroll
        for (int i = 0; i < kMax; i++)
            if (i < cnt && !found) { c += e[i]; if (us < c) { a = i; found = true; } }
    } else {
        for (int i = 1; i < cnt; i++) if (l[i] > l[a]) a = i;
    }
    aOut = a;
    return l[a] - lse;
}

/* Agent.forward + sampling + rollout bookkeeping for the 8 rows rw0 .. rw0 + 7 on the calling warp (quad (rb & 7) takes row
 * rw0 + (rb & 7); rb = the quad's row slot in S.hid / S.out; all 32 lanes must call it).  The body of bb_policy_kernel's row
 * loop, shared with bb_rollout_loop_kernel: the same instructions in the same order, hence the same bits. */
/* kImg (bb_rollout_loop_kernel): the rows are read from the step's observation IMAGE in shared memory (io.obsImg: the tile's
 * rows as the step staged them, 26 chunks of columns 0..103; columns 104..127 are zero in the rows and have no chunk) instead
 * of from the Observations column the step has just written through to L2 -- the same floats, without the L2 round trips. */
template <bool kImg, class Smem>
BBD void policy_rows8(const bb_policy &P, Smem &S, const PolicyIo &io, int64_t rw0, int rb, int q, int lane)
{
    const int H = P.hidden;
    const int64_t rr = rw0 + (rb & 7);
    const bool valid = rr < io.n;
    const int64_t r = valid ? rr : io.n - 1; /* tail quads shadow the last row and never store */
    /* ---- backbone.0: observation normaliser (agent.py:27-39) fused into Linear(128,H) ---- */
    const float4 *orow = kImg ? io.obsImg + (r - io.obsImgWorld0) * kImgWorld4 + io.obsImgRow * kImgRow4
                              : reinterpret_cast<const float4 *>(io.obs + r * io.obsStride);
    float4 *brow = (io.bufObs && valid) ? reinterpret_cast<float4 *>(io.bufObs + r * BB_OBS_DIM) : nullptr;
    float acc[kCh];
#pragma unroll
    for (int c = 0; c < kCh; c++) acc[c] = S.b1[4 * c + q];
#if BB_POL_MMA
    if (brow) { /* buffer.obs[step] = obs (ppo.py:128): this thread's 8 of the row's 32 chunks */
#pragma unroll
        for (int i = 0; i < 8; i++) {
            const int c = 4 * i + q;
            brow[c] = (!kImg || c < kObsChunks) ? orow[c] : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    {   /* 3xTF32 tensor-core product (see quad_mma_tf32x3): the A fragment comes straight from global memory, normalised on the fly */
        const float *xr = reinterpret_cast<const float *>(orow);
        const int g = lane >> 2;
        const float *wbase = &S.w1[0][0] + (4 * (g & 1) + (g >> 1)) * kW1Stride + q;
        float acc2[kCh];   /* odd k-steps accumulate separately: two independent MMA chains per tile (this kernel is latency bound) */
#pragma unroll
        for (int c = 0; c < kCh; c++) acc2[c] = 0.f;
#pragma unroll 4
        for (int k0 = 0; k0 < BB_OBS_DIM; k0 += 16) {
            float xv[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
                const int k = k0 + q + 4 * i;
                const float xraw = (!kImg || k < 4 * kObsChunks) ? xr[k] : 0.f;
                xv[i] = (xraw - S.mean[k]) * S.inv[k];
                if (P.obs_clamp > 0.f) xv[i] = fminf(fmaxf(xv[i], -P.obs_clamp), P.obs_clamp);
            }
            quad_mma_step<kW1Stride, kCh / 2>(xv[0], xv[1], wbase + k0, acc);
            quad_mma_step<kW1Stride, kCh / 2>(xv[2], xv[3], wbase + k0 + 8, acc2);
        }
#pragma unroll
        for (int c = 0; c < kCh; c++) acc[c] += acc2[c];
    }
#else
    static_assert(!kImg, "the image-fed rows exist for the tensor-core form only");
    float4 xq = orow[0];
#pragma unroll 2
    for (int k4 = 0; k4 < BB_OBS_DIM / 4; k4++) {
        const float4 x = xq;
        if (k4 + 1 < BB_OBS_DIM / 4) xq = orow[k4 + 1]; /* next quad of inputs in flight while this one is used */
        if (brow && (k4 & 3) == q) brow[k4] = x;         /* buffer.obs[step] = obs (ppo.py:128), one store per chunk */
        const float4 mu = *reinterpret_cast<const float4 *>(&S.mean[k4 * 4]);
        const float4 is = *reinterpret_cast<const float4 *>(&S.inv[k4 * 4]);
        float x0 = (x.x - mu.x) * is.x, x1 = (x.y - mu.y) * is.y, x2 = (x.z - mu.z) * is.z, x3 = (x.w - mu.w) * is.w;
        if (P.obs_clamp > 0.f) {
            x0 = fminf(fmaxf(x0, -P.obs_clamp), P.obs_clamp); x1 = fminf(fmaxf(x1, -P.obs_clamp), P.obs_clamp);
            x2 = fminf(fmaxf(x2, -P.obs_clamp), P.obs_clamp); x3 = fminf(fmaxf(x3, -P.obs_clamp), P.obs_clamp);
        }
#pragma unroll
        for (int c = 0; c < kCh; c++) {
            const float4 w = *reinterpret_cast<const float4 *>(&S.w1[4 * c + q][k4 * 4]);
            acc[c] = fmaf(w.w, x3, fmaf(w.z, x2, fmaf(w.y, x1, fmaf(w.x, x0, acc[c]))));
        }
    }
#endif
    layer_norm_relu(acc, q, H, S.g1, S.e1);              /* backbone.1-2 */
#pragma unroll
    for (int c = 0; c < kCh; c++) S.hid[rb][4 * c + q] = acc[c];
    __syncwarp();
    /* ---- backbone.3-5: Linear(H,H) + LayerNorm + ReLU ---- */
#pragma unroll
    for (int c = 0; c < kCh; c++) acc[c] = S.b2[4 * c + q];
#if BB_POL_MMA
    quad_mma_tf32x3<kH, kW2Stride, kCh / 2>(&S.hid[rb][0], &S.w2[0][0], lane, acc);
#else
#pragma unroll
    for (int k4 = 0; k4 < kH / 4; k4++) {
        const float4 h = *reinterpret_cast<const float4 *>(&S.hid[rb][k4 * 4]);
#pragma unroll
        for (int c = 0; c < kCh; c++) {
            const float4 w = *reinterpret_cast<const float4 *>(&S.w2[4 * c + q][k4 * 4]);
            acc[c] = fmaf(w.w, h.w, fmaf(w.z, h.z, fmaf(w.y, h.y, fmaf(w.x, h.x, acc[c]))));
        }
    }
#endif
    layer_norm_relu(acc, q, H, S.g2, S.e2);
    __syncwarp(); /* every read of the first hidden vector is done */
#pragma unroll
    for (int c = 0; c < kCh; c++) S.hid[rb][4 * c + q] = acc[c];
    __syncwarp();
    /* ---- heads: outputs q, q+4, .., q+16 of (19 actor logits | critic) ---- */
#if BB_POL_MMA
    float o[kHeadPerThread + 1];   /* o[5]: "output" 20 + q, rows of the next shared array: computed and dropped */
#pragma unroll
    for (int c = 0; c < kHeadPerThread; c++) o[c] = S.bh[4 * c + q];
    o[kHeadPerThread] = 0.f;
    quad_mma_tf32x3<kH, kW2Stride, 3>(&S.hid[rb][0], &S.wh[0][0], lane, o);
#else
    float o[kHeadPerThread];
#pragma unroll
    for (int c = 0; c < kHeadPerThread; c++) o[c] = S.bh[4 * c + q];
#pragma unroll
    for (int k4 = 0; k4 < kH / 4; k4++) {
        const float4 h = *reinterpret_cast<const float4 *>(&S.hid[rb][k4 * 4]);
#pragma unroll
        for (int c = 0; c < kHeadPerThread; c++) {
            const float4 w = *reinterpret_cast<const float4 *>(&S.wh[4 * c + q][k4 * 4]);
            o[c] = fmaf(w.w, h.w, fmaf(w.z, h.z, fmaf(w.y, h.y, fmaf(w.x, h.x, o[c]))));
        }
    }
#endif
#pragma unroll
    for (int c = 0; c < kHeadPerThread; c++) {
        const int j = 4 * c + q;
        S.out[rb][j] = o[c];
        if (io.logitsOut && valid && j < BB_NUM_LOGITS) io.logitsOut[r * BB_NUM_LOGITS + j] = o[c];
    }
    if (q == 3 && valid) { /* output 19 = critic */
        if (io.evaluateOnly) { if (io.nextValue) io.nextValue[r] = o[kHeadPerThread - 1]; }
        else if (io.valueOut) io.valueOut[r] = o[kHeadPerThread - 1];
    }
    __syncwarp();
    if (!io.evaluateOnly) {
        /* ---- bucketed categoricals (action.py:5-33): this thread's one or two buckets ----
         * Threefry call c gives the words of buckets 2c (x0) and 2c + 1 (x1).  The quad needs call 0 (q0: bucket 0, q1: bucket 1),
         * call 1 (q0: bucket 2, q2: bucket 3) and call 2 (q2: bucket 4, q3: bucket 5): q0 / q1 compute call 0, q2 call 1, q3 call 2
         * -- ONE Threefry per thread -- and the words of the second pass travel by shuffle (q0 takes call 1's x0 from q2, q2 call
         * 2's x0 from q3).  Same words, same buckets as one call per bucket. */
        float lp[2] = {0.f, 0.f};
        uint32_t wx0 = 0u, wx1 = 0u;
        if (io.stochastic) threefry2x32(io.seed, (uint32_t)(io.rowOffset + r), io.draw, (uint32_t)(q < 2 ? 0 : q - 1), wx0, wx1);
        const uint32_t nx0 = __shfl_sync(0xffffffffu, wx0, (lane & ~3) + (q == 0 ? 2 : 3));
#pragma unroll
        for (int s2 = 0; s2 < 2; s2++) {
            const int b = kQuadBuckets[q][s2];
            if (b < 0) continue;
            const float *l = &S.out[rb][kBucketOff[b]];
            int a;
            /* first pass: buckets of 2 / 8 / 2 / 2 entries over the quad; second pass: 3 / - / 2 / - */
            if (s2 == 0) lp[0] = sample_bucket<8>(l, kBucketCnt[b], io.stochastic != 0, (b & 1) ? wx1 : wx0, a);
            else lp[1] = sample_bucket<3>(l, kBucketCnt[b], io.stochastic != 0, nx0, a);
            if (valid) {
                if (io.actionsOut) io.actionsOut[r * io.actionStride + b] = a;
                if (io.bufActions) io.bufActions[r * BB_NUM_ACTION_DIMS + b] = (float)a;
            }
        }
        /* log_probs.sum(dim=-1), agent.py:152, in bucket order 0..5: owners q0.a, q1.a, q0.b, q2.a, q2.b, q3.a */
        const int qb = lane & ~3;
        float tot = __shfl_sync(0xffffffffu, lp[0], qb);
        tot += __shfl_sync(0xffffffffu, lp[0], qb + 1);
        tot += __shfl_sync(0xffffffffu, lp[1], qb);
        tot += __shfl_sync(0xffffffffu, lp[0], qb + 2);
        tot += __shfl_sync(0xffffffffu, lp[1], qb + 2);
        tot += __shfl_sync(0xffffffffu, lp[0], qb + 3);
        if (q == 0 && valid && io.logpOut) io.logpOut[r] = tot;
    }
    /* ---- bookkeeping of the previous step (ppo.py:84-85,128-134; ppo_stats.py:21-36) ---- */
    if (q == 1 && valid && io.bufRewPrev) {
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
    if (q == 2 && valid && io.resetCol) *reinterpret_cast<int2 *>(io.resetCol + r * 2) = make_int2(0, 0);
    __syncwarp(); /* S.hid / S.out of this quad are rewritten by the next row */
}

__global__ void __launch_bounds__(kPolThreads) bb_policy_kernel(const bb_policy P, const PolicyIo io)
{
    __shared__ PolicySmem S;
    const int H = P.hidden;
    const int tid = threadIdx.x, q = tid & 3, rb = tid >> 2, lane = tid & 31;
    /* ---- weights -> shared memory in their own layout (coalesced 16-byte reads); padded tails are zero ---- */
#pragma unroll
    for (int i = tid; i < kH * (BB_OBS_DIM / 4); i += kPolThreads) {
        const int j = i / (BB_OBS_DIM / 4), k4 = i % (BB_OBS_DIM / 4);
        const float4 v = j < H ? *reinterpret_cast<const float4 *>(P.w1 + j * BB_OBS_DIM + k4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(&S.w1[j][k4 * 4]) = v;
    }
#pragma unroll
    for (int i = tid; i < kH * kH; i += kPolThreads) {
        const int j = i / kH, k = i % kH;
        S.w2[j][k] = (j < H && k < H) ? P.w2[j * H + k] : 0.f;
    }
    for (int i = tid; i < kHeads * kH; i += kPolThreads) {
        const int j = i / kH, k = i % kH;
        S.wh[j][k] = k < H ? (j < BB_NUM_LOGITS ? P.wa[j * H + k] : P.wc[k]) : 0.f;
    }
    if (tid < BB_OBS_DIM) { S.mean[tid] = P.obs_mean[tid]; S.inv[tid] = P.obs_inv_std[tid]; }
    if (tid >= 128 && tid < 128 + kH) {
        const int j = tid - 128;
        const bool a = j < H;
        S.b1[j] = a ? P.b1[j] : 0.f; S.g1[j] = a ? P.ln1_g[j] : 0.f; S.e1[j] = a ? P.ln1_b[j] : 0.f;
        S.b2[j] = a ? P.b2[j] : 0.f; S.g2[j] = a ? P.ln2_g[j] : 0.f; S.e2[j] = a ? P.ln2_b[j] : 0.f;
    }
    if (tid >= 192 && tid < 192 + kHeads) S.bh[tid - 192] = (tid - 192) < BB_NUM_LOGITS ? P.ba[tid - 192] : P.bc[0];
    /* programmatic dependent launch (bb_rollout): the weights above do not depend on the step kernel that precedes this
     * launch in the stream; everything below does.  No-ops for an ordinary launch. */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (io.earlyTrigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();

    const int64_t rowStride = (int64_t)gridDim.x * kRowsPerBlock;
    for (int64_t rw0 = (int64_t)blockIdx.x * kRowsPerBlock + (rb & ~7); rw0 < io.n; rw0 += rowStride) { /* warp-uniform trip count */
        policy_rows8<false>(P, S, io, rw0, rb, q, lane);
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

/* ------------------------------------------------------------------------------------------------
 * bb_rollout_loop_kernel: ppo.py rollout() (ppo.py:60-141) as ONE launch.
 *
 * Worlds never interact, and a world's trainee / opponent rows are that world's own observation rows: the warp that steps
 * worlds [w0, w0 + 16) can also run Agent.forward for their rows.  So each warp takes its 16 worlds through env.reset() and
 * all T x { policy rows (two groups of 8) [+ the frozen opponent's rows]; step_tile } of the rollout, then the closing
 * critic evaluation, without a launch, a grid dependency or a weight re-staging between steps: the policy weights are
 * staged once per block (once per rollout instead of once per step), and the observation rows / actions / reward / done
 * a step hands to the next policy call are this SM's own stores (L1 / L2 hits).  Arithmetic per row = policy_rows8, per tile =
 * step_tile: bit-identical to the chain of bb_policy_kernel / bb_step_kernel launches bb_rollout enqueues otherwise
 * (tests/test_rollout_gpu.py runs both against the stepwise loop).  Between a tile's step and its rows' policy call the
 * warp's stores are ordered before its other lanes' loads by __syncwarp().
 * Block = ONE tile of 16 worlds and TWO warps: Agent.forward for 8 rows is a ~3000-instruction dependent chain (twice the
 * step's), so the tile's 16 rows go to the two warps in parallel (8 each, policy_rows8), the block meets at a barrier, warp 0
 * steps the tile, and the block meets again.  (One warp doing both row groups in turn was measured first: 0.85 ms per rollout
 * against 0.74 ms for the kernel chain -- the chain runs the row groups on 1024 warps, that form on 512.)
 * Dynamic shared memory = PolicySmemT<16> (trainee) [+ one for the frozen opponent] + one observation image.  Meant for
 * grids that are resident at once (bb_rollout falls back to the kernel chain otherwise: with more blocks than slots every
 * wave of blocks would run the whole rollout one after the other). */
struct RolloutLoop {
    bb_rollout_buffers buf;
    int32_t agentIdx, hasFrozen;
    uint32_t seed, drawBase;
    int64_t worldOffset;
};

template <class Smem>
BBD void stage_policy(const bb_policy &P, Smem &S, int tid, int nthreads)
{
    const int H = P.hidden;
    for (int i = tid; i < kH * (BB_OBS_DIM / 4); i += nthreads) {
        const int j = i / (BB_OBS_DIM / 4), k4 = i % (BB_OBS_DIM / 4);
        const float4 v = j < H ? *reinterpret_cast<const float4 *>(P.w1 + j * BB_OBS_DIM + k4 * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        *reinterpret_cast<float4 *>(&S.w1[j][k4 * 4]) = v;
    }
    for (int i = tid; i < kH * kH; i += nthreads) {
        const int j = i / kH, k = i % kH;
        S.w2[j][k] = (j < H && k < H) ? P.w2[j * H + k] : 0.f;
    }
    for (int i = tid; i < kHeads * kH; i += nthreads) {
        const int j = i / kH, k = i % kH;
        S.wh[j][k] = k < H ? (j < BB_NUM_LOGITS ? P.wa[j * H + k] : P.wc[k]) : 0.f;
    }
    for (int i = tid; i < BB_OBS_DIM; i += nthreads) { S.mean[i] = P.obs_mean[i]; S.inv[i] = P.obs_inv_std[i]; }
    for (int j = tid; j < kH; j += nthreads) {
        const bool a = j < H;
        S.b1[j] = a ? P.b1[j] : 0.f; S.g1[j] = a ? P.ln1_g[j] : 0.f; S.e1[j] = a ? P.ln1_b[j] : 0.f;
        S.b2[j] = a ? P.b2[j] : 0.f; S.g2[j] = a ? P.ln2_g[j] : 0.f; S.e2[j] = a ? P.ln2_b[j] : 0.f;
    }
    for (int j = tid; j < kHeads; j += nthreads) S.bh[j] = j < BB_NUM_LOGITS ? P.ba[j] : P.bc[0];
}

#ifndef BB_ROLLOUT_IMG
#define BB_ROLLOUT_IMG 1 /* the looping rollout's policy rows read the step's shared-memory observation image (0: the Observations column) */
#endif
constexpr bool kLoopImg = BB_ROLLOUT_IMG != 0 && BB_POL_MMA != 0;
constexpr int kLoopWarps = 2; /* warps per tile */
/* kTPB tiles per block (1 or 2): the policy weights are staged once per BLOCK, so two tiles per block halve their share of
 * shared memory -- with a frozen opponent (two weight sets) that is what lets 8192 envs be resident at once. */
template <int kTPB> using LoopSmemT = PolicySmemT<8 * kLoopWarps * kTPB>;
template <int kTPB>
__host__ __device__ constexpr size_t rollout_loop_smem_bytes(bool frozen)
{
    return sizeof(LoopSmemT<kTPB>) * (frozen ? 2 : 1) + (size_t)kTPB * kWorldsPerWarp * kImgWorld4 * 16;
}

template <int kTPB>
__global__ void __launch_bounds__(kLoopWarps * 32 * kTPB) bb_rollout_loop_kernel(const __grid_constant__ StepParams p, const bb_policy trainee,
                                                                                 const bb_policy frozen, const RolloutLoop rl)
{
    using LoopSmem = LoopSmemT<kTPB>;
    constexpr int kThreads = kLoopWarps * 32 * kTPB;
    extern __shared__ float4 smem4[];
    LoopSmem &St = *reinterpret_cast<LoopSmem *>(smem4);
    LoopSmem &Sf = *(reinterpret_cast<LoopSmem *>(smem4) + 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, q = tid & 3, rb = tid >> 2;
    const int tileInBlock = warp / kLoopWarps, half = warp % kLoopWarps;
    float4 *tileImg = reinterpret_cast<float4 *>(reinterpret_cast<LoopSmem *>(smem4) + (rl.hasFrozen ? 2 : 1)) + tileInBlock * kWorldsPerWarp * kImgWorld4;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    stage_policy(trainee, St, tid, kThreads);
    if (rl.hasFrozen) stage_policy(frozen, Sf, tid, kThreads);
    const Columns &C = p.col;
    const int64_t n = p.numWorlds;
    if ((int64_t)blockIdx.x * kTPB * kWorldsPerWarp >= n) return; /* block-uniform */
    const int64_t tileWorld0 = ((int64_t)blockIdx.x * kTPB + tileInBlock) * kWorldsPerWarp;
    const bool stepper = half == 0 && tileWorld0 < n; /* this warp steps the tile (a block's last tile may lie past the end: it only keeps the barriers) */
    const int agent = rl.agentIdx, T = rl.buf.horizon;
    if (stepper) { /* env.reset(), env.py:178-185: Reset flags = 1, trainee action = 0, one step (the flags are cleared by the first policy call) */
        const int64_t w = tileWorld0 + (lane >> 1);
        const int me = lane & 1;
        if (w < n) {
            C.reset[w * 2 + me] = 1;
            if (me == agent) {
                int2 *a = reinterpret_cast<int2 *>(C.action + (w * 2 + me) * 6);
                a[0] = make_int2(0, 0); a[1] = make_int2(0, 0); a[2] = make_int2(0, 0);
            }
        }
        __syncwarp();
        step_tile<false, kWorldsPerWarp>(p, p.io, tileImg, tileWorld0, lane);
    }
    __syncthreads(); /* the staged weights; the reset step's observation rows before the first policy call's loads */
    const int64_t rw0 = tileWorld0 + 8 * half; /* this warp's 8 rows */
    for (int t = 0; t <= T; t++) {
        PolicyIo io{};
        io.obs = C.obs + agent * BB_OBS_DIM; io.obsStride = 2 * BB_OBS_DIM; io.n = n; io.rowOffset = rl.worldOffset;
        io.stochastic = 1; io.seed = rl.seed; io.draw = rl.drawBase + (uint32_t)t;
        io.evaluateOnly = (t == T);
        if (t < T) {
            io.actionsOut = C.action + agent * 6; io.actionStride = 12;
            io.logpOut = rl.buf.log_probs + (int64_t)t * n;
            io.valueOut = rl.buf.values + (int64_t)t * n;
            io.bufObs = rl.buf.obs + (int64_t)t * n * BB_OBS_DIM;
            io.bufActions = rl.buf.actions + (int64_t)t * n * BB_NUM_ACTION_DIMS;
        } else {
            io.nextValue = rl.buf.next_value;
        }
        if (t > 0) {
            io.rew = C.reward + agent; io.done = C.done + agent; io.rdStride = 2;
            io.bufRewPrev = rl.buf.rewards + (int64_t)(t - 1) * n;
            io.bufNotDonePrev = rl.buf.not_dones + (int64_t)(t - 1) * n;
            io.curReward = rl.buf.cur_reward; io.epLength = rl.buf.ep_length; io.epSums = rl.buf.ep_sums;
        } else {
            io.resetCol = C.reset;
        }
        io.obsImg = tileImg; io.obsImgWorld0 = tileWorld0; io.obsImgRow = agent;
        if (rw0 < n) policy_rows8<kLoopImg>(trainee, St, io, rw0, rb, q, lane);
        if (t == T) break;
        if (rl.hasFrozen && rw0 < n) { /* env.py:126-143: the frozen policy drives the other agent from its own observation row */
            PolicyIo fo{};
            const int other = 1 - agent;
            fo.obs = C.obs + other * BB_OBS_DIM; fo.obsStride = 2 * BB_OBS_DIM; fo.n = n; fo.rowOffset = rl.worldOffset;
            fo.stochastic = 1; fo.seed = rl.seed; fo.draw = 0x80000000u + rl.drawBase + (uint32_t)t;
            fo.actionsOut = C.action + other * 6; fo.actionStride = 12;
            fo.obsImg = tileImg; fo.obsImgWorld0 = tileWorld0; fo.obsImgRow = other;
            policy_rows8<kLoopImg>(frozen, Sf, fo, rw0, rb, q, lane);
        }
        __syncthreads(); /* the tile's sampled actions (and cleared Reset flags) before the step's loads */
        if (stepper) step_tile<false, kWorldsPerWarp>(p, p.io, tileImg, tileWorld0, lane);
        __syncthreads(); /* the step's observation rows, reward and done before the next policy call's loads */
    }
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
