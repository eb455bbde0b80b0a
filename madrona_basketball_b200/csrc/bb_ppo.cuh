/*
 * bb_ppo.cuh — fused PPO minibatch step (include/bb_ppo.h): forward + clipped-PPO loss + backward of
 * scripts/agent.py Agent(num_channels = 32, num_layers = 2), gradient reduction, grad-norm clip + Adam.
 *
 * Reference semantics: scripts/ppo.py:180-216 (update_policy), scripts/agent.py:171-178 (get_stats),
 * scripts/action.py (bucketed categoricals: log-prob of the taken action and entropy per bucket).
 * In the reference this is ~200 small torch kernels per minibatch step (3.4 ms at 65 536 rows, launch
 * bound); here the whole network fits one kernel: 5 972 parameters live in shared memory, a row's
 * activations in registers / a private shared row, and the weight gradients are accumulated in registers
 * across the block's row tiles.
 *
 * bb_ppo_grad_kernel (256 threads, tiles of 64 rows, persistent grid of 2 blocks per SM):
 *   row phase   four threads ("quad") per row, thread q owns hidden channels q, q+4, .., q+28, exactly
 *               the mapping of bb_policy_kernel: broadcast 16-byte reads of the weights in torch's own
 *               [out][in] layout.  Forward keeps xhat / rstd of both LayerNorms in registers; the loss
 *               seeds d logits / d value; backward goes through transposed copies of the head and
 *               second-layer weights, and leaves X, A1, A2, dH1, dH2, dZ of the tile in shared memory.
 *   tile phase  the weight gradients are three small GEMMs over the tile's 64 rows (dW1 = dH1^T X,
 *               dW2 = dH2^T A1, dWh = dZ^T A2), each thread owning a register patch of the output.
 *   epilogue    every block writes its partial gradient (all 5 972 + 4 entries) to the workspace;
 *               bb_ppo_reduce_kernel adds the partials in block order.  No atomics: bitwise reproducible.
 * 26 kFLOP per row, 1.7 GFLOP per minibatch of 65 536 rows.  Agreement with torch autograd (fp32) is the parity
 * bar, so the matrix products run on the tensor cores as 3xTF32 (mma.sync m16n8k8, every operand split hi + lo,
 * three products, fp32 accumulation: fp32-level accuracy), see quad_mma_tf32x3 / mma_tf32x3 below: the plain FFMA
 * version was bound by shared-memory bandwidth (73 % of the LSU wavefront peak in ncu), the fragment layout of
 * mma.sync reads each shared value once per warp instead of once per quad.  129 -> 89 us per minibatch.
 */
#pragma once
#include <cub/device/device_radix_sort.cuh>
#include "../../include/bb_ppo.h"

namespace bb {

constexpr int kPpoThreads = 256;
constexpr int kPpoRows = kPpoThreads / 4;   /* rows per tile */
constexpr int kPpoPartial = BB_PPO_NUM_PARAMS + BB_PPO_NUM_STATS; /* floats per block partial */
constexpr int kXS = BB_OBS_DIM + 4;         /* 132 */
constexpr int kHS = kH + 4;                 /* 36  */
constexpr int kZS = 24;                     /* 20 head outputs, padded to 6 float4 */

struct __align__(16) PpoSmem {
    float w1[kH][kXS];
    float w2[kH][kHS];
    float w2t[kH][kHS];        /* w2t[k][j] = w2[j][k] */
    float wh[kHeads][kHS];     /* rows 0..18 actor, row 19 critic */
    float wht[kH][kZS];        /* wht[k][j] = wh[j][k] */
    float x[kPpoRows][kXS];    /* normalised observations of the tile */
    float a1[kPpoRows][kHS];
    float a2[kPpoRows][kHS];
    float dh1[kPpoRows][kHS];
    float dh2[kPpoRows][kHS];
    float z[kPpoRows][kZS];    /* head outputs, then their gradients */
    float mean[BB_OBS_DIM], inv[BB_OBS_DIM];
    float b1[kH], g1[kH], e1[kH], b2[kH], g2[kH], e2[kH], bh[kZS];
};

struct PpoArgs {
    bb_ppo_batch b;
    float invM;
    int64_t numTiles;
    int rowsPerTile;          /* multiple of 8, <= kPpoRows: chosen so that the resident blocks get equal shares */
    float *partials;
};

/* LayerNorm(32) forward over the quad: leaves xhat in v, returns rstd */
BBD float ppo_ln_fwd(float (&v)[kCh])
{
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kCh; c++) s += v[c];
    const float mean = quad_sum(s) * (1.f / kH);
    s = 0.f;
#pragma unroll
    for (int c = 0; c < kCh; c++) { v[c] -= mean; s = fmaf(v[c], v[c], s); }
    const float rstd = rsqrtf(quad_sum(s) * (1.f / kH) + 1e-5f);
#pragma unroll
    for (int c = 0; c < kCh; c++) v[c] *= rstd;
    return rstd;
}

/* LayerNorm backward: d (in: dL/dy, out: dL/dh), xhat, gamma of the thread's channels; accumulates dgamma / dbeta */
BBD void ppo_ln_bwd(float (&d)[kCh], const float (&xhat)[kCh], float rstd, const float *g, int q, float (&dg)[kCh], float (&de)[kCh])
{
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int c = 0; c < kCh; c++) {
        dg[c] = fmaf(d[c], xhat[c], dg[c]);
        de[c] += d[c];
        d[c] *= g[4 * c + q];
        s1 += d[c];
        s2 = fmaf(d[c], xhat[c], s2);
    }
    s1 = quad_sum(s1) * (1.f / kH);
    s2 = quad_sum(s2) * (1.f / kH);
#pragma unroll
    for (int c = 0; c < kCh; c++) d[c] = rstd * (d[c] - s1 - xhat[c] * s2);
}

/* tf32_split, mma_tf32_half and quad_mma_tf32x3 (the row-phase products) live in bb_rollout.cuh: bb_policy_kernel uses them too */
/* full-tile 3xTF32 product for the weight-gradient GEMMs: c += A(16x8) B(8x8), operands pre-split */
BBD void mma_tf32(float (&c)[4], const uint32_t (&a)[4], const uint32_t (&b)[2])
{
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}
BBD void mma_tf32x3(float (&c)[4], const uint32_t (&ah)[4], const uint32_t (&al)[4], const uint32_t (&bh)[2], const uint32_t (&bl)[2])
{
    mma_tf32(c, al, bh);
    mma_tf32(c, ah, bl);
    mma_tf32(c, ah, bh);
}
/* A fragment of G^T for the k-step whose eight rows are r0..r0+7: A[m][k] = G[row(k)][m0 + m] with row(t) = r0 + 2t,
 * row(t + 4) = r0 + 2t + 1 (any bijection works as long as B uses the same one; this one is bank-conflict free for
 * the 36- and 132-float row strides) */
template <int STR>
BBD void load_a_frag(const float *G, int ra, int m0, int g, uint32_t (&ah)[4], uint32_t (&al)[4])
{
    tf32_split(G[ra * STR + m0 + g], ah[0], al[0]);
    tf32_split(G[ra * STR + m0 + g + 8], ah[1], al[1]);
    tf32_split(G[(ra + 1) * STR + m0 + g], ah[2], al[2]);
    tf32_split(G[(ra + 1) * STR + m0 + g + 8], ah[3], al[3]);
}
template <int STR>
BBD void load_b_frag(const float *X, int ra, int n0, int g, uint32_t (&bh)[2], uint32_t (&bl)[2])
{
    tf32_split(X[ra * STR + n0 + g], bh[0], bl[0]);
    tf32_split(X[(ra + 1) * STR + n0 + g], bh[1], bl[1]);
}
__global__ void __launch_bounds__(kPpoThreads, 2) bb_ppo_grad_kernel(const bb_policy P, const PpoArgs A)
{
    extern __shared__ __align__(16) unsigned char ppoSmemRaw[];
    PpoSmem &S = *reinterpret_cast<PpoSmem *>(ppoSmemRaw);
    const int tid = threadIdx.x, q = tid & 3, rb = tid >> 2, lane = tid & 31, warp = tid >> 5;
    asm volatile("griddepcontrol.wait;" ::: "memory"); /* programmatic dependent launch: the previous Adam step wrote the weights */

    /* ---- parameters -> shared memory ---- */
    for (int i = tid; i < kH * (BB_OBS_DIM / 4); i += kPpoThreads) {
        const int j = i / (BB_OBS_DIM / 4), k4 = i % (BB_OBS_DIM / 4);
        *reinterpret_cast<float4 *>(&S.w1[j][k4 * 4]) = *reinterpret_cast<const float4 *>(P.w1 + j * BB_OBS_DIM + k4 * 4);
    }
    for (int i = tid; i < kH * kH; i += kPpoThreads) {
        const int j = i / kH, k = i % kH;
        const float w = P.w2[i];
        S.w2[j][k] = w;
        S.w2t[k][j] = w;
    }
    for (int i = tid; i < kHeads * kH; i += kPpoThreads) {
        const int j = i / kH, k = i % kH;
        const float w = j < BB_NUM_LOGITS ? P.wa[j * kH + k] : P.wc[k];
        S.wh[j][k] = w;
        S.wht[k][j] = w;
    }
    if (tid < kH) { S.wht[tid][20] = S.wht[tid][21] = S.wht[tid][22] = S.wht[tid][23] = 0.f; }
    if (tid < BB_OBS_DIM) { S.mean[tid] = P.obs_mean[tid]; S.inv[tid] = P.obs_inv_std[tid]; }
    if (tid >= 128 && tid < 128 + kH) {
        const int j = tid - 128;
        S.b1[j] = P.b1[j]; S.g1[j] = P.ln1_g[j]; S.e1[j] = P.ln1_b[j];
        S.b2[j] = P.b2[j]; S.g2[j] = P.ln2_g[j]; S.e2[j] = P.ln2_b[j];
    }
    if (tid >= 192 && tid < 192 + kZS) { const int j = tid - 192; S.bh[j] = j < BB_NUM_LOGITS ? P.ba[j] : (j == BB_NUM_LOGITS ? P.bc[0] : 0.f); }
    __syncthreads();

    /* ---- accumulators that live across the block's tiles ---- */
    float gW1[4][4], gW2[4], gWh[4], gB = 0.f;                     /* tile-phase patches */
    float dg1[kCh], de1[kCh], dg2[kCh], de2[kCh];                  /* LayerNorm parameters, per (row slot, channel) */
    float st0 = 0.f, st1 = 0.f, st2 = 0.f, st3 = 0.f;              /* loss statistics (q == 0 threads) */
#pragma unroll
    for (int i = 0; i < 4; i++) { gW2[i] = 0.f; gWh[i] = 0.f;
#pragma unroll
        for (int j = 0; j < 4; j++) gW1[i][j] = 0.f; }
#pragma unroll
    for (int c = 0; c < kCh; c++) { dg1[c] = de1[c] = dg2[c] = de2[c] = 0.f; }

    const bb_ppo_batch &B = A.b;
    const float clipLo = 1.f - B.clip_coef, clipHi = 1.f + B.clip_coef;

    for (int64_t tile = blockIdx.x; tile < A.numTiles; tile += gridDim.x) {
        /* =========================== row phase =========================== */
        const int64_t mrow = tile * A.rowsPerTile + rb;
        const bool valid = mrow < B.num_rows;
        if ((rb & ~7) < A.rowsPerTile) { /* warp-uniform: warps beyond the tile's rows only join the barriers */
        const int64_t mr = valid ? mrow : B.num_rows - 1;
        const int64_t src = B.indices ? B.indices[mr] : (B.indices32 ? (int64_t)B.indices32[mr] : mr);
        {   /* observation normaliser (agent.py:27-39): this thread's 8 of the row's 32 float4 chunks */
            const float4 *orow = reinterpret_cast<const float4 *>(B.obs + src * BB_OBS_DIM);
#pragma unroll
            for (int i = 0; i < 8; i++) {
                const int k4 = 4 * i + q;
                const float4 x = orow[k4];
                const float4 mu = *reinterpret_cast<const float4 *>(&S.mean[k4 * 4]);
                const float4 is = *reinterpret_cast<const float4 *>(&S.inv[k4 * 4]);
                float4 y = make_float4((x.x - mu.x) * is.x, (x.y - mu.y) * is.y, (x.z - mu.z) * is.z, (x.w - mu.w) * is.w);
                if (P.obs_clamp > 0.f) {
                    y.x = fminf(fmaxf(y.x, -P.obs_clamp), P.obs_clamp); y.y = fminf(fmaxf(y.y, -P.obs_clamp), P.obs_clamp);
                    y.z = fminf(fmaxf(y.z, -P.obs_clamp), P.obs_clamp); y.w = fminf(fmaxf(y.w, -P.obs_clamp), P.obs_clamp);
                }
                *reinterpret_cast<float4 *>(&S.x[rb][k4 * 4]) = y;
            }
        }
        __syncwarp();
        /* ---- backbone.0-2: Linear(128,32) + LayerNorm + ReLU ---- */
        float xh1[kCh], xh2[kCh];
        float rstd1, rstd2;
        unsigned mask1 = 0, mask2 = 0;
        {
#pragma unroll
            for (int c = 0; c < kCh; c++) xh1[c] = S.b1[4 * c + q];
#if BB_PPO_MMA & 1
            quad_mma_tf32x3<BB_OBS_DIM, kXS, kCh / 2>(&S.x[rb][0], &S.w1[0][0], lane, xh1);
#else
#pragma unroll 4
            for (int k4 = 0; k4 < BB_OBS_DIM / 4; k4++) {
                const float4 x = *reinterpret_cast<const float4 *>(&S.x[rb][k4 * 4]);
#pragma unroll
                for (int c = 0; c < kCh; c++) {
                    const float4 w = *reinterpret_cast<const float4 *>(&S.w1[4 * c + q][k4 * 4]);
                    xh1[c] = fmaf(w.w, x.w, fmaf(w.z, x.z, fmaf(w.y, x.y, fmaf(w.x, x.x, xh1[c]))));
                }
            }
#endif
            rstd1 = ppo_ln_fwd(xh1);
#pragma unroll
            for (int c = 0; c < kCh; c++) {
                const float y = fmaf(xh1[c], S.g1[4 * c + q], S.e1[4 * c + q]);
                if (y > 0.f) mask1 |= 1u << c;
                S.a1[rb][4 * c + q] = fmaxf(y, 0.f);
            }
        }
        __syncwarp();
        /* ---- backbone.3-5: Linear(32,32) + LayerNorm + ReLU ---- */
        {
#pragma unroll
            for (int c = 0; c < kCh; c++) xh2[c] = S.b2[4 * c + q];
#if BB_PPO_MMA & 2
            quad_mma_tf32x3<kH, kHS, kCh / 2>(&S.a1[rb][0], &S.w2[0][0], lane, xh2);
#else
#pragma unroll
            for (int k4 = 0; k4 < kH / 4; k4++) {
                const float4 h = *reinterpret_cast<const float4 *>(&S.a1[rb][k4 * 4]);
#pragma unroll
                for (int c = 0; c < kCh; c++) {
                    const float4 w = *reinterpret_cast<const float4 *>(&S.w2[4 * c + q][k4 * 4]);
                    xh2[c] = fmaf(w.w, h.w, fmaf(w.z, h.z, fmaf(w.y, h.y, fmaf(w.x, h.x, xh2[c]))));
                }
            }
#endif
            rstd2 = ppo_ln_fwd(xh2);
#pragma unroll
            for (int c = 0; c < kCh; c++) {
                const float y = fmaf(xh2[c], S.g2[4 * c + q], S.e2[4 * c + q]);
                if (y > 0.f) mask2 |= 1u << c;
                S.a2[rb][4 * c + q] = fmaxf(y, 0.f);
            }
        }
        __syncwarp();
        /* ---- heads: outputs q, q+4, .., q+16 of (19 actor logits | critic) ---- */
        {
#if BB_PPO_MMA & 4
            float o[6];   /* o[5] = "output" 20 + q: rows of the next array, computed and dropped */
#pragma unroll
            for (int c = 0; c < kHeadPerThread; c++) o[c] = S.bh[4 * c + q];
            o[5] = 0.f;
            quad_mma_tf32x3<kH, kHS, 3>(&S.a2[rb][0], &S.wh[0][0], lane, o);
#pragma unroll
            for (int c = 0; c < kHeadPerThread; c++) S.z[rb][4 * c + q] = o[c];
            S.z[rb][20 + q] = 0.f;   /* the padding the backward K = 24 product reads */
#else
            float o[kHeadPerThread];
#pragma unroll
            for (int c = 0; c < kHeadPerThread; c++) o[c] = S.bh[4 * c + q];
#pragma unroll
            for (int k4 = 0; k4 < kH / 4; k4++) {
                const float4 h = *reinterpret_cast<const float4 *>(&S.a2[rb][k4 * 4]);
#pragma unroll
                for (int c = 0; c < kHeadPerThread; c++) {
                    const float4 w = *reinterpret_cast<const float4 *>(&S.wh[4 * c + q][k4 * 4]);
                    o[c] = fmaf(w.w, h.w, fmaf(w.z, h.z, fmaf(w.y, h.y, fmaf(w.x, h.x, o[c]))));
                }
            }
#pragma unroll
            for (int c = 0; c < kHeadPerThread; c++) S.z[rb][4 * c + q] = o[c];
#endif
        }
        __syncwarp();
        /* ---- loss (ppo.py:192-208) and its gradient w.r.t. the head outputs; buckets spread over the quad ---- */
        {
            float lpPart = 0.f, entPart = 0.f;
            float Hb[2] = {0.f, 0.f};
            int act[2] = {0, 0};
#pragma unroll
            for (int s2 = 0; s2 < 2; s2++) { /* pass 1: the row's logits become log-probabilities in place */
                const int b = kQuadBuckets[q][s2];
                if (b < 0) continue;
                const int off = kBucketOff[b], cnt = kBucketCnt[b];
                float *l = &S.z[rb][off];
                float m = l[0];
                for (int i = 1; i < cnt; i++) m = fmaxf(m, l[i]);
                float s = 0.f;
                for (int i = 0; i < cnt; i++) s += expf(l[i] - m);
                const float lse = m + logf(s);
                int a = (int)B.actions[src * BB_NUM_ACTION_DIMS + b];
                a = a < 0 ? 0 : (a >= cnt ? cnt - 1 : a);
                act[s2] = a;
                float h = 0.f, la = 0.f;
                for (int i = 0; i < cnt; i++) {
                    const float lgp = l[i] - lse;
                    l[i] = lgp;
                    h = fmaf(-expf(lgp), lgp, h);
                    if (i == a) la = lgp;
                }
                lpPart += la;
                Hb[s2] = h;
                entPart += h;
            }
            const float lp = quad_sum(lpPart), ent = quad_sum(entPart);
            /* actor: ratio = exp(lp - lp_old); max(-adv * ratio, -adv * clamp(ratio)) */
            const float adv = B.advantages[src];
            const float ratio = expf(lp - B.log_probs[src]);
            const float surr1 = -adv * ratio;
            const float surr2 = -adv * fminf(fmaxf(ratio, clipLo), clipHi);
            const bool inside = ratio >= clipLo && ratio <= clipHi;
            /* d max / d ratio: inside the clip both branches carry -adv (torch splits a tie 0.5 + 0.5); outside only surr1 can */
            float dRatio = 0.f;
            if (inside) dRatio = -adv;
            else if (surr1 > surr2) dRatio = -adv;
            else if (surr1 == surr2) dRatio = -0.5f * adv;
            const float gLp = valid ? A.invM * dRatio * ratio : 0.f;
            const float gEnt = valid ? A.invM * B.ent_coef : 0.f;   /* d(-ent_coef * mean H)/dz_j = ent_coef/m * p_j (log p_j + H) */
#pragma unroll
            for (int s2 = 0; s2 < 2; s2++) {
                const int b = kQuadBuckets[q][s2];
                if (b < 0) continue;
                const int off = kBucketOff[b], cnt = kBucketCnt[b];
                float *l = &S.z[rb][off];
                for (int i = 0; i < cnt; i++) { /* pass 2: log-probabilities become gradients */
                    const float lgp = l[i], pr = expf(lgp);
                    const float oneHot = (i == act[s2]) ? 1.f : 0.f;
                    l[i] = fmaf(gLp, oneHot - pr, gEnt * pr * (lgp + Hb[s2]));
                }
            }
            if (q == 3) { /* critic (ppo.py:197-201): 0.5 * max((v - ret)^2, (v_old + clamp(v - v_old) - ret)^2) * vf_coef */
                const float v = S.z[rb][BB_NUM_LOGITS];
                const float vo = B.values[src], ret = B.returns[src];
                const float dvo = v - vo;
                const float vclip = vo + fminf(fmaxf(dvo, -B.clip_coef), B.clip_coef);
                const float e1 = v - ret, e2 = vclip - ret;
                const float l1 = e1 * e1, l2 = e2 * e2;
                const float passes = (dvo >= -B.clip_coef && dvo <= B.clip_coef) ? 1.f : 0.f;
                float dv;
                if (l1 > l2) dv = 2.f * e1;
                else if (l2 > l1) dv = 2.f * e2 * passes;
                else dv = e1 + e2 * passes;
                S.z[rb][BB_NUM_LOGITS] = valid ? A.invM * 0.5f * B.vf_coef * dv : 0.f;
                if (valid) st1 += 0.5f * fmaxf(l1, l2);
            }
            if (q == 0 && valid) { st0 += fmaxf(surr1, surr2); st2 += ent; st3 += inside ? 0.f : 1.f; }
        }
        __syncwarp();
        /* ---- backward through the heads, LayerNorm 2, Linear 2, LayerNorm 1 ---- */
        {
            float d[kCh];
#pragma unroll
            for (int c = 0; c < kCh; c++) d[c] = 0.f;
#if (BB_PPO_MMA & 12) == 12
            quad_mma_tf32x3<kZS, kZS, kCh / 2>(&S.z[rb][0], &S.wht[0][0], lane, d);
#else
#pragma unroll
            for (int j4 = 0; j4 < kHeads / 4; j4++) {
                const float4 g = *reinterpret_cast<const float4 *>(&S.z[rb][j4 * 4]);
#pragma unroll
                for (int c = 0; c < kCh; c++) {
                    const float4 w = *reinterpret_cast<const float4 *>(&S.wht[4 * c + q][j4 * 4]);
                    d[c] = fmaf(w.w, g.w, fmaf(w.z, g.z, fmaf(w.y, g.y, fmaf(w.x, g.x, d[c]))));
                }
            }
#endif
#pragma unroll
            for (int c = 0; c < kCh; c++) if (!((mask2 >> c) & 1u)) d[c] = 0.f;
            ppo_ln_bwd(d, xh2, rstd2, S.g2, q, dg2, de2);
#pragma unroll
            for (int c = 0; c < kCh; c++) S.dh2[rb][4 * c + q] = d[c];
            __syncwarp();
#pragma unroll
            for (int c = 0; c < kCh; c++) d[c] = 0.f;
#if BB_PPO_MMA & 16
            quad_mma_tf32x3<kH, kHS, kCh / 2>(&S.dh2[rb][0], &S.w2t[0][0], lane, d);
#else
#pragma unroll
            for (int k4 = 0; k4 < kH / 4; k4++) {
                const float4 g = *reinterpret_cast<const float4 *>(&S.dh2[rb][k4 * 4]);
#pragma unroll
                for (int c = 0; c < kCh; c++) {
                    const float4 w = *reinterpret_cast<const float4 *>(&S.w2t[4 * c + q][k4 * 4]);
                    d[c] = fmaf(w.w, g.w, fmaf(w.z, g.z, fmaf(w.y, g.y, fmaf(w.x, g.x, d[c]))));
                }
            }
#endif
#pragma unroll
            for (int c = 0; c < kCh; c++) if (!((mask1 >> c) & 1u)) d[c] = 0.f;
            ppo_ln_bwd(d, xh1, rstd1, S.g1, q, dg1, de1);
#pragma unroll
            for (int c = 0; c < kCh; c++) S.dh1[rb][4 * c + q] = d[c];
        }
        } /* row phase */
        __syncthreads();
        /* =========================== tile phase: weight gradients =========================== */
#if BB_PPO_MMA & 32
        {   /* dW1 = dH1^T X, dW2 = dH2^T A1, dWh = dZ^T A2 as 3xTF32 tensor-core products over the tile's rows (k = rows, 8 per step).
             * Warp w owns dW1[all 32 channels][inputs 16w .. 16w+15] (2 x 2 accumulator tiles = the 16 registers the FFMA patch had),
             * one 16 x 8 tile of dW2 and one of dWh (m-tile w >> 2, n-tile w & 3). */
            const int g = lane >> 2, t = lane & 3, mt2 = warp >> 2, nt2 = warp & 3;
            for (int r0 = 0; r0 < A.rowsPerTile; r0 += 8) {
                const int ra = r0 + 2 * t;
                uint32_t bh[2][2], bl[2][2];
#pragma unroll
                for (int nt = 0; nt < 2; nt++) load_b_frag<kXS>(&S.x[0][0], ra, 16 * warp + 8 * nt, g, bh[nt], bl[nt]);
#pragma unroll
                for (int mt = 0; mt < 2; mt++) {
                    uint32_t ah[4], al[4];
                    load_a_frag<kHS>(&S.dh1[0][0], ra, 16 * mt, g, ah, al);
#pragma unroll
                    for (int nt = 0; nt < 2; nt++) mma_tf32x3(gW1[2 * mt + nt], ah, al, bh[nt], bl[nt]);
                }
                {
                    uint32_t ah[4], al[4], b2h[2], b2l[2];
                    load_a_frag<kHS>(&S.dh2[0][0], ra, 16 * mt2, g, ah, al);
                    load_b_frag<kHS>(&S.a1[0][0], ra, 8 * nt2, g, b2h, b2l);
                    mma_tf32x3(gW2, ah, al, b2h, b2l);
                    /* heads: m-tile 1 covers outputs 16..31 of which 20..23 are the zero padding and 24..31 (the next row's entries) are dropped */
                    load_a_frag<kZS>(&S.z[0][0], ra, 16 * mt2, g, ah, al);
                    load_b_frag<kHS>(&S.a2[0][0], ra, 8 * nt2, g, b2h, b2l);
                    mma_tf32x3(gWh, ah, al, b2h, b2l);
                }
                if (tid >= 160 && tid < 160 + 2 * kH + kHeads) { /* biases: b1 | b2 | head biases */
                    const int tb = tid - 160;
                    const float *col = tb < kH ? &S.dh1[0][tb] : (tb < 2 * kH ? &S.dh2[0][tb - kH] : &S.z[0][tb - 2 * kH]);
                    const int str = tb < 2 * kH ? kHS : kZS;
                    float v[8];
#pragma unroll
                    for (int rr = 0; rr < 8; rr++) v[rr] = col[(r0 + rr) * str];
#pragma unroll
                    for (int rr = 0; rr < 8; rr++) gB += v[rr];
                }
            }
        }
#else
        {
            const int cg = warp, kg = lane;            /* dW1 patch: channels 4cg.., inputs 4kg.. */
            const int j2 = tid >> 3, k42 = tid & 7;    /* dW2 / dWh patch: output j2, inputs 4k42.. */
#pragma unroll 4
            for (int r = 0; r < A.rowsPerTile; r++) {
                const float4 dd = *reinterpret_cast<const float4 *>(&S.dh1[r][cg * 4]);
                const float4 xx = *reinterpret_cast<const float4 *>(&S.x[r][kg * 4]);
                gW1[0][0] = fmaf(dd.x, xx.x, gW1[0][0]); gW1[0][1] = fmaf(dd.x, xx.y, gW1[0][1]);
                gW1[0][2] = fmaf(dd.x, xx.z, gW1[0][2]); gW1[0][3] = fmaf(dd.x, xx.w, gW1[0][3]);
                gW1[1][0] = fmaf(dd.y, xx.x, gW1[1][0]); gW1[1][1] = fmaf(dd.y, xx.y, gW1[1][1]);
                gW1[1][2] = fmaf(dd.y, xx.z, gW1[1][2]); gW1[1][3] = fmaf(dd.y, xx.w, gW1[1][3]);
                gW1[2][0] = fmaf(dd.z, xx.x, gW1[2][0]); gW1[2][1] = fmaf(dd.z, xx.y, gW1[2][1]);
                gW1[2][2] = fmaf(dd.z, xx.z, gW1[2][2]); gW1[2][3] = fmaf(dd.z, xx.w, gW1[2][3]);
                gW1[3][0] = fmaf(dd.w, xx.x, gW1[3][0]); gW1[3][1] = fmaf(dd.w, xx.y, gW1[3][1]);
                gW1[3][2] = fmaf(dd.w, xx.z, gW1[3][2]); gW1[3][3] = fmaf(dd.w, xx.w, gW1[3][3]);
                const float d2 = S.dh2[r][j2];
                const float4 aa = *reinterpret_cast<const float4 *>(&S.a1[r][k42 * 4]);
                gW2[0] = fmaf(d2, aa.x, gW2[0]); gW2[1] = fmaf(d2, aa.y, gW2[1]);
                gW2[2] = fmaf(d2, aa.z, gW2[2]); gW2[3] = fmaf(d2, aa.w, gW2[3]);
                if (j2 < kHeads) { /* warps 0..4 */
                    const float dz = S.z[r][j2];
                    const float4 bb2 = *reinterpret_cast<const float4 *>(&S.a2[r][k42 * 4]);
                    gWh[0] = fmaf(dz, bb2.x, gWh[0]); gWh[1] = fmaf(dz, bb2.y, gWh[1]);
                    gWh[2] = fmaf(dz, bb2.z, gWh[2]); gWh[3] = fmaf(dz, bb2.w, gWh[3]);
                } else if (tid >= 160 && tid < 160 + 2 * kH + kHeads) { /* biases: b1 | b2 | head biases */
                    const int t = tid - 160;
                    gB += t < kH ? S.dh1[r][t] : (t < 2 * kH ? S.dh2[r][t - kH] : S.z[r][t - 2 * kH]);
                }
            }
        }
#endif
        __syncthreads(); /* the next tile rewrites the shared activations */
    }

    /* =========================== epilogue: this block's partial =========================== */
    float *out = A.partials + (int64_t)blockIdx.x * kPpoPartial;
    {
#if BB_PPO_MMA & 32
        const int g = lane >> 2, t = lane & 3, mt2 = warp >> 2, nt2 = warp & 3;
#pragma unroll
        for (int mt = 0; mt < 2; mt++)
#pragma unroll
            for (int nt = 0; nt < 2; nt++) { /* accumulator fragment: rows g, g + 8; columns 2t, 2t + 1 */
                float *dst = out + BB_PPO_OFF_W1 + (16 * mt + g) * BB_OBS_DIM + 16 * warp + 8 * nt + 2 * t;
                *reinterpret_cast<float2 *>(dst) = make_float2(gW1[2 * mt + nt][0], gW1[2 * mt + nt][1]);
                *reinterpret_cast<float2 *>(dst + 8 * BB_OBS_DIM) = make_float2(gW1[2 * mt + nt][2], gW1[2 * mt + nt][3]);
            }
        {
            float *dst = out + BB_PPO_OFF_W2 + (16 * mt2 + g) * kH + 8 * nt2 + 2 * t;
            *reinterpret_cast<float2 *>(dst) = make_float2(gW2[0], gW2[1]);
            *reinterpret_cast<float2 *>(dst + 8 * kH) = make_float2(gW2[2], gW2[3]);
        }
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int j = 16 * mt2 + g + 8 * h;
            if (j < kHeads) {
                float *dst = out + (j < BB_NUM_LOGITS ? BB_PPO_OFF_WA + j * kH : BB_PPO_OFF_WC) + 8 * nt2 + 2 * t; /* WC is not 8-byte aligned */
                dst[0] = gWh[2 * h]; dst[1] = gWh[2 * h + 1];
            }
        }
        if (tid >= 160 && tid < 160 + 2 * kH + kHeads) {
#else
        const int cg = warp, kg = lane, j2 = tid >> 3, k42 = tid & 7;
#pragma unroll
        for (int i = 0; i < 4; i++)
            *reinterpret_cast<float4 *>(out + BB_PPO_OFF_W1 + (4 * cg + i) * BB_OBS_DIM + 4 * kg) = make_float4(gW1[i][0], gW1[i][1], gW1[i][2], gW1[i][3]);
        *reinterpret_cast<float4 *>(out + BB_PPO_OFF_W2 + j2 * kH + 4 * k42) = make_float4(gW2[0], gW2[1], gW2[2], gW2[3]);
        if (j2 < kHeads) {
            float *dst = out + (j2 < BB_NUM_LOGITS ? BB_PPO_OFF_WA + j2 * kH : BB_PPO_OFF_WC) + 4 * k42; /* WC is not 16-byte aligned */
            dst[0] = gWh[0]; dst[1] = gWh[1]; dst[2] = gWh[2]; dst[3] = gWh[3];
        } else if (tid >= 160 && tid < 160 + 2 * kH + kHeads) {
#endif
            const int t = tid - 160;
            const int off = t < kH ? BB_PPO_OFF_B1 + t : (t < 2 * kH ? BB_PPO_OFF_B2 + (t - kH)
                          : (t - 2 * kH < BB_NUM_LOGITS ? BB_PPO_OFF_BA + (t - 2 * kH) : BB_PPO_OFF_BC));
            out[off] = gB;
        }
    }
    /* LayerNorm parameter gradients: sum the 64 row slots per channel through shared memory (fixed order) */
#pragma unroll
    for (int c = 0; c < kCh; c++) {
        S.a1[rb][4 * c + q] = dg1[c]; S.a2[rb][4 * c + q] = de1[c];
        S.dh1[rb][4 * c + q] = dg2[c]; S.dh2[rb][4 * c + q] = de2[c];
    }
    if (q == 0) { S.z[rb][0] = st0; S.z[rb][2] = st2; S.z[rb][3] = st3; }
    if (q == 3) S.z[rb][1] = st1;
    __syncthreads();
    if (tid < 4 * kH) {
        const int arr = tid >> 5, ch = tid & 31;
        const float(*src)[kHS] = arr == 0 ? S.a1 : (arr == 1 ? S.a2 : (arr == 2 ? S.dh1 : S.dh2));
        float s = 0.f;
        for (int r = 0; r < kPpoRows; r++) s += src[r][ch];
        const int off = arr == 0 ? BB_PPO_OFF_G1 : (arr == 1 ? BB_PPO_OFF_E1 : (arr == 2 ? BB_PPO_OFF_G2 : BB_PPO_OFF_E2));
        out[off + ch] = s;
    } else if (tid < 4 * kH + BB_PPO_NUM_STATS) {
        const int k = tid - 4 * kH;
        float s = 0.f;
        for (int r = 0; r < kPpoRows; r++) s += S.z[r][k];
        out[BB_PPO_NUM_PARAMS + k] = s;
    }
}

/* grad[j] = sum over blocks of partial[b][j]; the last BB_PPO_NUM_STATS entries go to stats.  32 columns x 8 slices per
 * block: slice s adds the partials b = s, s + 8, .. in order, the eight slice sums are added in slice order (a fixed
 * summation tree: bitwise reproducible), and eight loads per column are in flight instead of one dependent chain. */
constexpr int kRedCols = 32, kRedSlices = 8;
__global__ void __launch_bounds__(kRedCols * kRedSlices) bb_ppo_reduce_kernel(const float *partials, int numBlocks, float *grad, float *stats)
{
    __shared__ float sh[kRedSlices][kRedCols];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int c = threadIdx.x & (kRedCols - 1), sl = threadIdx.x / kRedCols;
    const int j = blockIdx.x * kRedCols + c;
    float s0 = 0.f, s1 = 0.f;
    if (j < kPpoPartial) {
        int b = sl;
        for (; b + kRedSlices < numBlocks; b += 2 * kRedSlices) {
            s0 += partials[(int64_t)b * kPpoPartial + j];
            s1 += partials[(int64_t)(b + kRedSlices) * kPpoPartial + j];
        }
        if (b < numBlocks) s0 += partials[(int64_t)b * kPpoPartial + j];
    }
    sh[sl][c] = s0 + s1;
    __syncthreads();
    if (sl == 0 && j < kPpoPartial) {
        float s = 0.f;
#pragma unroll
        for (int k = 0; k < kRedSlices; k++) s += sh[k][c];
        if (j < BB_PPO_NUM_PARAMS) grad[j] = s;
        else if (stats) stats[j - BB_PPO_NUM_PARAMS] = s;
    }
}

/* clip_grad_norm_ + Adam on a flat vector; one block (the policy has 5 972 parameters) */
constexpr int kAdamThreads = 1024;
__global__ void __launch_bounds__(kAdamThreads) bb_ppo_adam_kernel(float *params, const float *grad, float *m, float *v, int32_t *stepDev,
                                                                    int64_t n, float gradScale, float maxNorm, float lr, float beta1,
                                                                    float beta2, float eps, float *normOut)
{
    __shared__ float red[kAdamThreads / 32];
    __shared__ float sNorm;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    float s = 0.f;
    for (int64_t i = threadIdx.x; i < n; i += kAdamThreads) { const float g = grad[i] * gradScale; s = fmaf(g, g, s); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float t = red[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) sNorm = sqrtf(t);
    }
    __syncthreads();
    const float norm = sNorm;
    /* torch: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1 */
    const float clip = maxNorm > 0.f ? fminf(maxNorm / (norm + 1e-6f), 1.f) : 1.f;
    const int32_t step = stepDev[0] + 1;
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    const float stepSize = (float)((double)lr / bc1);
    const float bc2Sqrt = (float)sqrt(bc2);
    for (int64_t i = threadIdx.x; i < n; i += kAdamThreads) {
        const float g = grad[i] * gradScale * clip;
        const float mi = m[i] + (g - m[i]) * (1.f - beta1);       /* exp_avg.lerp_(grad, 1 - beta1) */
        const float vi = fmaf(beta2, v[i], (1.f - beta2) * g * g); /* exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2) */
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2Sqrt + eps;
        params[i] -= stepSize * (mi / denom);
    }
    __syncthreads();
    if (threadIdx.x == 0) { stepDev[0] = step; if (normOut) normOut[0] = norm; }
}

/* ------------------------------------------------------------------------------------------------
 * compute_advantages after the GAE scan (ppo.py:167-176): normaliser updates and advantage / value
 * normalisation.  The reference (and ppo_dist.finish_advantages) does this with ~40 torch kernels, a
 * float64 copy of the 262 144 x 128 observation batch and two host synchronisations (1.3 ms); here one
 * pass computes every column's sum and sum of squares in float64 (bb_ppo_moments_kernel + a block-order
 * reduction), the sums of all ranks are added with ONE all-reduce, a single block applies the
 * RunningMeanStd updates (agent.py:41-51), and one elementwise kernel normalises the three buffers.
 * ------------------------------------------------------------------------------------------------ */
constexpr int kMomCols = BB_OBS_DIM + 3;      /* 128 observation columns, values, returns, advantages */
constexpr int kMomThreads = 256;

__global__ void __launch_bounds__(kMomThreads) bb_ppo_moments_kernel(const float *obs, const float *vals, const float *rets, const float *advs,
                                                                     int64_t rows, double *partials)
{
    __shared__ double sh[kMomThreads / 32][BB_OBS_DIM][2];
    __shared__ double sc[kMomThreads / 32][6];
    const int tid = threadIdx.x, c4 = tid & 31, rsub = tid >> 5;
    double s[4] = {0, 0, 0, 0}, q[4] = {0, 0, 0, 0};
    for (int64_t r = (int64_t)blockIdx.x * 8 + rsub; r < rows; r += (int64_t)gridDim.x * 8) {
        const float4 x = *reinterpret_cast<const float4 *>(obs + r * BB_OBS_DIM + c4 * 4);
        s[0] += x.x; s[1] += x.y; s[2] += x.z; s[3] += x.w;
        q[0] = fma((double)x.x, (double)x.x, q[0]); q[1] = fma((double)x.y, (double)x.y, q[1]);
        q[2] = fma((double)x.z, (double)x.z, q[2]); q[3] = fma((double)x.w, (double)x.w, q[3]);
    }
#pragma unroll
    for (int i = 0; i < 4; i++) { sh[rsub][c4 * 4 + i][0] = s[i]; sh[rsub][c4 * 4 + i][1] = q[i]; }
    double t[6] = {0, 0, 0, 0, 0, 0};
    for (int64_t i = (int64_t)blockIdx.x * kMomThreads + tid; i < rows; i += (int64_t)gridDim.x * kMomThreads) {
        const double v = vals[i], r = rets[i], a = advs[i];
        t[0] += v; t[1] += r; t[2] += a;
        t[3] = fma(v, v, t[3]); t[4] = fma(r, r, t[4]); t[5] = fma(a, a, t[5]);
    }
#pragma unroll
    for (int k = 0; k < 6; k++) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t[k] += __shfl_xor_sync(0xffffffffu, t[k], o);
    }
    if ((tid & 31) == 0) for (int k = 0; k < 6; k++) sc[tid >> 5][k] = t[k];
    __syncthreads();
    double *out = partials + (int64_t)blockIdx.x * 2 * kMomCols;
    if (tid < BB_OBS_DIM) {
        double a = 0, b = 0;
        for (int w = 0; w < kMomThreads / 32; w++) { a += sh[w][tid][0]; b += sh[w][tid][1]; }
        out[tid] = a; out[kMomCols + tid] = b;
    } else if (tid < BB_OBS_DIM + 3) {
        const int k = tid - BB_OBS_DIM;
        double a = 0, b = 0;
        for (int w = 0; w < kMomThreads / 32; w++) { a += sc[w][k]; b += sc[w][3 + k]; }
        out[tid] = a; out[kMomCols + tid] = b;
    }
}

__global__ void bb_ppo_moments_reduce_kernel(const double *partials, int numBlocks, double *sums)
{
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= 2 * kMomCols) return;
    double s = 0;
    for (int b = 0; b < numBlocks; b++) s += partials[(int64_t)b * 2 * kMomCols + j];
    sums[j] = s;
}

struct NormState { double *mean, *var, *count; };

/* RunningMeanStd.update (agent.py:41-51) for one column given the batch's sum / sum of squares / count */
BBD void rms_update(double sum, double sumsq, double n, double &mean, double &var, double &count)
{
    const double bmean = sum / n;
    const double bvar = fmax(sumsq - n * bmean * bmean, 0.0) / fmax(n - 1.0, 1.0); /* unbiased, torch.var_mean */
    const double tot = n + count;
    const double delta = bmean - mean;
    const double m = var * count + bvar * n + delta * delta * count * n / tot;
    mean = mean + delta * n / tot;
    var = m / tot;
    count = tot;
}

/* one block: sums are GLOBAL (all ranks), n the global row count.  scal[4] = {adv mean, adv std + 1e-8, value mean f32, value rstd f32} */
__global__ void __launch_bounds__(BB_OBS_DIM) bb_ppo_finish_stats_kernel(const double *sums, double n, NormState obsN, NormState valN,
                                                                         float *obsMeanF, float *obsInvStdF, float *scal)
{
    const int c = threadIdx.x;
    {
        double mean = obsN.mean[c], var = obsN.var[c], count = obsN.count[0];
        rms_update(sums[c], sums[kMomCols + c], n, mean, var, count);
        obsN.mean[c] = mean; obsN.var[c] = var;
        /* agent.py:28-30,36: float32 casts, then + epsilon, rsqrt */
        obsMeanF[c] = (float)mean;
        obsInvStdF[c] = rsqrtf((float)var + 1e-5f);
        __syncthreads(); /* every thread has read the old count */
        if (c == 0) obsN.count[0] = count;
    }
    if (c == 0) {
        double mean = valN.mean[0], var = valN.var[0], count = valN.count[0];
        rms_update(sums[BB_OBS_DIM + 0], sums[kMomCols + BB_OBS_DIM + 0], n, mean, var, count);   /* value_norm.update(values) */
        rms_update(sums[BB_OBS_DIM + 1], sums[kMomCols + BB_OBS_DIM + 1], n, mean, var, count);   /* value_norm.update(returns) */
        valN.mean[0] = mean; valN.var[0] = var; valN.count[0] = count;
        const double am = sums[BB_OBS_DIM + 2] / n;
        const double av = fmax(sums[kMomCols + BB_OBS_DIM + 2] - n * am * am, 0.0) / fmax(n - 1.0, 1.0);
        scal[0] = (float)am;
        scal[1] = (float)sqrt(av) + 1e-8f;
        scal[2] = (float)mean;
        scal[3] = rsqrtf((float)var + 1e-5f);
    }
}

/* ppo.py:173-176 elementwise: advantages normalised; values / returns through the UPDATED value normaliser (clamped) */
__global__ void bb_ppo_finish_apply_kernel(const float *scal, float clampV, float *vals, float *rets, float *advs, int64_t rows)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= rows) return;
    const float am = scal[0], as = scal[1], vm = scal[2], vr = scal[3];
    advs[i] = (advs[i] - am) / as;
    float v = (vals[i] - vm) * vr, r = (rets[i] - vm) * vr;
    if (clampV > 0.f) { v = fminf(fmaxf(v, -clampV), clampV); r = fminf(fmaxf(r, -clampV), clampV); }
    vals[i] = v; rets[i] = r;
}

/* keys / identity values for bb_randperm */
__global__ void bb_randperm_keys_kernel(uint32_t *keys, int32_t *vals, int64_t n, uint32_t seed, uint32_t stream)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t a, b;
    threefry2x32(seed, stream, (uint32_t)i, 0u, a, b);
    keys[i] = a;
    vals[i] = (int32_t)i;
}

/* ------------------------------------------------------------------------------------------------
 * bb_ppo_adam_p2p_kernel: one-shot all-reduce of the flat gradient over NVLink peer memory + clip + Adam.
 * Push protocol, flag in the data (no fences, no separate flag round trip): the exchange buffer of a rank (device
 * memory, mapped by every peer through CUDA IPC) is
 *     uint64_t slot[2 parities][world_size writers][stride]       word = { float gradient element, uint32 sequence number }
 * A rank stores its gradient, element by element as single 8-byte words tagged with the call's sequence number, into
 * slot[seq & 1][its rank] of EVERY peer (remote stores are fire-and-forget and pipeline over NVLink), then polls its
 * OWN copy of the slots (local L2) until every word carries the sequence number, and adds the ranks' values in rank
 * order, so every rank computes the same bits.  8-byte stores are single-copy atomic: a word with the right tag holds
 * the right value.  Two parities: a rank writes step s + 2 into a slot only after it finished step s + 1, which needed
 * the owner's step s + 1 words, which the owner sends after its step-s kernel (the reader of that slot) has ended.
 * ------------------------------------------------------------------------------------------------ */
constexpr int kP2pMaxRanks = 16;
constexpr int kP2pPer = 6;   /* elements per thread: up to 6 * 1024 parameters */
struct P2pPeers { unsigned long long *buf[kP2pMaxRanks]; };
BBD unsigned long long ld_volatile_u64(const unsigned long long *p) { unsigned long long v; asm volatile("ld.volatile.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory"); return v; }
BBD void st_volatile_u64(unsigned long long *p, unsigned long long v) { asm volatile("st.volatile.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory"); }
BBD uint64_t p2p_timer_ns() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

__global__ void __launch_bounds__(kAdamThreads) bb_ppo_adam_p2p_kernel(const P2pPeers peers, int rank, int ws, int64_t stride, uint32_t seq,
                                                                        float *params, const float *grad, float *m, float *v,
                                                                        int32_t *stepDev, int64_t n, float maxNorm, float lr, float beta1,
                                                                        float beta2, float eps, float *normOut, int32_t *errFlag,
                                                                        uint64_t timeoutNs)
{
    __shared__ float red[kAdamThreads / 32];
    __shared__ float sNorm;
    __shared__ int sBad;
    if (threadIdx.x == 0) sBad = 0;
    asm volatile("griddepcontrol.wait;" ::: "memory");
    /* sticky: once a call has given up on a peer this rank neither pushes nor applies again, so its peers run into
     * their own timeout on the next call and every rank raises -- no replica keeps training on diverged parameters */
    if (errFlag && *errFlag != 0) return;
    const int64_t parOff = (int64_t)(seq & 1u) * ws * stride;
    /* 1. push this rank's gradient into its slot at every peer */
    float own[kP2pPer];
#pragma unroll
    for (int k = 0; k < kP2pPer; k++) {
        const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
        own[k] = i < n ? grad[i] : 0.f;
    }
    for (int d = 1; d < ws; d++) {
        int r = rank + d; if (r >= ws) r -= ws;     /* staggered: the ranks do not all hit the same peer first */
        unsigned long long *dst = peers.buf[r] + parOff + (int64_t)rank * stride;
#pragma unroll
        for (int k = 0; k < kP2pPer; k++) {
            const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
            if (i < n) st_volatile_u64(dst + i, ((unsigned long long)seq << 32) | (unsigned long long)__float_as_uint(own[k]));
        }
    }
    __syncthreads();
    /* 2. gather from the local slots, two writers at a time, summing in rank order (bounded wait) */
    const unsigned long long *mine = peers.buf[rank] + parOff;
    float g[kP2pPer];
#pragma unroll
    for (int k = 0; k < kP2pPer; k++) g[k] = 0.f;
    const uint64_t t0 = p2p_timer_ns();
    for (int r0 = 0; r0 < ws; r0 += 2) {
        unsigned long long w[2][kP2pPer];
        unsigned spins = 0;
        bool pending = true;
        while (pending) {
            pending = false;
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int r = r0 + j;
                if (r >= ws || r == rank) continue;
#pragma unroll
                for (int k = 0; k < kP2pPer; k++) {
                    const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
                    if (i < n) w[j][k] = ld_volatile_u64(mine + (int64_t)r * stride + i);
                }
            }
#pragma unroll
            for (int j = 0; j < 2; j++) {
                const int r = r0 + j;
                if (r >= ws || r == rank) continue;
#pragma unroll
                for (int k = 0; k < kP2pPer; k++) {
                    const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
                    if (i < n && (uint32_t)(w[j][k] >> 32) != seq) pending = true;
                }
            }
            if (pending && (++spins & 63u) == 0 && p2p_timer_ns() - t0 > timeoutNs) { sBad = 1; break; }
        }
        if (pending) break;
#pragma unroll
        for (int j = 0; j < 2; j++) {
            const int r = r0 + j;
            if (r >= ws) continue;
#pragma unroll
            for (int k = 0; k < kP2pPer; k++) {
                const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
                if (i < n) g[k] += (r == rank) ? own[k] : __uint_as_float((uint32_t)w[j][k]);
            }
        }
    }
    __syncthreads();
    if (sBad) { if (threadIdx.x == 0 && errFlag) *errFlag = 1; return; }
    /* 3. clip + Adam (arithmetic of bb_ppo_adam_kernel with gradScale = 1 / ws) */
    const float scale = 1.f / (float)ws;
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < kP2pPer; k++) { g[k] *= scale; s = fmaf(g[k], g[k], s); }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float t = red[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
        if (threadIdx.x == 0) sNorm = sqrtf(t);
    }
    __syncthreads();
    const float norm = sNorm;
    const float clip = maxNorm > 0.f ? fminf(maxNorm / (norm + 1e-6f), 1.f) : 1.f;
    const int32_t step = stepDev[0] + 1;
    const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
    const float stepSize = (float)((double)lr / bc1);
    const float bc2Sqrt = (float)sqrt(bc2);
#pragma unroll
    for (int k = 0; k < kP2pPer; k++) {
        const int64_t i = threadIdx.x + (int64_t)k * kAdamThreads;
        if (i >= n) continue;
        const float gi = g[k] * clip;
        const float mi = m[i] + (gi - m[i]) * (1.f - beta1);
        const float vi = fmaf(beta2, v[i], (1.f - beta2) * gi * gi);
        m[i] = mi; v[i] = vi;
        const float denom = sqrtf(vi) / bc2Sqrt + eps;
        params[i] -= stepSize * (mi / denom);
    }
    __syncthreads();
    if (threadIdx.x == 0) { stepDev[0] = step; if (normOut) normOut[0] = norm; }
}

} // namespace bb
