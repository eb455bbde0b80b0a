/*
 * bb_umma.cuh — 5th-generation tensor-core building blocks (tcgen05 / TMEM) for the policy contraction, sm_100a.
 *
 * The policy network (scripts/agent.py:108-154) is three small dense layers over a tile of observation rows:
 * [rows x 128] . W1^T[128 x 32], [rows x 32] . W2^T, [rows x 32] . Wh^T.  With M = 128 rows per CTA each layer is one
 * tcgen05.mma shape (M = 128, N = 32, kind::tf32, fp32 accumulator in tensor memory): a single thread issues the MMAs,
 * the accumulator tile lives in TMEM with row i in TMEM lane i, and tcgen05.ld 32x32b hands every thread the 32 channels of
 * ITS row -- so LayerNorm, ReLU, softmax and sampling are thread-local, with no quad shuffles.
 *
 * fp32-level accuracy (the parity bar against torch fp32) comes from 3xTF32: every operand is split x = hi + lo with
 * hi = x & 0xffffe000 (what the tensor core reads) and lo = x - hi (exact), and each product is issued as
 * hi.lo + lo.hi + hi.hi into the same accumulator.
 *
 * Shared-memory operand layout: the canonical K-major, no-swizzle UMMA layout (cute: "INTERLEAVE", ((8,n),2):((1,SBO),LBO) in
 * 16-byte units): core matrices of 8 rows x 16 bytes (four tf32), 128 contiguous bytes each.  Element (row r, k) of an
 * [R x K] operand sits at byte   (k / 4) * LBO + (r / 8) * 128 + (r % 8) * 16 + (k % 4) * 4,   LBO = (R / 8) * 128, SBO = 128.
 * A thread that owns row r writes four consecutive k as one 16-byte store; the 32 lanes of a warp (32 consecutive rows)
 * write 512 contiguous bytes: conflict-free.  One MMA instruction consumes K = 8 (two core matrices along K), so k-step s
 * starts at byte 2 s LBO.
 *
 * Descriptor bit layouts follow cute/arch/mma_sm100_desc.hpp (SmemDescriptor, InstrDescriptor) of the CUTLASS tree vendored in
 * this image; the PTX strings follow cute/arch/{mma_sm100_umma,copy_sm100,tmem_allocator_sm100}.hpp and cutlass/arch/barrier.h.
 */
#pragma once
#include <cstdint>

namespace bb {

#define BBU __device__ __forceinline__

BBU uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

/* 64-bit shared-memory matrix descriptor, SWIZZLE_NONE, version 1 (sm_100) */
BBU uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes)
{
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3fffu);          /* start address, 16-byte units, bits [0,14)  */
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;    /* leading (K) byte offset, bits [16,30)      */
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;    /* stride (M/N) byte offset, bits [32,46)     */
    d |= (uint64_t)1 << 46;                               /* descriptor version = 1 (Blackwell)         */
    return d;                                             /* base offset 0, layout type 0 = no swizzle  */
}

/* 32-bit instruction descriptor: D f32, A and B tf32, both K-major, dense */
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N)
{
    return (1u << 4)                      /* c_format = F32  */
         | (2u << 7) | (2u << 10)         /* a_format = b_format = TF32 */
         | ((uint32_t)(N >> 3) << 17)     /* n_dim */
         | ((uint32_t)(M >> 4) << 24);    /* m_dim */
}

/* D[tmem] (+)= A[smem] . B[smem]^T, issued by ONE thread for the CTA */
BBU void umma_tf32_ss(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate)
{
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}\n"
                 :: "r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
/* mbarrier arrive when every MMA issued so far by this thread has completed (implies tcgen05.fence::before_thread_sync) */
BBU void umma_commit(uint64_t *mbar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(mbar)) : "memory");
}
BBU void mbar_init(uint64_t *mbar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(mbar)), "r"(count) : "memory");
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
/* bounded: a tensor-core operation that never completes (a malformed descriptor) must not hang the GPU; returns false after
 * ~2^24 polls (tens of milliseconds), the callers then publish an error value instead of a result */
BBU bool mbar_wait(uint64_t *mbar, uint32_t parity)
{
    const uint32_t a = smem_u32(mbar);
    uint32_t done = 0;
    for (uint32_t spins = 0; !done && spins < (1u << 24); spins++) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n"
                     : "=r"(done) : "r"(a), "r"(parity) : "memory");
    }
    return done != 0;
}
BBU void tc_fence_before_sync() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
BBU void tc_fence_after_sync() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
BBU void fence_smem_to_async_proxy() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

/* whole warp: allocate `cols` TMEM columns (power of two >= 32); the base address lands in *dst (shared memory) */
BBU void tmem_alloc(uint32_t *dst, uint32_t cols)
{
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(dst)), "r"(cols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
BBU void tmem_free(uint32_t addr, uint32_t cols)
{
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(addr), "r"(cols) : "memory");
}
/* whole warp w: lane i receives columns col .. col + 31 of TMEM lane 32 w + i (one accumulator row) */
BBU void tmem_ld_row32(uint32_t tmem_base, int warp, int col, float (&v)[32])
{
    uint32_t r[32];
    const uint32_t addr = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)col;
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(addr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}

/* ---- operand staging: 3xTF32 split of four consecutive k of one row, into the hi and lo copies of a K-major operand ---- */
BBU void umma_split4(float4 x, uint4 &hi, uint4 &lo)
{
    hi.x = __float_as_uint(x.x) & 0xffffe000u; lo.x = __float_as_uint(x.x - __uint_as_float(hi.x));
    hi.y = __float_as_uint(x.y) & 0xffffe000u; lo.y = __float_as_uint(x.y - __uint_as_float(hi.y));
    hi.z = __float_as_uint(x.z) & 0xffffe000u; lo.z = __float_as_uint(x.z - __uint_as_float(hi.z));
    hi.w = __float_as_uint(x.w) & 0xffffe000u; lo.w = __float_as_uint(x.w - __uint_as_float(hi.w));
}
/* byte offset of (row r, k-chunk kc) in an [R x K] operand (see the header comment) */
BBU uint32_t umma_off(int r, int kc, int R) { return (uint32_t)(kc * (R / 8) * 128 + (r >> 3) * 128 + (r & 7) * 16); }

/* One thread: D[128 x 32] (+)= A[128 x K] . B[32 x K]^T as 3xTF32: for every k-step hi.lo, lo.hi, hi.hi.
 * aHi / aLo / bHi / bLo: shared addresses of the operand copies; `first` clears the accumulator on the very first MMA. */
BBU void umma_issue_3xtf32(uint32_t tmem_d, uint32_t aHi, uint32_t aLo, uint32_t bHi, uint32_t bLo, int K, bool first)
{
    constexpr uint32_t idesc = umma_idesc_tf32(128, 32);
    const uint32_t lboA = (128 / 8) * 128, lboB = (32 / 8) * 128;
    uint32_t acc = first ? 0u : 1u;
    for (int s = 0; s < K / 8; s++) {
        const uint32_t oa = (uint32_t)(2 * s) * lboA, ob = (uint32_t)(2 * s) * lboB;
        const uint64_t dAh = umma_smem_desc(aHi + oa, lboA, 128), dAl = umma_smem_desc(aLo + oa, lboA, 128);
        const uint64_t dBh = umma_smem_desc(bHi + ob, lboB, 128), dBl = umma_smem_desc(bLo + ob, lboB, 128);
        umma_tf32_ss(tmem_d, dAh, dBl, idesc, acc);
        umma_tf32_ss(tmem_d, dAl, dBh, idesc, 1u);
        umma_tf32_ss(tmem_d, dAh, dBh, idesc, 1u);
        acc = 1u;
    }
}

/* ---- self-test kernel (bb_umma_selftest): D[128 x 32] = A[128 x K] . B[32 x K]^T, one CTA of 128 threads ---- */
__global__ void __launch_bounds__(128) bb_umma_selftest_kernel(const float *A, const float *B, float *D, int K)
{
    extern __shared__ __align__(128) unsigned char usm[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t tmemBase;
    const int tid = threadIdx.x, warp = tid >> 5;
    unsigned char *aHi = usm, *aLo = aHi + 128 * K * 4, *bHi = aLo + 128 * K * 4, *bLo = bHi + 32 * K * 4;
    if (warp == 0) tmem_alloc(&tmemBase, 32);
    if (tid == 0) mbar_init(&mbar, 1);
    for (int kc = 0; kc < K / 4; kc++) {
        uint4 hi, lo;
        umma_split4(*reinterpret_cast<const float4 *>(A + (int64_t)tid * K + kc * 4), hi, lo);
        *reinterpret_cast<uint4 *>(aHi + umma_off(tid, kc, 128)) = hi;
        *reinterpret_cast<uint4 *>(aLo + umma_off(tid, kc, 128)) = lo;
        if (tid < 32) {
            umma_split4(*reinterpret_cast<const float4 *>(B + (int64_t)tid * K + kc * 4), hi, lo);
            *reinterpret_cast<uint4 *>(bHi + umma_off(tid, kc, 32)) = hi;
            *reinterpret_cast<uint4 *>(bLo + umma_off(tid, kc, 32)) = lo;
        }
    }
    fence_smem_to_async_proxy();
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tmemBase;
    if (tid == 0) {
        umma_issue_3xtf32(tmem, smem_u32(aHi), smem_u32(aLo), smem_u32(bHi), smem_u32(bLo), K, true);
        umma_commit(&mbar);
    }
    const bool ok = mbar_wait(&mbar, 0);
    tc_fence_after_sync();
    float v[32];
    if (ok) tmem_ld_row32(tmem, warp, 0, v);   /* warp-uniform */
#pragma unroll
    for (int i = 0; i < 32; i++) D[tid * 32 + i] = ok ? v[i] : -12345.f;
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_free(tmem, 32);
}

/* ------------------------------------------------------------------------------------------------
 * bb_policy_umma_kernel — Agent.forward (scripts/agent.py:134-154, action.py:20-33) on the 5th-generation tensor cores.
 *
 * One CTA of 128 threads = one tile of 128 observation rows, one row per thread (grid-strided over the tiles).
 *   stage    the thread reads its 512-byte row (32 x LDG.128), copies it to the rollout buffer, normalises it
 *            (agent.py:27-39), splits it hi / lo and writes the two K-major operand copies (64 KB each);
 *   layer 1  one thread issues 16 k-steps x 3 tcgen05.mma (M 128, N 32, K 8, kind::tf32) into TMEM columns 0..31,
 *            commits to an mbarrier; every thread then reads ITS row's 32 channels with tcgen05.ld, adds the bias and does
 *            LayerNorm + ReLU on registers, and writes the row as the next operand (16 KB per copy, over the old one);
 *   layer 2  4 k-steps x 3 MMAs into columns 32..63, same epilogue;
 *   heads    4 k-steps x 3 MMAs into columns 64..95 (19 logits + value, 12 zero columns);
 *   tail     the six bucketed categoricals, the summed log-probability and the rollout bookkeeping, thread per row, with the
 *            arithmetic of bb_policy_kernel's tail (same softmax / inverse-CDF expressions, same Threefry words).
 * Weights are staged once per CTA as hi / lo K-major operands (W1 2 x 16 KB, W2 and heads 2 x 4 KB each) before
 * griddepcontrol.wait (they do not depend on the preceding step kernel).  Layer 1 is issued as two K = 64 passes through one
 * 2 x 32 KB operand buffer (the row waits in registers), which keeps the CTA at 114 KB of shared memory: in the rollout chain
 * its blocks must become resident NEXT TO the step kernel's (programmatic dependent launch overlaps this kernel's weight
 * staging with the step's tail and the next step's launch with this kernel's tail); at 178 KB they could not, and the rollout
 * lost 7 us per step.
 * ------------------------------------------------------------------------------------------------ */
constexpr int kUmRows = 128;
constexpr int kUmKHalf = BB_OBS_DIM / 2;           /* layer 1 runs as two K = 64 passes over one 32 KB operand buffer per copy */
constexpr int kUmA = kUmRows * kUmKHalf * 4;       /* 32768: one copy of (half of) the layer-1 A operand */
constexpr int kUmW1 = kH * BB_OBS_DIM * 4;          /* 16384 */
constexpr int kUmW2 = kH * kH * 4;                  /* 4096  */
constexpr int kUmOffW1 = 2 * kUmA;
constexpr int kUmOffW2 = kUmOffW1 + 2 * kUmW1;
constexpr int kUmOffWh = kUmOffW2 + 2 * kUmW2;
constexpr int kUmOffPar = kUmOffWh + 2 * kUmW2;
constexpr int kUmParFloats = 2 * BB_OBS_DIM + 7 * kH;   /* mean, inv, b1 g1 e1 b2 g2 e2, bh(32) */
constexpr int kUmSmemBytes = kUmOffPar + kUmParFloats * 4;

/* nn.LayerNorm(H) + ReLU over one thread's 32 channels (channels >= H hold 0 and stay 0); biased variance, eps 1e-5 */
BBU void um_layer_norm_relu(float (&v)[kH], int H, const float *g, const float *b)
{
    float s = 0.f;
#pragma unroll
    for (int c = 0; c < kH; c++) s += v[c];
    const float inv = 1.f / (float)H;
    const float mean = s * inv;
    s = 0.f;
#pragma unroll
    for (int c = 0; c < kH; c++) {
        v[c] = (c < H) ? v[c] - mean : 0.f;
        s = fmaf(v[c], v[c], s);
    }
    const float rstd = rsqrtf(s * inv + 1e-5f);
#pragma unroll
    for (int c = 0; c < kH; c++) v[c] = (c < H) ? fmaxf(fmaf(v[c] * rstd, g[c], b[c]), 0.f) : 0.f;
}

/* the row's 32 activations -> hi / lo copies of a [128 x 32] K-major operand */
BBU void um_store_row32(unsigned char *hiBase, unsigned char *loBase, int row, const float (&v)[kH])
{
#pragma unroll
    for (int kc = 0; kc < kH / 4; kc++) {
        uint4 hi, lo;
        umma_split4(make_float4(v[4 * kc], v[4 * kc + 1], v[4 * kc + 2], v[4 * kc + 3]), hi, lo);
        const uint32_t off = umma_off(row, kc, kUmRows);
        *reinterpret_cast<uint4 *>(hiBase + off) = hi;
        *reinterpret_cast<uint4 *>(loBase + off) = lo;
    }
}

__global__ void __launch_bounds__(kUmRows, 1) bb_policy_umma_kernel(const bb_policy P, const PolicyIo io)
{
    extern __shared__ __align__(128) unsigned char usm[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t tmemBase;
    const int H = P.hidden;
    const int tid = threadIdx.x, warp = tid >> 5;
    unsigned char *aHi = usm, *aLo = usm + kUmA;
    unsigned char *w1Hi = usm + kUmOffW1, *w1Lo = w1Hi + kUmW1;
    unsigned char *w2Hi = usm + kUmOffW2, *w2Lo = w2Hi + kUmW2;
    unsigned char *whHi = usm + kUmOffWh, *whLo = whHi + kUmW2;
    float *par = reinterpret_cast<float *>(usm + kUmOffPar);
    float *sMean = par, *sInv = par + BB_OBS_DIM, *sB1 = sInv + BB_OBS_DIM, *sG1 = sB1 + kH, *sE1 = sG1 + kH;
    float *sB2 = sE1 + kH, *sG2 = sB2 + kH, *sE2 = sG2 + kH, *sBh = sE2 + kH;

    if (warp == 0) tmem_alloc(&tmemBase, 128);
    if (tid == 0) mbar_init(&mbar, 1);
    /* ---- weights -> hi / lo K-major operands (consecutive threads = consecutive output rows: 512 contiguous bytes per chunk) ---- */
    for (int i = tid; i < kH * (BB_OBS_DIM / 4); i += kUmRows) {
        const int kc = i / kH, j = i % kH;
        const float4 w = j < H ? *reinterpret_cast<const float4 *>(P.w1 + j * BB_OBS_DIM + kc * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
        uint4 hi, lo;
        umma_split4(w, hi, lo);
        *reinterpret_cast<uint4 *>(w1Hi + umma_off(j, kc, kH)) = hi;
        *reinterpret_cast<uint4 *>(w1Lo + umma_off(j, kc, kH)) = lo;
    }
    for (int i = tid; i < kH * (kH / 4); i += kUmRows) {
        const int kc = i / kH, j = i % kH;
        float w2v[4], whv[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int k = kc * 4 + e;
            w2v[e] = (j < H && k < H) ? P.w2[j * H + k] : 0.f;
            whv[e] = (k < H && j < kHeads) ? (j < BB_NUM_LOGITS ? P.wa[j * H + k] : P.wc[k]) : 0.f;
        }
        uint4 hi, lo;
        umma_split4(make_float4(w2v[0], w2v[1], w2v[2], w2v[3]), hi, lo);
        *reinterpret_cast<uint4 *>(w2Hi + umma_off(j, kc, kH)) = hi;
        *reinterpret_cast<uint4 *>(w2Lo + umma_off(j, kc, kH)) = lo;
        umma_split4(make_float4(whv[0], whv[1], whv[2], whv[3]), hi, lo);
        *reinterpret_cast<uint4 *>(whHi + umma_off(j, kc, kH)) = hi;
        *reinterpret_cast<uint4 *>(whLo + umma_off(j, kc, kH)) = lo;
    }
    if (tid < BB_OBS_DIM) { sMean[tid] = P.obs_mean[tid]; sInv[tid] = P.obs_inv_std[tid]; }
    if (tid < kH) {
        const bool a = tid < H;
        sB1[tid] = a ? P.b1[tid] : 0.f; sG1[tid] = a ? P.ln1_g[tid] : 0.f; sE1[tid] = a ? P.ln1_b[tid] : 0.f;
        sB2[tid] = a ? P.b2[tid] : 0.f; sG2[tid] = a ? P.ln2_g[tid] : 0.f; sE2[tid] = a ? P.ln2_b[tid] : 0.f;
        sBh[tid] = tid < BB_NUM_LOGITS ? P.ba[tid] : (tid == BB_NUM_LOGITS ? P.bc[0] : 0.f);
    }
    /* programmatic dependent launch (bb_rollout): everything below depends on the step kernel that precedes this launch */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem = tmemBase;
    const uint32_t aHiS = smem_u32(aHi), aLoS = smem_u32(aLo);
    uint32_t phase = 0;
    bool ok = true;

    for (int64_t row0 = (int64_t)blockIdx.x * kUmRows; row0 < io.n; row0 += (int64_t)gridDim.x * kUmRows) {
        const int64_t rr = row0 + tid;
        const bool valid = rr < io.n;
        const int64_t r = valid ? rr : io.n - 1; /* tail threads shadow the last row and never store */
        /* ---- the row: requested up front (32 independent 16-byte loads per thread in flight: with one thread per row the
         * loads of a warp are strided by the row pitch, so their latency is paid once, not per chunk), copied to the rollout
         * buffer and normalised in registers (agent.py:27-39) ---- */
        float4 xr[BB_OBS_DIM / 4];
        {
            const float4 *orow = reinterpret_cast<const float4 *>(io.obs + r * io.obsStride);
            float4 *brow = (io.bufObs && valid) ? reinterpret_cast<float4 *>(io.bufObs + r * BB_OBS_DIM) : nullptr;
#pragma unroll
            for (int kc = 0; kc < BB_OBS_DIM / 4; kc++) xr[kc] = orow[kc];
#pragma unroll
            for (int kc = 0; kc < BB_OBS_DIM / 4; kc++) {
                const float4 x = xr[kc];
                if (brow) brow[kc] = x;   /* buffer.obs[step] = obs (ppo.py:128) */
                const float4 mu = *reinterpret_cast<const float4 *>(sMean + kc * 4);
                const float4 is = *reinterpret_cast<const float4 *>(sInv + kc * 4);
                float4 xn = make_float4((x.x - mu.x) * is.x, (x.y - mu.y) * is.y, (x.z - mu.z) * is.z, (x.w - mu.w) * is.w);
                if (P.obs_clamp > 0.f) {
                    xn.x = fminf(fmaxf(xn.x, -P.obs_clamp), P.obs_clamp); xn.y = fminf(fmaxf(xn.y, -P.obs_clamp), P.obs_clamp);
                    xn.z = fminf(fmaxf(xn.z, -P.obs_clamp), P.obs_clamp); xn.w = fminf(fmaxf(xn.w, -P.obs_clamp), P.obs_clamp);
                }
                xr[kc] = xn;
            }
        }
        float v[kH];
        /* ---- four tensor-core passes: layer 1 (k 0..63), layer 1 (k 64..127, accumulating), layer 2, heads.  Each pass: the
         * threads write their row's operand chunks, make them visible to the async proxy, one thread issues, everyone waits on
         * the mbarrier (so the operand buffer is free again) ---- */
#pragma unroll 1
        for (int pass = 0; pass < 4; pass++) {
            if (pass < 2) {
#pragma unroll
                for (int kc = 0; kc < kUmKHalf / 4; kc++) {
                    uint4 hi, lo;
                    const float4 x = pass == 0 ? xr[kc] : xr[kc + kUmKHalf / 4];
                    umma_split4(x, hi, lo);
                    const uint32_t off = umma_off(tid, kc, kUmRows);
                    *reinterpret_cast<uint4 *>(aHi + off) = hi;
                    *reinterpret_cast<uint4 *>(aLo + off) = lo;
                }
            }
            fence_smem_to_async_proxy();
            tc_fence_before_sync();
            __syncthreads();
            tc_fence_after_sync();
            const int layer = pass == 0 ? 0 : pass - 1;
            if (tid == 0) {
                if (pass < 2) {   /* W1's k-chunks 16 p .. 16 p + 15: chunk kc of the [32 x 128] operand starts at kc * LBO_B */
                    const uint32_t wOff = (uint32_t)pass * (uint32_t)(kUmKHalf / 4) * (uint32_t)((kH / 8) * 128);
                    umma_issue_3xtf32(tmem, aHiS, aLoS, smem_u32(w1Hi) + wOff, smem_u32(w1Lo) + wOff, kUmKHalf, pass == 0);
                } else {
                    umma_issue_3xtf32(tmem + 32u * (uint32_t)layer, aHiS, aLoS, smem_u32(layer == 1 ? w2Hi : whHi),
                                      smem_u32(layer == 1 ? w2Lo : whLo), kH, true);
                }
                umma_commit(&mbar);
            }
            ok = mbar_wait(&mbar, phase) && ok;
            phase ^= 1u;
            tc_fence_after_sync();
            if (pass == 0) continue;
            tmem_ld_row32(tmem, warp, 32 * layer, v);
            if (layer < 2) {
                const float *bb = layer == 0 ? sB1 : sB2;
#pragma unroll
                for (int c = 0; c < kH; c++) v[c] += bb[c];
                um_layer_norm_relu(v, H, layer == 0 ? sG1 : sG2, layer == 0 ? sE1 : sE2);
                um_store_row32(aHi, aLo, tid, v);   /* the operand buffer is free: the MMAs that read it have completed */
            }
        }
        /* ---- heads: v[0..18] actor logits, v[19] critic ---- */
#pragma unroll
        for (int c = 0; c < kHeads; c++) v[c] += sBh[c];
        if (!ok) {   /* a tensor-core batch never completed: make the failure visible instead of publishing garbage */
#pragma unroll
            for (int c = 0; c < kHeads; c++) v[c] = __int_as_float(0x7fc00000);
        }
        if (valid) {
            if (io.logitsOut) {
#pragma unroll
                for (int c = 0; c < BB_NUM_LOGITS; c++) io.logitsOut[r * BB_NUM_LOGITS + c] = v[c];
            }
            if (io.evaluateOnly) { if (io.nextValue) io.nextValue[r] = v[BB_NUM_LOGITS]; }
            else if (io.valueOut) io.valueOut[r] = v[BB_NUM_LOGITS];
        }
        if (!io.evaluateOnly) {
            /* ---- bucketed categoricals (action.py:5-33), the expressions of bb_policy_kernel's tail, bucket by bucket ---- */
            float tot = 0.f;
            uint32_t x0 = 0, x1 = 0;
#pragma unroll
            for (int b = 0; b < BB_NUM_ACTION_DIMS; b++) {
                constexpr int kOff[BB_NUM_ACTION_DIMS] = {0, 2, 10, 13, 15, 17};
                constexpr int kCnt[BB_NUM_ACTION_DIMS] = {2, 8, 3, 2, 2, 2};
                const int off = kOff[b], cnt = kCnt[b];
                float m = v[off];
#pragma unroll
                for (int i = 1; i < cnt; i++) m = fmaxf(m, v[off + i]);
                float e[8];
                float s = 0.f;
#pragma unroll
                for (int i = 0; i < cnt; i++) { e[i] = expf(v[off + i] - m); s += e[i]; }
                const float lse = m + logf(s);
                int a = 0;
                float la = v[off];
                if (io.stochastic) { /* inverse CDF: smallest a with u * s < e_0 + .. + e_a; word (b & 1) of Threefry call b >> 1 */
                    if ((b & 1) == 0) threefry2x32(io.seed, (uint32_t)(io.rowOffset + r), io.draw, (uint32_t)(b >> 1), x0, x1);
                    const float us = ((float)(((b & 1) ? x1 : x0) >> 8) * 0x1p-24f) * s;
                    float c = 0.f;
                    a = cnt - 1;
                    la = v[off + cnt - 1];
                    bool found = false;
#pragma unroll
                    for (int i = 0; i < cnt; i++) {
                        c += e[i];
                        if (!found && us < c) { a = i; la = v[off + i]; found = true; }
                    }
                } else {
#pragma unroll
                    for (int i = 1; i < cnt; i++) if (v[off + i] > la) { a = i; la = v[off + i]; }
                }
                tot += la - lse;
                if (valid) {
                    if (io.actionsOut) io.actionsOut[r * io.actionStride + b] = a;
                    if (io.bufActions) io.bufActions[r * BB_NUM_ACTION_DIMS + b] = (float)a;
                }
            }
            if (valid && io.logpOut) io.logpOut[r] = tot;
        }
        /* ---- bookkeeping of the previous step (ppo.py:84-85,128-134; ppo_stats.py:21-36) ---- */
        if (valid && io.bufRewPrev) {
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
        if (valid && io.resetCol) *reinterpret_cast<int2 *>(io.resetCol + r * 2) = make_int2(0, 0);
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 0) tmem_free(tmem, 128);
}

} // namespace bb
