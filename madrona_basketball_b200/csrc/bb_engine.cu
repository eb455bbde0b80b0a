/*
 * bb_engine.cu — kernels + C ABI (include/bb_engine.h) of the B200 step engine.
 *
 * Replaces Manager / Impl / MWCudaExecutor of the reference
 * (/root/reference/src/mgr.cpp) for the step path: one fused kernel per
 * Manager::step() instead of Madrona's 19 ParallelForNodes + 3 sorts
 * (game.cpp:1467-1523, sim.cpp:100-136).
 *
 * Data layout: one device column per component, in exactly the AoS-per-component
 * layout the reference exports (mgr.cpp:317-445), so exported pointers ARE the
 * simulation state (zero-copy torch views, writable from Python like the
 * reference's).  World slots are fixed (world w owns rows w*k..w*k+k-1), which is
 * why the reference's GPU-only sort-by-world pass has no equivalent here.
 *
 * Build: nvcc -gencode arch=compute_100a,code=sm_100a -fmad=false (bit-exact
 * float parity with the CPU oracle needs unfused multiply-add).
 */
#include <cuda_runtime.h>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "../../include/bb_engine.h"
#include "bb_pair.cuh"

namespace bb {

/* ------------------------------ columns ----------------------------------- */
struct Columns {
    int32_t *reset;        /* [N,2]     */
    float *gameState;      /* [N,14] (int fields stored as bit patterns) */
    int32_t *action;       /* [N,2,6]   */
    int32_t *actionMask;   /* [N,2,4]   */
    float *agentPos;       /* [N,2,3]   */
    float *obs;            /* [N,2,128] */
    float *reward;         /* [N,2]     */
    float *done;           /* [N,2]     */
    int32_t *agentEntity;  /* [N,2,2]  {gen,id} */
    int32_t *possession;   /* [N,2,3]   */
    float *orientation;    /* [N,2,4]   */
    int32_t *team;         /* [N,2,5]   */
    float *stats;          /* [N,2,2]   */
    float *ballPos;        /* [N,3]     */
    int32_t *ballPhysics;  /* [N,7]     */
    int32_t *ballEntity;   /* [N,2]     */
    int32_t *grabbed;      /* [N,2]     */
    float *ballVel;        /* [N,3]     */
    float *hoopPos;        /* [N,2,3]   */
    float *agentVel;       /* [N,2,3]   */
    int32_t *inbounding;   /* [N,2,2]   */
    float *cooldown;       /* [N,2]     */
    float *attributes;     /* [N,2,9]   */
    uint32_t *curStep;     /* [N,2]     */
    int32_t *worldClock;   /* [N]       */
    uint32_t *rngCount;    /* [N]       */
};

/* bytes per world of each column, in canonical (checksum) order */
static const int kColBytes[26] = {8, 56, 48, 32, 24, 1024, 8, 8, 16, 24, 32, 40, 16, 12, 28, 8, 8, 12, 24,
                                  24, 16, 8, 72, 8, 4, 4};

/* Per-step I/O redirection (all optional). */
struct StepIo {
    const int32_t *actionSrc; /* external Action source, [N,2,6] (device, or mapped pinned host) */
    float *rewardOut;         /* mirrors of Reward / Done, [N,2] (mapped pinned host): host-buffer steps */
    float *doneOut;
    float *obsOut;            /* mirror of Observations, [N,2,128] (mapped pinned host) */
    int actionsHost;          /* actionSrc is mapped pinned HOST memory: a warp fetches its tile's rows with 16-byte
                               * uncached loads (few, wide PCIe reads) and hands them out through shared memory */
    int actionsU8;            /* actionSrc holds packed bytes, u8 [N,2,6] (bb_step_host_u8 / bb_serve_step_u8) */
};

struct StepParams {
    Columns col;
    SimConst sc;
    StepIo io;
    int64_t numWorlds;
    int prefetchBlocks;       /* > 0: block b pulls the state rows of block b + prefetchBlocks into L2 */
    int prefetchSelf;         /* ahead of griddepcontrol.wait, pull this warp's own rows into L2: bit 0 state rows, bit 1 action rows */
    int earlyTrigger;         /* 1: release the stream's next programmatic launch as soon as this grid has started (step chains
                               * of one-warp blocks: the next grid's blocks sit resident at their griddepcontrol.wait) */
};

/* ------------------------------ obs image --------------------------------- *
 * Observation rows (game.cpp:1256-1460, layout SURVEY App. A) are 55 % of the step's bytes.  Both 512-byte rows of a
 * world are assembled in shared memory IN OUTPUT ORDER as 16-byte chunks (chunk c = columns 4c..4c+3; columns 104..127
 * are zero and have no chunk), every chunk written with one 16-byte shared store and read back with one 16-byte shared
 * load by the lane that stores it to global memory (lane L owns columns 4L..4L+3 of every row: each store instruction
 * writes 512 contiguous bytes).  Lane (world, agent a) writes what it holds in registers: chunks 0..15 and 25 of its
 * OWN row (game / ball / hoop header and its self block, game.cpp:1257-1322) and chunks 16..24 of its PARTNER's row
 * (the partner's opponent block describes this agent, game.cpp:1378-1422; the block sits 38 columns further, i.e. two
 * words off the self block's chunk phase, which is why it is written a second time rather than re-read misaligned).
 * Same code in both lanes: no divergence, no table lookups.
 *
 * Image geometry (float4 units): world stride 58, row stride 27.  A quarter-warp of a 16-byte store is 4 worlds x 2
 * rows -> float4 indices 58 j + 27 a + c -> bank groups (2 j + 3 a + c) mod 8, all different: conflict-free stores;
 * loads read 26 consecutive float4.  kWPR worlds of the warp are staged per round (16: one round, 14.5 KB per warp,
 * the latency-bound launch shapes; 8: two rounds, 7.3 KB per warp, the 16-warps-per-SM shape). */
constexpr int kWorldsPerWarp = 16;
constexpr int kImgWorld4 = 58;  /* float4 per world slot ( = 2 mod 8) */
constexpr int kImgRow4 = 27;    /* row 1 of a world starts here (odd) */
constexpr int kObsChunks = 26;  /* chunks with content: columns 0..103 */

BBD float4 f4(float a, float b, float c, float d) { return make_float4(a, b, c, d); }

/* Derived floats of one agent's block (game.cpp:1293-1312), each computed once per step. */
struct ObsDerived {
    Vec3 vn; float vdot, accel;      /* velocity direction, dot with facing, acceleration factor */
    Vec3 hoopDir; float hoopDist;    /* towards the hoop this agent attacks */
    Vec3 ballDir; float ballDist;
    Vec3 toMe; float dist;           /* the PARTNER's vec_to_agent / distance (game.cpp:1334,1381-1383): unit vector from the
                                      * partner to this agent, and their distance (the same float in both directions) */
};

BBD ObsDerived obs_derive(const Lane &L, Vec3 opos, const SimConst &c)
{
    const AgentS &A = L.A;
    ObsDerived d;
    const Vec3 zero = v3(0.f, 0.f, 0.f);
    const bool moving = L.vel2 > 1e-6f;
    /* normalize(v) = v * (1 / sqrt(length2(v))) with the sqrt shared with length(v) */
    d.vn = moving ? A.vel * f_div(1.f, L.speed) : zero;
    d.vdot = moving ? dot(d.vn, L.facing) : 0.f;
    d.accel = (d.vdot <= 0.8f) ? 0.1f : 1.f;
    const Vec3 dh = v3(L.me ? c.hoopX[0] : c.hoopX[1], c.hoopY, 0.f) - A.pos;
    d.hoopDist = f_sqrt(len2(dh));
    d.hoopDir = (d.hoopDist > 1e-6f) ? dh * f_div(1.f, d.hoopDist) : zero;
    const Vec3 db = L.B.pos - A.pos;
    d.ballDist = f_sqrt(len2(db));
    d.ballDir = (d.ballDist > 1e-6f) ? db * f_div(1.f, d.ballDist) : zero;
    const Vec3 v = A.pos - opos;
    const float l2 = len2(v);
    d.dist = f_sqrt(l2);
    d.toMe = (l2 > 1e-6f) ? v * f_div(1.f, d.dist) : zero;
    return d;
}

/* System 18 (fillObservationsSystem, game.cpp:1175-1461) for the lane's world: 26 chunk stores.
 * own / other: float4 base of this agent's row / of the partner's row in the image. */
BBD void write_obs_chunks(float4 *own, float4 *other, const Lane &L, const ObsDerived &d, Vec3 opos, int32_t oInb, const SimConst &c)
{
    const AgentS &A = L.A;
    const BallS &B = L.B;
    const GameS &g = L.G;
    const bool me = L.me != 0;
    const Vec3 f = L.facing;
    /* one-hots (game.cpp:1441-1452): ball holder; inbounder = last agent (index order) with imInbounding */
    const int32_t id0 = L.idBase + 3, id1 = L.idBase + 4;
    const int32_t inb0 = me ? oInb : A.imInb, inb1 = me ? A.imInb : oInb;
    int32_t inbounder = -1;
    if ((float)inb0 > 0.5f) inbounder = id0;
    if ((float)inb1 > 0.5f) inbounder = id1;
    const float oh99 = (id0 == B.holder) ? 1.f : 0.f, oh100 = (id1 == B.holder) ? 1.f : 0.f;
    const float oh101 = (id0 == inbounder) ? 1.f : 0.f, oh102 = (id1 == inbounder) ? 1.f : 0.f;
    const float atkX = me ? c.hoopX[0] : c.hoopX[1], defX = me ? c.hoopX[1] : c.hoopX[0];
    const float imInb = (float)A.imInb, pw = (float)A.pointsWorth, has = (float)A.hasBall;
    /* own row: header (columns 0..22) and self block (23..60) */
    own[0] = f4(g.gameClock, g.shotClock, g.period, (float)g.inbInProgress);
    own[1] = f4(g.inboundClock, me ? g.team1Score : g.team0Score, me ? g.team0Score : g.team1Score, B.pos.x);
    own[2] = f4(B.pos.y, B.pos.z, B.vel.x, B.vel.y);
    own[3] = f4(B.vel.z, (float)B.isGrabbed, (float)B.inFlight, (float)B.spv);
    own[4] = f4((float)B.ltTeam, atkX, c.hoopY, 0.f);
    own[5] = f4(defX, c.hoopY, 0.f, A.pos.x);
    own[6] = f4(A.pos.y, A.pos.z, 0.f, 0.f);
    own[7] = f4(0.f, 0.f, A.q.w, A.q.x);
    own[8] = f4(A.q.y, A.q.z, f.x, f.y);
    own[9] = f4(f.z, d.vn.x, d.vn.y, d.vn.z);
    own[10] = f4(L.speed, d.vdot, d.accel, d.hoopDir.x);
    own[11] = f4(d.hoopDir.y, d.hoopDir.z, d.hoopDist, d.ballDir.x);
    own[12] = f4(d.ballDir.y, d.ballDir.z, d.ballDist, imInb);
    own[13] = f4(A.cooldown, A.maxSpeed, A.quickness, A.shooting);
    own[14] = f4(A.ftPct, A.reaction, A.curShot, pw);
    own[15] = f4(has, opos.x, opos.y, opos.z);            /* 60 | opponent block [0..2] = the partner's position */
    own[25] = f4(oh100, oh101, oh102, 0.f);
    /* partner's row: its opponent block (columns 64..98) describes this agent; column 99 */
    other[16] = f4(d.toMe.x, d.toMe.y, d.toMe.z, d.dist);
    other[17] = f4(A.q.w, A.q.x, A.q.y, A.q.z);
    other[18] = f4(f.x, f.y, f.z, d.vn.x);
    other[19] = f4(d.vn.y, d.vn.z, L.speed, d.vdot);
    other[20] = f4(d.accel, d.hoopDir.x, d.hoopDir.y, d.hoopDir.z);
    other[21] = f4(d.hoopDist, d.ballDir.x, d.ballDir.y, d.ballDir.z);
    other[22] = f4(d.ballDist, imInb, A.cooldown, A.maxSpeed);
    other[23] = f4(A.quickness, A.shooting, A.ftPct, A.reaction);
    other[24] = f4(A.curShot, pw, has, oh99);
}

/* Coalesced emission of both observation rows of `count` staged worlds (image slots 0..count-1) to the rows of worlds
 * world0 .. world0+count-1.  kCount > 0: compile-time count, fully unrolled (one LDS.128 + one STG.128 per row). */
template <int kCount>
BBD void emit_obs_rows(const float4 *img, float *obs, int64_t world0, int count, int lane)
{
    float4 *dst = reinterpret_cast<float4 *>(obs + world0 * 256) + lane;
    const bool has = lane < kObsChunks;
    const float4 *src = img + (has ? lane : 0);
    const float4 zero = f4(0.f, 0.f, 0.f, 0.f);
    if (kCount > 0) {
#pragma unroll
        for (int j = 0; j < kCount; j++) {
#pragma unroll
            for (int r = 0; r < 2; r++) {
                float4 v = zero;
                if (has) v = src[j * kImgWorld4 + r * kImgRow4];
                dst[j * 64 + r * 32] = v;
            }
        }
    } else {
        for (int j = 0; j < count; j++) {
            float4 v0 = zero, v1 = zero;
            if (has) { v0 = src[j * kImgWorld4]; v1 = src[j * kImgWorld4 + kImgRow4]; }
            dst[j * 64] = v0;
            dst[j * 64 + 32] = v1;
        }
    }
}

/* ------------------------------ load / store ------------------------------ */

/* Lane (w, me): its agent's rows + the world's ball / GameState / clock / RNG rows
 * (the pair reads the same world addresses: one broadcast transaction). */
/* staged: the action rows of the warp's 16 worlds in shared memory (row = lane), or NULL: read them in place */
BBD void load_lane(Lane &L, const StepParams &p, const StepIo &io, int64_t w, int me, const uint32_t *staged, int lane)
{
    const Columns &C = p.col;
    const int64_t r = w * 2 + me; /* agent row */
    AgentS &A = L.A;
    L.idBase = (int32_t)(w * 5);
    L.me = me;
    L.didReset = false;
    if (io.actionsU8) { /* six bytes per agent, 2-byte aligned */
        const unsigned short *ap = staged ? reinterpret_cast<const unsigned short *>(staged) + lane * 3
                                          : reinterpret_cast<const unsigned short *>(reinterpret_cast<const uint8_t *>(io.actionSrc) + r * 6);
        const unsigned h0 = ap[0], h1 = ap[1], h2 = ap[2];
        A.move = (int32_t)(h0 & 0xffu); A.moveAngle = (int32_t)(h0 >> 8); A.rotate = (int32_t)(h1 & 0xffu);
        A.grab = (int32_t)(h1 >> 8); A.pass = (int32_t)(h2 & 0xffu); A.shoot = (int32_t)(h2 >> 8);
    } else {
        const int2 *ap = staged ? reinterpret_cast<const int2 *>(staged) + lane * 3
                                : reinterpret_cast<const int2 *>((io.actionSrc ? io.actionSrc : C.action) + r * 6);
        const int2 a0 = ap[0], a1 = ap[1], a2 = ap[2];
        A.move = a0.x; A.moveAngle = a0.y; A.rotate = a1.x; A.grab = a1.y; A.pass = a2.x; A.shoot = a2.y;
    }
    A.reset = C.reset[r];
    {
        const float *pp = C.agentPos + r * 3;
        A.pos = v3(pp[0], pp[1], pp[2]);
        const float *vp = C.agentVel + r * 3;
        A.vel = v3(vp[0], vp[1], vp[2]);
        float4 q = *reinterpret_cast<const float4 *>(C.orientation + r * 4);
        A.q.w = q.x; A.q.x = q.y; A.q.y = q.z; A.q.z = q.w;
    }
    {
        const int32_t *ip = C.possession + r * 3;
        A.hasBall = ip[0]; A.ballId = ip[1]; A.pointsWorth = ip[2];
        int2 ib = *reinterpret_cast<const int2 *>(C.inbounding + r * 2);
        A.imInb = ib.x; A.allowedToMove = ib.y;
        A.cooldown = C.cooldown[r];
        float2 st = *reinterpret_cast<const float2 *>(C.stats + r * 2);
        A.points = st.x; A.fouls = st.y;
        A.curStep = C.curStep[r];
    }
    {
        const float *tp = C.attributes + r * 9;
        A.maxSpeed = tp[0]; A.quickness = tp[1]; A.shooting = tp[2]; A.ftPct = tp[3]; A.reaction = tp[4];
        A.target = v3(tp[5], tp[6], tp[7]); A.curShot = tp[8];
    }
    {
        BallS &B = L.B;
        const float *bp = C.ballPos + w * 3;
        B.pos = v3(bp[0], bp[1], bp[2]);
        const float *bv = C.ballVel + w * 3;
        B.vel = v3(bv[0], bv[1], bv[2]);
        const int32_t *ph = C.ballPhysics + w * 7;
        B.inFlight = ph[0]; B.ltAgent = ph[1]; B.ltTeam = ph[2]; B.sbAgent = ph[3];
        B.sbTeam = ph[4]; B.spv = ph[5]; B.goingIn = ph[6];
        int2 gr = *reinterpret_cast<const int2 *>(C.grabbed + w * 2);
        B.isGrabbed = gr.x; B.holder = gr.y;
    }
    {
        const float2 *gp = reinterpret_cast<const float2 *>(C.gameState + w * 14);
        float2 g0 = gp[0], g1 = gp[1], g2 = gp[2], g3 = gp[3], g4 = gp[4], g5 = gp[5], g6 = gp[6];
        GameS &g = L.G;
        g.inbInProgress = __float_as_int(g0.x); g.liveBall = __float_as_int(g0.y);
        g.period = g1.x; g.teamInPoss = g1.y;
        g.team0Hoop = __float_as_int(g2.x); g.team0Score = g2.y;
        g.team1Hoop = __float_as_int(g3.x); g.team1Score = g3.y;
        g.gameClock = g4.x; g.shotClock = g4.y;
        g.scoredBaskets = g5.x; g.oobCount = g5.y;
        g.inboundClock = g6.x; g.isOneOnOne = __float_as_int(g6.y);
    }
    L.resetNow = C.worldClock[w];
    L.rng = C.rngCount[w];
    L.dirty = 0;
    L.cooldownOld = A.cooldown; L.curShotOld = A.curShot; L.pointsWorthOld = A.pointsWorth;
    L.resetNowOld = L.resetNow; L.rngOld = L.rng;
    /* mask / reward / done are fully rewritten by tick and actionMask */
}

BBD void store_lane(const Lane &L, const StepParams &p, const StepIo &io, int64_t w)
{
    const Columns &C = p.col;
    const int me = L.me;
    const int64_t r = w * 2 + me;
    const AgentS &A = L.A;
    {
        int2 *ap = reinterpret_cast<int2 *>(C.action + r * 6);
        ap[0] = make_int2(A.move, A.moveAngle); ap[1] = make_int2(A.rotate, A.grab); ap[2] = make_int2(A.pass, A.shoot);
        *reinterpret_cast<int4 *>(C.actionMask + r * 4) = make_int4(A.canMove, A.canGrab, A.canPass, A.canShoot);
    }
    if (L.didReset) { /* resetWorld clears Reset (gen.cpp:273) and rewrites Team (gen.cpp:258,283) */
        C.reset[r] = A.reset;
        int32_t *tp = C.team + r * 5;
        tp[0] = me;
        tp[1] = __float_as_int(me ? 255.f : 0.f); tp[2] = __float_as_int(me ? 0.f : 100.f); tp[3] = __float_as_int(me ? 100.f : 255.f);
        tp[4] = L.idBase + me;
    }
    const uint32_t dirty = L.dirty;
    auto bits = [](float x) { return __float_as_int(x); };
    {
        float *pp = C.agentPos + r * 3;
        pp[0] = A.pos.x; pp[1] = A.pos.y; pp[2] = A.pos.z;
        float *vp = C.agentVel + r * 3;
        vp[0] = A.vel.x; vp[1] = A.vel.y; vp[2] = A.vel.z;
        if (dirty & kDirtyQ) *reinterpret_cast<float4 *>(C.orientation + r * 4) = make_float4(A.q.w, A.q.x, A.q.y, A.q.z);
    }
    {
        if ((dirty & kDirtyPoss) || A.pointsWorth != L.pointsWorthOld) {
            int32_t *ip = C.possession + r * 3;
            ip[0] = A.hasBall; ip[1] = A.ballId; ip[2] = A.pointsWorth;
        }
        if (dirty & kDirtyInb) *reinterpret_cast<int2 *>(C.inbounding + r * 2) = make_int2(A.imInb, A.allowedToMove);
        if (bits(A.cooldown) != bits(L.cooldownOld)) C.cooldown[r] = A.cooldown;
        if (dirty & kDirtyStats) *reinterpret_cast<float2 *>(C.stats + r * 2) = make_float2(A.points, A.fouls);
        C.curStep[r] = A.curStep;
        C.done[r] = A.done;
        if (io.doneOut) io.doneOut[r] = A.done;
    }
    {
        float *tp = C.attributes + r * 9;
        if (dirty & kDirtyAttr) { tp[0] = A.maxSpeed; tp[1] = A.quickness; tp[2] = A.shooting; tp[3] = A.ftPct; tp[4] = A.reaction; }
        if (dirty & kDirtyTarget) { tp[5] = A.target.x; tp[6] = A.target.y; tp[7] = A.target.z; }
        if (bits(A.curShot) != bits(L.curShotOld)) tp[8] = A.curShot;
    }
    if (me == 0) { /* replicated world state: lane 0 writes the ball, lane 1 the singletons */
        const BallS &B = L.B;
        float *bp = C.ballPos + w * 3;
        bp[0] = B.pos.x; bp[1] = B.pos.y; bp[2] = B.pos.z;
        if (dirty & kDirtyBall) {
            float *bv = C.ballVel + w * 3;
            bv[0] = B.vel.x; bv[1] = B.vel.y; bv[2] = B.vel.z;
            int32_t *ph = C.ballPhysics + w * 7;
            ph[0] = B.inFlight; ph[1] = B.ltAgent; ph[2] = B.ltTeam; ph[3] = B.sbAgent;
            ph[4] = B.sbTeam; ph[5] = B.spv; ph[6] = B.goingIn;
            *reinterpret_cast<int2 *>(C.grabbed + w * 2) = make_int2(B.isGrabbed, B.holder);
        }
    } else {
        float2 *gp = reinterpret_cast<float2 *>(C.gameState + w * 14);
        const GameS &g = L.G;
        gp[4] = make_float2(g.gameClock, g.shotClock);
        if (dirty & kDirtyGame) {
            gp[0] = make_float2(__int_as_float(g.inbInProgress), __int_as_float(g.liveBall));
            gp[1] = make_float2(g.period, g.teamInPoss);
            gp[2] = make_float2(__int_as_float(g.team0Hoop), g.team0Score);
            gp[3] = make_float2(__int_as_float(g.team1Hoop), g.team1Score);
            gp[5] = make_float2(g.scoredBaskets, g.oobCount);
            gp[6] = make_float2(g.inboundClock, __int_as_float(g.isOneOnOne));
        }
        if (L.resetNow != L.resetNowOld) C.worldClock[w] = L.resetNow;
        if (L.rng != L.rngOld) C.rngCount[w] = L.rng;
    }
}

/* Reward is final only after system 19; everything else is stored right after system 17 so that the stores drain
 * while the observation values are computed */
BBD void store_reward(const Lane &L, const StepParams &p, const StepIo &io, int64_t w)
{
    const int64_t r = w * 2 + L.me;
    p.col.reward[r] = L.A.reward;
    if (io.rewardOut) io.rewardOut[r] = L.A.reward;
}

/* thread-per-world store of a freshly generated world (init kernel only) */
BBD void store_world(const World &W, const StepParams &p, int64_t w)
{
    Lane L;
    L.B = W.b; L.G = W.g; L.resetNow = W.resetNow; L.rng = W.rng; L.idBase = W.idBase; L.didReset = false;
    L.dirty = kDirtyAll; L.resetNowOld = ~W.resetNow; L.rngOld = ~W.rng; /* a fresh world: every row is written */
#pragma unroll
    for (int i = 0; i < 2; i++) {
        L.A = W.a[i];
        L.me = i;
        L.pointsWorthOld = ~W.a[i].pointsWorth;
        L.cooldownOld = __int_as_float(~__float_as_int(W.a[i].cooldown)); L.curShotOld = __int_as_float(~__float_as_int(W.a[i].curShot));
        store_lane(L, p, p.io, w);
        store_reward(L, p, p.io, w);
        const Columns &C = p.col;
        const int64_t r = w * 2 + i;
        C.reset[r] = W.a[i].reset;
        int32_t *tp = C.team + r * 5; /* generateWorld colours: constants.hpp:18-19 */
        tp[0] = i;
        tp[1] = __float_as_int(i ? 128.f : 0.f); tp[2] = __float_as_int(i ? 0.f : 100.f); tp[3] = __float_as_int(i ? 128.f : 255.f);
        tp[4] = W.idBase + i;
    }
}

/* L2 prefetch of the rows a later block will load (HBM-bound regime).  Every column is contiguous
 * over worlds, so the rows of `worlds` consecutive worlds are one 16-byte-aligned segment per column:
 * lane c of the calling warp issues one cp.async.bulk.prefetch.L2 for column c.  By the time the
 * block that owns those worlds starts, its up-front loads would hit L2 instead of DRAM.  Tuning knob
 * only (BB_PREFETCH_BLOCKS): on B200 it costs more than it hides, see bb_create. */
BBD void prefetch_state_rows(const StepParams &p, int64_t world0, int worlds, int lane, bool withActions = true)
{
    if (lane == 0 && !withActions) return;
    const Columns &C = p.col;
    const void *base = nullptr;
    int bpw = 0; /* bytes per world */
    switch (lane) {
    case 0: base = p.io.actionSrc ? (const void *)p.io.actionSrc : (const void *)C.action; bpw = 48; break;
    case 1: base = C.reset; bpw = 8; break;
    case 2: base = C.gameState; bpw = 56; break;
    case 3: base = C.agentPos; bpw = 24; break;
    case 4: base = C.agentVel; bpw = 24; break;
    case 5: base = C.orientation; bpw = 32; break;
    case 6: base = C.possession; bpw = 24; break;
    case 7: base = C.inbounding; bpw = 16; break;
    case 8: base = C.cooldown; bpw = 8; break;
    case 9: base = C.stats; bpw = 16; break;
    case 10: base = C.curStep; bpw = 8; break;
    case 11: base = C.attributes; bpw = 72; break;
    case 12: base = C.ballPos; bpw = 12; break;
    case 13: base = C.ballVel; bpw = 12; break;
    case 14: base = C.ballPhysics; bpw = 28; break;
    case 15: base = C.grabbed; bpw = 8; break;
    case 16: base = C.worldClock; bpw = 4; break;
    case 17: base = C.rngCount; bpw = 4; break;
    default: return;
    }
    const char *src = (const char *)base + world0 * bpw;
    const uint32_t bytes = (uint32_t)(worlds * bpw);
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(src), "r"(bytes) : "memory");
}

/* ------------------------------ kernels ----------------------------------- */
constexpr int kMaxThreads = 128;
constexpr int kInitThreads = 128;
/* dynamic shared memory of a launch: one image of kWPR world slots per warp */
__host__ __device__ constexpr int smem_bytes_for(int threads, int wpr) { return (threads / 32) * wpr * kImgWorld4 * 16; }

#ifndef BB_MIN_BLOCKS_128
#define BB_MIN_BLOCKS_128 4 /* resident 128-thread blocks per SM the register budget is cut for */
#endif
/* One Manager::step() for the 16 worlds [warpWorld0, warpWorld0 + 16) on the calling warp:
 * lane = (world (lane >> 1), agent (lane & 1)).  warpImg: this warp's observation image (kWPR world slots). */
template <bool kSync, int kWPR>
BBD void step_tile(const StepParams &p, const StepIo &io, float4 *warpImg, int64_t warpWorld0, int lane)
{
    const int64_t w = warpWorld0 + (lane >> 1);
    const bool valid = w < p.numWorlds;
    const int me = lane & 1;
    Lane L;
    /* host-resident action rows (bb_step_host*, bb_serve_*): the tile's rows are one contiguous, 16-byte aligned segment
     * (768 B of int32 rows, 192 B packed); 48 / 12 lanes fetch 16 bytes each past the caches and the pair reads its six
     * fields from shared memory -- instead of three 2- or 8-byte loads per lane over PCIe */
    const uint32_t *staged = nullptr;
    if (io.actionsHost && warpWorld0 + kWorldsPerWarp <= p.numWorlds) {
        const int bytesPerWorld = io.actionsU8 ? 12 : 48;
        const uint4 *src = reinterpret_cast<const uint4 *>(reinterpret_cast<const uint8_t *>(io.actionSrc) + warpWorld0 * bytesPerWorld);
        uint4 *dst = reinterpret_cast<uint4 *>(warpImg);
        for (int i = lane; i < bytesPerWorld; i += 32) dst[i] = __ldcv(src + i); /* 16 worlds x bytesPerWorld / 16 B */
        __syncwarp();
        staged = reinterpret_cast<const uint32_t *>(warpImg);
    }
    /* tail lanes shadow the last world so that every shuffle stays warp-wide; they never store */
    load_lane(L, p, io, valid ? w : p.numWorlds - 1, me, staged, lane);
    __syncwarp(); /* all of the pair's loads precede any of its stores */
    const bool tag = (p.sc.flags & BB_FLAG_TAG_MASK) != 0;
    const bool fast = __all_sync(kFullMask, tag && L.G.isOneOnOne == 1 && L.G.inbInProgress == 0);
    if (fast) pair_step_fast<kSync>(L, p.sc);                /* systems 1-17 */
    else { Lane T = L; slow_step(&T, &p.sc, tag, kSync); L = T; finish_derived(L); L.dirty = kDirtyAll; } /* only the copy's address escapes */
    __syncwarp();
    const Vec3 opos = xorv(L.A.pos);
    const int32_t oInb = xori(L.A.imInb);
    if (valid) store_lane(L, p, io, w);                      /* state after systems 1-17 (18 and 19 write only obs / reward) */
    const ObsDerived d = obs_derive(L, opos, p.sc);
    pair_reward(L, d.dist);                                  /* system 19 (reads nothing system 18 writes) */
    if (valid) store_reward(L, p, io, w);
    if (warpWorld0 >= p.numWorlds) return;
    const int64_t remain = p.numWorlds - warpWorld0;
    const int nValid = remain >= kWorldsPerWarp ? kWorldsPerWarp : (int)remain;
    /* system 18: stage kWPR worlds' rows in the image, emit them, next round */
#pragma unroll
    for (int rd = 0; rd < kWorldsPerWarp / kWPR; rd++) {
        if (kWPR == kWorldsPerWarp || (lane >> 4) == rd) {
            float4 *slot = warpImg + ((lane >> 1) & (kWPR - 1)) * kImgWorld4;
            write_obs_chunks(slot + me * kImgRow4, slot + (me ^ 1) * kImgRow4, L, d, opos, oInb, p.sc);
        }
        __syncwarp();
        const int left = nValid - rd * kWPR; /* staged worlds of this round that exist */
        if (left >= kWPR) {
            emit_obs_rows<kWPR>(warpImg, p.col.obs, warpWorld0 + rd * kWPR, kWPR, lane);
            if (io.obsOut) emit_obs_rows<kWPR>(warpImg, io.obsOut, warpWorld0 + rd * kWPR, kWPR, lane);
        } else if (left > 0) {
            emit_obs_rows<0>(warpImg, p.col.obs, warpWorld0 + rd * kWPR, left, lane);
            if (io.obsOut) emit_obs_rows<0>(warpImg, io.obsOut, warpWorld0 + rd * kWPR, left, lane);
        }
        if (rd + 1 < kWorldsPerWarp / kWPR) __syncwarp(); /* the next round overwrites the slots */
    }
}

/* One Manager::step(): thread t = (world t/2, agent t%2).  blockDim.x is any multiple of 32 up to kMaxT; dynamic shared
 * memory = smem_bytes_for(blockDim.x, kWPR).  kSync: see phase_sync; kWPR: worlds per observation round (16 or 8). */
/* kMinB: resident blocks per SM the register budget is cut for.  kMinB == 1 with kMaxT == 128 is the "lone warps" shape of the
 * latency-bound regime (at most 4 x SMs warps): four warps per block, ONE block per SM, so that every warp has a scheduler
 * (SMSP) to itself and no register cap -- one-warp blocks are spread by the block scheduler unevenly (ncu of 512 one-warp blocks:
 * some SMSPs hold two warps for the whole kernel while others hold none, SM active time 1.6x apart). */
template <int kMaxT, bool kSync, int kWPR, int kMinB = (kMaxT == 32 ? 8 : BB_MIN_BLOCKS_128 * 128 / kMaxT)>
__global__ void __launch_bounds__(kMaxT, kMinB) bb_step_kernel(const __grid_constant__ StepParams p)
{
    extern __shared__ float4 smem4[];
    constexpr bool kLat = true; /* early release / self prefetch are the host's choice per launch shape (bb_create) */
    if (kLat && p.prefetchSelf) {
        /* A block released early by its predecessor (earlyTrigger) sits here until that grid has drained.  L2 prefetches of
         * the rows it is about to load cost nothing meanwhile and are safe before the wait: the L2 is the point of coherence,
         * so a line the predecessor still writes is simply written in place; nothing is read into registers or L1. */
        const int64_t w0 = ((int64_t)blockIdx.x * blockDim.x + (threadIdx.x & ~31)) >> 1;
        if (w0 + kWorldsPerWarp <= p.numWorlds) prefetch_state_rows(p, w0, kWorldsPerWarp, threadIdx.x & 31, (p.prefetchSelf & 2) != 0);
    }
    /* programmatic dependent launch (bb_rollout): let the next kernel of the chain be scheduled early, and wait here for
     * the previous one to complete; both are no-ops for an ordinary launch */
    asm volatile("griddepcontrol.wait;" ::: "memory");
    if (kLat && p.earlyTrigger) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 *warpImg = smem4 + warp * kWPR * kImgWorld4;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t warpWorld0 = (t - lane) >> 1;
    if (!kSync && warpWorld0 >= p.numWorlds) return; /* whole warp past the end (with barriers it shadows the last world) */
    if (p.prefetchBlocks > 0 && warp == 0) { /* worlds per block = blockDim.x / 2, a multiple of 16 */
        const int wpb = blockDim.x >> 1;
        const int64_t w0 = ((int64_t)blockIdx.x + p.prefetchBlocks) * wpb;
        if (w0 + wpb <= p.numWorlds) prefetch_state_rows(p, w0, wpb, lane);
    }
    step_tile<kSync, kWPR>(p, p.io, warpImg, warpWorld0, lane);
}

/* n x Manager::step() in ONE launch (bb_step_n, bb_step_n_with_actions): worlds never interact and a tile's rows are read and
 * written by its own warp only, so the warp simply takes n steps in a row -- no grid-wide dependency, no launch, no drain and
 * no cold instruction fetch between steps, and the tile's rows stay in this SM's L1 / the L2 from one step to the next (every
 * step still stores its full state, observation rows, reward and done, exactly like n launches: the columns ARE the exported
 * tensors).  Between two steps the warp's stores are ordered before its partner lanes' loads by __syncwarp() (with phase
 * barriers: the block's).  actionStride != 0: step s reads its Action rows from io.actionSrc + s * actionStride (int32
 * elements), a recorded action stream [n,N,2,6]; otherwise every step reads the Action column the previous one left, like
 * scripts/run.py:16-20. */
template <int kMaxT, bool kSync, int kWPR, int kMinB = (kMaxT == 32 ? 8 : BB_MIN_BLOCKS_128 * 128 / kMaxT)>
__global__ void __launch_bounds__(kMaxT, kMinB) bb_step_loop_kernel(const __grid_constant__ StepParams p, int32_t n, int64_t actionStride)
{
    extern __shared__ float4 smem4[];
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float4 *warpImg = smem4 + warp * kWPR * kImgWorld4;
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t warpWorld0 = (t - lane) >> 1;
    if (!kSync && warpWorld0 >= p.numWorlds) return;
    StepIo io = p.io;
    for (int32_t s = 0; s < n; s++) {
        step_tile<kSync, kWPR>(p, io, warpImg, warpWorld0, lane);
        __syncwarp(); /* this step's stores (state rows, and the image the observation rows were emitted from) before the next step's loads */
        if (io.actionSrc) io.actionSrc += actionStride;
    }
}

/* ------------------------------ persistent server -------------------------- *
 * Host-driven stepping without a kernel launch or a stream wait per step (bb_serve_*): the kernel
 * stays resident, block 0 polls a doorbell in mapped pinned host memory, every block then runs
 * step_tile() for its tiles (reading the action rows from, and mirroring reward / done / observations
 * into, host memory over PCIe), and the last block to finish publishes the sequence number back to
 * the host.  All blocks must be co-resident (grid <= occupancy x SMs; tiles are strided over it).
 * The server exits by itself after idleTimeoutNs without a doorbell, so a vanished host cannot wedge
 * the GPU. */
struct ServeHostCtl { /* mapped pinned host memory */
    /* host -> device doorbell.  word[0] is the ONLY word block 0 polls (one 8-byte PCIe read per sample): the low 16 bits of
     * the sequence number to run in its top 16 bits, kServeU8Bit, kServeLongBit and, in its low 6 bits, the slot of the
     * device-side table that holds the step's buffer combination {actions, reward, done, obs} (device aliases).
     * Long form (first use of a combination, or a re-used slot): word[2..5] carry the four aliases; the host writes them
     * BEFORE word[0], block 0 reads them once AFTER it has acquired word[0] and stores them in the table slot.  Short form:
     * the slot is already filled.  word[0] == kServeExitWord: leave. */
    unsigned long long word[8];
    int32_t complete;        /* device -> host: last completed sequence number (its own 64-byte line: the host spins on it) */
    int32_t alive;           /* device -> host: 1 while the server runs */
    unsigned long long tSeen, tDone; /* %globaltimer when block 0 saw the doorbell / the last block finished (tuning) */
    /* tuning marks (%globaltimer, ns) of the last step: [0] doorbell seen by block 0, [1] after its acquire fence,
     * [2] release to the other blocks done, [3] block 0 starts its tile, [4] block 0 done with its tile, [5] last block
     * done with its tile, [6] last block past its counter increment, [7] just before the completion word is written */
    unsigned long long mark[8];
};
constexpr unsigned long long kServeExitWord = ~0ull;
constexpr unsigned long long kServePtrMask = (1ull << 47) - 1;
constexpr unsigned long long kServeU8Bit = 1ull << 47;   /* word[0]: the action rows are packed bytes */
constexpr unsigned long long kServeLongBit = 1ull << 46; /* word[0]: word[2..5] carry the buffer combination of the slot */
constexpr int kServeSlots = 64;
struct ServeDevCtl { /* device memory; the polled word, the counter and the flag on different 128-byte lines: blocks that are
                      * done poll `bell` for the next step while the last blocks still count themselves done */
    unsigned long long bell;  /* block 0 -> other blocks: the doorbell word of the step to run (tag | flags | slot) */
    unsigned long long pad0[15];
    unsigned int doneCount;
    unsigned int pad1[31];
    int32_t stepDone;         /* last sequence number whose completion word has been sent */
    int32_t pad2[31];
    unsigned long long combo[kServeSlots][4]; /* buffer combinations: device aliases of {actions, reward, done, obs} */
};
constexpr int32_t kServeExit = -1;

BBD int32_t ld_acquire_sys(const int32_t *p) { int32_t v; asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
BBD int32_t ld_acquire_gpu(const int32_t *p) { int32_t v; asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v; }
BBD void st_release_gpu(int32_t *p, int32_t v) { asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
BBD void st_release_sys(int32_t *p, int32_t v) { asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
BBD uint64_t global_timer_ns() { uint64_t t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }

constexpr int kServeThreads = 128; /* upper bound; launched with 32 (one tile per block) or 128 (four tiles, one block per SM) */
__global__ void __launch_bounds__(kServeThreads, 1) bb_serve_kernel(const __grid_constant__ StepParams p, ServeHostCtl *hc, ServeDevCtl *dc,
                                                                  int32_t startSeq, uint64_t idleTimeoutNs, int flags)
{
    extern __shared__ float4 smem4[];
    __shared__ int32_t sSeq;
    __shared__ StepIo sIo;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int warpsPerBlock = blockDim.x >> 5;
    float4 *warpImg = smem4 + warp * kWorldsPerWarp * kImgWorld4;
    const int64_t numTiles = (p.numWorlds + kWorldsPerWarp - 1) / kWorldsPerWarp;
    int32_t seq = startSeq;
    uint64_t tSeen = 0;
    const bool marks = (flags & 4) != 0; /* BB_SERVE_FLAGS bit 2: write the tuning marks (a few extra posted writes per step) */
    for (;;) {
        const int32_t want = seq + 1;
        if (threadIdx.x == 0) {
            int32_t got;
            const unsigned long long tag = (unsigned long long)((unsigned)want & 0xffffu);
            unsigned long long w0;
            if (blockIdx.x == 0) {
                const uint64_t t0 = global_timer_ns();
                unsigned polls = 0;
                /* one 8-byte read in flight at a time: several outstanding reads of the same host line serialise on PCIe and
                 * delay everything else on the link (measured: 4 in flight cost 5 us per step) */
                for (;;) {
                    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(w0) : "l"(&hc->word[0]) : "memory");
                    if (w0 == kServeExitWord) { got = kServeExit; break; }
                    if ((w0 >> 48) == tag) { got = want; break; }
                    if ((++polls & 31u) == 0 && global_timer_ns() - t0 > idleTimeoutNs) { got = kServeExit; w0 = kServeExitWord; break; }
                }
                tSeen = global_timer_ns();
                if (marks) { hc->mark[0] = tSeen; hc->tSeen = tSeen; }
                /* acquire: the host's action rows (and word[2..5]) precede its doorbell.  A relaxed load followed by an ACQUIRE
                 * fence is the acquire pattern of the PTX memory model; it compiles to CCTL.IVALL (drop this SM's L1 lines).
                 * The acq_rel fence used before (flag bit 3 brings it back for A/B runs) adds MEMBAR.ALL.SYS -- the release
                 * half, which waits for a PCIe round trip although this thread has written nothing the host is waiting for:
                 * 17.8 -> 16.4 us per served step with packed rows, 23.4 -> 22.0 with int32 rows
                 * (profiles/r02j_ab_serve_fences.txt). */
                if (flags & 8) asm volatile("fence.acq_rel.sys;" ::: "memory");
                else asm volatile("fence.acquire.sys;" ::: "memory");
                if (marks) hc->mark[1] = global_timer_ns();
                if (got != kServeExit && (w0 & kServeLongBit)) { /* a new buffer combination: into its table slot */
                    unsigned long long p0, p1, p2, p3;
                    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(p0), "=l"(p1) : "l"(&hc->word[2]) : "memory");
                    asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(p2), "=l"(p3) : "l"(&hc->word[4]) : "memory");
                    unsigned long long *c = dc->combo[w0 & (kServeSlots - 1)];
                    c[0] = p0 & kServePtrMask; c[1] = p1 & kServePtrMask; c[2] = p2 & kServePtrMask; c[3] = p3 & kServePtrMask;
                }
                asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(&dc->bell), "l"(w0) : "memory");
                if (marks) hc->mark[2] = global_timer_ns();
            } else {
                for (;;) {
                    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(w0) : "l"(&dc->bell) : "memory");
                    if (w0 == kServeExitWord) { got = kServeExit; break; }
                    if ((w0 >> 48) == tag) { got = want; break; }
                    __nanosleep(20);
                }
            }
            sSeq = got;
            if (got != kServeExit) {
                const volatile unsigned long long *c = dc->combo[w0 & (kServeSlots - 1)];
                const unsigned long long a0 = c[0];
                sIo.actionSrc = (const int32_t *)a0; sIo.rewardOut = (float *)c[1]; sIo.doneOut = (float *)c[2]; sIo.obsOut = (float *)c[3];
                sIo.actionsU8 = (w0 & kServeU8Bit) ? 1 : 0;
                sIo.actionsHost = (a0 != 0 && (a0 & 15ull) == 0) ? 1 : 0;
            }
        }
        __syncthreads();
        if (sSeq == kServeExit) break;
        const StepIo io = sIo;
        if (marks && blockIdx.x == 0 && threadIdx.x == 0) hc->mark[3] = global_timer_ns();
        for (int64_t tile = (int64_t)blockIdx.x * warpsPerBlock + warp; tile < numTiles; tile += (int64_t)gridDim.x * warpsPerBlock)
            step_tile<false, kWorldsPerWarp>(p, io, warpImg, tile * kWorldsPerWarp, lane);
        __syncthreads();
        if (threadIdx.x == 0) {
            const uint64_t tTile = marks ? global_timer_ns() : 0;
            if (marks && blockIdx.x == 0) hc->mark[4] = tTile;
            /* release (gpu scope): this block's writes before the count; the last block's system-scope release below is
             * cumulative over everything it has observed.  (A system-scope fence in every block, so that the last block's
             * release finds the host writes already performed, was measured: 16.5 -> 22.8 us per step.) */
            __threadfence();
            const unsigned old = atomicAdd(&dc->doneCount, 1u);
            if (old == gridDim.x - 1) { /* last block of this step */
                dc->doneCount = 0;      /* nobody touches it again before the host rings the next doorbell */
                if (marks) { hc->mark[5] = tTile; hc->mark[6] = global_timer_ns(); hc->tDone = hc->mark[6]; hc->mark[7] = hc->mark[6]; }
                st_release_sys(&hc->complete, want); /* release at system scope: one fence, inside the store */
                st_release_gpu(&dc->stepDone, want);
            }
            /* block 0 goes back to polling the host only once the step is complete: doorbell reads that share the link
             * with the last blocks' reward / done writes and the completion word delay them */
            if (blockIdx.x == 0 && !(flags & 2))
                while (ld_acquire_gpu(&dc->stepDone) != want) __nanosleep(100);
        }
        seq = want;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) st_release_sys(&hc->alive, 0);
}

/* Sim::Sim -> generateWorld (sim.cpp:86-96) for every world + the constant columns. */
__global__ void __launch_bounds__(kInitThreads) bb_init_kernel(const __grid_constant__ StepParams p)
{
    const int64_t w = (int64_t)blockIdx.x * kInitThreads + threadIdx.x;
    if (w >= p.numWorlds) return;
    const Columns &C = p.col;
    World W;
    W.idBase = (int32_t)(w * 5);
    W.rng = 0;
    W.didReset = false;
    generate_world(W, p.sc);
    store_world(W, p, w);
    /* entity columns {gen = 0, id} and hoops */
    C.agentEntity[w * 4 + 0] = 0; C.agentEntity[w * 4 + 1] = agent_id(W, 0);
    C.agentEntity[w * 4 + 2] = 0; C.agentEntity[w * 4 + 3] = agent_id(W, 1);
    C.ballEntity[w * 2 + 0] = 0;  C.ballEntity[w * 2 + 1] = ball_id(W);
#pragma unroll
    for (int h = 0; h < 2; h++) {
        Vec3 hp = hoop_pos(p.sc, h);
        C.hoopPos[w * 6 + h * 3 + 0] = hp.x; C.hoopPos[w * 6 + h * 3 + 1] = hp.y; C.hoopPos[w * 6 + h * 3 + 2] = hp.z;
    }
}

/* SURVEY 8(d) synthetic action stream; ranges scripts/env.py:96-102 */
__global__ void bb_fill_actions_kernel(int32_t *actions, int64_t numWorlds, uint32_t seed, uint32_t step, int64_t worldOffset)
{
    int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; /* (world, agent) */
    if (idx >= numWorlds * 2) return;
    uint32_t world = (uint32_t)(idx / 2 + worldOffset), agent = (uint32_t)(idx & 1);
    uint32_t a, b;
    threefry2x32(seed, agent, world, step, a, b);
    int2 *dst = reinterpret_cast<int2 *>(actions + idx * 6);
    dst[0] = make_int2((int32_t)(a & 1u), (int32_t)((a >> 1) & 7u));
    dst[1] = make_int2((int32_t)((a >> 8) % 3u), (int32_t)(b & 1u));
    dst[2] = make_int2((int32_t)((b >> 1) & 1u), (int32_t)((b >> 2) & 1u));
}

/* test hook: the spec's libm call sites evaluated on the device (bb_math_eval) */
__global__ void bb_math_eval_kernel(int fn, const float *x, const float *y, float *out, int64_t n)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    float r = 0.f;
    switch (fn) {
    case 0: r = bb_sinf(x[i]); break;
    case 1: r = bb_cosf(x[i]); break;
    case 2: r = bb_atanf(x[i]); break;
    case 3: r = bb_atan2f(x[i], y[i]); break;
    case 4: r = bb_acosf(x[i]); break;
    case 5: r = bb_expf(x[i]); break;
    case 6: r = bb_erff(x[i]); break;
    }
    out[i] = r;
}

struct ChecksumParams {
    const uint32_t *col[26];
    int words[26];
};

__global__ void bb_checksum_kernel(const __grid_constant__ ChecksumParams cp, int64_t numWorlds, uint64_t *out)
{
    int64_t w = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (w >= numWorlds) return;
    uint64_t h = 0xcbf29ce484222325ull;
    for (int c = 0; c < 26; c++) {
        const uint32_t *ptr = cp.col[c] + w * cp.words[c];
        for (int i = 0; i < cp.words[c]; i++) h = (h ^ ptr[i]) * 0x100000001b3ull;
    }
    out[w] = h;
}

} // namespace bb

#include "bb_rollout.cuh"
#include "bb_ppo.cuh"
#include "bb_umma.cuh"

/* ================================ host side ================================ */

static thread_local std::string g_last_error;

static void set_error(const std::string &msg) { g_last_error = msg; }

#define BB_CUDA_OK(expr)                                                                           \
    do {                                                                                           \
        cudaError_t err__ = (expr);                                                                \
        if (err__ != cudaSuccess) {                                                                \
            set_error(std::string(#expr) + ": " + cudaGetErrorString(err__));                      \
            return -1;                                                                             \
        }                                                                                          \
    } while (0)

struct bb_engine {
    bb_config cfg;
    int64_t N;        /* worlds */
    int64_t Npad;     /* allocation, multiple of 32 worlds */
    void *slab;       /* one allocation for every column */
    size_t slabBytes;
    void *colPtr[26];
    bb::StepParams params;
    cudaStream_t hostStream; /* bb_step_host / setters */
    int blockThreads, phaseSync, prefetchBlocks, obsRound, pdlDefault, pdlEarly, prefetchSelf, stepLoop;
    int pdlSync;      /* programmatic launches (+ early release, self prefetch) for the phase-barrier shape too (BB_PDL_SYNC) */
    int lone;         /* the "lone warps" launch shape (bb_step_kernel): 128-thread blocks, one per SM, no register cap */
    int serveThreads; /* block size of the persistent server: 32 or 128 */
    int64_t launches;
    /* bb_step_n: graph of stepGraphN step kernels */
    cudaGraphExec_t stepGraphExec;
    int32_t stepGraphN;
    cudaStream_t captureStream;
    /* persistent server (bb_serve_*) */
    bb::ServeHostCtl *serveHost; /* mapped pinned */
    bb::ServeDevCtl *serveDev;
    cudaStream_t serveStream;
    int32_t serveSeq;
    bool serving;
    /* device aliases of the pinned buffers seen during this serve session (cudaPointerGetAttributes costs about a
     * microsecond per pointer, four pointers per 18 us step).  Contract (include/bb_engine.h): a buffer passed to
     * bb_serve_step stays pinned until bb_serve_end; the cache is dropped by bb_serve_begin. */
    const void *aliasHost[64];
    void *aliasDev[64];
    int aliasCount;
    /* buffer combinations {actions, reward, done, obs} the server's device table holds (host pointers; slot = index) */
    const void *comboHost[bb::kServeSlots][4];
    int comboCount, comboNext;
};

namespace {

/* Host-side evaluation of the finite-domain trig of SURVEY App. E with the same
 * spec the device uses (include/bb_math.h is bit-identical on both sides). */
bb::Quat host_angle_axis_z(float angle)
{
    float c = bb_cosf(angle / 2.f), s = bb_sinf(angle / 2.f);
    bb::Quat q;
    q.w = c; q.x = 0.f * s; q.y = 0.f * s; q.z = 1.f * s;
    return q;
}

void fill_sim_const(bb::SimConst &sc, const bb_config &cfg)
{
    /* bindings.cpp:28-51 */
    const int32_t cells_per_meter = 1;
    sc.startX = cfg.start_x;
    sc.startY = cfg.start_y;
    sc.width = (float)cfg.discrete_x / (float)cells_per_meter;
    sc.height = (float)cfg.discrete_y / (float)cells_per_meter;
    /* gen.cpp:96-141 */
    volatile float court_start_x = (sc.width - bb::kCourtLen) / 2.0f;
    volatile float hx0 = court_start_x + bb::kHoopFromBaseline;
    volatile float t = court_start_x + bb::kCourtLen;
    volatile float hx1 = t - bb::kHoopFromBaseline;
    sc.hoopX[0] = hx0;
    sc.hoopX[1] = hx1;
    sc.hoopY = sc.height / 2.0f;
    volatile float deg = bb::kPi / 180.f;
    sc.turnPos = host_angle_axis_z(deg * 6);
    sc.turnNeg = host_angle_axis_z(deg * -6);
    sc.spawnQ[0] = host_angle_axis_z(-bb::kPi / 2.0f);
    sc.spawnQ[1] = host_angle_axis_z(bb::kPi / 2.0f);
    sc.piZ = host_angle_axis_z(bb::kPi);
    for (int k = 0; k < 8; k++) {
        volatile float ang = (float)k * bb::kAngleBetweenDirs;
        sc.moveSin[k] = bb_sinf(ang);
        sc.moveCos[k] = bb_cosf(ang);
    }
    volatile float two = 1.f * 1.f + 1.f * 1.f;
    sc.invSqrt2 = 1.f / __builtin_sqrtf(two);
    sc.flags = cfg.flags;
}

/* host Threefry for the world key (mgr.cpp:155 initKey, sim.cpp:89 split_i(key,0,0)) */
void host_threefry(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t out[2])
{
    static const int R[8] = {13, 15, 26, 6, 17, 29, 16, 24};
    uint32_t ks[3] = {k0, k1, 0x1BD11BDAu ^ k0 ^ k1};
    uint32_t x0 = c0 + ks[0], x1 = c1 + ks[1];
    for (int r = 0; r < 20; r++) {
        x0 += x1;
        x1 = (x1 << R[r & 7]) | (x1 >> (32 - R[r & 7]));
        x1 ^= x0;
        if ((r & 3) == 3) {
            int s = r / 4 + 1;
            x0 += ks[s % 3];
            x1 += ks[(s + 1) % 3] + (uint32_t)s;
        }
    }
    out[0] = x0;
    out[1] = x1;
}

/* Launch shape.  Small world counts are latency-bound: spread the warps over all 148 SMs x 4
 * schedulers (1 warp per block) until every SM holds a few.  Large counts are instruction-fetch /
 * HBM bound: big blocks whose warps re-align at phase barriers.  BB_BLOCK_THREADS (32..512) and
 * BB_PHASE_SYNC (0/1) override for tuning. */
/* kLen2Le* / kLen2Ge* (bb_world.cuh) against this host's correctly rounded sqrtf: T is the last float whose root is
 * still <= c (resp. the first whose root is >= c). */
bool check_sqrt_thresholds()
{
    auto le_ok = [](float T, float c) { return sqrtf(T) <= c && !(sqrtf(nextafterf(T, INFINITY)) <= c); };
    auto ge_ok = [](float T, float c) { return sqrtf(T) >= c && !(sqrtf(nextafterf(T, -INFINITY)) >= c); };
    return le_ok(bb::kLen2LeHoopZone, bb::kHoopZone) && le_ok(bb::kLen2LeAgentSize, bb::kAgentSize) &&
           ge_ok(bb::kLen2GeArc, bb::kArcRadius);
}

void pick_launch_shape(int64_t n_worlds, int *threads, int *sync, int *obs_round)
{
    const int64_t warps = (n_worlds * 2 + 31) / 32;
    /* obs_round: worlds staged per observation round (image size per warp: 14.5 KB at 16, 7.3 KB at 8) */
    if (warps <= 148 * 8) { *threads = 32; *sync = 0; *obs_round = 16; }
    else if (warps <= 148 * 32) { *threads = 64; *sync = 0; *obs_round = 16; }
    else { *threads = 128; *sync = 1; *obs_round = 8; }
    if (const char *s = getenv("BB_BLOCK_THREADS")) {
        int v = atoi(s);
        if (v >= 32 && v <= 512 && v % 32 == 0) *threads = v;
    }
    if (const char *s = getenv("BB_PHASE_SYNC")) *sync = atoi(s) != 0;
    if (const char *s = getenv("BB_OBS_ROUND")) { int v = atoi(s); if (v == 8 || v == 16) *obs_round = v; }
    if (*threads == 32) *sync = 0;
    if (*threads > 128) *obs_round = 8;
}

/* launch with programmaticStreamSerialization: the kernel may start while its predecessor in the stream is still
 * running; it must griddepcontrol.wait before touching anything the predecessor writes (our step / policy kernels do) */
template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kernel)(KArgs...), int blocks, int threads, size_t smem, cudaStream_t stream, bool pdl, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)blocks); cfg.blockDim = dim3((unsigned)threads);
    cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

/* loop_n > 0: ONE launch of bb_step_loop_kernel taking loop_n steps (action rows of step s at actions_dev + s * action_stride);
 * returns 1 without launching when the engine's launch shape has no looping instantiation (tuning overrides) */
int launch_step(bb_engine *e, const int32_t *actions_dev, cudaStream_t stream, float *reward_out = nullptr,
                float *done_out = nullptr, bool host_mapped = false, bool pdl = false, bool actions_u8 = false, bool early = false,
                int32_t loop_n = 0, int64_t action_stride = 0)
{
    bb::StepParams p = e->params;
    p.io.actionSrc = actions_dev;
    p.io.actionsU8 = actions_u8 ? 1 : 0;
    p.io.rewardOut = reward_out;
    p.io.doneOut = done_out;
    p.io.obsOut = nullptr;
    p.io.actionsHost = (host_mapped && actions_dev && ((uintptr_t)actions_dev & 15u) == 0) ? 1 : 0;
    p.prefetchBlocks = host_mapped ? 0 : e->prefetchBlocks; /* no L2 prefetch of host-mapped rows */
    const int threads = e->blockThreads;
    const int blocks = (int)((e->N * 2 + threads - 1) / threads);
    const int wpr = e->obsRound;
    const int smem = bb::smem_bytes_for(threads, wpr);
    const bool lat = threads <= 128 && (!e->phaseSync || e->pdlSync); /* shapes whose launches are programmatic (bb_create decides) */
    pdl = pdl && lat;
    p.earlyTrigger = (pdl && early && lat) ? 1 : 0;
    p.prefetchSelf = 0;
    if (pdl && lat && e->prefetchSelf && !host_mapped)
        p.prefetchSelf = 1 | ((!actions_dev || ((uintptr_t)actions_dev & 15u) == 0) ? 2 : 0);
#define BB_LAUNCH_STEP(MAXT, SYNC, WPR)                                                                          \
    do {                                                                                                         \
        if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_step_kernel<MAXT, SYNC, WPR>, blocks, threads, (size_t)smem, stream, true, p)); \
        else bb::bb_step_kernel<MAXT, SYNC, WPR><<<blocks, threads, smem, stream>>>(p);                          \
    } while (0)
#ifndef BB_NO_T32
#define BB_NO_T32 0
#endif
    if (loop_n > 0) {
        p.earlyTrigger = 0; p.prefetchSelf = 0; p.prefetchBlocks = 0;
#define BB_LAUNCH_LOOP(MAXT, SYNC, WPR)                                                                          \
    do {                                                                                                         \
        if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_step_loop_kernel<MAXT, SYNC, WPR>, blocks, threads, (size_t)smem, stream, true, p, loop_n, action_stride)); \
        else bb::bb_step_loop_kernel<MAXT, SYNC, WPR><<<blocks, threads, smem, stream>>>(p, loop_n, action_stride); \
    } while (0)
        if (e->lone) {
            if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_step_loop_kernel<128, false, 16, 1>, blocks, threads, (size_t)smem, stream, true, p, loop_n, action_stride));
            else bb::bb_step_loop_kernel<128, false, 16, 1><<<blocks, threads, smem, stream>>>(p, loop_n, action_stride);
        } else if (!BB_NO_T32 && threads == 32 && wpr == 16) BB_LAUNCH_LOOP(32, false, 16);
        else if (threads <= 128 && !e->phaseSync && wpr == 16) BB_LAUNCH_LOOP(128, false, 16);
        else if (threads <= 128 && e->phaseSync && wpr == 8) BB_LAUNCH_LOOP(128, true, 8);
        else return 1;
#undef BB_LAUNCH_LOOP
        e->launches += loop_n; /* passes of the step code over every world */
        BB_CUDA_OK(cudaGetLastError());
        return 0;
    }
    if (e->lone) { /* four lone warps per block, one block per SM */
        if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_step_kernel<128, false, 16, 1>, blocks, threads, (size_t)smem, stream, true, p));
        else bb::bb_step_kernel<128, false, 16, 1><<<blocks, threads, smem, stream>>>(p);
    } else if (!BB_NO_T32 && threads == 32 && wpr == 16) { /* the latency-bound shape: one-warp blocks, no register cap worth having */
        BB_LAUNCH_STEP(32, false, 16);
    } else if (threads <= 128) {
        if (e->phaseSync) { if (wpr == 8) BB_LAUNCH_STEP(128, true, 8); else BB_LAUNCH_STEP(128, true, 16); }
        else              { if (wpr == 8) BB_LAUNCH_STEP(128, false, 8); else BB_LAUNCH_STEP(128, false, 16); }
    } else {
        if (e->phaseSync) BB_LAUNCH_STEP(512, true, 8); else BB_LAUNCH_STEP(512, false, 8);
    }
#undef BB_LAUNCH_STEP
    e->launches++;
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

} // namespace

extern "C" {

const char *bb_last_error(void) { return g_last_error.c_str(); }

bb_engine *bb_create(const bb_config *cfg)
{
    g_last_error.clear();
    if (!cfg || cfg->num_worlds <= 0) { set_error("bb_create: num_worlds must be positive"); return nullptr; }
    int ndev = 0;
    cudaError_t err = cudaGetDeviceCount(&ndev);
    if (err != cudaSuccess || ndev == 0) {
        set_error(std::string("bb_create: no CUDA device (") + cudaGetErrorString(err) +
                  "); this engine has no CPU path");
        return nullptr;
    }
    int dev = cfg->gpu_id < 0 ? 0 : cfg->gpu_id;
    if (dev >= ndev) { set_error("bb_create: gpu_id out of range"); return nullptr; }
    if (cudaSetDevice(dev) != cudaSuccess) { set_error("bb_create: cudaSetDevice failed"); return nullptr; }
    cudaDeviceProp prop;
    cudaGetDeviceProperties(&prop, dev);
    if (prop.major != 10) {
        set_error("bb_create: device is sm_" + std::to_string(prop.major * 10 + prop.minor) +
                  "; libbball.so is built for sm_100a only");
        return nullptr;
    }

    bb_engine *e = new bb_engine();
    e->cfg = *cfg;
    e->cfg.gpu_id = dev;
    e->N = cfg->num_worlds;
    e->Npad = (e->N + 31) / 32 * 32;
    e->launches = 0;
    e->stepGraphExec = nullptr;
    e->stepGraphN = 0;
    e->captureStream = nullptr;
    e->serveHost = nullptr;
    e->serveDev = nullptr;
    e->serveStream = nullptr;
    e->serveSeq = 0;
    e->serving = false;
    e->aliasCount = 0;

    /* one slab, every column 256-byte aligned */
    size_t offs[26], total = 0;
    for (int c = 0; c < 26; c++) {
        offs[c] = total;
        size_t bytes = (size_t)bb::kColBytes[c] * (size_t)e->Npad;
        total += (bytes + 255) / 256 * 256;
    }
    e->slabBytes = total;
    if (cudaMalloc(&e->slab, total) != cudaSuccess) {
        set_error("bb_create: cudaMalloc of " + std::to_string(total) + " bytes failed");
        delete e;
        return nullptr;
    }
    cudaMemset(e->slab, 0, total);
    for (int c = 0; c < 26; c++) e->colPtr[c] = (char *)e->slab + offs[c];

    bb::Columns &C = e->params.col;
    C.reset = (int32_t *)e->colPtr[0];        C.gameState = (float *)e->colPtr[1];
    C.action = (int32_t *)e->colPtr[2];       C.actionMask = (int32_t *)e->colPtr[3];
    C.agentPos = (float *)e->colPtr[4];       C.obs = (float *)e->colPtr[5];
    C.reward = (float *)e->colPtr[6];         C.done = (float *)e->colPtr[7];
    C.agentEntity = (int32_t *)e->colPtr[8];  C.possession = (int32_t *)e->colPtr[9];
    C.orientation = (float *)e->colPtr[10];   C.team = (int32_t *)e->colPtr[11];
    C.stats = (float *)e->colPtr[12];         C.ballPos = (float *)e->colPtr[13];
    C.ballPhysics = (int32_t *)e->colPtr[14]; C.ballEntity = (int32_t *)e->colPtr[15];
    C.grabbed = (int32_t *)e->colPtr[16];     C.ballVel = (float *)e->colPtr[17];
    C.hoopPos = (float *)e->colPtr[18];       C.agentVel = (float *)e->colPtr[19];
    C.inbounding = (int32_t *)e->colPtr[20];  C.cooldown = (float *)e->colPtr[21];
    C.attributes = (float *)e->colPtr[22];    C.curStep = (uint32_t *)e->colPtr[23];
    C.worldClock = (int32_t *)e->colPtr[24];  C.rngCount = (uint32_t *)e->colPtr[25];

    fill_sim_const(e->params.sc, e->cfg);
    uint32_t key[2];
    host_threefry(cfg->rand_seed, 0u, 0u, 0u, key);
    e->params.sc.key0 = key[0];
    e->params.sc.key1 = key[1];
    e->params.io = bb::StepIo{nullptr, nullptr, nullptr, nullptr, 0};
    e->params.numWorlds = e->N;
    e->params.prefetchBlocks = 0;

    pick_launch_shape(e->N, &e->blockThreads, &e->phaseSync, &e->obsRound);
    {   /* lone warps (see bb_step_kernel): when the warps fit one per scheduler, four per block and one block per SM */
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->cfg.gpu_id);
        const int64_t warps = (e->N * 2 + 31) / 32;
        e->lone = 0;
        if (const char *s = getenv("BB_LONE")) e->lone = atoi(s) != 0 && warps <= 4 * (int64_t)sms;
        if (e->lone) { e->blockThreads = 128; e->phaseSync = 0; e->obsRound = 16; }
        e->serveThreads = 32;
        if (const char *s = getenv("BB_SERVE_THREADS")) { const int v = atoi(s); if ((v == 32 || v == 128) && (v == 32 || warps <= 4 * (int64_t)sms)) e->serveThreads = v; }
    }
    e->pdlDefault = 1;
    if (const char *s = getenv("BB_PDL")) e->pdlDefault = atoi(s) != 0;
    {   /* Early release of the next programmatic launch (every block executes griddepcontrol.launch_dependents once it is past
         * its own griddepcontrol.wait, so the next grid is scheduled as soon as this grid's LAST block has started) together
         * with the L2 self-prefetch (a block that sits at its griddepcontrol.wait has already asked for the rows it is about
         * to load): the next step's blocks fill the slots this grid's tail frees, rows on their way.  On by default for every
         * launch shape without phase barriers.  Same-box A/B, cold rows, us per step (profiles r02i, r02r): 2048 worlds
         * 9.4 -> 6.6, 8192 10.0 -> 8.9, 16 384 10.95 -> 10.1, 32 768 16.4 -> 14.4, 65 536 30.0 -> 28.8; either half alone
         * does nothing (early release) or harm (the prefetch on the critical path: 16 384 worlds 12.4).
         * BB_PDL_EARLY / BB_PREFETCH_SELF = 0/1 override; BB_PDL_SYNC=1 extends all three to the phase-barrier shape. */
        e->pdlSync = 0;
        if (const char *s = getenv("BB_PDL_SYNC")) e->pdlSync = atoi(s) != 0;
        const int on = (e->blockThreads <= 128 && (!e->phaseSync || e->pdlSync)) ? 1 : 0;
        e->pdlEarly = on;
        if (const char *s = getenv("BB_PDL_EARLY")) e->pdlEarly = atoi(s) != 0;
        e->prefetchSelf = on;
        if (const char *s = getenv("BB_PREFETCH_SELF")) e->prefetchSelf = atoi(s) != 0;
    }
    if (!check_sqrt_thresholds()) {
        set_error("bb_create: the squared-length thresholds of bb_world.cuh do not match this host's sqrtf");
        cudaFree(e->slab);
        delete e;
        return nullptr;
    }
    /* L2 prefetch of a later block's rows: measured SLOWER on B200 (2^20 worlds: 340 us -> 380-390 us at
     * distances of 296..2368 blocks, gpurun exp1), so it is off unless BB_PREFETCH_BLOCKS asks for it. */
    e->prefetchBlocks = 0;
    if (const char *s = getenv("BB_PREFETCH_BLOCKS")) e->prefetchBlocks = atoi(s);
    { /* the 16-world image of a 128-thread block is 58 KB: above the 48 KB default */
        const int big = bb::smem_bytes_for(512, 8);
        cudaFuncSetAttribute(bb::bb_step_kernel<128, true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<128, true, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<128, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<128, false, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<512, true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<512, false, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_kernel<128, false, 16, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_loop_kernel<128, false, 16, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_serve_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_loop_kernel<128, true, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
        cudaFuncSetAttribute(bb::bb_step_loop_kernel<128, false, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, big);
    }
    e->stepLoop = 1; /* bb_step_n = one launch of the looping kernel */
    if (const char *s = getenv("BB_STEP_N_GRAPH")) e->stepLoop = atoi(s) != 0 ? 0 : 1;
    if (cudaStreamCreateWithFlags(&e->hostStream, cudaStreamNonBlocking) != cudaSuccess) {
        set_error("bb_create: cudaStreamCreate failed");
        cudaFree(e->slab);
        delete e;
        return nullptr;
    }

    const int blocks = (int)((e->N + bb::kInitThreads - 1) / bb::kInitThreads);
    bb::bb_init_kernel<<<blocks, bb::kInitThreads>>>(e->params);
    e->launches++;
    err = cudaDeviceSynchronize();
    if (err != cudaSuccess) {
        set_error(std::string("bb_create: init kernel failed: ") + cudaGetErrorString(err));
        cudaStreamDestroy(e->hostStream);
        cudaFree(e->slab);
        delete e;
        return nullptr;
    }
    return e;
}

void bb_destroy(bb_engine *e)
{
    if (!e) return;
    cudaSetDevice(e->cfg.gpu_id);
    bb_serve_end(e);
    if (e->serveStream) cudaStreamDestroy(e->serveStream);
    if (e->stepGraphExec) cudaGraphExecDestroy(e->stepGraphExec);
    if (e->captureStream) cudaStreamDestroy(e->captureStream);
    if (e->serveHost) cudaFreeHost(e->serveHost);
    if (e->serveDev) cudaFree(e->serveDev);
    cudaDeviceSynchronize();
    cudaStreamDestroy(e->hostStream);
    cudaFree(e->slab);
    delete e;
}

/* Ordinary (uncaptured) step launches carry the programmatic-stream-serialization attribute: the kernel's blocks may be
 * scheduled while the previous kernel of the stream drains, and wait at their first instruction (griddepcontrol.wait)
 * until that kernel has completed and flushed -- whatever kernel it is -- so only launch latency overlaps, never data.
 * Inside a caller's stream capture the edge needs a kernel node to hang on: taken only when the capture's single pending
 * dependency is a kernel node (a captured first node, or one behind a memcpy / event node, gets a plain edge).  BB_PDL=0 disables. */
static bool pdl_for(bb_engine *e, cudaStream_t s)
{
    if (!e->pdlDefault) return false;
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    const cudaGraphNode_t *deps = nullptr;
    size_t numDeps = 0;
    if (cudaStreamGetCaptureInfo(s, &st, nullptr, nullptr, &deps, &numDeps) != cudaSuccess) { cudaGetLastError(); return false; }
    if (st == cudaStreamCaptureStatusNone) return true;
    if (st != cudaStreamCaptureStatusActive || numDeps != 1 || !deps) return false;
    cudaGraphNodeType ty;
    if (cudaGraphNodeGetType(deps[0], &ty) != cudaSuccess) { cudaGetLastError(); return false; }
    return ty == cudaGraphNodeTypeKernel;
}

int bb_step(bb_engine *e, void *cuda_stream)
{
    if (e->serving) { set_error("bb_step: the persistent server is running (bb_serve_end first)"); return -1; }
    return launch_step(e, nullptr, (cudaStream_t)cuda_stream, nullptr, nullptr, false, pdl_for(e, (cudaStream_t)cuda_stream), false, e->pdlEarly != 0);
}

int bb_step_n(bb_engine *e, int32_t n, void *cuda_stream)
{
    if (n <= 0) return 0;
    if (e->serving) { set_error("bb_step_n: the persistent server is running (bb_serve_end first)"); return -1; }
    if (e->stepLoop) { /* one launch of the looping kernel; BB_STEP_N_GRAPH=1 keeps the graph of n launches below */
        const int rc = launch_step(e, nullptr, (cudaStream_t)cuda_stream, nullptr, nullptr, false, pdl_for(e, (cudaStream_t)cuda_stream), false, false, n, 0);
        if (rc <= 0) return rc;
    }
    if (!e->stepGraphExec || e->stepGraphN != n) {
        if (e->stepGraphExec) { cudaGraphExecDestroy(e->stepGraphExec); e->stepGraphExec = nullptr; }
        if (!e->captureStream) BB_CUDA_OK(cudaStreamCreateWithFlags(&e->captureStream, cudaStreamNonBlocking));
        const int64_t launches0 = e->launches;
        cudaGraph_t graph = nullptr;
        for (int attempt = 0; attempt < 2 && !e->stepGraphExec; attempt++) {
            const bool pdl = attempt == 0; /* second attempt: plain edges, should a driver refuse programmatic ones in capture */
            BB_CUDA_OK(cudaStreamBeginCapture(e->captureStream, cudaStreamCaptureModeThreadLocal));
            int rc = 0;
            for (int32_t i = 0; i < n && rc == 0; i++) rc = launch_step(e, nullptr, e->captureStream, nullptr, nullptr, false, pdl && i > 0, false, e->pdlEarly != 0);
            cudaError_t err = cudaStreamEndCapture(e->captureStream, &graph);
            if (rc == 0 && err == cudaSuccess && cudaGraphInstantiate(&e->stepGraphExec, graph, 0) != cudaSuccess) e->stepGraphExec = nullptr;
            if (graph) { cudaGraphDestroy(graph); graph = nullptr; }
            cudaGetLastError();
        }
        e->launches = launches0;
        if (!e->stepGraphExec) { set_error("bb_step_n: graph capture failed"); return -1; }
        e->stepGraphN = n;
    }
    BB_CUDA_OK(cudaGraphLaunch(e->stepGraphExec, (cudaStream_t)cuda_stream));
    e->launches += n;
    return 0;
}

int bb_step_n_with_actions(bb_engine *e, int32_t n, const int32_t *actions_dev, void *cuda_stream)
{
    if (n <= 0) return 0;
    if (e->serving) { set_error("bb_step_n_with_actions: the persistent server is running (bb_serve_end first)"); return -1; }
    if (!actions_dev) { set_error("bb_step_n_with_actions: actions_dev is NULL"); return -1; }
    cudaStream_t s = (cudaStream_t)cuda_stream;
    const int64_t stride = e->N * 12; /* int32 elements per step */
    if (e->stepLoop) {
        const int rc = launch_step(e, actions_dev, s, nullptr, nullptr, false, pdl_for(e, s), false, false, n, stride);
        if (rc <= 0) return rc;
    }
    for (int32_t i = 0; i < n; i++) /* launch shapes without a looping instantiation: n ordinary launches */
        if (launch_step(e, actions_dev + (int64_t)i * stride, s, nullptr, nullptr, false, pdl_for(e, s), false, e->pdlEarly != 0) != 0) return -1;
    return 0;
}

int bb_step_with_actions(bb_engine *e, const int32_t *actions_dev, void *cuda_stream)
{
    if (e->serving) { set_error("bb_step_with_actions: the persistent server is running (bb_serve_end first)"); return -1; }
    return launch_step(e, actions_dev, (cudaStream_t)cuda_stream, nullptr, nullptr, false, pdl_for(e, (cudaStream_t)cuda_stream), false, e->pdlEarly != 0);
}

/* Device alias of a pinned (page-locked, mapped) host buffer, or NULL for pageable memory.
 * Queried every call (about a microsecond): a cached answer could outlive the buffer. */
static void *mapped_alias(const void *host)
{
    if (!host) return nullptr;
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, host) == cudaSuccess && at.type == cudaMemoryTypeHost) return at.devicePointer;
    cudaGetLastError();
    return nullptr;
}

static int step_host_impl(bb_engine *e, const void *actions_host, bool u8, float *reward_host, float *done_host, float *obs_host);

int bb_step_host(bb_engine *e, const int32_t *actions_host, float *reward_host, float *done_host, float *obs_host)
{
    return step_host_impl(e, actions_host, false, reward_host, done_host, obs_host);
}

int bb_step_host_u8(bb_engine *e, const uint8_t *actions_host, float *reward_host, float *done_host, float *obs_host)
{
    return step_host_impl(e, actions_host, true, reward_host, done_host, obs_host);
}

static int step_host_impl(bb_engine *e, const void *actions_host, bool u8, float *reward_host, float *done_host, float *obs_host)
{
    if (e->serving) { set_error("bb_step_host: the persistent server is running (bb_serve_end first)"); return -1; }
    cudaStream_t s = e->hostStream;
    const bb::Columns &C = e->params.col;
    /* Pinned buffers: the kernel itself moves the bytes over PCIe (reads the action rows from, and
     * mirrors reward / done into, the mapped host pages) -- one launch + one wait per step instead of
     * three DMA operations around it.  Pageable buffers: staged copies. */
    const int32_t *act_dev = (const int32_t *)mapped_alias(actions_host);
    float *rew_dev = (float *)mapped_alias(reward_host);
    float *done_dev = (float *)mapped_alias(done_host);
    if (actions_host && !act_dev) {
        if (u8) { set_error("bb_step_host_u8: the packed action rows must be pinned (page-locked, mapped) host memory"); return -1; }
        BB_CUDA_OK(cudaMemcpyAsync(C.action, actions_host, (size_t)e->N * 48, cudaMemcpyHostToDevice, s));
    }
    if (launch_step(e, act_dev, s, rew_dev, done_dev, act_dev != nullptr, false, u8 && act_dev != nullptr) != 0) return -1;
    if (reward_host && !rew_dev) BB_CUDA_OK(cudaMemcpyAsync(reward_host, C.reward, (size_t)e->N * 8, cudaMemcpyDeviceToHost, s));
    if (done_host && !done_dev) BB_CUDA_OK(cudaMemcpyAsync(done_host, C.done, (size_t)e->N * 8, cudaMemcpyDeviceToHost, s));
    if (obs_host) BB_CUDA_OK(cudaMemcpyAsync(obs_host, C.obs, (size_t)e->N * 1024, cudaMemcpyDeviceToHost, s));
    BB_CUDA_OK(cudaStreamSynchronize(s));
    return 0;
}

/* ---- persistent server: host-driven steps without a launch or a stream wait per step ---- */

static int serve_launch(bb_engine *e)
{
    const int sthreads = e->serveThreads;
    const int smem = bb::smem_bytes_for(sthreads, bb::kWorldsPerWarp);
    int perSm = 0, sms = 0;
    BB_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm, bb::bb_serve_kernel, sthreads, smem));
    BB_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->cfg.gpu_id));
    const int64_t tiles = (e->N + bb::kWorldsPerWarp - 1) / bb::kWorldsPerWarp;
    const int64_t blocksWanted = (tiles * 32 + sthreads - 1) / sthreads;
    int64_t cap = (int64_t)perSm * sms;
    if (cap > 8 * (int64_t)sms) cap = 8 * (int64_t)sms; /* leave room for other kernels on every SM */
    if (sthreads == 128 && cap > sms) cap = sms;        /* four lone warps per block: one block per SM */
    const int grid = (int)(blocksWanted < cap ? blocksWanted : cap);
    if (grid <= 0) { set_error("bb_serve_begin: server kernel cannot be resident"); return -1; }
    uint64_t idle_ns = 2000000000ull; /* the server leaves after 2 s without a doorbell */
    if (const char *s = getenv("BB_SERVE_IDLE_MS")) idle_ns = (uint64_t)atoll(s) * 1000000ull;
    for (int i = 0; i < 8; i++) e->serveHost->word[i] = 0;
    e->serveHost->word[0] = (unsigned long long)((unsigned)e->serveSeq & 0xffffu) << 48;
    e->comboCount = 0; e->comboNext = 0; /* the device table is cleared below */
    e->serveHost->complete = e->serveSeq;
    e->serveHost->alive = 1;
    BB_CUDA_OK(cudaMemsetAsync(e->serveDev, 0, sizeof(bb::ServeDevCtl), e->serveStream));
    {   /* the forwarded doorbell starts with the tag of the last completed sequence, like the host's */
        const unsigned long long bell = (unsigned long long)((unsigned)e->serveSeq & 0xffffu) << 48;
        BB_CUDA_OK(cudaMemcpyAsync(&e->serveDev->bell, &bell, sizeof bell, cudaMemcpyHostToDevice, e->serveStream));
        BB_CUDA_OK(cudaStreamSynchronize(e->serveStream)); /* `bell` is on this stack frame */
    }
    bb::StepParams p = e->params;
    p.prefetchBlocks = 0;
    bb::ServeHostCtl *hc_dev = nullptr;
    BB_CUDA_OK(cudaHostGetDevicePointer((void **)&hc_dev, e->serveHost, 0));
    int flags = 0;
    if (const char *s = getenv("BB_SERVE_FLAGS")) flags = atoi(s);
    bb::bb_serve_kernel<<<grid, sthreads, smem, e->serveStream>>>(p, hc_dev, e->serveDev, e->serveSeq, idle_ns, flags);
    e->launches++;
    BB_CUDA_OK(cudaGetLastError());
    e->serving = true;
    return 0;
}

int bb_serve_begin(bb_engine *e)
{
    if (e->serving && e->serveHost->alive) return 0;
    cudaSetDevice(e->cfg.gpu_id);
    if (!e->serveStream) BB_CUDA_OK(cudaStreamCreateWithFlags(&e->serveStream, cudaStreamNonBlocking));
    if (!e->serveHost) {
        BB_CUDA_OK(cudaHostAlloc((void **)&e->serveHost, sizeof(bb::ServeHostCtl), cudaHostAllocMapped));
        memset(e->serveHost, 0, sizeof(bb::ServeHostCtl));
    }
    if (!e->serveDev) BB_CUDA_OK(cudaMalloc((void **)&e->serveDev, sizeof(bb::ServeDevCtl)));
    BB_CUDA_OK(cudaDeviceSynchronize()); /* earlier steps / tensor writes on other streams are complete */
    e->aliasCount = 0;
    return serve_launch(e);
}

static int serve_step_impl(bb_engine *e, const void *actions_host, bool u8, float *reward_host, float *done_host, float *obs_host);

int bb_serve_step(bb_engine *e, const int32_t *actions_host, float *reward_host, float *done_host, float *obs_host)
{
    return serve_step_impl(e, actions_host, false, reward_host, done_host, obs_host);
}

int bb_serve_step_u8(bb_engine *e, const uint8_t *actions_host, float *reward_host, float *done_host, float *obs_host)
{
    return serve_step_impl(e, actions_host, true, reward_host, done_host, obs_host);
}

static int serve_step_impl(bb_engine *e, const void *actions_host, bool u8, float *reward_host, float *done_host, float *obs_host)
{
    if (!e->serving) { set_error("bb_serve_step: call bb_serve_begin first"); return -1; }
    bb::ServeHostCtl *hc = e->serveHost;
    if (!__atomic_load_n(&hc->alive, __ATOMIC_ACQUIRE)) { /* idle timeout: the server left; start it again */
        BB_CUDA_OK(cudaStreamSynchronize(e->serveStream));
        if (serve_launch(e) != 0) return -1;
    }
    const void *hostPtrs[4] = {actions_host, reward_host, done_host, obs_host};
    const void *ptrs[4] = {nullptr, nullptr, nullptr, nullptr}; /* device aliases: resolved only for a combination the table lacks */
    auto resolve = [&]() -> int {
        auto alias = [e](const void *host) -> void * {
            if (!host) return nullptr;
            for (int i = 0; i < e->aliasCount; i++)
                if (e->aliasHost[i] == host) return e->aliasDev[i];
            void *dev = mapped_alias(host);
            if (dev && e->aliasCount < 64) { e->aliasHost[e->aliasCount] = host; e->aliasDev[e->aliasCount++] = dev; }
            return dev;
        };
        for (int i = 0; i < 4; i++) {
            ptrs[i] = alias(hostPtrs[i]);
            if (hostPtrs[i] && !ptrs[i]) { set_error("bb_serve_step: buffers must be pinned (page-locked, mapped) host memory"); return -1; }
            if ((unsigned long long)(uintptr_t)ptrs[i] & ~bb::kServePtrMask) { set_error("bb_serve_step: buffer alias beyond 47 bits"); return -1; }
        }
        return 0;
    };
    /* ring: ONE 8-byte word (the only one block 0 polls).  A buffer combination the device table does not hold yet is
     * validated (pinned, mapped) and travels once, in word[2..5], written before the doorbell word; a combination the
     * table holds costs a scan of at most 64 x 4 host pointers. */
    int slot = -1;
    for (int i = 0; i < e->comboCount && slot < 0; i++)
        if (e->comboHost[i][0] == hostPtrs[0] && e->comboHost[i][1] == hostPtrs[1] && e->comboHost[i][2] == hostPtrs[2] &&
            e->comboHost[i][3] == hostPtrs[3]) slot = i;
    if (slot < 0 && resolve() != 0) return -1;
    const int32_t seq = ++e->serveSeq;
    e->launches++; /* one pass of the step code over every world, like a launch */
    const unsigned long long tag = (unsigned long long)((unsigned)seq & 0xffffu) << 48;
    auto ring = [&](int useSlot) -> int {
        unsigned long long w0 = tag | ((u8 && actions_host) ? bb::kServeU8Bit : 0ull);
        if (useSlot < 0) { /* a new combination, or the server was relaunched (its table is empty) */
            if (!ptrs[0] && !ptrs[1] && !ptrs[2] && !ptrs[3] && resolve() != 0) return -1;
            useSlot = e->comboCount < bb::kServeSlots ? e->comboCount++ : (e->comboNext++ % bb::kServeSlots);
            for (int i = 0; i < 4; i++) {
                e->comboHost[useSlot][i] = hostPtrs[i];
                __atomic_store_n(&hc->word[2 + i], (unsigned long long)(uintptr_t)ptrs[i], __ATOMIC_RELAXED);
            }
            w0 |= bb::kServeLongBit;
        }
        __atomic_store_n(&hc->word[0], w0 | (unsigned long long)useSlot, __ATOMIC_RELEASE);
        return 0;
    };
    if (ring(slot) != 0) return -1;
    /* spin on the completion word the last block writes into this page */
    uint64_t spins = 0;
    int relaunches = 0;
    while (__atomic_load_n(&hc->complete, __ATOMIC_ACQUIRE) != seq) {
        __builtin_ia32_pause();
        if ((++spins & 0xfffff) == 0) { /* every ~1M spins: is the server still there? */
            if (!__atomic_load_n(&hc->alive, __ATOMIC_ACQUIRE) || cudaStreamQuery(e->serveStream) != cudaErrorNotReady) {
                cudaError_t err = cudaStreamSynchronize(e->serveStream);
                if (err == cudaSuccess && __atomic_load_n(&hc->complete, __ATOMIC_ACQUIRE) != seq && relaunches++ == 0) {
                    /* the server idled out between the alive check above and the doorbell (it saw no step for
                     * BB_SERVE_IDLE_MS): start it again for THIS sequence number and ring once more */
                    e->serveSeq = seq - 1;
                    const int rc = serve_launch(e);
                    e->serveSeq = seq;
                    if (rc != 0) return -1;
                    if (ring(-1) != 0) return -1; /* serve_launch cleared the device table */
                    continue;
                }
                e->serving = false;
                set_error(std::string("bb_serve_step: server kernel stopped: ") + cudaGetErrorString(err));
                return -1;
            }
        }
    }
    return 0;
}

/* tuning aid: GPU-side nanoseconds between "doorbell seen" and "last block done" of the last step */
long long bb_serve_last_gpu_ns(bb_engine *e) { return e->serveHost ? (long long)(e->serveHost->tDone - e->serveHost->tSeen) : -1; }
/* tuning aid: the eight %globaltimer marks of the last served step (see ServeHostCtl::mark), relative to mark 0 */
int bb_serve_last_marks(bb_engine *e, long long out[8])
{
    if (!e->serveHost) return -1;
    for (int i = 0; i < 8; i++) out[i] = (long long)(e->serveHost->mark[i] - e->serveHost->mark[0]);
    return 0;
}

int bb_serve_end(bb_engine *e)
{
    if (!e || !e->serving) return 0;
    __atomic_store_n(&e->serveHost->word[0], bb::kServeExitWord, __ATOMIC_RELEASE);
    cudaError_t err = cudaStreamSynchronize(e->serveStream);
    e->serving = false;
    if (err != cudaSuccess) { set_error(std::string("bb_serve_end: ") + cudaGetErrorString(err)); return -1; }
    return 0;
}

int bb_export(bb_engine *e, int slot, void **dev_ptr, int *dtype, int64_t dims[4], int *ndim, int *device)
{
    struct Desc { int slot, col, dtype, ndim; int64_t d1, d2; };
    static const Desc table[] = {
        {BB_SLOT_RESET, 0, BB_DTYPE_I32, 3, 2, 1},           {BB_SLOT_GAME_STATE, 1, BB_DTYPE_F32, 2, 14, 0},
        {BB_SLOT_ACTION, 2, BB_DTYPE_I32, 3, 2, 6},          {BB_SLOT_ACTION_MASK, 3, BB_DTYPE_I32, 3, 2, 4},
        {BB_SLOT_AGENT_POS, 4, BB_DTYPE_F32, 3, 2, 3},       {BB_SLOT_OBSERVATIONS, 5, BB_DTYPE_F32, 3, 2, 128},
        {BB_SLOT_REWARD, 6, BB_DTYPE_F32, 2, 2, 0},          {BB_SLOT_DONE, 7, BB_DTYPE_F32, 2, 2, 0},
        {BB_SLOT_AGENT_ENTITY_ID, 8, BB_DTYPE_I32, 2, 2, 0}, {BB_SLOT_AGENT_POSSESSION, 9, BB_DTYPE_I32, 3, 2, 3},
        {BB_SLOT_ORIENTATION, 10, BB_DTYPE_F32, 3, 2, 4},    {BB_SLOT_TEAM_DATA, 11, BB_DTYPE_I32, 3, 2, 5},
        {BB_SLOT_AGENT_STATS, 12, BB_DTYPE_I32, 3, 2, 2},    {BB_SLOT_BASKETBALL_POS, 13, BB_DTYPE_F32, 3, 1, 3},
        {BB_SLOT_BALL_PHYSICS, 14, BB_DTYPE_I32, 3, 1, 7},   {BB_SLOT_BALL_ENTITY_ID, 15, BB_DTYPE_I32, 2, 1, 0},
        {BB_SLOT_BALL_GRABBED, 16, BB_DTYPE_I32, 3, 1, 2},   {BB_SLOT_BALL_VELOCITY, 17, BB_DTYPE_F32, 3, 1, 3},
        {BB_SLOT_HOOP_POS, 18, BB_DTYPE_F32, 3, 2, 3},
        {BB_SLOT_AGENT_VELOCITY, 19, BB_DTYPE_F32, 3, 2, 3}, {BB_SLOT_INBOUNDING, 20, BB_DTYPE_I32, 3, 2, 2},
        {BB_SLOT_GRAB_COOLDOWN, 21, BB_DTYPE_F32, 2, 2, 0},  {BB_SLOT_ATTRIBUTES, 22, BB_DTYPE_F32, 3, 2, 9},
        {BB_SLOT_CUR_STEP, 23, BB_DTYPE_I32, 2, 2, 0},       {BB_SLOT_WORLD_CLOCK, 24, BB_DTYPE_I32, 1, 0, 0},
        {BB_SLOT_RNG_COUNT, 25, BB_DTYPE_I32, 1, 0, 0},
    };
    for (const Desc &d : table) {
        if (d.slot != slot) continue;
        *dev_ptr = e->colPtr[d.col];
        *dtype = d.dtype;
        *ndim = d.ndim;
        dims[0] = e->N; dims[1] = d.d1; dims[2] = d.d2; dims[3] = 0;
        *device = e->cfg.gpu_id;
        return 0;
    }
    set_error("bb_export: unknown slot " + std::to_string(slot));
    return -1;
}

int bb_set_action(bb_engine *e, int32_t world_idx, int32_t agent_idx, int32_t move_speed, int32_t move_angle,
                  int32_t rotate, int32_t grab, int32_t pass, int32_t shoot)
{
    if (world_idx >= 0 && world_idx < (int32_t)e->N && agent_idx >= 0 && agent_idx < 2) {
        int32_t host[6] = {move_speed, move_angle, rotate, grab, pass, shoot};
        int32_t *dst = e->params.col.action + ((int64_t)world_idx * 2 + agent_idx) * 6;
        BB_CUDA_OK(cudaMemcpy(dst, host, sizeof host, cudaMemcpyHostToDevice));
    } else {
        /* mgr.cpp:288-292 */
        printf("ERROR: Invalid indices! world=%d (max=%d), agent=%d (max=%d)\n", world_idx, (int32_t)e->N - 1, agent_idx, 1);
    }
    return 0;
}

int bb_trigger_reset(bb_engine *e, int32_t world_idx)
{
    if (world_idx >= 0 && world_idx < (int32_t)e->N) {
        int32_t host[2] = {1, 1};
        BB_CUDA_OK(cudaMemcpy(e->params.col.reset + (int64_t)world_idx * 2, host, sizeof host, cudaMemcpyHostToDevice));
    }
    return 0;
}

int bb_fill_actions(bb_engine *e, int32_t *actions_dev, uint32_t seed, int64_t step, int64_t world_offset, void *cuda_stream)
{
    int32_t *dst = actions_dev ? actions_dev : e->params.col.action;
    const int64_t n = e->N * 2;
    bb::bb_fill_actions_kernel<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(dst, e->N, seed, (uint32_t)step, world_offset);
    e->launches++;
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int bb_checksum(bb_engine *e, uint64_t *out_dev, void *cuda_stream)
{
    bb::ChecksumParams cp;
    for (int c = 0; c < 26; c++) {
        cp.col[c] = (const uint32_t *)e->colPtr[c];
        cp.words[c] = bb::kColBytes[c] / 4;
    }
    bb::bb_checksum_kernel<<<(int)((e->N + 127) / 128), 128, 0, (cudaStream_t)cuda_stream>>>(cp, e->N, out_dev);
    e->launches++;
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int bb_umma_selftest(const float *a_dev, const float *b_dev, float *d_dev, int k, void *cuda_stream)
{
    if (!a_dev || !b_dev || !d_dev || k < 8 || k > 128 || k % 8) { set_error("bb_umma_selftest: bad arguments"); return -1; }
    const int smem = (128 + 128 + 32 + 32) * k * 4;
    BB_CUDA_OK(cudaFuncSetAttribute(bb::bb_umma_selftest_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    bb::bb_umma_selftest_kernel<<<1, 128, smem, (cudaStream_t)cuda_stream>>>(a_dev, b_dev, d_dev, k);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int bb_math_eval(int fn, const float *x_dev, const float *y_dev, float *out_dev, int64_t n, void *cuda_stream)
{
    if (fn < 0 || fn > 6 || n < 0) { set_error("bb_math_eval: bad arguments"); return -1; }
    if (n == 0) return 0;
    bb::bb_math_eval_kernel<<<(int)((n + 255) / 256), 256, 0, (cudaStream_t)cuda_stream>>>(fn, x_dev, y_dev ? y_dev : x_dev, out_dev, n);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int64_t bb_launch_count(const bb_engine *e) { return e->launches; }
int64_t bb_num_worlds(const bb_engine *e) { return e->N; }

/* ---- fused rollout path (include/bb_rollout.h) ---- */

static int policy_ok(const bb_policy *p, const char *who)
{
    if (!p || p->hidden < 1 || p->hidden > BB_POLICY_MAX_HIDDEN) { set_error(std::string(who) + ": policy.hidden must be 1..32"); return -1; }
    if (((uintptr_t)p->w1 & 15) != 0) { set_error(std::string(who) + ": policy.w1 must be 16-byte aligned"); return -1; }
    return 0;
}

/* BB_POLICY_UMMA=1 (read once) selects the tcgen05 policy kernel; BB_POLICY_UMMA_FORCE=0/1 overrides it for as long as it is
 * set (read per call: the tests run both kernels in one process) */
static bool policy_uses_umma()
{
    static const int dflt = [] { const char *s = getenv("BB_POLICY_UMMA"); return (s && atoi(s) != 0) ? 1 : 0; }();
    if (const char *s = getenv("BB_POLICY_UMMA_FORCE")) return atoi(s) != 0;
    return dflt != 0;
}

static int launch_policy(const bb_policy *pol, const bb::PolicyIo &io, cudaStream_t stream, bool pdl = false)
{
    if (io.n <= 0) return 0;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    {   /* BB_POLICY_UMMA=1: the tcgen05 / TMEM kernel (csrc/bb_umma.cuh): 128 rows per CTA, one CTA per SM */
        if (policy_uses_umma()) {
            static bool attr = false;
            if (!attr) {
                BB_CUDA_OK(cudaFuncSetAttribute(bb::bb_policy_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bb::kUmSmemBytes));
                attr = true;
            }
            int64_t tiles = (io.n + bb::kUmRows - 1) / bb::kUmRows;
            if (tiles > sms) tiles = sms;
            if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_policy_umma_kernel, (int)tiles, bb::kUmRows, (size_t)bb::kUmSmemBytes, stream, true, *pol, io));
            else bb::bb_policy_umma_kernel<<<(int)tiles, bb::kUmRows, bb::kUmSmemBytes, stream>>>(*pol, io);
            BB_CUDA_OK(cudaGetLastError());
            return 0;
        }
    }
    int64_t blocks = (io.n + bb::kRowsPerBlock - 1) / bb::kRowsPerBlock;
    if (blocks > 4 * (int64_t)sms) blocks = 4 * (int64_t)sms; /* rows are strided over the resident blocks */
    if (pdl) BB_CUDA_OK(launch_pdl(bb::bb_policy_kernel, (int)blocks, bb::kPolThreads, (size_t)0, stream, true, *pol, io));
    else bb::bb_policy_kernel<<<(int)blocks, bb::kPolThreads, 0, stream>>>(*pol, io);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int bb_policy_forward(const bb_policy *pol, const float *obs, int64_t obs_stride, int64_t n, int stochastic, uint32_t seed,
                      uint32_t draw, int64_t row_offset, int32_t *actions_out, int64_t action_stride, float *logp_out,
                      float *value_out, float *logits_out, void *cuda_stream)
{
    if (policy_ok(pol, "bb_policy_forward") != 0) return -1;
    bb::PolicyIo io{};
    io.obs = obs; io.obsStride = obs_stride; io.n = n; io.rowOffset = row_offset;
    io.stochastic = stochastic; io.seed = seed; io.draw = draw;
    io.actionsOut = actions_out; io.actionStride = action_stride;
    io.logpOut = logp_out; io.valueOut = value_out; io.logitsOut = logits_out;
    return launch_policy(pol, io, (cudaStream_t)cuda_stream);
}

int bb_rollout(bb_engine *e, const bb_policy *trainee, const bb_policy *frozen, const bb_rollout_buffers *buf, int32_t agent_idx,
               uint32_t seed, uint32_t draw_base, int64_t world_offset, void *cuda_stream)
{
    if (policy_ok(trainee, "bb_rollout") != 0) return -1;
    if (frozen && policy_ok(frozen, "bb_rollout(frozen)") != 0) return -1;
    if (!buf || buf->horizon < 1 || agent_idx < 0 || agent_idx > 1) { set_error("bb_rollout: bad buffers / agent_idx"); return -1; }
    if (e->serving) { set_error("bb_rollout: the persistent server is running (bb_serve_end first)"); return -1; }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    const bb::Columns &C = e->params.col;
    const int64_t n = e->N;
    const int T = buf->horizon;
    /* Default: the whole rollout as ONE launch of bb_rollout_loop_kernel (every warp takes its 16 worlds through env.reset() and
     * the T x { policy rows; step } in a row).  BB_ROLLOUT_LOOP=0: round 2's chain of 2-3 kernels per step below. */
    const char *loopEnv = getenv("BB_ROLLOUT_LOOP"); /* read per call: the tests run both forms in one process */
    bool loop = !(loopEnv && atoi(loopEnv) == 0) && !policy_uses_umma(); /* the looping kernel carries the mma.sync policy rows */
    const int64_t tiles = (n + bb::kWorldsPerWarp - 1) / bb::kWorldsPerWarp;
    int tpb = 0; /* tiles per block of the looping kernel: 1, or 2 when only that makes the grid resident (frozen opponent) */
    if (loop) { /* only when every block of the looping kernel is resident at once (see the kernel) */
        static bool attr = false;
        if (!attr) {
            BB_CUDA_OK(cudaFuncSetAttribute(bb::bb_rollout_loop_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bb::rollout_loop_smem_bytes<1>(true)));
            BB_CUDA_OK(cudaFuncSetAttribute(bb::bb_rollout_loop_kernel<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bb::rollout_loop_smem_bytes<2>(true)));
            attr = true;
        }
        int perSm1 = 0, perSm2 = 0, sms = 148;
        BB_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm1, bb::bb_rollout_loop_kernel<1>, bb::kLoopWarps * 32, bb::rollout_loop_smem_bytes<1>(frozen != nullptr)));
        BB_CUDA_OK(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSm2, bb::bb_rollout_loop_kernel<2>, bb::kLoopWarps * 64, bb::rollout_loop_smem_bytes<2>(frozen != nullptr)));
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, e->cfg.gpu_id);
        const int forced = loopEnv ? atoi(loopEnv) : 1; /* 2 / 3: force one / two tiles per block whatever the residency (tests, A/B) */
        if (forced == 2) tpb = 1;
        else if (forced == 3) tpb = 2;
        else if ((tiles + 1) / 2 <= (int64_t)perSm2 * sms) tpb = 2; /* measured 2 % faster than one tile per block where both fit */
        else if (tiles <= (int64_t)perSm1 * sms) tpb = 1;
        else loop = false;
    }
    if (loop) {
        bb::RolloutLoop rl{};
        rl.buf = *buf; rl.agentIdx = agent_idx; rl.hasFrozen = frozen ? 1 : 0;
        rl.seed = seed; rl.drawBase = draw_base; rl.worldOffset = world_offset;
        bb::StepParams p = e->params;
        p.io = bb::StepIo{nullptr, nullptr, nullptr, nullptr, 0, 0};
        p.prefetchBlocks = 0; p.prefetchSelf = 0; p.earlyTrigger = 0;
        if (tpb == 1)
            BB_CUDA_OK(launch_pdl(bb::bb_rollout_loop_kernel<1>, (int)tiles, bb::kLoopWarps * 32, bb::rollout_loop_smem_bytes<1>(frozen != nullptr), st,
                                  pdl_for(e, st), p, *trainee, frozen ? *frozen : *trainee, rl));
        else
            BB_CUDA_OK(launch_pdl(bb::bb_rollout_loop_kernel<2>, (int)((tiles + 1) / 2), bb::kLoopWarps * 64, bb::rollout_loop_smem_bytes<2>(frozen != nullptr), st,
                                  pdl_for(e, st), p, *trainee, frozen ? *frozen : *trainee, rl));
        e->launches += T + 1; /* passes of the step code over every world */
        BB_CUDA_OK(cudaGetLastError());
        return 0;
    }
    /* env.reset(), env.py:178-185 */
    bb::bb_reset_prologue_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(C.reset, C.action, n, agent_idx);
    e->launches++;
    BB_CUDA_OK(cudaGetLastError());
    static const bool pdl = [] { const char *s = getenv("BB_ROLLOUT_PDL"); return !(s && atoi(s) == 0); }();
    /* bit 0: the step kernel releases the policy kernel early (its blocks stage the weights while the step runs: rollout
     * 0.715 -> 0.689 ms); bit 1: the policy kernel releases the step kernel early (measured slower, 0.746 ms: off) */
    static const int early = [] { const char *s = getenv("BB_ROLLOUT_EARLY"); return s ? atoi(s) : 1; }();
    if (launch_step(e, nullptr, st) != 0) return -1;
    for (int t = 0; t <= T; t++) {
        bb::PolicyIo io{};
        io.obs = C.obs + agent_idx * BB_OBS_DIM; io.obsStride = 2 * BB_OBS_DIM; io.n = n; io.rowOffset = world_offset;
        io.stochastic = 1; io.seed = seed; io.draw = draw_base + (uint32_t)t;
        io.evaluateOnly = (t == T);
        if (t < T) {
            io.actionsOut = C.action + agent_idx * 6; io.actionStride = 12;
            io.logpOut = buf->log_probs + (int64_t)t * n;
            io.valueOut = buf->values + (int64_t)t * n;
            io.bufObs = buf->obs + (int64_t)t * n * BB_OBS_DIM;
            io.bufActions = buf->actions + (int64_t)t * n * BB_NUM_ACTION_DIMS;
        } else {
            io.nextValue = buf->next_value;
        }
        if (t > 0) { /* reward / done of step t-1 are in the columns now */
            io.rew = C.reward + agent_idx; io.done = C.done + agent_idx; io.rdStride = 2;
            io.bufRewPrev = buf->rewards + (int64_t)(t - 1) * n;
            io.bufNotDonePrev = buf->not_dones + (int64_t)(t - 1) * n;
            io.curReward = buf->cur_reward; io.epLength = buf->ep_length; io.epSums = buf->ep_sums;
        } else {
            io.resetCol = C.reset; /* env.py:182 resets.fill_(0) */
        }
        io.earlyTrigger = (pdl && (early & 2) && !frozen) ? 1 : 0;
        if (launch_policy(trainee, io, st, pdl) != 0) return -1;
        e->launches++;
        if (t == T) break;
        if (frozen) { /* env.py:126-143: the frozen policy drives the other agent from its own observation row */
            bb::PolicyIo fo{};
            const int other = 1 - agent_idx;
            fo.obs = C.obs + other * BB_OBS_DIM; fo.obsStride = 2 * BB_OBS_DIM; fo.n = n; fo.rowOffset = world_offset;
            fo.stochastic = 1; fo.seed = seed; fo.draw = 0x80000000u + draw_base + (uint32_t)t;
            fo.actionsOut = C.action + other * 6; fo.actionStride = 12;
            fo.earlyTrigger = (pdl && (early & 2)) ? 1 : 0;
            if (launch_policy(frozen, fo, st, pdl) != 0) return -1;
            e->launches++;
        }
        if (launch_step(e, nullptr, st, nullptr, nullptr, false, pdl, false, (early & 1) != 0) != 0) return -1;
    }
    return 0;
}

int bb_gae(const bb_rollout_buffers *buf, int64_t n, double gamma, double gae_lambda, const double *value_mean,
           const double *value_var, float value_clamp, float *values_unnorm_out, float *advantages_out, float *returns_out,
           void *cuda_stream)
{
    if (!buf || buf->horizon < 1 || n <= 0 || !advantages_out || !returns_out) { set_error("bb_gae: bad arguments"); return -1; }
    bb::bb_gae_kernel<<<(int)((n + 127) / 128), 128, 0, (cudaStream_t)cuda_stream>>>(
        buf->values, buf->next_value, buf->rewards, buf->not_dones, buf->horizon, n, (float)gamma, (float)(gamma * gae_lambda),
        value_mean, value_var, value_clamp, values_unnorm_out, advantages_out, returns_out);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

/* ---- fused PPO update (include/bb_ppo.h) ---- */

static int ppo_grid(int *blocks_out)
{
    int dev = 0, sms = 148;
    BB_CUDA_OK(cudaGetDevice(&dev));
    BB_CUDA_OK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    *blocks_out = 2 * sms; /* __launch_bounds__(256, 2): every block resident, tiles strided over them */
    return 0;
}

int64_t bb_ppo_workspace_bytes(void)
{
    int blocks = 0;
    if (ppo_grid(&blocks) != 0) return -1;
    return (int64_t)blocks * bb::kPpoPartial * (int64_t)sizeof(float);
}

int bb_ppo_grad(const bb_policy *pol, const bb_ppo_batch *batch, void *workspace, float *grad_out, float *stats_out,
                void *cuda_stream)
{
    if (!pol || pol->hidden != BB_PPO_HIDDEN) { set_error("bb_ppo_grad: policy.hidden must be 32 (Agent(num_channels=32, num_layers=2))"); return -1; }
    if (((uintptr_t)pol->w1 & 15) != 0) { set_error("bb_ppo_grad: policy.w1 must be 16-byte aligned"); return -1; }
    if (!batch || batch->num_rows <= 0 || !batch->obs || !batch->actions || !batch->log_probs || !batch->values ||
        !batch->advantages || !batch->returns || !workspace || !grad_out) { set_error("bb_ppo_grad: bad arguments"); return -1; }
    if (((uintptr_t)batch->obs & 15) != 0 || ((uintptr_t)workspace & 15) != 0) { set_error("bb_ppo_grad: obs / workspace must be 16-byte aligned"); return -1; }
    static bool attrSet = false;
    if (!attrSet) {
        BB_CUDA_OK(cudaFuncSetAttribute(bb::bb_ppo_grad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(bb::PpoSmem)));
        attrSet = true;
    }
    int maxBlocks = 0;
    if (ppo_grid(&maxBlocks) != 0) return -1;
    bb::PpoArgs a{};
    a.b = *batch;
    a.invM = 1.0f / (float)batch->num_rows;
    /* equal shares: with t tiles per resident block, rows per tile = ceil(m / (blocks * t)) rounded up to a warp's 8 rows
     * (65 536 rows on 296 blocks: 4 tiles of 56 rows instead of 3-or-4 tiles of 64) */
    const int64_t tilesPerBlock = (batch->num_rows + (int64_t)maxBlocks * bb::kPpoRows - 1) / ((int64_t)maxBlocks * bb::kPpoRows);
    int64_t rpt = (batch->num_rows + maxBlocks * tilesPerBlock - 1) / (maxBlocks * tilesPerBlock);
    rpt = (rpt + 7) / 8 * 8;
    if (rpt > bb::kPpoRows) rpt = bb::kPpoRows;
    a.rowsPerTile = (int)rpt;
    a.numTiles = (batch->num_rows + rpt - 1) / rpt;
    a.partials = (float *)workspace;
    const int blocks = (int)(a.numTiles < maxBlocks ? a.numTiles : maxBlocks);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    static const bool pdl = [] { const char *s = getenv("BB_PPO_PDL"); return !(s && atoi(s) == 0); }();
    BB_CUDA_OK(launch_pdl(bb::bb_ppo_grad_kernel, blocks, bb::kPpoThreads, sizeof(bb::PpoSmem), st, pdl, *pol, a));
    BB_CUDA_OK(launch_pdl(bb::bb_ppo_reduce_kernel, (bb::kPpoPartial + bb::kRedCols - 1) / bb::kRedCols, bb::kRedCols * bb::kRedSlices, (size_t)0, st, pdl,
                          (const float *)a.partials, blocks, grad_out, stats_out));
    return 0;
}

int bb_ppo_adam(float *params, const float *grad, float *exp_avg, float *exp_avg_sq, int32_t *step_dev, int64_t n,
                float grad_scale, float max_grad_norm, float lr, float beta1, float beta2, float eps, float *grad_norm_out,
                void *cuda_stream)
{
    if (!params || !grad || !exp_avg || !exp_avg_sq || !step_dev || n <= 0) { set_error("bb_ppo_adam: bad arguments"); return -1; }
    static const bool pdl = [] { const char *s = getenv("BB_PPO_PDL"); return !(s && atoi(s) == 0); }();
    BB_CUDA_OK(launch_pdl(bb::bb_ppo_adam_kernel, 1, bb::kAdamThreads, (size_t)0, (cudaStream_t)cuda_stream, pdl, params, grad, exp_avg,
                          exp_avg_sq, step_dev, n, grad_scale, max_grad_norm, lr, beta1, beta2, eps, grad_norm_out));
    return 0;
}

/* ---- gradient all-reduce over NVLink peer memory fused into the optimiser step (bb_ppo.h) ---- */
struct bb_p2p {
    int rank = 0, ws = 1, dev = 0;
    int64_t stride = 0;           /* floats per parity slot */
    uint32_t seq = 0;             /* calls made so far */
    unsigned long long *local = nullptr; /* this rank's exchange buffer (cudaMalloc): [2][ws][stride] tagged words */
    bb::P2pPeers peers{};         /* every rank's buffer as seen from this device */
    bool connected = false;
};

bb_p2p *bb_p2p_create(int32_t rank, int32_t world_size, int64_t num_floats, unsigned char handle_out[64])
{
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "CUDA IPC handle size");
    if (world_size < 1 || world_size > bb::kP2pMaxRanks || rank < 0 || rank >= world_size || num_floats <= 0 || !handle_out) {
        set_error("bb_p2p_create: bad arguments (world size 1..16)");
        return nullptr;
    }
    if (num_floats > bb::kP2pPer * bb::kAdamThreads) { set_error("bb_p2p_create: at most 6144 floats"); return nullptr; }
    bb_p2p *c = new bb_p2p();
    c->rank = rank; c->ws = world_size;
    c->stride = (num_floats + 31) / 32 * 32;
    const size_t bytes = (size_t)(2 * world_size * c->stride) * sizeof(unsigned long long);
    cudaIpcMemHandle_t h;
    if (cudaGetDevice(&c->dev) != cudaSuccess || cudaMalloc(&c->local, bytes) != cudaSuccess || cudaMemset(c->local, 0, bytes) != cudaSuccess ||
        cudaDeviceSynchronize() != cudaSuccess || cudaIpcGetMemHandle(&h, c->local) != cudaSuccess) {
        set_error(std::string("bb_p2p_create: ") + cudaGetErrorString(cudaGetLastError()));
        if (c->local) cudaFree(c->local);
        delete c;
        return nullptr;
    }
    memcpy(handle_out, &h, 64);
    c->peers.buf[rank] = c->local;
    return c;
}

int bb_p2p_connect(bb_p2p *c, const unsigned char *handles)
{
    if (!c || !handles) { set_error("bb_p2p_connect: bad arguments"); return -1; }
    if (c->connected) return 0;
    for (int r = 0; r < c->ws; r++) {
        if (r == c->rank) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)r * 64, 64);
        void *p = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
        if (e != cudaSuccess) {
            cudaGetLastError();
            set_error(std::string("bb_p2p_connect: cudaIpcOpenMemHandle(rank ") + std::to_string(r) + "): " + cudaGetErrorString(e));
            return -1;
        }
        c->peers.buf[r] = (unsigned long long *)p;
    }
    c->connected = true;
    return 0;
}

void bb_p2p_destroy(bb_p2p *c)
{
    if (!c) return;
    cudaDeviceSynchronize();
    for (int r = 0; r < c->ws; r++)
        if (r != c->rank && c->peers.buf[r]) cudaIpcCloseMemHandle(c->peers.buf[r]);
    if (c->local) cudaFree(c->local);
    delete c;
}

int bb_ppo_adam_p2p(bb_p2p *c, float *params, const float *grad, float *exp_avg, float *exp_avg_sq, int32_t *step_dev, int64_t n,
                    float max_grad_norm, float lr, float beta1, float beta2, float eps, float *grad_norm_out, int32_t *error_flag_dev,
                    int32_t timeout_ms, void *cuda_stream)
{
    if (!c || !params || !grad || !exp_avg || !exp_avg_sq || !step_dev || n <= 0 || n > c->stride) { set_error("bb_ppo_adam_p2p: bad arguments"); return -1; }
    if (!c->connected && c->ws > 1) { set_error("bb_ppo_adam_p2p: bb_p2p_connect has not been called"); return -1; }
    if (timeout_ms <= 0) timeout_ms = 2000;
    static const bool pdl = [] { const char *s = getenv("BB_PPO_PDL"); return !(s && atoi(s) == 0); }();
    c->seq += 1;
    BB_CUDA_OK(launch_pdl(bb::bb_ppo_adam_p2p_kernel, 1, bb::kAdamThreads, (size_t)0, (cudaStream_t)cuda_stream, pdl, c->peers, c->rank, c->ws,
                          c->stride, c->seq, params, grad, exp_avg, exp_avg_sq, step_dev, n, max_grad_norm, lr, beta1, beta2, eps,
                          grad_norm_out, error_flag_dev, (uint64_t)timeout_ms * 1000000ull));
    return 0;
}

int64_t bb_ppo_moments_workspace_bytes(void)
{
    int blocks = 0;
    if (ppo_grid(&blocks) != 0) return -1;
    return (int64_t)blocks * 2 * bb::kMomCols * (int64_t)sizeof(double);
}

int bb_ppo_moments(const float *obs, const float *values_unnorm, const float *returns, const float *advantages, int64_t rows,
                   void *workspace, double *sums_out, void *cuda_stream)
{
    if (!obs || !values_unnorm || !returns || !advantages || rows <= 0 || !workspace || !sums_out) { set_error("bb_ppo_moments: bad arguments"); return -1; }
    if (((uintptr_t)obs & 15) != 0) { set_error("bb_ppo_moments: obs must be 16-byte aligned"); return -1; }
    int maxBlocks = 0;
    if (ppo_grid(&maxBlocks) != 0) return -1;
    const int64_t want = (rows + 7) / 8;
    const int blocks = (int)(want < maxBlocks ? want : maxBlocks);
    cudaStream_t st = (cudaStream_t)cuda_stream;
    bb::bb_ppo_moments_kernel<<<blocks, bb::kMomThreads, 0, st>>>(obs, values_unnorm, returns, advantages, rows, (double *)workspace);
    BB_CUDA_OK(cudaGetLastError());
    bb::bb_ppo_moments_reduce_kernel<<<(2 * bb::kMomCols + 63) / 64, 64, 0, st>>>((const double *)workspace, blocks, sums_out);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

/* ---- bb_ppo_moments for callers that may not allocate (graph capture, serving loops): the per-block partial sums live in
 * a file-scope device array instead of a caller-provided workspace.  The array belongs to the library image: every GPU
 * that loads libbball.so holds these bytes until the process exits.  Its NAME is only a host-side stand-in here -- the
 * device address is looked up once per device with cudaGetSymbolAddress (that device current) and handed to the kernels. */
} // extern "C" (device storage and its address table have C++ linkage)

namespace bb {
constexpr int kMomStaticDoubles = BB_PPO_MOMENTS_STATIC_MAX_BLOCKS * 2 * kMomCols;
static __device__ __align__(256) double g_momentsScratch[kMomStaticDoubles];
} // namespace bb

static constexpr int kMomStaticMaxDevices = 64;
static double *g_momentsScratchAddr[kMomStaticMaxDevices]; /* per device, nullptr until resolved */

static double *moments_static_scratch()
{
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMomStaticMaxDevices) {
        set_error("bb_ppo_moments_static: no current CUDA device (or device ordinal beyond " + std::to_string(kMomStaticMaxDevices - 1) + ")");
        return nullptr;
    }
    double *p = __atomic_load_n(&g_momentsScratchAddr[dev], __ATOMIC_ACQUIRE);
    if (p) return p;
    void *sym = nullptr;
    const cudaError_t err = cudaGetSymbolAddress(&sym, bb::g_momentsScratch);
    if (err != cudaSuccess || !sym) {
        set_error(std::string("bb_ppo_moments_static: cudaGetSymbolAddress(g_momentsScratch) failed: ") + cudaGetErrorString(err));
        return nullptr;
    }
    if (((uintptr_t)sym & 255) != 0) { set_error("bb_ppo_moments_static: scratch array is not 256-byte aligned"); return nullptr; }
    __atomic_store_n(&g_momentsScratchAddr[dev], (double *)sym, __ATOMIC_RELEASE); /* racing threads resolve the same address */
    return (double *)sym;
}

extern "C" {

int bb_ppo_moments_static(const float *obs, const float *values_unnorm, const float *returns, const float *advantages, int64_t rows,
                          double *sums_out, void *cuda_stream)
{
    if (!obs || !values_unnorm || !returns || !advantages || rows <= 0 || !sums_out) { set_error("bb_ppo_moments_static: bad arguments"); return -1; }
    if (((uintptr_t)obs & 15) != 0) { set_error("bb_ppo_moments_static: obs must be 16-byte aligned"); return -1; }
    if (rows > (int64_t)BB_PPO_MOMENTS_STATIC_MAX_ROWS) {
        set_error("bb_ppo_moments_static: " + std::to_string(rows) + " rows exceed the static limit of " +
                  std::to_string((int64_t)BB_PPO_MOMENTS_STATIC_MAX_ROWS) + " (BB_PPO_MOMENTS_STATIC_MAX_ROWS); use bb_ppo_moments with a workspace");
        return -1;
    }
    int maxBlocks = 0;
    if (ppo_grid(&maxBlocks) != 0) return -1;
    const int64_t want = (rows + 7) / 8;
    const int blocks = (int)(want < maxBlocks ? want : maxBlocks); /* the grid bb_ppo_moments uses: same partials, same bits */
    if (blocks > BB_PPO_MOMENTS_STATIC_MAX_BLOCKS) {
        set_error("bb_ppo_moments_static: this device's grid of " + std::to_string(blocks) + " blocks exceeds the static scratch of " +
                  std::to_string(BB_PPO_MOMENTS_STATIC_MAX_BLOCKS) + " block partials (BB_PPO_MOMENTS_STATIC_MAX_BLOCKS); use bb_ppo_moments");
        return -1;
    }
    double *scratch = moments_static_scratch();
    if (!scratch) return -1;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    /* nothing has to start at zero: blocks 0 .. blocks-1 overwrite every entry of their partial, and the reduction reads
     * exactly those; whatever an earlier (larger) call left beyond them is never read */
    bb::bb_ppo_moments_kernel<<<blocks, bb::kMomThreads, 0, st>>>(obs, values_unnorm, returns, advantages, rows, scratch);
    BB_CUDA_OK(cudaGetLastError());
    bb::bb_ppo_moments_reduce_kernel<<<(2 * bb::kMomCols + 63) / 64, 64, 0, st>>>((const double *)scratch, blocks, sums_out);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

int bb_ppo_finish_advantages(const double *sums, double global_rows, const bb_running_norm *obs_norm, const bb_running_norm *value_norm,
                             float *obs_mean_f32, float *obs_inv_std_f32, float value_clamp, float *values_unnorm, float *returns,
                             float *advantages, int64_t rows, float *scratch, void *cuda_stream)
{
    if (!sums || global_rows < 2.0 || !obs_norm || !value_norm || !obs_mean_f32 || !obs_inv_std_f32 || !values_unnorm || !returns ||
        !advantages || rows <= 0 || !scratch) { set_error("bb_ppo_finish_advantages: bad arguments"); return -1; }
    cudaStream_t st = (cudaStream_t)cuda_stream;
    bb::NormState on{obs_norm->mean, obs_norm->var, obs_norm->count}, vn{value_norm->mean, value_norm->var, value_norm->count};
    bb::bb_ppo_finish_stats_kernel<<<1, BB_OBS_DIM, 0, st>>>(sums, global_rows, on, vn, obs_mean_f32, obs_inv_std_f32, scratch);
    BB_CUDA_OK(cudaGetLastError());
    bb::bb_ppo_finish_apply_kernel<<<(int)((rows + 255) / 256), 256, 0, st>>>(scratch, value_clamp, values_unnorm, returns, advantages, rows);
    BB_CUDA_OK(cudaGetLastError());
    return 0;
}

static int64_t randperm_sort_bytes(int64_t n)
{
    size_t temp = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, temp, (const uint32_t *)nullptr, (uint32_t *)nullptr, (const int32_t *)nullptr,
                                    (int32_t *)nullptr, (int)n, 0, 32, (cudaStream_t)0);
    return (int64_t)((temp + 255) / 256 * 256);
}

int64_t bb_randperm_workspace_bytes(int64_t n)
{
    if (n <= 0 || n > 0x7fffffff) return -1;
    const int64_t arr = (n * 4 + 255) / 256 * 256; /* keys in, keys out, identity values */
    return 3 * arr + randperm_sort_bytes(n);
}

int bb_randperm(int64_t n, uint32_t seed, uint32_t stream_id, void *workspace, int64_t workspace_bytes, int32_t *perm_out, void *cuda_stream)
{
    if (n <= 0 || n > 0x7fffffff || !workspace || !perm_out) { set_error("bb_randperm: bad arguments"); return -1; }
    const int64_t arr = (n * 4 + 255) / 256 * 256;
    const int64_t sortBytes = randperm_sort_bytes(n);
    if (workspace_bytes < 3 * arr + sortBytes || ((uintptr_t)workspace & 255) != 0) { set_error("bb_randperm: workspace too small or misaligned"); return -1; }
    char *w = (char *)workspace;
    uint32_t *keysIn = (uint32_t *)w, *keysOut = (uint32_t *)(w + arr);
    int32_t *valsIn = (int32_t *)(w + 2 * arr);
    void *temp = w + 3 * arr;
    cudaStream_t st = (cudaStream_t)cuda_stream;
    bb::bb_randperm_keys_kernel<<<(int)((n + 255) / 256), 256, 0, st>>>(keysIn, valsIn, n, seed, stream_id);
    BB_CUDA_OK(cudaGetLastError());
    size_t tb = (size_t)sortBytes;
    BB_CUDA_OK(cub::DeviceRadixSort::SortPairs(temp, tb, keysIn, keysOut, valsIn, perm_out, (int)n, 0, 32, st));
    return 0;
}

} // extern "C"
