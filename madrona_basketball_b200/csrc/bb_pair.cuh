/*
 * bb_pair.cuh — one world on a lane PAIR (sub-warp of 2): lane 2k owns agent 0,
 * lane 2k+1 owns agent 1 of the same world.
 *
 * Each lane keeps its own agent in registers; ball, GameState, WorldClock and the
 * RNG counter are REPLICATED in both lanes and advanced identically, so world
 * state never needs a broadcast.  What one agent needs of the other (position,
 * orientation, possession flag, a few predicates) crosses with __shfl_xor(.., 1).
 * Systems run in the task-graph order of /root/reference/src/game.cpp:1463-1526;
 * where the reference's agent-0-then-agent-1 order is observable (shootSystem:
 * RNG draw order, ball hand-off) the pair takes two "turns" in that order.
 *
 * This fast path covers what the reference actually runs: 1-v-1 worlds
 * (GameState.isOneOnOne == 1), no inbound in progress, and the FOR-TAG action
 * mask (game.cpp:525-528).  Any other world (the 5-v-5 / inbound / grab / pass
 * code that is dead at the reference's constants but still part of its source)
 * sends its whole WARP through slow_step(): both lanes gather the partner's agent
 * and run the serial thread-per-world restatement of bb_world.cuh redundantly,
 * then keep their own agent.  Same results, one code path to verify.
 */
#pragma once
#include "bb_world.cuh"

namespace bb {

constexpr unsigned kFullMask = 0xffffffffu;

struct Lane {
    AgentS A;
    BallS B;
    GameS G;
    int32_t resetNow;
    uint32_t rng;
    int32_t idBase;
    int32_t me;      /* 0 / 1: agent index == team index == defended hoop */
    bool didReset;
    /* derived values of the lane's agent, valid from the end of systems 1-17 on (set by pair_step_fast, or by
     * finish_derived after the slow path): what collision, defence, observations and reward would each recompute */
    Vec3 facing;     /* q.rotateVec({0,1,0}) */
    float vel2, speed; /* length2 / length of the velocity */
    bool qDirty;     /* the orientation changed after `facing` was taken (shot, reset) */
    /* Which state rows this step actually changed (store_lane skips the others: most rows are untouched on most steps,
     * and an unchanged row costs neither a store slot nor HBM write-back).  A bit is set at every site that assigns the
     * row; rows with a single cheap field are compared against the value loaded instead (the *Old copies). */
    uint32_t dirty;
    float cooldownOld, curShotOld;
    int32_t pointsWorthOld, resetNowOld;
    uint32_t rngOld;
};
enum : uint32_t {
    kDirtyQ = 1u << 0,       /* Orientation */
    kDirtyPoss = 1u << 1,    /* InPossession.hasBall / ballEntityID */
    kDirtyInb = 1u << 2,     /* Inbounding */
    kDirtyStats = 1u << 3,   /* Stats */
    kDirtyTarget = 1u << 4,  /* Attributes.currentTargetPosition */
    kDirtyBall = 1u << 5,    /* ball Velocity, BallPhysics, Grabbed */
    kDirtyGame = 1u << 6,    /* GameState fields other than the two clocks */
    kDirtyAttr = 1u << 7,    /* Attributes[0..4] (written by resetWorld only) */
    kDirtyAll = 0xffu
};

BBD void finish_derived(Lane &L)
{
    L.facing = rotate_vec(L.A.q, v3(0.f, 1.f, 0.f));
    L.vel2 = len2(L.A.vel);
    L.speed = f_sqrt(L.vel2);
    L.qDirty = false;
}

/* Partner exchange.  Always warp-wide: the fast / slow choice is made per WARP
 * (the slow path is valid for every world), so all 32 lanes reach every exchange. */
BBD float xorf(float v) { return __shfl_xor_sync(kFullMask, v, 1); }
BBD int32_t xori(int32_t v) { return __shfl_xor_sync(kFullMask, v, 1); }
BBD Vec3 xorv(Vec3 v) { return v3(xorf(v.x), xorf(v.y), xorf(v.z)); }
BBD Quat xorq(Quat q) { Quat r; r.w = xorf(q.w); r.x = xorf(q.x); r.y = xorf(q.y); r.z = xorf(q.z); return r; }

BBD int32_t my_id(const Lane &L) { return L.idBase + 3 + L.me; }
BBD Vec3 sel(bool c, Vec3 a, Vec3 b) { return v3(c ? a.x : b.x, c ? a.y : b.y, c ? a.z : b.z); }
BBD Quat selq(bool c, Quat a, Quat b) { Quat r; r.w = c ? a.w : b.w; r.x = c ? a.x : b.x; r.y = c ? a.y : b.y; r.z = c ? a.z : b.z; return r; }

/* RNG::sampleUniform + helper.cpp:8-11, on the replicated counter */
BBD float pair_uniform(Lane &L, const SimConst &c, float lo, float hi)
{
    float u = rng_uniform01(c.key0, c.key1, L.rng++);
    return lo + (hi - lo) * u;
}

/* ------------------------------------------------------------------ slow path */

BBD AgentS xor_agent(const AgentS &A)
{
    AgentS o;
    o.pos = xorv(A.pos); o.vel = xorv(A.vel); o.q = xorq(A.q);
    o.hasBall = xori(A.hasBall); o.ballId = xori(A.ballId); o.pointsWorth = xori(A.pointsWorth);
    o.imInb = xori(A.imInb); o.allowedToMove = xori(A.allowedToMove);
    o.cooldown = xorf(A.cooldown);
    o.maxSpeed = xorf(A.maxSpeed); o.quickness = xorf(A.quickness); o.shooting = xorf(A.shooting);
    o.ftPct = xorf(A.ftPct); o.reaction = xorf(A.reaction);
    o.target = xorv(A.target); o.curShot = xorf(A.curShot);
    o.points = xorf(A.points); o.fouls = xorf(A.fouls);
    o.curStep = (uint32_t)xori((int32_t)A.curStep);
    o.move = xori(A.move); o.moveAngle = xori(A.moveAngle); o.rotate = xori(A.rotate);
    o.grab = xori(A.grab); o.pass = xori(A.pass); o.shoot = xori(A.shoot);
    o.canMove = xori(A.canMove); o.canGrab = xori(A.canGrab); o.canPass = xori(A.canPass); o.canShoot = xori(A.canShoot);
    o.reward = xorf(A.reward); o.done = xorf(A.done);
    o.reset = xori(A.reset);
    return o;
}

/* Systems 1-17 through the serial restatement, for a whole warp (cold path, kept out
 * of line so that it costs the fast path no registers and no instruction-cache lines):
 * every lane gathers its partner's agent, runs the thread-per-world code on the
 * complete world, and keeps its own agent.  Both lanes of a pair compute the same
 * world, so the replicated ball / GameState / RNG stay identical. */
__device__ __noinline__ void slow_step(Lane *Lp, const SimConst *c, bool tagMask, bool sync)
{
    if (sync) { /* the barriers the fast path of the other warps passes (kPhases) */
        __syncthreads(); __syncthreads(); __syncthreads(); __syncthreads(); __syncthreads(); __syncthreads();
    }
#ifndef BB_NO_SLOW
    Lane L = *Lp;
    AgentS other = xor_agent(L.A);
    World W;
    if (L.me == 0) { W.a[0] = L.A; W.a[1] = other; }
    else           { W.a[0] = other; W.a[1] = L.A; }
    W.b = L.B; W.g = L.G; W.resetNow = L.resetNow; W.rng = L.rng; W.idBase = L.idBase; W.didReset = false;
    if (tagMask) step_world_logic<true>(W, *c);
    else step_world_logic<false>(W, *c);
    L.A = (L.me == 0) ? W.a[0] : W.a[1];
    L.B = W.b; L.G = W.g; L.resetNow = W.resetNow; L.rng = W.rng; L.didReset = W.didReset;
    *Lp = L;
#endif
}

/* ------------------------------------------------------------------ fast path */

/* shootSystem (game.cpp:273-407) for actor k, executed identically by both lanes.
 * apos/dpos: actor / other agent position; amove, avlen: actor's Action.move, |vel|. */
BBD void pair_shoot_turn(Lane &L, const SimConst &c, int k, Vec3 apos, Vec3 dpos, int32_t amove, float avlen)
{
    const bool acting = (L.me == k);
    const Vec3 hoop = v3(k ? c.hoopX[0] : c.hoopX[1], c.hoopY, 0.f);
    const float radius = kHoopZone;
    Vec3 ideal = hoop - apos;
    float intended = bb_atan2f(ideal.x, ideal.y);
    float dist_hoop = len(ideal);
    float dist_sd = kDistDev * dist_hoop;
    float dev_dist = pair_uniform(L, c, -dist_sd, dist_sd);
    float dev_def = 0.f;
    float nearest = len(apos - dpos);
    if (nearest < 2.0f) {
        float def_sd = f_div(kDefDev, nearest + 0.1f);
        dev_def = pair_uniform(L, c, -def_sd, def_sd);
    }
    float dev_vel = 0.f;
    if (amove > 0) {
        float vel_sd = kVelDev * avlen;
        dev_vel = pair_uniform(L, c, -vel_sd, vel_sd);
    }
    float total = dev_dist + dev_def + dev_vel;
    float dir = intended + total;
    Vec3 shot = v3(bb_sinf(dir), bb_cosf(dir), 0.f);
    bool going_in = false;
    float along = dot(ideal, shot);
    if (!(along < 0.f)) {
        float closest_sq = len2(ideal) - along * along;
        going_in = closest_sq <= radius * radius;
    }
    Quat rot = rotation_between(v3(0.f, 1.f, 0.f), shot, c);
    if (acting) { L.A.q = rot; L.qDirty = true; L.dirty |= kDirtyQ; }
    BallS &B = L.B;
    const int32_t aid = L.idBase + 3 + k;
    if (B.holder == aid) {
        int32_t spv = shot_point_value(apos, hoop);
        if (going_in) { B.goingIn = 1; L.G.scoredBaskets += 1.f; }
        else if (acting) L.A.reward -= 1.f;
        B.isGrabbed = 0; B.holder = kPlaceholder;
        L.dirty |= kDirtyBall | kDirtyGame;
        if (acting) { L.A.hasBall = 0; L.A.ballId = kPlaceholder; L.A.imInb = 0; L.dirty |= kDirtyPoss | kDirtyInb; }
        B.vel = shot * .1f;
        B.inFlight = 1;
        B.sbAgent = aid; B.sbTeam = k; B.spv = spv;
        B.ltAgent = aid; B.ltTeam = k;
    }
}

/* updateCurrentShotPercentage (game.cpp:758-809) for the lane's agent */
BBD void pair_shot_pct(Lane &L, Vec3 opos, const SimConst &c)
{
    AgentS &A = L.A;
    if (A.hasBall == 0) { A.curShot = 0.f; return; }
    const Vec3 hoop = v3(L.me ? c.hoopX[0] : c.hoopX[1], c.hoopY, 0.f);
    float dist_hoop = len(hoop - A.pos);
    float nearest = len(A.pos - opos);
    float dist_sd = kDistDev * dist_hoop;
    float def_sd = f_div(kDefDev, nearest) + .0001f;
    float vel_sd = kVelDev * len(A.vel);
    float final_sd = f_sqrt((f_div(dist_sd * dist_sd, 3.f) + f_div(def_sd * def_sd, 3.f)) + f_div(vel_sd * vel_sd, 3.f));
    float max_ang = bb_atanf(f_div(kHoopZone, dist_hoop));
    float z = f_div(max_ang, final_sd);
    A.curShot = bb_erff(f_div(z, f_sqrt(2.f)));
}

/* resetWorld (gen.cpp:216-316) + setupAgentPositions (helper.cpp:108-160), 1-v-1 branch */
BBD void pair_reset_world(Lane &L, const SimConst &c)
{
    GameS &g = L.G;
    /* gameClock <= 0 && isOneOnOne == 0 is false in a 1-v-1 world: full GameState reset, ids / mode kept */
    g.inbInProgress = 0; g.liveBall = 1; g.period = 1.0f; g.teamInPoss = 0.0f;
    g.team0Score = 0.f; g.team1Score = 0.f; g.gameClock = kTimePerPeriod; g.shotClock = 24.0f;
    g.scoredBaskets = 0.f; g.oobCount = 0.f; g.inboundClock = 0.f;
    AgentS &A = L.A;
    const int i = L.me;
    A.move = A.moveAngle = A.rotate = A.grab = A.pass = A.shoot = 0;
    A.canMove = A.canGrab = A.canPass = A.canShoot = 0;
    A.reset = 0; A.imInb = 0; A.allowedToMove = 1;
    A.done = 1.f; A.curStep = 0;
    A.q = i ? c.spawnQ[1] : c.spawnQ[0];
    A.cooldown = 0.f; A.points = 0.f; A.fouls = 0.f;
    A.vel = v3(0.f, 0.f, 0.f);
    L.vel2 = 0.f; L.speed = 0.f; L.qDirty = true; /* sqrt(0) == 0 */
    L.dirty = kDirtyAll;
    /* the three draws of helper.cpp:121,122,132, replicated */
    Vec3 base = v3(c.startX + (0.f * 2.f), c.startY, 0.f);
    float xd = pair_uniform(L, c, -kStartDev, kStartDev);
    float yd = pair_uniform(L, c, -kStartDev, kStartDev);
    Vec3 p0 = base + v3(xd, yd, 0.f);
    p0.x = clampf(p0.x, 0.f, c.width);
    p0.y = clampf(p0.y, 0.f, c.height);
    float ang = pair_uniform(L, c, 0.f, 2.f * kPi);
    Vec3 off = v3(kDefenderSpawnRadius * bb_cosf(ang), kDefenderSpawnRadius * bb_sinf(ang), 0.f);
    Vec3 p1 = p0 + off;
    p1.x = clampf(p1.x, 0.f, c.width);
    p1.y = clampf(p1.y, 0.f, c.height);
    if (i == 0) { A.pos = p0; A.hasBall = 1; A.ballId = L.idBase + 2; A.pointsWorth = 2; }
    else        { A.pos = p1; A.hasBall = 0; A.ballId = kPlaceholder; A.pointsWorth = 2; }
    A.maxSpeed = kDefaultSpeed - (float)i * kDefenderSlowdown;
    A.quickness = 1.f; A.shooting = 0.f; A.ftPct = 0.f;
    A.reaction = (float)i * kDefenderReaction;
    A.target = A.pos;
    A.curShot = 0.f;
    BallS &B = L.B;
    B.pos = p0;
    B.inFlight = 0; B.ltAgent = kPlaceholder; B.ltTeam = kPlaceholder;
    B.sbAgent = kPlaceholder; B.sbTeam = kPlaceholder; B.spv = 2; B.goingIn = 0;
    B.vel = v3(0.f, 0.f, 0.f);
    B.isGrabbed = 1; B.holder = L.idBase + 3;
    L.didReset = true;
}

/* agentCollisionSystem (game.cpp:537-648) for the pair (agent 0, agent 1), split over the two lanes: each lane builds
 * its OWN agent's rectangle and tests the two separating axes that rectangle contributes (the reference's axis order is
 * a.right, a.forward, b.right, b.forward: lane 0 holds axes 0-1, lane 1 axes 2-3), the partner's four vertices and the
 * partial verdicts cross by shuffle.  The reference stops at the first separating axis and keeps the FIRST axis of
 * least overlap (strict <): lane 0's candidate wins unless lane 1's overlap is strictly smaller. */
BBD void pair_collision(Lane &L, Vec3 opos)
{
    const float hw = kShoulderWidth / 2.0f, hd = kAgentDepth / 2.0f;
    const int me = L.me;
    const Vec3 c = L.A.pos;
    const Vec3 f = L.facing;
    const Vec3 r = v3(f.y, -f.x, 0.f);
    /* Far apart: nothing to test.  A rectangle's corners lie within R = sqrt(hw^2 + hd^2) * |f| of its centre (hwv and hdv
     * are orthogonal and |r| <= |f|; |f| != 1 is covered, the orientation is never renormalised), so the two rectangles
     * cannot touch when the centres are further apart than R_a + R_b.  The warp skips the SAT only when EVERY pair is beyond
     * sqrt(2) (R_a + R_b) (tested without a root: d^2 > 4 (R_a^2 + R_b^2) >= 2 (R_a + R_b)^2): then one of the four axes
     * separates the projections by more than 0.1 (R_a + R_b), five orders of magnitude above float rounding, so the
     * reference's SAT finds `collision = false` too and writes nothing (game.cpp:598-617).  NaNs compare false -> SAT. */
    {
        const float R2 = (hw * hw + hd * hd) * len2(f);
        const float oR2 = xorf(R2);
        const bool apart = len2(opos - c) > 4.f * (R2 + oR2);
        if (__all_sync(kFullMask, apart)) return;
    }
    const Vec3 hwv = r * hw, hdv = f * hd;
    const Vec3 mv[4] = {c - hdv + hwv, c - hdv - hwv, c + hdv - hwv, c + hdv + hwv};
    const Vec3 ov[4] = {xorv(mv[0]), xorv(mv[1]), xorv(mv[2]), xorv(mv[3])};
    /* normalize(r), normalize(f): length2(r) = (f.y*f.y + f.x*f.x) + 0 and length2(f) = (f.x*f.x + f.y*f.y) + f.z*f.z are the
     * same float whenever f.z is +-0 (IEEE addition commutes) -- always, for the z-rotations the game produces -- so the
     * two axes share one 1/sqrt; a warp that holds any other orientation computes r's own. */
    const float invf = f_div(1.f, f_sqrt(len2(f)));
    float invr = invf;
    if (__any_sync(kFullMask, f.z != 0.f)) invr = f_div(1.f, f_sqrt(len2(r)));
    const Vec3 axes[2] = {r * invr, f * invf};
    bool coll = true;
    float minOv = FLT_MAX;
    Vec3 mtv = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 2; j++) {
        const Vec3 ax = axes[j];
        float amin = dot(mv[0], ax), amax = amin, bmin = dot(ov[0], ax), bmax = bmin;
#pragma unroll
        for (int k = 1; k < 4; k++) {
            const float pa = dot(mv[k], ax), pb = dot(ov[k], ax);
            if (pa < amin) amin = pa;
            if (pa > amax) amax = pa;
            if (pb < bmin) bmin = pb;
            if (pb > bmax) bmax = pb;
        }
        if (!(amax > bmin && bmax > amin)) {
            coll = false;
        } else {
            const float overlap = fminf(amax, bmax) - fmaxf(amin, bmin);
            if (coll && overlap < minOv) { minOv = overlap; mtv = ax; }
        }
    }
    const bool ocoll = xori(coll ? 1 : 0) != 0;
    const float oMin = xorf(minOv);
    const Vec3 oMtv = xorv(mtv);
    if (coll && ocoll) {
        const float m0 = me ? oMin : minOv, m1 = me ? minOv : oMin;
        const Vec3 t0 = sel(me != 0, oMtv, mtv), t1 = sel(me != 0, mtv, oMtv);
        const bool second = m1 < m0;
        const float min_overlap = second ? m1 : m0;
        Vec3 corr = second ? t1 : t0;
        if (L.G.teamInPoss == (float)0) {
            if (me == 0) L.A.reward -= 10.f; else L.A.reward += 10.f;
            L.resetNow = 1;
        }
        const Vec3 ca = sel(me == 0, c, opos), cb = sel(me == 0, opos, c);
        if (dot(cb - ca, corr) < 0.f) corr = -corr;
        const Vec3 shift = (corr * min_overlap) * 0.5f;
        if (me == 0) L.A.pos = L.A.pos - shift; else L.A.pos = L.A.pos + shift;
    }
}

/* hardCodeDefenseSystem (game.cpp:651-755) for the lane's agent */
BBD void pair_defense(Lane &L, Vec3 opos, int32_t ohas, const SimConst &c)
{
    AgentS &D = L.A;
    const int i = L.me;
    if (L.G.teamInPoss == (float)i) { D.move = 0; return; }
    D.grab = 1;
    /* first agent (index order) holding a ball */
    const int32_t has0 = i ? ohas : D.hasBall, has1 = i ? D.hasBall : ohas;
    const Vec3 pos0 = sel(i != 0, opos, D.pos), pos1 = sel(i != 0, D.pos, opos);
    if (has0 != 1 && has1 != 1) { D.move = 0; return; }
    Vec3 op = (has0 == 1) ? pos0 : pos1;
    Vec3 hdv = v3(i ? c.hoopX[1] : c.hoopX[0], c.hoopY, 0.f) - op;
    Vec3 guard = op;
    if (len2(hdv) > 1e-6f) guard = op + normalize(hdv) * kGuardDist;
    float f = D.reaction * kDt;
    D.target = D.target + (guard - D.target) * f;
    L.dirty |= kDirtyTarget;
    Vec3 mv = D.target - D.pos;
    if (len2(mv) < 0.01f) { D.move = 0; return; }
    Vec3 dir = normalize(mv);
    const float k = c.invSqrt2;
    const float dxs[8] = {0.f, 1.f * k, 1.f, 1.f * k, 0.f, -1.f * k, -1.f, -1.f * k};
    const float dys[8] = {-1.f, -1.f * k, 0.f, 1.f * k, 1.f, 1.f * k, 0.f, -1.f * k};
    float best = -2.f;
    int32_t best_k = 0;
#pragma unroll
    for (int m = 0; m < 8; m++) {
        float dd = dir.x * dxs[m] + dir.y * dys[m] + dir.z * 0.f;
        if (dd > best) { best = dd; best_k = m; }
    }
    D.move = 1;
    D.moveAngle = best_k;
    Vec3 ov = L.facing;
    float ang = bb_acosf(clampf(dot(ov, dir), -1.f, 1.f));
    if (ang > kPi / 8.f) {
        float cr = ov.x * mv.y - ov.y * mv.x;
        if (cr < 0.f) D.rotate = -1;
        else if (cr > 0.f) D.rotate = 1;
        else D.rotate = 0;
    } else {
        D.rotate = 0;
    }
}

/* rewardSystem (game.cpp:811-870) for the lane's agent */
/* dist_other: |other.pos - pos| (game.cpp:825); the caller passes the agent-to-agent distance the observation rows
 * already hold -- the squared differences are the same floats in either direction */
BBD void pair_reward(Lane &L, float dist_other)
{
    AgentS &A = L.A;
    if ((float)L.me == L.G.teamInPoss) {
        if (L.G.gameClock > 5.f) {
            const int32_t id = my_id(L);
            if (L.B.sbAgent == id && L.B.goingIn == 1) A.reward += (float)L.B.spv;
            else if (L.B.sbAgent == id && L.B.goingIn == 0 && L.B.inFlight == 1) A.reward -= 1.f;
            A.reward += A.curShot;
        }
    } else {
        A.reward -= 1.f;
        A.reward += bb_expf(-0.4f * dist_other);
    }
}

/* Phase alignment.  The step is ~3000 straight-line instructions executed once per warp,
 * more than the SM's 32 KB instruction cache holds; warps of one block that drift apart
 * each stream the whole kernel through that cache.  With kSync the block re-aligns at
 * kPhases points, so its warps fetch the same ~8 KB of code at the same time.  Every
 * warp of the block must pass the same number of barriers on either path. */
constexpr int kPhases = 6;
template <bool kSync> BBD void phase_sync() { if (kSync) __syncthreads(); }

/* Systems 1..17 (game.cpp:1467-1516) for a 1-v-1, tag-mask, no-inbound world. */
template <bool kSync>
BBD void pair_step_fast(Lane &L, const SimConst &c)
{
    AgentS &A = L.A;
    BallS &B = L.B;
    GameS &g = L.G;
    const int me = L.me;
    sys_tick(A);                       /* 1 */
    sys_action_mask<true>(A, g);       /* 2 */
    if (A.rotate != 0) L.dirty |= kDirtyQ;
    sys_move_agent(A, c, L.facing);    /* 3 */
    L.qDirty = false;
    L.vel2 = len2(A.vel);
    L.speed = f_sqrt(L.vel2);
    /* 4 grab, 5 pass: can_grab == can_pass == 0 under the tag mask -> both return at once */
    phase_sync<kSync>();
    Vec3 opos = xorv(A.pos);
    {                                  /* 6 shoot: agent 0's turn, then agent 1's */
        const int32_t wants = (A.canShoot != 0 && A.shoot != 0) ? 1 : 0;
        const float vlen = L.speed;
        const int32_t ovmove = xori(A.move), owants = xori(wants);
        const float ovlen = xorf(vlen);
#pragma unroll 1
        for (int k = 0; k < 2; k++) {
            const bool mine = (me == k);
            if (mine ? wants : owants)
                pair_shoot_turn(L, c, k, mine ? A.pos : opos, mine ? opos : A.pos, mine ? A.move : ovmove,
                                mine ? vlen : ovlen);
        }
    }
    phase_sync<kSync>();
    {                                  /* 7 moveBall (game.cpp:82-125) */
        const int32_t holds = (A.hasBall == 1 && B.isGrabbed == 1 && B.holder == my_id(L)) ? 1 : 0;
        const int32_t oholds = xori(holds);
        if (holds) B.pos = A.pos; else if (oholds) B.pos = opos;
        if (!(len2(B.vel) == 0.f || B.isGrabbed == 1)) { /* sqrt(x) == 0 <=> x == 0 */
            float nx = clampf(B.pos.x + B.vel.x, 0.f, c.width);
            float ny = clampf(B.pos.y + B.vel.y, 0.f, c.height);
            float nz = B.pos.z + B.vel.z;
            B.pos = v3(nx, ny, nz);
        }
    }
    pair_shot_pct(L, opos, c);         /* 8 */
    phase_sync<kSync>();
#pragma unroll
    for (int h = 0; h < 2; h++) {      /* 9 score (game.cpp:873-953), 1-v-1 branch */
        float dx = B.pos.x - c.hoopX[h], dy = B.pos.y - c.hoopY;
        if (dx * dx + dy * dy <= kLen2LeHoopZone && (float)B.inFlight == 1.f) { /* sqrt(..) <= kHoopZone */
            int32_t points = B.spv;
            if (my_id(L) == B.sbAgent) A.points += (float)((me == h) ? -B.spv : B.spv);
            if (L.idBase + h == g.team0Hoop) g.team1Score += (float)points;
            else g.team0Score += (float)points;
            g.scoredBaskets += 1.f;
            L.dirty |= kDirtyStats | kDirtyGame | kDirtyBall;
            B.inFlight = 0;
            B.vel = v3(0.f, 0.f, 0.f);
            B.sbAgent = kPlaceholder; B.sbTeam = kPlaceholder; B.spv = 2; B.goingIn = 0;
            L.resetNow = 1;
        }
    }
    const int off = ((float)1 == g.teamInPoss) ? 1 : 0; /* game.cpp:1012-1018, 1072-1078 */
    {                                  /* 10 outOfBounds (game.cpp:1055-1113), 1-v-1 branch */
        bool out = B.pos.x < kCourtMinX || B.pos.x > kCourtMaxX || B.pos.y < kCourtMinY || B.pos.y > kCourtMaxY;
        if (out) { /* inboundingInProgress == 0 on this path */
            if (me == off) A.reward -= 100.f;
            L.resetNow = 1;
        }
    }
    {                                  /* 11 updateLastTouch (game.cpp:1034-1051): the later agent wins */
        const int32_t touch = (len2(B.pos - A.pos) <= kLen2LeAgentSize) ? 1 : 0; /* length <= kAgentSize */
        const int32_t otouch = xori(touch);
        const int32_t t0 = me ? otouch : touch, t1 = me ? touch : otouch;
        const int32_t ltA = B.ltAgent, ltT = B.ltTeam;
        if (t1) { B.ltAgent = L.idBase + 4; B.ltTeam = 1; }
        else if (t0) { B.ltAgent = L.idBase + 3; B.ltTeam = 0; }
        if (B.ltAgent != ltA || B.ltTeam != ltT) L.dirty |= kDirtyBall;
    }
    {                                  /* 12 clock (game.cpp:992-1030) */
        bool live = (float)g.liveBall > 0.5f;
        if (live && g.gameClock > 0.f) { g.gameClock -= kDt; g.shotClock -= kDt; }
        if (g.gameClock <= 0.f && live) {
            if (me == off) A.reward += 10.f;
            L.resetNow = 1;
        }
        if (g.shotClock < 0.f) g.shotClock = 0.f;
    }
    /* 13 inboundViolation: needs inboundingInProgress > 0.5 -> nothing on this path */
    if (L.resetNow != 0) { pair_reset_world(L, c); L.resetNow = 0; } /* 14 */
    if (__any_sync(kFullMask, L.qDirty)) L.facing = rotate_vec(A.q, v3(0.f, 1.f, 0.f)); /* the orientation is final here */
    A.pointsWorth = shot_point_value(A.pos, v3(me ? c.hoopX[0] : c.hoopX[1], c.hoopY, 0.f)); /* 15 */
    phase_sync<kSync>();
    opos = xorv(A.pos);
    pair_collision(L, opos);           /* 16 collision */
    phase_sync<kSync>();
    opos = xorv(A.pos);
    pair_defense(L, opos, xori(A.hasBall), c); /* 17 */
    phase_sync<kSync>();
}

} // namespace bb
