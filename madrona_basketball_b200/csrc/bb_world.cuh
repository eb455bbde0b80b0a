/*
 * bb_world.cuh — one world's step, register resident, for sm_100a.
 *
 * Device restatement of the 19 systems the reference registers in
 * setupGameStepTasks (/root/reference/src/game.cpp:1463-1526) plus world
 * generation / reset (src/gen.cpp) and helpers (src/helper.cpp).  One thread owns
 * one world: two agents, one ball, two hoops, GameState and WorldClock live in
 * registers for the whole step, systems run in task-graph order with agent 0
 * before agent 1 and hoop 0 before hoop 1 (the Madrona CPU executor's order).
 *
 * Structural facts used (all follow from gen.cpp and are re-established by every
 * resetWorld): agent i is on team i and defends hoop i (gen.cpp:202-205,282-283);
 * entity ids of world w are 5w + {hoop0, hoop1, ball, agent0, agent1}
 * (creation order gen.cpp:101,131,167,187); hoop positions / scoring zones never
 * change after generateWorld (gen.cpp:96-156).
 *
 * Floating point: compiled with -fmad=false; every operation below is written
 * in the reference's evaluation order so results match the CPU oracle bit for
 * bit.  Transcendentals come from include/bb_math.h.
 */
#pragma once
#include <cstdint>
#include <cfloat>
#include "../../include/bb_math.h"

namespace bb {

/* ------------------------------ constants.hpp ------------------------------ */
constexpr int32_t kPlaceholder = INT32_MAX;            /* constants.hpp:8 */
constexpr float kPi = 3.14159265358979323846f;         /* madrona::math::pi */
constexpr float kSimHz = 62.0f;                        /* :11 */
constexpr float kDt = 1.0f / kSimHz;                   /* :12 */
constexpr float kTimePerPeriod = 10.f;                 /* :13 */
constexpr float kPixelsPerMeter = 110.f;               /* :17 */
constexpr float kHoopZone = 0.1f;                      /* :24 */
constexpr int32_t kOneOnOne = 1;                       /* :27 */
constexpr float kAgentSize = 0.2f;                     /* :39 */
constexpr float kShoulderWidth = 0.4290;               /* :40 */
constexpr float kAgentDepth = .1;                      /* :41 */
constexpr float kGuardDist = .2f;                      /* :45 */
constexpr float kStartDev = 5.f;                       /* :46 */
constexpr float kDefaultSpeed = 3.f;                   /* :47 */
constexpr float kDefenderSlowdown = 0.2f;              /* :48 */
constexpr float kDefenderReaction = 10.f;              /* :49 */
constexpr float kDefenderSpawnRadius = 8.f;            /* :50 */
constexpr float kAngleBetweenDirs = kPi / 4.0f;        /* :53 */
constexpr float kBallSlowdown = 0.9f;                  /* :55 */
constexpr float kDistDev = .008f, kDefDev = .002f, kVelDev = .001f; /* :59-61 */
constexpr float kCourtLen = 28.65f, kCourtWid = 15.24f;             /* :67-68 */
constexpr float kMargin = 1.1f;                                     /* :71 */
constexpr float kWorldW = kCourtLen * kMargin, kWorldH = kCourtWid * kMargin; /* :72-73 */
constexpr float kCourtMinX = (kWorldW - kCourtLen) / 2.0f;          /* :76 */
constexpr float kCourtMaxX = kCourtMinX + kCourtLen;                /* :77 */
constexpr float kCourtMinY = (kWorldH - kCourtWid) / 2.0f;          /* :78 */
constexpr float kCourtMaxY = kCourtMinY + kCourtWid;                /* :79 */
constexpr float kHoopFromBaseline = 1.575f;                         /* :84 */
constexpr float kArcRadius = 7.24f;                                 /* :91 */
constexpr float kCorner3Side = 0.91f, kCorner3Len = 4.27f;          /* :92-93 */

/* ------------------------------ small math -------------------------------- */
struct Vec3 { float x, y, z; };
struct Quat { float w, x, y, z; };

#define BBD __device__ __forceinline__

/* IEEE sqrt / divide of the reference's float code.  -DBB_SUB_DIVSQRT=1 turns the ~15-instruction
 * inline sequences into two shared subroutines (instruction-cache footprint vs call latency). */
#ifndef BB_SUB_DIVSQRT
#define BB_SUB_DIVSQRT 0
#endif
#if BB_SUB_DIVSQRT
__device__ __noinline__ float f_sqrt(float x) { return bb_sqrtf(x); }
__device__ __noinline__ float f_div(float a, float b) { return bb_divf(a, b); }
#else
BBD float f_sqrt(float x) { return bb_sqrtf(x); }
BBD float f_div(float a, float b) { return bb_divf(a, b); }
#endif

BBD Vec3 v3(float x, float y, float z) { Vec3 r; r.x = x; r.y = y; r.z = z; return r; }
BBD Vec3 operator+(Vec3 a, Vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
BBD Vec3 operator-(Vec3 a, Vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
BBD Vec3 operator-(Vec3 a) { return v3(-a.x, -a.y, -a.z); }
BBD Vec3 operator*(Vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
BBD float dot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
BBD Vec3 cross(Vec3 a, Vec3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
BBD float len2(Vec3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
BBD float len(Vec3 a) { return f_sqrt(len2(a)); }
/* Vector3::normalize = *this * (1 / sqrt(length2)) */
BBD Vec3 normalize(Vec3 a) { float inv = f_div(1.f, f_sqrt(len2(a))); return a * inv; }
BBD float clampf(float v, float lo, float hi) { return fminf(fmaxf(v, lo), hi); }

/* Distance tests without the square root.  IEEE sqrt is monotone, so for a constant c the set
 * { s >= 0 : sqrt_rn(s) <= c } is an interval [0, T] and { s : sqrt_rn(s) >= c } is [T', inf): comparing the squared
 * length with T / T' gives the SAME verdict as the reference's `length() <= c` for every float (NaN included: both sides
 * compare false).  T = largest float with sqrtf(T) <= c, T' = smallest with sqrtf(T') >= c; bb_create re-derives them
 * with the host's sqrtf and refuses to start on a mismatch (check_sqrt_thresholds), tests/test_abi_cpu.py sweeps 2e5
 * neighbouring floats on each side. */
constexpr float kLen2LeHoopZone = 0x1.47ae16p-7f;   /* length <= kHoopZone (0.1f)   */
constexpr float kLen2LeAgentSize = 0x1.47ae16p-5f;  /* length <= kAgentSize (0.2f)  */
constexpr float kLen2GeArc = 0x1.a3573cp+5f;        /* length >= kArcRadius (7.24f) */

BBD Quat qmul(Quat a, Quat b)
{
    Quat r;
    r.w = (a.w * b.w - a.x * b.x) - (a.y * b.y + a.z * b.z);
    r.x = (a.w * b.x + a.x * b.w) + (a.y * b.z - a.z * b.y);
    r.y = (a.w * b.y + a.y * b.w) + (a.z * b.x - a.x * b.z);
    r.z = (a.w * b.z + a.z * b.w) + (a.x * b.y - a.y * b.x);
    return r;
}
BBD Vec3 rotate_vec(Quat q, Vec3 v)
{
    Vec3 pure = v3(q.x, q.y, q.z);
    Vec3 pxv = cross(pure, v);
    Vec3 ppxv = cross(pure, pxv);
    return v + ((pxv * q.w) + ppxv) * 2.f;
}
BBD Quat angle_axis(float angle, Vec3 n)
{
    float c = bb_cosf(angle / 2.f), s = bb_sinf(angle / 2.f);
    Quat r; r.w = c; r.x = n.x * s; r.y = n.y * s; r.z = n.z * s; return r;
}
BBD Quat quat_id() { Quat r; r.w = 1.f; r.x = 0.f; r.y = 0.f; r.z = 0.f; return r; }

/* ------------------------------ per-engine constants ---------------------- */
struct SimConst {
    float startX, startY, width, height;     /* GridState, bindings.cpp:28-51 */
    float hoopX[2], hoopY;                   /* gen.cpp:96-141 */
    Quat turnPos, turnNeg;                   /* game.cpp:423-424: angleAxis(+-6deg, z) */
    Quat spawnQ[2];                          /* gen.cpp:196,277: angleAxis(-+pi/2, z) */
    Quat piZ;                                /* helper.cpp:30: angleAxis(pi, z) */
    float moveSin[8], moveCos[8];            /* game.cpp:432-435 for moveAngle 0..7 */
    float invSqrt2;                          /* (1,1,0).normalize() component, game.cpp:731 */
    uint32_t key0, key1;                     /* sim.cpp:89 world RNG key (same in every world) */
    uint32_t flags;
};

/* ------------------------------ world registers --------------------------- */
struct AgentS {
    Vec3 pos, vel;
    Quat q;
    int32_t hasBall, ballId, pointsWorth;            /* InPossession */
    int32_t imInb, allowedToMove;                    /* Inbounding */
    float cooldown;                                  /* GrabCooldown */
    float maxSpeed, quickness, shooting, ftPct, reaction; /* Attributes */
    Vec3 target;
    float curShot;
    float points, fouls;                             /* Stats */
    uint32_t curStep;
    int32_t move, moveAngle, rotate, grab, pass, shoot; /* Action */
    int32_t canMove, canGrab, canPass, canShoot;     /* ActionMask */
    float reward, done;
    int32_t reset;
};
struct BallS {
    Vec3 pos, vel;
    int32_t inFlight, ltAgent, ltTeam, sbAgent, sbTeam, spv, goingIn; /* BallPhysics */
    int32_t isGrabbed, holder;                                       /* Grabbed */
};
struct GameS { /* types.hpp:46-67 */
    int32_t inbInProgress, liveBall;
    float period, teamInPoss;
    int32_t team0Hoop;
    float team0Score;
    int32_t team1Hoop;
    float team1Score, gameClock, shotClock, scoredBaskets, oobCount, inboundClock;
    int32_t isOneOnOne;
};
struct World {
    AgentS a[2];
    BallS b;
    GameS g;
    int32_t resetNow;      /* WorldClock */
    uint32_t rng;          /* RNG draw counter */
    int32_t idBase;        /* 5 * world index */
    bool didReset;         /* resetWorld ran this step: Team / Reset columns need a store */
};

BBD int32_t agent_id(const World &W, int i) { return W.idBase + 3 + i; }
BBD int32_t ball_id(const World &W) { return W.idBase + 2; }
BBD int32_t hoop_id(const World &W, int h) { return W.idBase + h; }
BBD Vec3 hoop_pos(const SimConst &c, int h) { return v3(c.hoopX[h], c.hoopY, 0.f); }

/* ------------------------------ rand ---------------------- */
BBD uint32_t rotl32(uint32_t x, int r) { return __funnelshift_l(x, x, r); }

BBD void threefry2x32(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t &o0, uint32_t &o1)
{
    const uint32_t k2 = 0x1BD11BDAu ^ k0 ^ k1;
    uint32_t x0 = c0 + k0, x1 = c1 + k1;
#define BB_TF_R(r) x0 += x1; x1 = rotl32(x1, r); x1 ^= x0;
    BB_TF_R(13) BB_TF_R(15) BB_TF_R(26) BB_TF_R(6)   x0 += k1; x1 += k2 + 1u;
    BB_TF_R(17) BB_TF_R(29) BB_TF_R(16) BB_TF_R(24)  x0 += k2; x1 += k0 + 2u;
    BB_TF_R(13) BB_TF_R(15) BB_TF_R(26) BB_TF_R(6)   x0 += k0; x1 += k1 + 3u;
    BB_TF_R(17) BB_TF_R(29) BB_TF_R(16) BB_TF_R(24)  x0 += k1; x1 += k2 + 4u;
    BB_TF_R(13) BB_TF_R(15) BB_TF_R(26) BB_TF_R(6)   x0 += k2; x1 += k0 + 5u;
#undef BB_TF_R
    o0 = x0; o1 = x1;
}

/* RNG::sampleUniform for draw `count` of the world key: one shared subroutine */
__device__ __noinline__ float rng_uniform01(uint32_t k0, uint32_t k1, uint32_t count)
{
    uint32_t a, b;
    threefry2x32(k0, k1, count, 0u, a, b);
    return (float)((a ^ b) >> 8) * 0x1p-24f;
}

/* RNG::sampleUniform then helper.cpp:8-11 */
BBD float sample_uniform(World &W, const SimConst &c, float lo, float hi)
{
    float u = rng_uniform01(c.key0, c.key1, W.rng++);
    return lo + (hi - lo) * u;
}

/* ------------------------------ helper.cpp -------------------------------- */

/* helper.cpp:14-42 */
BBD Quat rotation_between(Vec3 start, Vec3 target, const SimConst &c)
{
    start = normalize(start);
    target = normalize(target);
    float d = dot(start, target);
    if (d > 0.999999f) return quat_id();
    if (d < -0.999999f) return c.piZ;
    Vec3 axis = normalize(cross(start, target));
    float ang = bb_acosf(d);
    return angle_axis(ang, axis);
}

/* helper.cpp:44-48 */
BBD Vec3 vector_to_center(Vec3 pos, const SimConst &c) { return normalize(v3(c.startX, c.startY, 0.f) - pos); }

/* helper.cpp:50-81 */
BBD int32_t shot_point_value(Vec3 shot_pos, Vec3 hoop)
{
    const float dist2 = len2(shot_pos - hoop); /* `dist >= kArcRadius` as a test on the squared length, see kLen2GeArc */
    bool corner = (shot_pos.y < kCourtMinY + kCorner3Side) || (shot_pos.y > kCourtMinY + kCourtWid - kCorner3Side);
    if (corner) {
        bool left = hoop.x < kWorldW / 2.0f;
        if (left) { if (shot_pos.x <= kCourtMinX + kCorner3Len) return 3; }
        else      { if (shot_pos.x >= kCourtMinX + kCourtLen - kCorner3Len) return 3; }
    }
    if (dist2 >= kLen2GeArc) return 3;
    return 2;
}

/* helper.cpp:108-160 */
BBD void setup_agent_positions(World &W, const SimConst &c, int32_t &offensive_id, Vec3 &pos_for_ball)
{
#pragma unroll
    for (int i = 0; i < 2; i++) {
        AgentS &A = W.a[i];
        if ((float)W.g.isOneOnOne == 1.f) {
            if (i == 0) {
                Vec3 base = v3(c.startX + ((float)i * 2.f), c.startY, 0.f);
                float xd = sample_uniform(W, c, -kStartDev, kStartDev);
                float yd = sample_uniform(W, c, -kStartDev, kStartDev);
                A.pos = base + v3(xd, yd, 0.f);
                A.pos.x = clampf(A.pos.x, 0.f, c.width);
                A.pos.y = clampf(A.pos.y, 0.f, c.height);
                pos_for_ball = A.pos;
                offensive_id = agent_id(W, i);
                A.hasBall = 1; A.ballId = ball_id(W); A.pointsWorth = 2;
            } else {
                float ang = sample_uniform(W, c, 0.f, 2.f * kPi);
                Vec3 off = v3(kDefenderSpawnRadius * bb_cosf(ang), kDefenderSpawnRadius * bb_sinf(ang), 0.f);
                A.pos = pos_for_ball + off;
                A.pos.x = clampf(A.pos.x, 0.f, c.width);
                A.pos.y = clampf(A.pos.y, 0.f, c.height);
                A.hasBall = 0; A.ballId = kPlaceholder; A.pointsWorth = 2;
            }
        } else {
            A.pos = v3(c.startX - 1.f - (float)(-2 * (i % 2)), c.startY - 2.f + (float)(i / 2), 0.f);
            if (i == 0) { offensive_id = agent_id(W, i); A.hasBall = 1; A.ballId = ball_id(W); A.pointsWorth = 2; }
            else        { A.hasBall = 0; A.ballId = kPlaceholder; A.pointsWorth = 2; }
        }
        A.maxSpeed = kDefaultSpeed - (float)i * kDefenderSlowdown;
        A.quickness = 1.f; A.shooting = 0.f; A.ftPct = 0.f;
        A.reaction = (float)i * kDefenderReaction;
        A.target = A.pos;
        A.curShot = 0.f;
    }
}

/* ------------------------------ gen.cpp ----------------------------------- */

BBD void reset_ball(World &W)
{
    W.b.inFlight = 0; W.b.ltAgent = kPlaceholder; W.b.ltTeam = kPlaceholder;
    W.b.sbAgent = kPlaceholder; W.b.sbTeam = kPlaceholder; W.b.spv = 2; W.b.goingIn = 0;
    W.b.vel = v3(0.f, 0.f, 0.f);
}

/* gen.cpp:13-214 */
BBD void generate_world(World &W, const SimConst &c)
{
    GameS &g = W.g;
    g.inbInProgress = 0; g.liveBall = 1; g.period = 1.0f; g.teamInPoss = 0.0f;
    g.team0Hoop = hoop_id(W, 0); g.team0Score = 0.f; g.team1Hoop = hoop_id(W, 1); g.team1Score = 0.f;
    g.gameClock = kTimePerPeriod; g.shotClock = 24.0f; g.scoredBaskets = 0.f; g.oobCount = 0.f;
    g.inboundClock = 0.f; g.isOneOnOne = kOneOnOne;
    W.resetNow = 0; /* quirk B9: column starts zeroed */
    W.b.pos = v3(c.startX, c.startY, 0.f);
    W.b.isGrabbed = 0; W.b.holder = kPlaceholder;
    reset_ball(W);
#pragma unroll
    for (int i = 0; i < 2; i++) {
        AgentS &A = W.a[i];
        A.move = A.moveAngle = A.rotate = A.grab = A.pass = A.shoot = 0;
        A.canMove = A.canGrab = A.canPass = A.canShoot = 0;
        A.reset = 0; A.imInb = 0; A.allowedToMove = 1;
        A.reward = 0.f; A.done = 0.f; A.curStep = 0;
        A.q = c.spawnQ[i];
        A.cooldown = 0.f; A.points = 0.f; A.fouls = 0.f;
        A.vel = v3(0.f, 0.f, 0.f);
    }
    int32_t off_id = kPlaceholder;
    Vec3 pfb = v3(c.startX, c.startY, 0.f);
    setup_agent_positions(W, c, off_id, pfb);
    if ((float)g.isOneOnOne == 1.f) { W.b.isGrabbed = 1; W.b.holder = off_id; }
}

/* gen.cpp:216-316 */
BBD void reset_world(World &W, const SimConst &c)
{
    GameS &g = W.g;
    if (g.gameClock <= 0.f && (float)g.isOneOnOne == 0.f) {
        if (g.period < 4.f || g.team0Score == g.team1Score) {
            g.period += 1.f;
            g.gameClock = kTimePerPeriod; g.shotClock = 24.0f; g.liveBall = 1; g.inbInProgress = 0;
        } else {
            g.liveBall = 0;
        }
    } else {
        g.inbInProgress = 0; g.liveBall = 1; g.period = 1.0f; g.teamInPoss = 0.0f;
        g.team0Score = 0.f; g.team1Score = 0.f; g.gameClock = kTimePerPeriod; g.shotClock = 24.0f;
        g.scoredBaskets = 0.f; g.oobCount = 0.f; g.inboundClock = 0.f;
        /* team0Hoop, team1Hoop, isOneOnOne preserved (gen.cpp:245-254) */
    }
#pragma unroll
    for (int i = 0; i < 2; i++) {
        AgentS &A = W.a[i];
        A.move = A.moveAngle = A.rotate = A.grab = A.pass = A.shoot = 0;
        A.canMove = A.canGrab = A.canPass = A.canShoot = 0;
        A.reset = 0; A.imInb = 0; A.allowedToMove = 1;
        A.done = 1.f; A.curStep = 0;
        A.q = c.spawnQ[i];
        A.cooldown = 0.f; A.points = 0.f; A.fouls = 0.f;
        A.vel = v3(0.f, 0.f, 0.f);
    }
    int32_t off_id = kPlaceholder;
    Vec3 pfb = v3(c.startX, c.startY, 0.f);
    setup_agent_positions(W, c, off_id, pfb);
    W.b.pos = pfb;
    reset_ball(W);
    if ((float)g.isOneOnOne == 1.f) { W.b.isGrabbed = 1; W.b.holder = off_id; }
    else { W.b.isGrabbed = 0; W.b.holder = kPlaceholder; }
    W.didReset = true;
}

/* ------------------------------ game.cpp ---------------------------------- */

/* game.cpp:14-53 */
BBD void assign_inbounder(World &W, Vec3 ball_pos, int32_t new_team, Quat orient, bool is_oob)
{
    bool assigned = false;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        if (i == new_team && !assigned) {
            assigned = true;
            AgentS &A = W.a[i];
            A.imInb = 1;
            A.pos = ball_pos;
            W.b.isGrabbed = 1; W.b.holder = agent_id(W, i);
            A.hasBall = 1; A.ballId = ball_id(W);
            A.q = orient;
        }
    }
    if (assigned) {
        W.g.teamInPoss = (float)new_team;
        W.g.inbInProgress = 1;
        W.g.inboundClock = 5.f;
        if (is_oob) W.g.oobCount += 1.f;
    }
}

/* game.cpp:969-988 */
BBD void sys_tick(AgentS &A)
{
    A.reward = 0.f;
    if (A.reset == 1) { A.done = 1.f; A.curStep = 0; }
    else { A.done = 0.f; A.curStep++; }
    A.cooldown = fmaxf(0.f, A.cooldown - 1.f);
}

/* game.cpp:489-533 */
template <bool kTagMask>
BBD void sys_action_mask(AgentS &A, const GameS &g)
{
    A.canMove = 1; A.canGrab = 1; A.canPass = 0; A.canShoot = 0;
    if (A.hasBall == 1) { A.canPass = 1; A.canShoot = 1; }
    if (g.inbInProgress == 1) {
        A.canShoot = 0; A.canGrab = 0;
        if (A.imInb == 1 && g.liveBall == 0) A.canMove = 0;
    }
    if (A.cooldown > 0.f) A.canGrab = 0;
    if (kTagMask) { A.canPass = 0; A.canGrab = 0; } /* game.cpp:525-528 */
}

/* game.cpp:410-486 */
/* facing: q.rotateVec({0,1,0}) of the orientation after the turn (game.cpp:439), handed back so the lane-pair
 * path can reuse it for collision / defence / observations while the orientation stays what it is here */
BBD void sys_move_agent(AgentS &A, const SimConst &c, Vec3 &facing)
{
    if (A.rotate != 0) {
        Quat t = (A.rotate == 1) ? c.turnPos : c.turnNeg;
        A.q = qmul(t, A.q);
    }
    Vec3 ov = rotate_vec(A.q, v3(0.f, 1.f, 0.f));
    facing = ov;
    if (A.canMove == 0) return;
    float s, co;
    if ((uint32_t)A.moveAngle < 8u) { s = c.moveSin[A.moveAngle]; co = c.moveCos[A.moveAngle]; }
    else { float ang = (float)A.moveAngle * kAngleBetweenDirs; s = bb_sinf(ang); co = bb_cosf(ang); }
    Vec3 delta = (v3(s, -co, 0.f) * A.quickness) * (float)A.move;
    float maxS = A.maxSpeed;
    float d = 0.f;
    if (len2(A.vel) > 1e-6f) d = dot(normalize(A.vel), ov);
    if (d < -0.1f) { maxS *= .1f; delta = delta * .1f; }
    else if (d <= 0.8f) { maxS *= .7f; delta = delta * .1f; }
    A.vel = A.vel + delta;
    if (A.hasBall == 1) maxS *= kBallSlowdown;
    float L = len(A.vel);
    if (L > maxS) A.vel = A.vel * f_div(maxS, L);
    float dx = A.vel.x * kDt, dy = A.vel.y * kDt;
    A.pos.x = clampf(A.pos.x + dx, 0.f, c.width);
    A.pos.y = clampf(A.pos.y + dy, 0.f, c.height);
    /* game.cpp:473-484: the cell table is all zero (bindings.cpp:7-12): no walls */
    A.vel = A.vel * .95f;
}

/* game.cpp:164-239 */
BBD void sys_grab(World &W, int i)
{
    AgentS &A = W.a[i];
    if (A.canGrab == 0 || A.grab == 0) return;
    A.cooldown = 10.f;
    A.grab = 0;
    BallS &B = W.b;
    if (B.inFlight == 1) return;
    bool holding = (A.hasBall == 1 && B.isGrabbed == 1 && B.holder == agent_id(W, i));
    if (holding) {
        A.ballId = kPlaceholder; A.hasBall = 0; B.holder = kPlaceholder; B.isGrabbed = 0;
        return;
    }
    float dist = len(B.pos - A.pos);
    if (dist <= 0.3f) {
        if ((float)W.g.isOneOnOne == 1.f && (float)i != W.g.teamInPoss) { W.resetNow = 1; return; }
#pragma unroll
        for (int j = 0; j < 2; j++) {
            if (W.a[j].ballId == ball_id(W)) {
                W.a[j].hasBall = 0; W.a[j].ballId = kPlaceholder; W.a[j].cooldown = kSimHz;
            }
        }
        A.hasBall = 1; A.ballId = ball_id(W);
        B.holder = agent_id(W, i); B.isGrabbed = 1; B.inFlight = 0;
        B.vel = v3(0.f, 0.f, 0.f);
        B.sbAgent = kPlaceholder; B.sbTeam = kPlaceholder; B.spv = 2;
        W.g.teamInPoss = (float)i;
        W.g.liveBall = 1;
    }
}

/* game.cpp:243-270 */
BBD void sys_pass(World &W, int i)
{
    AgentS &A = W.a[i];
    if (A.canPass == 0 || A.pass == 0) return;
    if (W.b.holder == agent_id(W, i)) {
        W.b.isGrabbed = 0; W.b.holder = kPlaceholder;
        A.hasBall = 0; A.ballId = kPlaceholder; A.imInb = 0;
        W.b.vel = rotate_vec(A.q, v3(0.f, 0.1f, 0.f));
        W.g.inbInProgress = 0;
    }
}

/* game.cpp:273-407 */
BBD void sys_shoot(World &W, int i, const SimConst &c)
{
    AgentS &A = W.a[i];
    if (A.canShoot == 0 || A.shoot == 0) return;
    const Vec3 hoop = hoop_pos(c, 1 - i); /* the hoop this team does not defend */
    const float radius = kHoopZone;
    Vec3 ideal = hoop - A.pos;
    float intended = bb_atan2f(ideal.x, ideal.y);
    float dist_hoop = len(ideal);
    float dist_sd = kDistDev * dist_hoop;
    float dev_dist = sample_uniform(W, c, -dist_sd, dist_sd);
    float dev_def = 0.f;
    float nearest = len(A.pos - W.a[1 - i].pos); /* the only agent on the other team */
    if (nearest < 2.0f) {
        float def_sd = f_div(kDefDev, nearest + 0.1f);
        dev_def = sample_uniform(W, c, -def_sd, def_sd);
    }
    float dev_vel = 0.f;
    if (A.move > 0) {
        float vel_sd = kVelDev * len(A.vel);
        dev_vel = sample_uniform(W, c, -vel_sd, vel_sd);
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
    A.q = rotation_between(v3(0.f, 1.f, 0.f), shot, c);
    BallS &B = W.b;
    if (B.holder == agent_id(W, i)) {
        int32_t spv = shot_point_value(A.pos, hoop);
        if (going_in) { B.goingIn = 1; W.g.scoredBaskets += 1.f; }
        else A.reward -= 1.f;
        B.isGrabbed = 0; B.holder = kPlaceholder;
        A.hasBall = 0; A.ballId = kPlaceholder; A.imInb = 0;
        B.vel = shot * .1f;
        B.inFlight = 1;
        B.sbAgent = agent_id(W, i); B.sbTeam = i; B.spv = spv;
        B.ltAgent = agent_id(W, i); B.ltTeam = i;
    }
}

/* game.cpp:82-125 */
BBD void sys_move_ball(World &W, const SimConst &c)
{
    BallS &B = W.b;
#pragma unroll
    for (int i = 0; i < 2; i++)
        if (W.a[i].hasBall == 1 && B.isGrabbed == 1 && B.holder == agent_id(W, i)) B.pos = W.a[i].pos;
    if (len(B.vel) == 0.f || B.isGrabbed == 1) return;
    float nx = clampf(B.pos.x + B.vel.x, 0.f, c.width);
    float ny = clampf(B.pos.y + B.vel.y, 0.f, c.height);
    float nz = B.pos.z + B.vel.z;
    B.pos = v3(nx, ny, nz);
}

/* game.cpp:758-809 */
BBD void sys_shot_pct(World &W, int i, const SimConst &c)
{
    AgentS &A = W.a[i];
    if (A.hasBall == 0) { A.curShot = 0.f; return; }
    float dist_hoop = len(hoop_pos(c, 1 - i) - A.pos);
    float nearest = len(A.pos - W.a[1 - i].pos);
    float dist_sd = kDistDev * dist_hoop;
    float def_sd = f_div(kDefDev, nearest) + .0001f;
    float vel_sd = kVelDev * len(A.vel);
    float final_sd = f_sqrt((f_div(dist_sd * dist_sd, 3.f) + f_div(def_sd * def_sd, 3.f)) + f_div(vel_sd * vel_sd, 3.f));
    float max_ang = bb_atanf(f_div(kHoopZone, dist_hoop));
    float z = f_div(max_ang, final_sd);
    A.curShot = bb_erff(f_div(z, f_sqrt(2.f)));
}

/* game.cpp:873-953 */
BBD void sys_score(World &W, int h, const SimConst &c)
{
    BallS &B = W.b;
    GameS &g = W.g;
    const Vec3 hp = hoop_pos(c, h);
    float dx = B.pos.x - hp.x, dy = B.pos.y - hp.y;
    float dist = f_sqrt(dx * dx + dy * dy);
    if (dist <= kHoopZone && (float)B.inFlight == 1.f) {
        int32_t points = B.spv;
        int32_t inb_team = h; /* agent h defends hoop h */
#pragma unroll
        for (int j = 0; j < 2; j++)
            if (agent_id(W, j) == B.sbAgent) W.a[j].points += (float)((j == h) ? -B.spv : B.spv);
        Vec3 spot;
        if (hoop_id(W, h) == g.team0Hoop) { g.team1Score += (float)points; spot = v3(kCourtMinX, hp.y + (kPixelsPerMeter / 60.f), 0.f); }
        else                              { g.team0Score += (float)points; spot = v3(kCourtMaxX, hp.y + (kPixelsPerMeter / 60.f), 0.f); }
        g.scoredBaskets += 1.f;
        B.inFlight = 0;
        B.vel = v3(0.f, 0.f, 0.f);
        B.sbAgent = kPlaceholder; B.sbTeam = kPlaceholder; B.spv = 2; B.goingIn = 0;
        if ((float)g.isOneOnOne == 0.f) {
            B.pos = spot;
            Quat o = rotation_between(v3(0.f, 1.f, 0.f), vector_to_center(B.pos, c), c);
            assign_inbounder(W, spot, inb_team, o, false);
        } else {
            W.resetNow = 1;
        }
    }
}

/* offence = agent 0 unless a later agent's team is in possession (game.cpp:1012-1018, 1072-1078) */
BBD int offence_agent(const World &W) { return ((float)1 == W.g.teamInPoss) ? 1 : 0; }

/* game.cpp:1055-1113 */
BBD void sys_out_of_bounds(World &W, const SimConst &c)
{
    BallS &B = W.b;
    GameS &g = W.g;
    bool out = B.pos.x < kCourtMinX || B.pos.x > kCourtMaxX || B.pos.y < kCourtMinY || B.pos.y > kCourtMaxY;
    if (out && (float)g.inbInProgress == 0.f) {
        if ((float)g.isOneOnOne == 1.f) {
            int off = offence_agent(W);
            if (off == 0) W.a[0].reward -= 100.f; else W.a[1].reward -= 100.f;
            W.resetNow = 1;
        } else {
            B.inFlight = 0;
            B.vel = v3(0.f, 0.f, 0.f);
            g.liveBall = 0;
            int32_t new_team = 1 - B.ltTeam;
#pragma unroll
            for (int i = 0; i < 2; i++) {
                AgentS &A = W.a[i];
                if (A.hasBall == 1 && A.ballId == ball_id(W)) {
                    A.pos = A.pos + vector_to_center(A.pos, c);
                    A.hasBall = 0; A.ballId = kPlaceholder;
                }
            }
            Quat o = rotation_between(v3(0.f, 1.f, 0.f), vector_to_center(B.pos, c), c);
            assign_inbounder(W, B.pos, new_team, o, true);
        }
    }
}

/* game.cpp:1034-1051 */
BBD void sys_last_touch(World &W)
{
#pragma unroll
    for (int i = 0; i < 2; i++) {
        float d = len(W.b.pos - W.a[i].pos);
        if (d <= kAgentSize) { W.b.ltAgent = agent_id(W, i); W.b.ltTeam = i; }
    }
}

/* game.cpp:992-1030 */
BBD void sys_clock(World &W)
{
    GameS &g = W.g;
    bool live = (float)g.liveBall > 0.5f;
    if (live && g.gameClock > 0.f) { g.gameClock -= kDt; g.shotClock -= kDt; }
    if ((float)g.inbInProgress > 0.5f) g.inboundClock -= kDt;
    if (g.gameClock <= 0.f && live) {
        int off = offence_agent(W);
        if (off == 0) W.a[0].reward += 10.f; else W.a[1].reward += 10.f;
        W.resetNow = 1;
    }
    if (g.shotClock < 0.f) g.shotClock = 0.f;
}

/* game.cpp:1116-1157 */
BBD void sys_inbound_violation(World &W, const SimConst &c)
{
    GameS &g = W.g;
    if (!((float)g.inbInProgress > 0.5f && g.inboundClock <= 0.f)) return;
    int32_t cur_team = (int32_t)g.teamInPoss;
    int32_t new_team = 1 - cur_team;
    int32_t turnover_ball = kPlaceholder;
    g.liveBall = 0;
#pragma unroll
    for (int i = 0; i < 2; i++) {
        AgentS &A = W.a[i];
        if ((float)A.imInb > 0.5f) {
            turnover_ball = A.ballId;
            A.imInb = 0; A.hasBall = 0; A.ballId = kPlaceholder;
            A.pos = A.pos + vector_to_center(A.pos, c);
        }
    }
    if (turnover_ball != kPlaceholder && ball_id(W) == turnover_ball) {
        W.b.isGrabbed = 0; W.b.holder = kPlaceholder;
        Quat o = rotation_between(v3(0.f, 1.f, 0.f), vector_to_center(W.b.pos, c), c);
        assign_inbounder(W, W.b.pos, new_team, o, true);
    }
}

/* game.cpp:129-161 */
BBD void sys_points_worth(World &W, int i, const SimConst &c)
{
    W.a[i].pointsWorth = shot_point_value(W.a[i].pos, hoop_pos(c, 1 - i));
}

/* game.cpp:537-648 + helper.cpp:85-105, the single pair (agent0, agent1) */
BBD void sys_agent_collision(World &W)
{
    AgentS &A = W.a[0];
    AgentS &Bg = W.a[1];
    const float hw = kShoulderWidth / 2.0f, hd = kAgentDepth / 2.0f;
    Vec3 ca = A.pos;
    Vec3 fa = rotate_vec(A.q, v3(0.f, 1.f, 0.f));
    Vec3 ra = v3(fa.y, -fa.x, 0.f);
    Vec3 hwa = ra * hw, hda = fa * hd;
    Vec3 va[4] = {ca - hda + hwa, ca - hda - hwa, ca + hda - hwa, ca + hda + hwa};
    Vec3 cb = Bg.pos;
    Vec3 fb = rotate_vec(Bg.q, v3(0.f, 1.f, 0.f));
    Vec3 rb = v3(fb.y, -fb.x, 0.f);
    Vec3 hwb = rb * hw, hdb = fb * hd;
    Vec3 vb[4] = {cb - hdb + hwb, cb - hdb - hwb, cb + hdb - hwb, cb + hdb + hwb};
    Vec3 axes[4] = {normalize(ra), normalize(fa), normalize(rb), normalize(fb)};

    bool colliding = true;
    float min_overlap = FLT_MAX;
    Vec3 mtv = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 4; j++) {
        if (!colliding) continue; /* reference breaks at the first separating axis */
        Vec3 ax = axes[j];
        float amin = dot(va[0], ax), amax = amin, bmin = dot(vb[0], ax), bmax = bmin;
#pragma unroll
        for (int k = 1; k < 4; k++) {
            float pa = dot(va[k], ax), pb = dot(vb[k], ax);
            if (pa < amin) amin = pa;
            if (pa > amax) amax = pa;
            if (pb < bmin) bmin = pb;
            if (pb > bmax) bmax = pb;
        }
        if (!(amax > bmin && bmax > amin)) {
            colliding = false;
        } else {
            float overlap = fminf(amax, bmax) - fmaxf(amin, bmin);
            if (overlap < min_overlap) { min_overlap = overlap; mtv = ax; }
        }
    }
    if (colliding) {
        if (W.g.teamInPoss == (float)0) { /* agent a = agent 0 is on team 0 */
            A.reward -= 10.f;
            Bg.reward += 10.f;
            W.resetNow = 1;
        }
        Vec3 corr = mtv;
        if (dot(cb - ca, corr) < 0.f) corr = -corr;
        Vec3 shift = (corr * min_overlap) * 0.5f;
        A.pos = A.pos - shift;
        Bg.pos = Bg.pos + shift;
    }
}

/* game.cpp:651-755 */
BBD void sys_hardcode_defense(World &W, int i, const SimConst &c)
{
    AgentS &D = W.a[i];
    if (W.g.teamInPoss == (float)i) { D.move = 0; return; }
    D.grab = 1;
    /* first agent holding a ball; the defender's hoop (hoop i) always exists */
    bool found = false;
    Vec3 guard = v3(0.f, 0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 2; j++) {
        if (W.a[j].hasBall == 1 && !found) {
            Vec3 op = W.a[j].pos;
            Vec3 hd = hoop_pos(c, i) - op;
            if (len2(hd) > 1e-6f) guard = op + normalize(hd) * kGuardDist;
            else guard = op;
            found = true;
        }
    }
    if (!found) { D.move = 0; return; }
    float f = D.reaction * kDt;
    D.target = D.target + (guard - D.target) * f;
    Vec3 mv = D.target - D.pos;
    if (len2(mv) < 0.01f) { D.move = 0; return; }
    Vec3 dir = normalize(mv);
    /* dot with the 8 normalised compass directions of game.cpp:713-722; the
     * z terms are (0 * dir.z) and kept so signed zeros match */
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
    Vec3 ov = rotate_vec(D.q, v3(0.f, 1.f, 0.f));
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

/* game.cpp:811-870 */
BBD void sys_reward(World &W, int i)
{
    AgentS &A = W.a[i];
    float dist_other = len(W.a[1 - i].pos - A.pos);
    if ((float)i == W.g.teamInPoss) {
        if (W.g.gameClock > 5.f) {
            if (W.b.sbAgent == agent_id(W, i) && W.b.goingIn == 1) A.reward += (float)W.b.spv;
            else if (W.b.sbAgent == agent_id(W, i) && W.b.goingIn == 0 && W.b.inFlight == 1) A.reward -= 1.f;
            A.reward += A.curShot;
        }
    } else {
        A.reward -= 1.f;
        A.reward += bb_expf(-0.4f * dist_other);
    }
}

/* Systems 1..17 of the task graph (game.cpp:1467-1516); observation fill (18)
 * and reward (19) follow in the kernel because 18 writes through shared memory. */
template <bool kTagMask>
BBD void step_world_logic(World &W, const SimConst &c)
{
    W.didReset = false;
#pragma unroll
    for (int i = 0; i < 2; i++) sys_tick(W.a[i]);
#pragma unroll
    for (int i = 0; i < 2; i++) sys_action_mask<kTagMask>(W.a[i], W.g);
#pragma unroll
    for (int i = 0; i < 2; i++) { Vec3 facing; sys_move_agent(W.a[i], c, facing); }
    if (!kTagMask) {
#pragma unroll
        for (int i = 0; i < 2; i++) sys_grab(W, i);
#pragma unroll
        for (int i = 0; i < 2; i++) sys_pass(W, i);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) sys_shoot(W, i, c);
    sys_move_ball(W, c);
#pragma unroll
    for (int i = 0; i < 2; i++) sys_shot_pct(W, i, c);
#pragma unroll
    for (int h = 0; h < 2; h++) sys_score(W, h, c);
    sys_out_of_bounds(W, c);
    sys_last_touch(W);
    sys_clock(W);
    sys_inbound_violation(W, c);
    if (W.resetNow != 0) { reset_world(W, c); W.resetNow = 0; } /* game.cpp:957-967 */
#pragma unroll
    for (int i = 0; i < 2; i++) sys_points_worth(W, i, c);
    sys_agent_collision(W);
#pragma unroll
    for (int i = 0; i < 2; i++) sys_hardcode_defense(W, i, c);
}

} // namespace bb
