/*
 * bb_oracle.cpp — CPU restatement of the reference's per-world step.
 *
 * TEST INFRASTRUCTURE ONLY (see bb_oracle.h).  PARITY UNPINNED (Madrona absent,
 * reference has no tests).  One world at a time, 19 systems in the task-graph
 * order of /root/reference/src/game.cpp:1463-1526, agent 0 before agent 1,
 * hoop 0 before hoop 1 (Madrona CPU ParallelForNode order).  Every function
 * cites the reference lines it restates.  Components live in per-component
 * columns exactly as the reference exports them (src/mgr.cpp:317-445), so the
 * columns double as the exported tensors.
 *
 * Build: g++ -O2 -ffp-contract=off -mfma (oracle/Makefile).  -DBB_USE_LIBM
 * swaps the deterministic math spec (include/bb_math.h) for glibc.
 */
#include "bb_oracle.h"
#include "madrona_prims.h"

#include <cstdlib>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

namespace orc {

/* ======================= src/constants.hpp:5-98 ========================== */
constexpr int32_t NUM_AGENTS = 2;
constexpr int32_t NUM_BASKETBALLS = 1;
constexpr int32_t NUM_HOOPS = 2;
constexpr int32_t ENTITY_ID_PLACEHOLDER = INT32_MAX;
constexpr float SIMULATION_HZ = 62.0f;
constexpr float TIMESTEPS_TO_SECONDS_FACTOR = 1.0f / SIMULATION_HZ;
constexpr float TIME_PER_PERIOD = 10.f;
constexpr float PIXELS_PER_METER = 110.f;
constexpr float HOOP_SCORE_ZONE_SIZE = 0.1f;
constexpr int32_t ONE_ON_ONE = 1;
constexpr float AGENT_SIZE_M = 0.2f;
constexpr float AGENT_SHOULDER_WIDTH = 0.4290;
constexpr float AGENT_DEPTH = .1;
constexpr float GUARDING_DISTANCE = .2f;
constexpr float START_POS_STDDEV = 5.f;
constexpr float DEFAULT_SPEED = 3.f;
constexpr float DEFENDER_SLOWDOWN = 0.2f;
constexpr float DEFENDER_REACTION = 10.f;
constexpr float DEFENDER_SPAWN_RADIUS = 8.f;
constexpr float ANGLE_BETWEEN_DIRECTIONS = pi / 4.0f;
constexpr Vector3 AGENT_BASE_FORWARD = {0, 1, 0};
constexpr float BALL_AGENT_SLOWDOWN = 0.9f;
constexpr float DIST_DEVIATION_PER_METER = .008f;
constexpr float DEF_DEVIATION_PER_METER = .002f;
constexpr float VEL_DEVIATION_FACTOR = .001f;
constexpr float COURT_LENGTH_M = 28.65f;
constexpr float COURT_WIDTH_M = 15.24f;
constexpr float WORLD_MARGIN_FACTOR = 1.1f;
constexpr float WORLD_WIDTH_M = COURT_LENGTH_M * WORLD_MARGIN_FACTOR;
constexpr float WORLD_HEIGHT_M = COURT_WIDTH_M * WORLD_MARGIN_FACTOR;
constexpr float COURT_MIN_X = (WORLD_WIDTH_M - COURT_LENGTH_M) / 2.0f;
constexpr float COURT_MAX_X = COURT_MIN_X + COURT_LENGTH_M;
constexpr float COURT_MIN_Y = (WORLD_HEIGHT_M - COURT_WIDTH_M) / 2.0f;
constexpr float COURT_MAX_Y = COURT_MIN_Y + COURT_WIDTH_M;
constexpr float HOOP_FROM_BASELINE_M = 1.575f;
constexpr float ARC_RADIUS_M = 7.24f;
constexpr float CORNER_3_FROM_SIDELINE_M = 0.91f;
constexpr float CORNER_3_LENGTH_FROM_BASELINE_M = 4.27f;

/* ========================= src/types.hpp:46-215 =========================== */
struct GameState {
    int32_t inboundingInProgress;
    int32_t liveBall;
    float period;
    float teamInPossession;
    int32_t team0Hoop;
    float team0Score;
    int32_t team1Hoop;
    float team1Score;
    float gameClock;
    float shotClock;
    float scoredBaskets;
    float outOfBoundsCount;
    float inboundClock;
    int32_t isOneOnOne;
};
struct WorldClock { int32_t resetNow; };
struct Reset { int32_t resetNow; };
struct Position { Vector3 position; };
struct Velocity { Vector3 velocity; };
struct Action { int32_t move, moveAngle, rotate, grab, pass, shoot; };
struct ActionMask { int32_t can_move, can_grab, can_pass, can_shoot; };
struct Orientation { Quat orientation; };
struct Reward { float r; };
struct Done { float episodeDone; };
struct InPossession { int32_t hasBall, ballEntityID, pointsWorth; };
struct Inbounding { int32_t imInbounding, allowedToMove; };
struct Team { int32_t teamIndex; Vector3 teamColor; int32_t defendingHoopID; };
struct GrabCooldown { float cooldown; };
struct Observations { float observationsArray[128]; };
struct Stats { float points, fouls; };
struct Attributes {
    float maxSpeed, quickness, shooting, freeThrowPercentage, reactionSpeed;
    Vector3 currentTargetPosition;
    float currentShotPercentage;
};
struct Grabbed { int32_t isGrabbed, holderEntityID; };
struct BallPhysics {
    int32_t inFlight, lastTouchedByAgentID, lastTouchedByTeamID, shotByAgentID, shotByTeamID,
        shotPointValue, shotIsGoingIn;
};
struct ScoringZone { float radius, height; Vector3 center; };
struct CurStep { uint32_t step; };
struct Entity { uint32_t gen; int32_t id; }; /* Madrona Entity, SURVEY App. C */

static_assert(sizeof(GameState) == 56 && sizeof(Action) == 24 && sizeof(Team) == 20 &&
              sizeof(Attributes) == 36 && sizeof(BallPhysics) == 28 && sizeof(Observations) == 512, "layout");

/* src/grid.hpp:16-28 (cell table dropped: bindings.cpp:7-12 zero-fills it and nothing sets a wall) */
struct GridState {
    float startX, startY, width, height;
    int32_t cellsPerMeter, discreteWidth, discreteHeight;
};

} // namespace orc

using namespace orc;

/* Columns: one array per component, [num_worlds * entities_per_world]. */
struct orc_sim {
    orc_config cfg;
    GridState grid;
    RandKey worldKey; /* sim.cpp:89 — identical in every world */
    int64_t N;

    std::vector<Reset> reset;
    std::vector<GameState> gameState;
    std::vector<Action> action;
    std::vector<ActionMask> actionMask;
    std::vector<Position> agentPos;
    std::vector<Observations> obs;
    std::vector<Reward> reward;
    std::vector<Done> done;
    std::vector<Entity> agentEntity;
    std::vector<InPossession> possession;
    std::vector<Orientation> orientation;
    std::vector<Team> team;
    std::vector<Stats> stats;
    std::vector<Position> ballPos;
    std::vector<BallPhysics> ballPhysics;
    std::vector<Entity> ballEntity;
    std::vector<Grabbed> grabbed;
    std::vector<Velocity> ballVel;
    std::vector<Position> hoopPos;
    /* not exported by the reference */
    std::vector<Velocity> agentVel;
    std::vector<Inbounding> inbounding;
    std::vector<GrabCooldown> cooldown;
    std::vector<Attributes> attributes;
    std::vector<CurStep> curStep;
    std::vector<WorldClock> worldClock;
    std::vector<uint32_t> rngCount;
    std::vector<ScoringZone> scoringZone;
};

namespace orc {

/* View of one world: the "Engine &ctx" of the reference. */
struct Ctx {
    orc_sim *s;
    int64_t w;
    const GridState *grid;
    RNG rng;
    /* Sim::agents / balls / hoops (sim.hpp:39-41): entity ids of this world */
    int32_t agentId[NUM_AGENTS], ballId[NUM_BASKETBALLS], hoopId[NUM_HOOPS];

    GameState &gameState() { return s->gameState[w]; }
    WorldClock &worldClock() { return s->worldClock[w]; }
    int64_t ai(int i) const { return w * NUM_AGENTS + i; }
    int64_t hi(int i) const { return w * NUM_HOOPS + i; }
};

static inline Ctx makeCtx(orc_sim *s, int64_t w)
{
    Ctx c;
    c.s = s;
    c.w = w;
    c.grid = &s->grid;
    c.rng = RNG{s->worldKey, s->rngCount[w]};
    /* ids handed out sequentially, worlds built serially:
     * hoop0, hoop1, ball, agent0, agent1 (creation order gen.cpp:101,131,167,187) */
    int32_t base = (int32_t)(w * 5);
    c.hoopId[0] = base + 0;
    c.hoopId[1] = base + 1;
    c.ballId[0] = base + 2;
    c.agentId[0] = base + 3;
    c.agentId[1] = base + 4;
    return c;
}

/* ============================ src/helper.cpp ============================== */

/* helper.cpp:8-11 */
static float sampleUniform(Ctx &ctx, const float min, const float max)
{
    return min + (max - min) * ctx.rng.sampleUniform();
}

/* helper.cpp:14-42 */
static Quat findRotationBetweenVectors(Vector3 start, Vector3 target)
{
    start = start.normalize();
    target = target.normalize();
    float dot_product = dot(start, target);
    if (dot_product > 0.999999f) return Quat::id();
    if (dot_product < -0.999999f) return Quat::angleAxis(pi, Vector3{0.f, 0.f, 1.f});
    Vector3 rotation_axis = cross(start, target);
    rotation_axis = rotation_axis.normalize();
    float rotation_angle = m_acosf(dot_product);
    return Quat::angleAxis(rotation_angle, rotation_axis);
}

/* helper.cpp:44-48 */
static Vector3 findVectorToCenter(Ctx &ctx, Position pos)
{
    return (Vector3{ctx.grid->startX, ctx.grid->startY, 0.f} - pos.position).normalize();
}

/* helper.cpp:50-81 */
static int32_t getShotPointValue(Position shot_pos, Vector3 hoop_score_zone)
{
    float distance_to_hoop = (shot_pos.position - hoop_score_zone).length();
    bool isInCornerLane = (shot_pos.position.y < COURT_MIN_Y + CORNER_3_FROM_SIDELINE_M ||
                           shot_pos.position.y > COURT_MIN_Y + COURT_WIDTH_M - CORNER_3_FROM_SIDELINE_M);
    if (isInCornerLane) {
        bool isShootingAtLeftHoop = hoop_score_zone.x < WORLD_WIDTH_M / 2.0f;
        if (isShootingAtLeftHoop) {
            if (shot_pos.position.x <= COURT_MIN_X + CORNER_3_LENGTH_FROM_BASELINE_M) return 3;
        } else {
            if (shot_pos.position.x >= COURT_MIN_X + COURT_LENGTH_M - CORNER_3_LENGTH_FROM_BASELINE_M) return 3;
        }
    }
    if (distance_to_hoop >= ARC_RADIUS_M) return 3;
    return 2;
}

/* helper.cpp:85-105 */
struct Projection { float min, max; };
static Projection projectRectangle(const Vector3 *vertices, const Vector3 &axis)
{
    Projection p;
    p.min = dot(vertices[0], axis);
    p.max = p.min;
    for (int i = 1; i < 4; i++) {
        float proj = dot(vertices[i], axis);
        if (proj < p.min) p.min = proj;
        if (proj > p.max) p.max = proj;
    }
    return p;
}
static bool projectionsOverlap(const Projection &p1, const Projection &p2)
{
    return p1.max > p2.min && p2.max > p1.min;
}

/* helper.cpp:108-160 */
static void setupAgentPositions(Ctx &ctx, int32_t basketball_id, int32_t &offensive_agent_id,
                                Position &agent_pos_for_ball)
{
    const GridState *grid = ctx.grid;
    GameState &gameState = ctx.gameState();
    for (int i = 0; i < NUM_AGENTS; i++) {
        Position &pos = ctx.s->agentPos[ctx.ai(i)];
        InPossession &in_pos = ctx.s->possession[ctx.ai(i)];
        if (gameState.isOneOnOne == 1.f) {
            if (i == 0) {
                Vector3 base_pos = {grid->startX + (i * 2.f), grid->startY, 0.f};
                float x_dev = sampleUniform(ctx, -START_POS_STDDEV, START_POS_STDDEV);
                float y_dev = sampleUniform(ctx, -START_POS_STDDEV, START_POS_STDDEV);
                pos.position = base_pos + Vector3{x_dev, y_dev, 0.f};
                pos.position.x = clampf(pos.position.x, 0.f, grid->width);
                pos.position.y = clampf(pos.position.y, 0.f, grid->height);
                agent_pos_for_ball = pos;
                offensive_agent_id = ctx.agentId[i];
                in_pos = {1, basketball_id, 2};
            } else {
                float random_angle = sampleUniform(ctx, 0.f, 2.f * pi);
                Vector3 offset = {DEFENDER_SPAWN_RADIUS * m_cosf(random_angle),
                                  DEFENDER_SPAWN_RADIUS * m_sinf(random_angle), 0.f};
                pos.position = agent_pos_for_ball.position + offset;
                pos.position.x = clampf(pos.position.x, 0.f, grid->width);
                pos.position.y = clampf(pos.position.y, 0.f, grid->height);
                in_pos = {0, ENTITY_ID_PLACEHOLDER, 2};
            }
        } else {
            pos = Position{Vector3{grid->startX - 1 - (-2 * (i % 2)), grid->startY - 2 + i / 2, 0.f}};
            if (i == 0) {
                offensive_agent_id = ctx.agentId[i];
                in_pos = {1, basketball_id, 2};
            } else {
                in_pos = {0, ENTITY_ID_PLACEHOLDER, 2};
            }
        }
        ctx.s->attributes[ctx.ai(i)] = {DEFAULT_SPEED - i * DEFENDER_SLOWDOWN, 1.f, 0.f, 0.f,
                                        i * DEFENDER_REACTION, pos.position, 0.f};
    }
}

/* ============================== src/gen.cpp =============================== */

static Orientation spawnOrientation(int i)
{
    /* gen.cpp:196 / gen.cpp:277 */
    return (i % 2 == 0) ? Orientation{Quat::angleAxis(-pi / 2.0f, {0.f, 0.f, 1.f})}
                        : Orientation{Quat::angleAxis(pi / 2.0f, {0.f, 0.f, 1.f})};
}

/* gen.cpp:13-214 */
static void generateWorld(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    ctx.gameState() = GameState{0, 1, 1.0f, 0.0f, 0, 0.0f, 1, 0.0f, TIME_PER_PERIOD, 24.0f, 0.f, 0.f, 0.0f, ONE_ON_ONE};
    /* gen.cpp:33-34 zeroes a COPY of WorldClock (quirk B9); the column starts zeroed */
    const GridState *grid = ctx.grid;
    GameState &gameState = ctx.gameState();

    float court_start_x = (grid->width - COURT_LENGTH_M) / 2.0f;
    float court_center_y = grid->height / 2.0f;

    /* hoop 0: gen.cpp:101-127 */
    gameState.team0Hoop = ctx.hoopId[0];
    Position hoop0_pos = Position{Vector3{court_start_x + HOOP_FROM_BASELINE_M, court_center_y, 0.f}};
    s->hoopPos[ctx.hi(0)] = hoop0_pos;
    s->scoringZone[ctx.hi(0)] = ScoringZone{HOOP_SCORE_ZONE_SIZE, .1f, hoop0_pos.position};
    /* hoop 1: gen.cpp:131-156 */
    gameState.team1Hoop = ctx.hoopId[1];
    Position hoop1_pos = Position{Vector3{court_start_x + COURT_LENGTH_M - HOOP_FROM_BASELINE_M, court_center_y, 0.f}};
    s->hoopPos[ctx.hi(1)] = hoop1_pos;
    s->scoringZone[ctx.hi(1)] = ScoringZone{HOOP_SCORE_ZONE_SIZE, .1f, hoop1_pos.position};

    /* ball: gen.cpp:167-176 */
    s->ballEntity[ctx.w] = Entity{0, ctx.ballId[0]};
    s->ballPos[ctx.w] = Position{Vector3{grid->startX, grid->startY, 0.f}};
    s->grabbed[ctx.w] = Grabbed{0, ENTITY_ID_PLACEHOLDER};
    s->ballPhysics[ctx.w] = BallPhysics{0, ENTITY_ID_PLACEHOLDER, ENTITY_ID_PLACEHOLDER, ENTITY_ID_PLACEHOLDER,
                                        ENTITY_ID_PLACEHOLDER, 2, 0};
    s->ballVel[ctx.w].velocity = Vector3::zero();

    /* agents: gen.cpp:182-206 */
    int32_t offensive_agent_id = ENTITY_ID_PLACEHOLDER;
    Position agent_pos_for_ball = {grid->startX, grid->startY, 0.f};
    const Vector3 TEAM0_COLOR = {0, 100, 255}, TEAM1_COLOR = {128, 0, 128}; /* constants.hpp:18-19 */
    for (int i = 0; i < NUM_AGENTS; i++) {
        int64_t a = ctx.ai(i);
        s->agentEntity[a] = Entity{0, ctx.agentId[i]};
        s->action[a] = Action{0, 0, 0, 0, 0, 0};
        s->actionMask[a] = ActionMask{0, 0, 0, 0};
        s->reset[a] = Reset{0};
        s->inbounding[a] = Inbounding{0, 1};
        s->reward[a].r = 0.f;
        s->done[a].episodeDone = 0.f;
        s->curStep[a].step = 0;
        s->orientation[a] = spawnOrientation(i);
        s->cooldown[a] = GrabCooldown{0.f};
        s->stats[a] = {0.f, 0.f};
        s->agentVel[a].velocity = Vector3::zero();
        int32_t defending_hoop_id = (i % 2 == 0) ? gameState.team0Hoop : gameState.team1Hoop;
        const Vector3 team_color = (i % 2 == 0) ? TEAM0_COLOR : TEAM1_COLOR;
        s->team[a] = Team{i % 2, team_color, defending_hoop_id};
    }
    setupAgentPositions(ctx, ctx.ballId[0], offensive_agent_id, agent_pos_for_ball);
    if (gameState.isOneOnOne == 1.f) s->grabbed[ctx.w] = {1, offensive_agent_id}; /* gen.cpp:211-213 */
}

/* gen.cpp:216-316 */
static void resetWorld(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    const GridState *grid = ctx.grid;

    if (gameState.gameClock <= 0.f && gameState.isOneOnOne == 0.f) {
        if (gameState.period < 4 || gameState.team0Score == gameState.team1Score) {
            gameState.period++;
            gameState.gameClock = TIME_PER_PERIOD;
            gameState.shotClock = 24.0f;
            gameState.liveBall = 1.0f;
            gameState.inboundingInProgress = 0.f;
        } else {
            gameState.liveBall = 0.f;
        }
    } else {
        gameState = GameState{0, 1, 1.0f, 0.0f, gameState.team0Hoop, 0.0f, gameState.team1Hoop, 0.0f,
                              TIME_PER_PERIOD, 24.0f, 0.f, 0.f, 0.0f, gameState.isOneOnOne};
    }

    Vector3 team_colors[2] = {Vector3{0, 100, 255}, Vector3{255, 0, 100}};
    int32_t basketball_id = ctx.ballId[NUM_BASKETBALLS - 1];
    int32_t offensive_agent_id = ENTITY_ID_PLACEHOLDER;
    Position agent_pos_for_ball = {grid->startX, grid->startY, 0.f};

    for (int i = 0; i < NUM_AGENTS; i++) {
        int64_t a = ctx.ai(i);
        s->action[a] = Action{0, 0, 0, 0, 0, 0};
        s->actionMask[a] = ActionMask{0, 0, 0, 0};
        s->reset[a] = Reset{0};
        s->inbounding[a] = Inbounding{0, 1};
        s->done[a].episodeDone = 1.f;
        s->curStep[a].step = 0;
        s->orientation[a] = spawnOrientation(i);
        s->cooldown[a] = GrabCooldown{0.f};
        s->stats[a] = Stats{0.f, 0.f};
        s->agentVel[a].velocity = Vector3::zero();
        int32_t defending_hoop_id = (i % 2 == 0) ? gameState.team0Hoop : gameState.team1Hoop;
        s->team[a] = Team{static_cast<int32_t>(i % 2), team_colors[i % 2], defending_hoop_id};
    }

    setupAgentPositions(ctx, basketball_id, offensive_agent_id, agent_pos_for_ball);

    for (int i = 0; i < NUM_BASKETBALLS; i++) {
        s->ballPos[ctx.w] = agent_pos_for_ball;
        s->ballPhysics[ctx.w] = BallPhysics{0, ENTITY_ID_PLACEHOLDER, ENTITY_ID_PLACEHOLDER, ENTITY_ID_PLACEHOLDER,
                                            ENTITY_ID_PLACEHOLDER, 2, 0};
        s->ballVel[ctx.w].velocity = Vector3::zero();
        Grabbed &grabbed = s->grabbed[ctx.w];
        if (gameState.isOneOnOne == 1.f) grabbed = Grabbed{1, offensive_agent_id};
        else grabbed = Grabbed{0, ENTITY_ID_PLACEHOLDER};
    }
    /* hoops' Reset/Done/CurStep (gen.cpp:310-315) are not observable */
}

/* ============================== src/game.cpp ============================== */

/* game.cpp:14-53 */
static void assignInbounder(Ctx &ctx, Position ball_pos, int32_t new_team_idx, Quat new_orientation, bool is_oob)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    float inbounder_assigned = 0.0f;
    for (int i = 0; i < NUM_AGENTS; i++) {
        int64_t a = ctx.ai(i);
        if (s->team[a].teamIndex == new_team_idx && inbounder_assigned == 0) {
            inbounder_assigned = 1;
            s->inbounding[a].imInbounding = 1;
            s->agentPos[a] = ball_pos;
            s->grabbed[ctx.w] = {1, ctx.agentId[i]};
            s->possession[a].hasBall = 1;
            s->possession[a].ballEntityID = ctx.ballId[0];
            s->orientation[a].orientation = new_orientation;
        }
    }
    if (inbounder_assigned > 0) {
        gameState.teamInPossession = (float)new_team_idx;
        gameState.inboundingInProgress = 1;
        gameState.inboundClock = 5.f;
        if (is_oob) gameState.outOfBoundsCount++;
    }
}

/* game.cpp:969-988 */
static void tick(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    s->reward[a].r = 0.f;
    if (s->reset[a].resetNow == 1) {
        s->done[a].episodeDone = 1.f;
        s->curStep[a].step = 0;
    } else {
        s->done[a].episodeDone = 0.f;
        s->curStep[a].step++;
    }
    s->cooldown[a].cooldown = fmaxf(0.f, s->cooldown[a].cooldown - 1.f);
}

/* game.cpp:489-533 */
static void actionMaskSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    ActionMask &action_mask = s->actionMask[a];
    action_mask.can_move = 1;
    action_mask.can_grab = 1;
    action_mask.can_pass = 0;
    action_mask.can_shoot = 0;
    if (s->possession[a].hasBall == 1) {
        action_mask.can_pass = 1;
        action_mask.can_shoot = 1;
    }
    if (gameState.inboundingInProgress == 1) {
        action_mask.can_shoot = 0;
        action_mask.can_grab = 0;
        if (s->inbounding[a].imInbounding == 1 && gameState.liveBall == 0) action_mask.can_move = 0;
    }
    if (s->cooldown[a].cooldown > 0.f) action_mask.can_grab = 0;
    if (s->cfg.flags & ORC_FLAG_TAG_MASK) { /* game.cpp:525-528 "FOR TAG" */
        action_mask.can_pass = 0;
        action_mask.can_grab = 0;
    }
}

/* game.cpp:410-486 */
static void moveAgentSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    const GridState *grid = ctx.grid;
    Action &action = s->action[a];
    ActionMask &action_mask = s->actionMask[a];
    Position &agent_pos = s->agentPos[a];
    InPossession &in_possession = s->possession[a];
    Inbounding &inbounding = s->inbounding[a];
    Orientation &agent_orientation = s->orientation[a];
    Attributes &attributes = s->attributes[a];
    Velocity &agent_vel = s->agentVel[a];

    if (action.rotate != 0) {
        float turn_angle = (action.rotate == 1) ? (pi / 180.f) * 6 : (pi / 180.f) * -6;
        Quat turn = Quat::angleAxis(turn_angle, Vector3{0, 0, 1});
        agent_orientation.orientation = turn * agent_orientation.orientation;
    }
    if (action_mask.can_move == 0) return;

    float move_angle = action.moveAngle * ANGLE_BETWEEN_DIRECTIONS;
    Vector3 delta_vel = Vector3{m_sinf(move_angle), -m_cosf(move_angle), 0} * attributes.quickness * action.move;

    float maximum_speed = attributes.maxSpeed;
    Vector3 agent_orientation_as_vec = agent_orientation.orientation.rotateVec(AGENT_BASE_FORWARD);
    float dot_between_orientation_and_velocity = 0.0f;
    if (agent_vel.velocity.length2() > 1e-6f)
        dot_between_orientation_and_velocity = agent_vel.velocity.normalize().dot(agent_orientation_as_vec);

    if (dot_between_orientation_and_velocity < -0.1f) {
        maximum_speed *= .1f;
        delta_vel *= .1f;
    } else if (dot_between_orientation_and_velocity <= 0.8f) {
        maximum_speed *= .7f;
        delta_vel *= .1f;
    }
    agent_vel.velocity += delta_vel;
    if (inbounding.imInbounding == 1) delta_vel.x = 0.f; /* no effect: quirk B17 */
    if (in_possession.hasBall == 1) maximum_speed *= BALL_AGENT_SLOWDOWN;
    if (agent_vel.velocity.length() > maximum_speed) agent_vel.velocity *= maximum_speed / agent_vel.velocity.length();

    float dx = agent_vel.velocity.x * TIMESTEPS_TO_SECONDS_FACTOR;
    float dy = agent_vel.velocity.y * TIMESTEPS_TO_SECONDS_FACTOR;
    float new_x = agent_pos.position.x + dx;
    float new_y = agent_pos.position.y + dy;
    new_x = clampf(new_x, 0.f, grid->width);
    new_y = clampf(new_y, 0.f, grid->height);
    /* game.cpp:473-484: wall lookup in an all-zero cell table -> always passable */
    agent_pos.position.x = new_x;
    agent_pos.position.y = new_y;
    agent_vel.velocity *= .95f;
}

/* game.cpp:164-239 */
static void grabSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    Action &action = s->action[a];
    InPossession &in_possession = s->possession[a];
    Team &team = s->team[a];
    if (s->actionMask[a].can_grab == 0 || action.grab == 0) return;
    s->cooldown[a].cooldown = 10.f;
    action.grab = 0;

    for (int b = 0; b < NUM_BASKETBALLS; b++) {
        BallPhysics &ball_physics = s->ballPhysics[ctx.w];
        Grabbed &grabbed = s->grabbed[ctx.w];
        Velocity &ball_velocity = s->ballVel[ctx.w];
        if (ball_physics.inFlight == 1) continue;

        int32_t agent_is_holding_this_ball =
            (in_possession.hasBall == 1 && grabbed.isGrabbed == 1 && grabbed.holderEntityID == ctx.agentId[i]) ? 1 : 0;
        if (agent_is_holding_this_ball == 1) {
            in_possession.ballEntityID = ENTITY_ID_PLACEHOLDER;
            in_possession.hasBall = 0;
            grabbed.holderEntityID = ENTITY_ID_PLACEHOLDER;
            grabbed.isGrabbed = 0;
            continue;
        }
        Position basketball_pos = s->ballPos[ctx.w];
        float distance_between_ball_and_player = (basketball_pos.position - s->agentPos[a].position).length();
        if (distance_between_ball_and_player <= 0.3f) {
            if (gameState.isOneOnOne == 1.f && team.teamIndex != gameState.teamInPossession) {
                ctx.worldClock().resetNow = 1.0f;
                continue;
            }
            for (int j = 0; j < NUM_AGENTS; j++) {
                InPossession &other_in_possession = s->possession[ctx.ai(j)];
                if (other_in_possession.ballEntityID == ctx.ballId[b]) {
                    other_in_possession.hasBall = 0;
                    other_in_possession.ballEntityID = ENTITY_ID_PLACEHOLDER;
                    s->cooldown[ctx.ai(j)].cooldown = SIMULATION_HZ;
                }
            }
            in_possession.hasBall = 1;
            in_possession.ballEntityID = ctx.ballId[b];
            grabbed.holderEntityID = ctx.agentId[i];
            grabbed.isGrabbed = 1;
            ball_physics.inFlight = 0;
            ball_velocity.velocity = Vector3::zero();
            ball_physics.shotByAgentID = ENTITY_ID_PLACEHOLDER;
            ball_physics.shotByTeamID = ENTITY_ID_PLACEHOLDER;
            ball_physics.shotPointValue = 2;
            gameState.teamInPossession = (float)team.teamIndex;
            gameState.liveBall = 1;
        }
    }
}

/* game.cpp:243-270 */
static void passSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    if (s->actionMask[a].can_pass == 0 || s->action[a].pass == 0) return;
    GameState &gameState = ctx.gameState();
    Grabbed &grabbed = s->grabbed[ctx.w];
    if (grabbed.holderEntityID == ctx.agentId[i]) {
        grabbed.isGrabbed = 0;
        grabbed.holderEntityID = ENTITY_ID_PLACEHOLDER;
        s->possession[a].hasBall = 0;
        s->possession[a].ballEntityID = ENTITY_ID_PLACEHOLDER;
        s->inbounding[a].imInbounding = 0;
        s->ballVel[ctx.w].velocity = s->orientation[a].orientation.rotateVec(Vector3{0.f, 0.1f, 0.f});
        gameState.inboundingInProgress = 0;
    }
}

/* game.cpp:273-407 */
static void shootSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    Action &action = s->action[a];
    Position agent_pos = s->agentPos[a]; /* by value */
    Team &team = s->team[a];
    if (s->actionMask[a].can_shoot == 0 || action.shoot == 0) return;

    Vector3 attacking_hoop_score_zone = {0.f, 0.f, 0.f};
    float scoring_radius = 0.f;
    for (int h = 0; h < NUM_HOOPS; h++) {
        if (ctx.hoopId[h] != team.defendingHoopID) {
            attacking_hoop_score_zone = s->scoringZone[ctx.hi(h)].center;
            scoring_radius = s->scoringZone[ctx.hi(h)].radius;
        }
    }
    Vector3 ideal_shot_vector = attacking_hoop_score_zone - agent_pos.position;
    float intended_direction = m_atan2f(ideal_shot_vector.x, ideal_shot_vector.y);

    float distance_to_hoop = ideal_shot_vector.length();
    float dist_stddev = DIST_DEVIATION_PER_METER * distance_to_hoop;
    float deviation_from_distance = sampleUniform(ctx, -dist_stddev, dist_stddev);

    float deviation_from_defender = 0.0f;
    float distance_to_nearest_defender = std::numeric_limits<float>::infinity();
    for (int j = 0; j < NUM_AGENTS; j++) {
        if (s->team[ctx.ai(j)].teamIndex != team.teamIndex) {
            float dist_to_def = (agent_pos.position - s->agentPos[ctx.ai(j)].position).length();
            if (dist_to_def < distance_to_nearest_defender) distance_to_nearest_defender = dist_to_def;
        }
    }
    if (distance_to_nearest_defender < 2.0f) {
        float def_stddev = (DEF_DEVIATION_PER_METER) / (distance_to_nearest_defender + 0.1f);
        deviation_from_defender = sampleUniform(ctx, -def_stddev, def_stddev);
    }
    float deviation_from_velocity = 0.0f;
    if (action.move > 0) {
        float vel_stddev = VEL_DEVIATION_FACTOR * s->agentVel[a].velocity.length();
        deviation_from_velocity = sampleUniform(ctx, -vel_stddev, vel_stddev);
    }
    float total_deviation = deviation_from_distance + deviation_from_defender + deviation_from_velocity;
    float shot_direction = intended_direction + total_deviation;
    Vector3 final_shot_vec = {m_sinf(shot_direction), m_cosf(shot_direction), 0.f};

    float shot_is_going_in = 0.0f;
    float how_far = ideal_shot_vector.dot(final_shot_vec);
    if (how_far < 0) {
        shot_is_going_in = 0.0f;
    } else {
        float closest_distance_to_hoop_sq = ideal_shot_vector.length2() - how_far * how_far;
        shot_is_going_in = (closest_distance_to_hoop_sq <= scoring_radius * scoring_radius) ? 1.0f : 0.0f;
    }

    const Vector3 base_forward = {0.0f, 1.0f, 0.0f};
    s->orientation[a].orientation = findRotationBetweenVectors(base_forward, final_shot_vec);

    GameState &gameState = ctx.gameState();
    for (int b = 0; b < NUM_BASKETBALLS; b++) {
        BallPhysics &ball_physics = s->ballPhysics[ctx.w];
        Grabbed &grabbed = s->grabbed[ctx.w];
        if (grabbed.holderEntityID == ctx.agentId[i]) {
            int32_t shot_point_value = getShotPointValue(agent_pos, attacking_hoop_score_zone);
            if (shot_is_going_in == 1) {
                ball_physics.shotIsGoingIn = 1;
                gameState.scoredBaskets++;
            } else {
                s->reward[a].r -= 1.f;
            }
            grabbed.isGrabbed = 0;
            grabbed.holderEntityID = ENTITY_ID_PLACEHOLDER;
            s->possession[a].hasBall = 0;
            s->possession[a].ballEntityID = ENTITY_ID_PLACEHOLDER;
            s->inbounding[a].imInbounding = 0.0f;
            s->ballVel[ctx.w].velocity = final_shot_vec * .1f;
            ball_physics.inFlight = 1;
            ball_physics.shotByAgentID = ctx.agentId[i];
            ball_physics.shotByTeamID = team.teamIndex;
            ball_physics.shotPointValue = shot_point_value;
            ball_physics.lastTouchedByAgentID = ctx.agentId[i];
            ball_physics.lastTouchedByTeamID = team.teamIndex;
        }
    }
}

/* game.cpp:82-125 */
static void moveBallSystem(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    Position &ball_pos = s->ballPos[ctx.w];
    Grabbed &grabbed = s->grabbed[ctx.w];
    Velocity &ball_velocity = s->ballVel[ctx.w];
    for (int i = 0; i < NUM_AGENTS; i++) {
        InPossession &in_possession = s->possession[ctx.ai(i)];
        int32_t holding =
            (in_possession.hasBall == 1 && grabbed.isGrabbed == 1 && grabbed.holderEntityID == ctx.agentId[i]) ? 1 : 0;
        if (holding == 1) ball_pos = s->agentPos[ctx.ai(i)];
    }
    if (ball_velocity.velocity.length() == 0 || grabbed.isGrabbed == 1) return;
    const GridState *grid = ctx.grid;
    float new_x = ball_pos.position.x + ball_velocity.velocity[0];
    float new_y = ball_pos.position.y + ball_velocity.velocity[1];
    float new_z = ball_pos.position.z + ball_velocity.velocity[2];
    new_x = clampf(new_x, 0.f, grid->width);
    new_y = clampf(new_y, 0.f, grid->height);
    /* game.cpp:113-120 wall lookup: always passable */
    ball_pos.position.x = new_x;
    ball_pos.position.y = new_y;
    ball_pos.position.z = new_z;
}

/* game.cpp:758-809 */
static void updateCurrentShotPercentage(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    Attributes &attributes = s->attributes[a];
    Team &agent_team = s->team[a];
    if (s->possession[a].hasBall == 0) {
        attributes.currentShotPercentage = 0.f;
        return;
    }
    Position attacking_hoop_pos = (ctx.hoopId[0] != agent_team.defendingHoopID) ? s->hoopPos[ctx.hi(0)] : s->hoopPos[ctx.hi(1)];
    float dist_to_hoop = (attacking_hoop_pos.position - s->agentPos[a].position).length();
    float distance_to_nearest_defender = std::numeric_limits<float>::infinity();
    for (int j = 0; j < NUM_AGENTS; j++) {
        if (s->team[ctx.ai(j)].teamIndex != agent_team.teamIndex) {
            float dist_to_def = (s->agentPos[a].position - s->agentPos[ctx.ai(j)].position).length();
            if (dist_to_def < distance_to_nearest_defender) distance_to_nearest_defender = dist_to_def;
        }
    }
    float dist_stddev = DIST_DEVIATION_PER_METER * dist_to_hoop;
    float def_stddev = DEF_DEVIATION_PER_METER / distance_to_nearest_defender + .0001f;
    float vel_stddev = VEL_DEVIATION_FACTOR * s->agentVel[a].velocity.length();
    float final_stddev = m_sqrtf((dist_stddev * dist_stddev / 3.f) + (def_stddev * def_stddev / 3) + (vel_stddev * vel_stddev / 3));
    float max_make_angle = m_atanf(HOOP_SCORE_ZONE_SIZE / dist_to_hoop);
    float z_score = max_make_angle / final_stddev;
    attributes.currentShotPercentage = m_erff(z_score / m_sqrtf(2.f));
}

/* game.cpp:873-953 */
static void scoreSystem(Ctx &ctx, int h)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    Position &hoop_pos = s->hoopPos[ctx.hi(h)];
    ScoringZone &scoring_zone = s->scoringZone[ctx.hi(h)];
    int32_t hoop_id = ctx.hoopId[h];

    Position &ball_pos = s->ballPos[ctx.w];
    BallPhysics &ball_physics = s->ballPhysics[ctx.w];
    Velocity &ball_velocity = s->ballVel[ctx.w];
    float distance_to_hoop = m_sqrtf((ball_pos.position.x - hoop_pos.position.x) * (ball_pos.position.x - hoop_pos.position.x) +
                                     (ball_pos.position.y - hoop_pos.position.y) * (ball_pos.position.y - hoop_pos.position.y));
    if (distance_to_hoop <= scoring_zone.radius && ball_physics.inFlight == 1.f) {
        int32_t points_scored = ball_physics.shotPointValue;
        int32_t inbounding_team_idx = 0;
        for (int j = 0; j < NUM_AGENTS; j++) {
            Team &team = s->team[ctx.ai(j)];
            Stats &agent_stats = s->stats[ctx.ai(j)];
            if (team.defendingHoopID == hoop_id) inbounding_team_idx = team.teamIndex;
            if (ctx.agentId[j] == ball_physics.shotByAgentID)
                agent_stats.points += (team.defendingHoopID == hoop_id) ? -ball_physics.shotPointValue : ball_physics.shotPointValue;
        }
        Position inbound_spot;
        if (hoop_id == gameState.team0Hoop) {
            gameState.team1Score += points_scored;
            inbound_spot = Position{{COURT_MIN_X, hoop_pos.position.y + (PIXELS_PER_METER / 60), 0.f}};
        } else {
            gameState.team0Score += points_scored;
            inbound_spot = Position{{COURT_MAX_X, hoop_pos.position.y + (PIXELS_PER_METER / 60), 0.f}};
        }
        gameState.scoredBaskets++;
        ball_physics.inFlight = 0.0f;
        ball_velocity.velocity = Vector3::zero();
        ball_physics.shotByAgentID = ENTITY_ID_PLACEHOLDER;
        ball_physics.shotByTeamID = ENTITY_ID_PLACEHOLDER;
        ball_physics.shotPointValue = 2;
        ball_physics.shotIsGoingIn = 0.0f;
        if (gameState.isOneOnOne == 0.f) {
            ball_pos = inbound_spot;
            Quat inbound_orientation = findRotationBetweenVectors(AGENT_BASE_FORWARD, findVectorToCenter(ctx, ball_pos));
            assignInbounder(ctx, inbound_spot, inbounding_team_idx, inbound_orientation, false);
        } else {
            ctx.worldClock().resetNow = 1.0f;
        }
    }
}

/* game.cpp:1055-1113 */
static void outOfBoundsSystem(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    Position &ball_pos = s->ballPos[ctx.w];
    BallPhysics &ball_physics = s->ballPhysics[ctx.w];
    if ((ball_pos.position.x < COURT_MIN_X || ball_pos.position.x > COURT_MAX_X ||
         ball_pos.position.y < COURT_MIN_Y || ball_pos.position.y > COURT_MAX_Y) &&
        gameState.inboundingInProgress == 0.f) {
        if (gameState.isOneOnOne == 1.f) {
            int off_agent = 0;
            for (int i = 1; i < NUM_AGENTS; i++)
                if (s->team[ctx.ai(i)].teamIndex == gameState.teamInPossession) off_agent = i;
            s->reward[ctx.ai(off_agent)].r -= 100.f;
            ctx.worldClock().resetNow = 1.0f;
        } else {
            ball_physics.inFlight = 0.0f;
            s->ballVel[ctx.w].velocity = Vector3::zero();
            gameState.liveBall = 0;
            int32_t new_team_idx = 1 - ball_physics.lastTouchedByTeamID;
            for (int i = 0; i < NUM_AGENTS; i++) {
                InPossession &in_possession = s->possession[ctx.ai(i)];
                Position &agent_pos = s->agentPos[ctx.ai(i)];
                if (in_possession.hasBall == 1 && in_possession.ballEntityID == ctx.ballId[0]) {
                    agent_pos.position += findVectorToCenter(ctx, agent_pos);
                    in_possession.hasBall = 0;
                    in_possession.ballEntityID = ENTITY_ID_PLACEHOLDER;
                }
            }
            Quat inbound_orientation = findRotationBetweenVectors(AGENT_BASE_FORWARD, findVectorToCenter(ctx, ball_pos));
            assignInbounder(ctx, ball_pos, new_team_idx, inbound_orientation, true);
        }
    }
}

/* game.cpp:1034-1051 */
static void updateLastTouchSystem(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    Position &ball_pos = s->ballPos[ctx.w];
    BallPhysics &ball_physics = s->ballPhysics[ctx.w];
    for (int i = 0; i < NUM_AGENTS; i++) {
        float distance = (ball_pos.position - s->agentPos[ctx.ai(i)].position).length();
        if (distance <= AGENT_SIZE_M) {
            ball_physics.lastTouchedByAgentID = ctx.agentId[i];
            ball_physics.lastTouchedByTeamID = s->team[ctx.ai(i)].teamIndex;
        }
    }
}

/* game.cpp:992-1030 */
static void clockSystem(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    if (gameState.liveBall > 0.5f && gameState.gameClock > 0.f) {
        gameState.gameClock -= TIMESTEPS_TO_SECONDS_FACTOR;
        gameState.shotClock -= TIMESTEPS_TO_SECONDS_FACTOR;
    }
    if (gameState.inboundingInProgress > 0.5f) gameState.inboundClock -= TIMESTEPS_TO_SECONDS_FACTOR;
    if (gameState.gameClock <= 0.f && gameState.liveBall > 0.5f) {
        int off_agent = 0;
        for (int i = 1; i < NUM_AGENTS; i++)
            if (s->team[ctx.ai(i)].teamIndex == gameState.teamInPossession) off_agent = i;
        s->reward[ctx.ai(off_agent)].r += 10.f;
        ctx.worldClock().resetNow = 1.0f;
    }
    if (gameState.shotClock < 0.f) gameState.shotClock = 0.f;
}

/* game.cpp:1116-1157 */
static void inboundViolationSystem(Ctx &ctx)
{
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    if (!(gameState.inboundingInProgress > 0.5f && gameState.inboundClock <= 0.f)) return;
    int32_t current_team_idx = gameState.teamInPossession;
    int32_t new_team_idx = 1 - current_team_idx;
    int32_t ball_to_turnover_id = ENTITY_ID_PLACEHOLDER;
    gameState.liveBall = 0;
    for (int i = 0; i < NUM_AGENTS; i++) {
        Inbounding &inb = s->inbounding[ctx.ai(i)];
        Position &agent_pos = s->agentPos[ctx.ai(i)];
        InPossession &poss = s->possession[ctx.ai(i)];
        if (inb.imInbounding > 0.5f) {
            ball_to_turnover_id = poss.ballEntityID;
            inb.imInbounding = 0.0f;
            poss.hasBall = 0.0f;
            poss.ballEntityID = ENTITY_ID_PLACEHOLDER;
            agent_pos.position += findVectorToCenter(ctx, agent_pos);
        }
    }
    if (ball_to_turnover_id != ENTITY_ID_PLACEHOLDER) {
        for (int b = 0; b < NUM_BASKETBALLS; b++) {
            Position &ball_pos = s->ballPos[ctx.w];
            if (ctx.ballId[b] == (int32_t)ball_to_turnover_id) {
                s->grabbed[ctx.w] = {0, ENTITY_ID_PLACEHOLDER};
                Quat inbound_orientation = findRotationBetweenVectors(AGENT_BASE_FORWARD, findVectorToCenter(ctx, ball_pos));
                assignInbounder(ctx, ball_pos, new_team_idx, inbound_orientation, true);
            }
        }
    }
}

/* game.cpp:957-967 */
static void resetSystem(Ctx &ctx)
{
    if (ctx.worldClock().resetNow == 0) return;
    resetWorld(ctx);
    ctx.worldClock().resetNow = 0.0f;
}

/* game.cpp:129-161 */
static void updatePointsWorthSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    Vector3 target_hoop_score_zone{};
    float found_target_hoop = 0.0f;
    for (int h = 0; h < NUM_HOOPS; h++) {
        if (ctx.hoopId[h] != s->team[a].defendingHoopID) {
            target_hoop_score_zone = s->scoringZone[ctx.hi(h)].center;
            found_target_hoop = 1.0f;
            break;
        }
    }
    if (found_target_hoop == 1) s->possession[a].pointsWorth = getShotPointValue(s->agentPos[a], target_hoop_score_zone);
    else s->possession[a].pointsWorth = 2;
}

/* game.cpp:537-648 */
static void agentCollisionSystem(Ctx &ctx, int ia)
{
    orc_sim *s = ctx.s;
    Position &pos_a = s->agentPos[ctx.ai(ia)];
    Orientation &orient_a = s->orientation[ctx.ai(ia)];
    Reward &reward = s->reward[ctx.ai(ia)];
    Team &team = s->team[ctx.ai(ia)];
    for (int ib = 0; ib < NUM_AGENTS; ib++) {
        if (ctx.agentId[ia] >= ctx.agentId[ib]) continue;
        Position &pos_b = s->agentPos[ctx.ai(ib)];
        Orientation &orient_b = s->orientation[ctx.ai(ib)];

        Vector3 center_a = pos_a.position;
        Vector3 fwd_a = orient_a.orientation.rotateVec({0, 1, 0});
        Vector3 right_a = {fwd_a.y, -fwd_a.x, 0};
        Vector3 half_width_a = right_a * (AGENT_SHOULDER_WIDTH / 2.0f);
        Vector3 half_depth_a = fwd_a * (AGENT_DEPTH / 2.0f);
        Vector3 vertices_a[4] = {center_a - half_depth_a + half_width_a, center_a - half_depth_a - half_width_a,
                                 center_a + half_depth_a - half_width_a, center_a + half_depth_a + half_width_a};

        Vector3 center_b = pos_b.position;
        Vector3 fwd_b = orient_b.orientation.rotateVec({0, 1, 0});
        Vector3 right_b = {fwd_b.y, -fwd_b.x, 0};
        Vector3 half_width_b = right_b * (AGENT_SHOULDER_WIDTH / 2.0f);
        Vector3 half_depth_b = fwd_b * (AGENT_DEPTH / 2.0f);
        Vector3 vertices_b[4] = {center_b - half_depth_b + half_width_b, center_b - half_depth_b - half_width_b,
                                 center_b + half_depth_b - half_width_b, center_b + half_depth_b + half_width_b};

        Vector3 axes[4] = {right_a.normalize(), fwd_a.normalize(), right_b.normalize(), fwd_b.normalize()};

        bool is_colliding = true;
        float min_overlap = std::numeric_limits<float>::max();
        Vector3 mtv_axis = {0, 0, 0};
        for (int j = 0; j < 4; j++) {
            const Vector3 &axis = axes[j];
            Projection p_a = projectRectangle(vertices_a, axis);
            Projection p_b = projectRectangle(vertices_b, axis);
            if (!projectionsOverlap(p_a, p_b)) {
                is_colliding = false;
                break;
            } else {
                float overlap = fminf(p_a.max, p_b.max) - fmaxf(p_a.min, p_b.min);
                if (overlap < min_overlap) {
                    min_overlap = overlap;
                    mtv_axis = axis;
                }
            }
        }
        if (is_colliding) {
            GameState gameState = ctx.gameState();
            Reward &entity_b_reward = s->reward[ctx.ai(ib)];
            if (gameState.teamInPossession == team.teamIndex) {
                reward.r -= 10;
                entity_b_reward.r += 10;
                ctx.worldClock().resetNow = 1.0f;
            }
            float penetration_depth = min_overlap;
            Vector3 correction_vec = mtv_axis;
            Vector3 center_to_center = center_b - center_a;
            if (dot(center_to_center, correction_vec) < 0) correction_vec = -correction_vec;
            pos_a.position -= correction_vec * penetration_depth * 0.5f;
            pos_b.position += correction_vec * penetration_depth * 0.5f;
        }
    }
}

/* game.cpp:651-755 */
static void hardCodeDefenseSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    Team &defender_team = s->team[a];
    Position &defender_pos = s->agentPos[a];
    Action &defender_action = s->action[a];
    Attributes &defender_attributes = s->attributes[a];

    if (gameState.teamInPossession == defender_team.teamIndex) {
        defender_action.move = 0;
        return;
    }
    defender_action.grab = 1.f;
    Vector3 guarding_pos{};
    int32_t found_offender = 0;
    for (int j = 0; j < NUM_AGENTS; j++) {
        InPossession &offender_in_possession = s->possession[ctx.ai(j)];
        Position &offender_pos = s->agentPos[ctx.ai(j)];
        if (offender_in_possession.hasBall == 1 && found_offender == 0) {
            for (int h = 0; h < NUM_HOOPS; h++) {
                if (defender_team.defendingHoopID == ctx.hoopId[h]) {
                    Position &hoop_pos = s->hoopPos[ctx.hi(h)];
                    Vector3 hoop_direction = hoop_pos.position - offender_pos.position;
                    if (hoop_direction.length2() > 1e-6f)
                        guarding_pos = offender_pos.position + GUARDING_DISTANCE * hoop_direction.normalize();
                    else
                        guarding_pos = offender_pos.position;
                    found_offender = 1;
                }
            }
        }
    }
    if (found_offender == 0) {
        defender_action.move = 0;
        return;
    }
    Vector3 current_target = defender_attributes.currentTargetPosition;
    Vector3 ideal_target = guarding_pos;
    float interpolation_factor = defender_attributes.reactionSpeed * TIMESTEPS_TO_SECONDS_FACTOR;
    defender_attributes.currentTargetPosition = current_target + (ideal_target - current_target) * interpolation_factor;

    Vector3 move_vector = defender_attributes.currentTargetPosition - defender_pos.position;
    if (move_vector.length2() < 0.01f) {
        defender_action.move = 0;
        return;
    }
    const Vector3 move_directions[] = {{0.f, -1.f, 0.f}, {1.f, -1.f, 0.f}, {1.f, 0.f, 0.f},  {1.f, 1.f, 0.f},
                                       {0.f, 1.f, 0.f},  {-1.f, 1.f, 0.f}, {-1.f, 0.f, 0.f}, {-1.f, -1.f, 0.f}};
    Vector3 desired_dir = move_vector.normalize();
    float max_dot = -2.f;
    int32_t best_move_angle = 0;
    for (int32_t k = 0; k < 8; ++k) {
        float current_dot = dot(desired_dir, move_directions[k].normalize());
        if (current_dot > max_dot) {
            max_dot = current_dot;
            best_move_angle = k;
        }
    }
    defender_action.move = 1;
    defender_action.moveAngle = best_move_angle;

    Vector3 defender_orientation_as_vec = s->orientation[a].orientation.rotateVec(AGENT_BASE_FORWARD);
    float angle_between_vectors = m_acosf(clampf(defender_orientation_as_vec.dot(move_vector.normalize()), -1.f, 1.f));
    if (angle_between_vectors > pi / 8.f) {
        float direction_cross = defender_orientation_as_vec.x * move_vector.y - defender_orientation_as_vec.y * move_vector.x;
        if (direction_cross < 0) defender_action.rotate = -1.f;
        else if (direction_cross > 0) defender_action.rotate = 1.f;
        else defender_action.rotate = 0.f;
    } else {
        defender_action.rotate = 0.f;
    }
}

/* game.cpp:1162-1172 */
struct AgentObservationData {
    int32_t id, teamID;
    Position pos;
    Orientation orient;
    Velocity velocity;
    InPossession in_pos;
    Inbounding inb;
    GrabCooldown cooldown;
    Attributes attributes;
};

/* game.cpp:1175-1461 */
static void fillObservationsSystem(Ctx &ctx, int me)
{
    orc_sim *s = ctx.s;
    int64_t a = ctx.ai(me);
    float *obs = s->obs[a].observationsArray;
    const GameState &gameState = ctx.gameState();
    int32_t idx = 0;
    auto fill_vec3 = [&](const Vector3 &v) { obs[idx++] = v.x; obs[idx++] = v.y; obs[idx++] = v.z; };
    auto fill_quat = [&](const Quat &q) { obs[idx++] = q.w; obs[idx++] = q.x; obs[idx++] = q.y; obs[idx++] = q.z; };

    Position &agent_pos = s->agentPos[a];
    Orientation &agent_orientation = s->orientation[a];
    InPossession &in_possession = s->possession[a];
    Inbounding &inbounding = s->inbounding[a];
    Team &agent_team = s->team[a];
    GrabCooldown &grab_cooldown = s->cooldown[a];
    Velocity &agent_vel = s->agentVel[a];
    Attributes &agent_attributes = s->attributes[a];

    Position ball_pos = s->ballPos[ctx.w];
    BallPhysics ball_phys = s->ballPhysics[ctx.w];
    Grabbed ball_grabbed = s->grabbed[ctx.w];
    Velocity ball_velocity = s->ballVel[ctx.w];

    Position hoop_positions[NUM_HOOPS];
    int32_t hoop_ids[NUM_HOOPS];
    for (int h = 0; h < NUM_HOOPS; h++) {
        hoop_positions[h] = s->hoopPos[ctx.hi(h)];
        hoop_ids[h] = ctx.hoopId[h];
    }

    AgentObservationData all_agents[NUM_AGENTS];
    int agent_idx = 0;
    int32_t inbounder_id = -1;
    for (int i = 0; i < NUM_AGENTS; i++) {
        int64_t o = ctx.ai(i);
        all_agents[agent_idx++] = {ctx.agentId[i], s->team[o].teamIndex, s->agentPos[o], s->orientation[o], s->agentVel[o],
                                   s->possession[o], s->inbounding[o], s->cooldown[o], s->attributes[o]};
        if (s->inbounding[o].imInbounding > 0.5f) inbounder_id = ctx.agentId[i];
    }

    obs[idx++] = gameState.gameClock;
    obs[idx++] = gameState.shotClock;
    obs[idx++] = gameState.period;
    obs[idx++] = gameState.inboundingInProgress;
    obs[idx++] = gameState.inboundClock;
    if (agent_team.teamIndex == 0) {
        obs[idx++] = gameState.team0Score;
        obs[idx++] = gameState.team1Score;
    } else {
        obs[idx++] = gameState.team1Score;
        obs[idx++] = gameState.team0Score;
    }
    fill_vec3(ball_pos.position);
    fill_vec3(ball_velocity.velocity);
    obs[idx++] = ball_grabbed.isGrabbed;
    obs[idx++] = ball_phys.inFlight;
    obs[idx++] = (float)ball_phys.shotPointValue;
    obs[idx++] = (float)ball_phys.lastTouchedByTeamID;

    Position attacking_hoop_pos = (hoop_ids[0] != agent_team.defendingHoopID) ? hoop_positions[0] : hoop_positions[1];
    Position defending_hoop_pos = (hoop_ids[0] == agent_team.defendingHoopID) ? hoop_positions[0] : hoop_positions[1];
    fill_vec3(attacking_hoop_pos.position);
    fill_vec3(defending_hoop_pos.position);

    fill_vec3(agent_pos.position);
    fill_vec3(Vector3::zero());
    obs[idx++] = 0.f;
    fill_quat(agent_orientation.orientation);
    Vector3 agent_orient_as_vec = agent_orientation.orientation.rotateVec(AGENT_BASE_FORWARD);
    fill_vec3(agent_orient_as_vec);
    if (agent_vel.velocity.length2() > 1e-6f) fill_vec3(agent_vel.velocity.normalize());
    else fill_vec3(Vector3::zero());
    obs[idx++] = agent_vel.velocity.length();
    float dot_between_orient_and_vel = 0.f;
    if (agent_vel.velocity.length2() > 1e-6f) dot_between_orient_and_vel = agent_vel.velocity.normalize().dot(agent_orient_as_vec);
    obs[idx++] = dot_between_orient_and_vel;
    obs[idx++] = (dot_between_orient_and_vel <= 0.8f) ? 0.1f : 1.f;
    Vector3 dir_to_hoop = (attacking_hoop_pos.position - agent_pos.position);
    float dist_to_hoop = dir_to_hoop.length();
    if (dist_to_hoop > 1e-6f) fill_vec3(dir_to_hoop.normalize());
    else fill_vec3(Vector3{0.f, 0.f, 0.f});
    obs[idx++] = dist_to_hoop;
    Vector3 dir_to_ball = (ball_pos.position - agent_pos.position);
    float dist_to_ball = dir_to_ball.length();
    if (dist_to_ball > 1e-6f) fill_vec3(dir_to_ball.normalize());
    else fill_vec3(Vector3{0.f, 0.f, 0.f});
    obs[idx++] = dist_to_ball;
    obs[idx++] = inbounding.imInbounding;
    obs[idx++] = grab_cooldown.cooldown;
    obs[idx++] = agent_attributes.maxSpeed;
    obs[idx++] = agent_attributes.quickness;
    obs[idx++] = agent_attributes.shooting;
    obs[idx++] = agent_attributes.freeThrowPercentage;
    obs[idx++] = agent_attributes.reactionSpeed;
    obs[idx++] = agent_attributes.currentShotPercentage;
    obs[idx++] = (float)in_possession.pointsWorth;
    obs[idx++] = in_possession.hasBall;

    int teammate_count = 0, opponent_count = 0;
    const int max_teammates = (NUM_AGENTS / 2) - 1;
    const int max_opponents = NUM_AGENTS / 2;
    auto fill_other = [&](const AgentObservationData &o, Vector3 vec_to_agent, const Position &hoop_for_dir) {
        /* game.cpp:1339-1372 (teammate) and 1380-1421 (opponent) are the same
         * sequence; only the hoop differs (attacking vs observer's defending) */
        fill_vec3(o.pos.position);
        if (vec_to_agent.length2() > 1e-6f) fill_vec3(vec_to_agent.normalize());
        else fill_vec3(Vector3{0.f, 0.f, 0.f});
        obs[idx++] = vec_to_agent.length();
        fill_quat(o.orient.orientation);
        Vector3 o_orient_as_vec = o.orient.orientation.rotateVec(AGENT_BASE_FORWARD);
        fill_vec3(o_orient_as_vec);
        if (o.velocity.velocity.length2() > 1e-6f) fill_vec3(o.velocity.velocity.normalize());
        else fill_vec3(Vector3::zero());
        obs[idx++] = o.velocity.velocity.length();
        float o_dot = 0.f;
        if (o.velocity.velocity.length2() > 1e-6f) o_dot = o.velocity.velocity.normalize().dot(o_orient_as_vec);
        obs[idx++] = o_dot;
        obs[idx++] = (o_dot <= 0.8f) ? 0.1f : 1.f;
        Vector3 o_dir_to_hoop = (hoop_for_dir.position - o.pos.position);
        float o_dist_to_hoop = o_dir_to_hoop.length();
        if (o_dist_to_hoop > 1e-6f) fill_vec3(o_dir_to_hoop.normalize());
        else fill_vec3(Vector3{0.f, 0.f, 0.f});
        obs[idx++] = o_dist_to_hoop;
        Vector3 o_dir_to_ball = (ball_pos.position - o.pos.position);
        float o_dist_to_ball = o_dir_to_ball.length();
        if (o_dist_to_ball > 1e-6f) fill_vec3(o_dir_to_ball.normalize());
        else fill_vec3(Vector3{0.f, 0.f, 0.f});
        obs[idx++] = o_dist_to_ball;
        obs[idx++] = o.inb.imInbounding;
        obs[idx++] = o.cooldown.cooldown;
        obs[idx++] = o.attributes.maxSpeed;
        obs[idx++] = o.attributes.quickness;
        obs[idx++] = o.attributes.shooting;
        obs[idx++] = o.attributes.freeThrowPercentage;
        obs[idx++] = o.attributes.reactionSpeed;
        obs[idx++] = o.attributes.currentShotPercentage;
        obs[idx++] = (float)o.in_pos.pointsWorth;
        obs[idx++] = o.in_pos.hasBall;
    };
    for (int i = 0; i < agent_idx; i++) {
        if (all_agents[i].id == ctx.agentId[me]) continue;
        Vector3 vec_to_agent = all_agents[i].pos.position - agent_pos.position;
        if (all_agents[i].teamID == agent_team.teamIndex) {
            if (teammate_count < max_teammates) {
                fill_other(all_agents[i], vec_to_agent, attacking_hoop_pos);
                teammate_count++;
            }
        } else {
            if (opponent_count < max_opponents) {
                fill_other(all_agents[i], vec_to_agent, defending_hoop_pos);
                opponent_count++;
            }
        }
    }
    constexpr int agent_feature_size = 3 + 3 + 1 + 4 + 3 + 3 + 1 + 1 + 1 + 3 + 1 + 3 + 11;
    for (int i = teammate_count; i < max_teammates; i++)
        for (int j = 0; j < agent_feature_size; j++) obs[idx++] = 0.f;
    for (int i = opponent_count; i < max_opponents; i++)
        for (int j = 0; j < agent_feature_size; j++) obs[idx++] = 0.f;

    for (int i = 0; i < agent_idx; i++) obs[idx++] = (all_agents[i].id == (int32_t)ball_grabbed.holderEntityID) ? 1.f : 0.f;
    for (int i = agent_idx; i < NUM_AGENTS; i++) obs[idx++] = 0.f;
    for (int i = 0; i < agent_idx; i++) obs[idx++] = (all_agents[i].id == inbounder_id) ? 1.f : 0.f;
    for (int i = agent_idx; i < NUM_AGENTS; i++) obs[idx++] = 0.f;

    if (!(idx < 128)) abort(); /* game.cpp:1455 assert */
    for (; idx < 128; idx++) obs[idx] = 0.f;
}

/* game.cpp:811-870 */
static void rewardSystem(Ctx &ctx, int i)
{
    int64_t a = ctx.ai(i);
    orc_sim *s = ctx.s;
    GameState &gameState = ctx.gameState();
    Reward &reward = s->reward[a];
    Team &team = s->team[a];
    int other = 0;
    for (int j = 0; j < NUM_AGENTS; j++)
        if (ctx.agentId[j] != ctx.agentId[i]) other = j;
    float dist_to_other_agent = (s->agentPos[ctx.ai(other)].position - s->agentPos[a].position).length();
    if (team.teamIndex == gameState.teamInPossession) {
        if (gameState.gameClock > 5.f) {
            BallPhysics &ball_physics = s->ballPhysics[ctx.w];
            if (ball_physics.shotByAgentID == ctx.agentId[i] && ball_physics.shotIsGoingIn == 1)
                reward.r += ball_physics.shotPointValue;
            else if (ball_physics.shotByAgentID == ctx.agentId[i] && ball_physics.shotIsGoingIn == 0 && ball_physics.inFlight == 1)
                reward.r -= 1;
            reward.r += s->attributes[a].currentShotPercentage;
        }
    } else {
        reward.r -= 1.f;
        reward.r += m_expf(-0.4f * dist_to_other_agent);
    }
}

/* game.cpp:1463-1526: the task graph, one world */
static void stepWorld(orc_sim *s, int64_t w)
{
    Ctx ctx = makeCtx(s, w);
    for (int i = 0; i < NUM_AGENTS; i++) tick(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) actionMaskSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) moveAgentSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) grabSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) passSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) shootSystem(ctx, i);
    moveBallSystem(ctx);
    for (int i = 0; i < NUM_AGENTS; i++) updateCurrentShotPercentage(ctx, i);
    for (int h = 0; h < NUM_HOOPS; h++) scoreSystem(ctx, h);
    outOfBoundsSystem(ctx);
    updateLastTouchSystem(ctx);
    clockSystem(ctx);
    inboundViolationSystem(ctx);
    resetSystem(ctx);
    for (int i = 0; i < NUM_AGENTS; i++) updatePointsWorthSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) agentCollisionSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) hardCodeDefenseSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) fillObservationsSystem(ctx, i);
    for (int i = 0; i < NUM_AGENTS; i++) rewardSystem(ctx, i);
    s->rngCount[w] = ctx.rng.count;
}

/* SURVEY 8(d) synthetic action stream: Threefry keyed by (seed, agent),
 * counter (global world, step).  Ranges: scripts/env.py:96-102. */
static inline Action syntheticAction(uint32_t seed, uint32_t world, uint32_t step, uint32_t agent)
{
    RandKey r = threefry2x32(RandKey{seed, agent}, world, step);
    Action act;
    act.move = (int32_t)(r.a & 1u);
    act.moveAngle = (int32_t)((r.a >> 1) & 7u);
    act.rotate = (int32_t)((r.a >> 8) % 3u);
    act.grab = (int32_t)(r.b & 1u);
    act.pass = (int32_t)((r.b >> 1) & 1u);
    act.shoot = (int32_t)((r.b >> 2) & 1u);
    return act;
}

template <typename F>
static void parallelFor(int64_t n, int num_threads, F fn)
{
    if (num_threads <= 1 || n < 2) {
        fn(0, n);
        return;
    }
    if (num_threads > n) num_threads = (int)n;
    std::vector<std::thread> th;
    int64_t chunk = (n + num_threads - 1) / num_threads;
    for (int t = 0; t < num_threads; t++) {
        int64_t lo = t * chunk, hi = lo + chunk > n ? n : lo + chunk;
        if (lo >= hi) break;
        th.emplace_back([=] { fn(lo, hi); });
    }
    for (auto &t : th) t.join();
}

} // namespace orc

extern "C" {

orc_sim *orc_create(const orc_config *cfg)
{
    orc_sim *s = new orc_sim();
    s->cfg = *cfg;
    int64_t N = s->N = cfg->num_worlds;
    /* bindings.cpp:28-51 */
    int32_t cells_per_meter = 1;
    s->grid.startX = cfg->start_x;
    s->grid.startY = cfg->start_y;
    s->grid.width = (float)cfg->discrete_x / (float)cells_per_meter;
    s->grid.height = (float)cfg->discrete_y / (float)cells_per_meter;
    s->grid.cellsPerMeter = cells_per_meter;
    s->grid.discreteWidth = (int32_t)cfg->discrete_x;
    s->grid.discreteHeight = (int32_t)cfg->discrete_y;
    /* mgr.cpp:155 initKey(randSeed); sim.cpp:89 rng(split_i(initRandKey, 0, 0)) */
    s->worldKey = split_i(initKey(cfg->rand_seed), 0, 0);

    s->reset.assign(N * 2, Reset{0});
    s->gameState.assign(N, GameState{});
    s->action.assign(N * 2, Action{});
    s->actionMask.assign(N * 2, ActionMask{});
    s->agentPos.assign(N * 2, Position{});
    s->obs.assign(N * 2, Observations{});
    s->reward.assign(N * 2, Reward{});
    s->done.assign(N * 2, Done{});
    s->agentEntity.assign(N * 2, Entity{});
    s->possession.assign(N * 2, InPossession{});
    s->orientation.assign(N * 2, Orientation{});
    s->team.assign(N * 2, Team{});
    s->stats.assign(N * 2, Stats{});
    s->ballPos.assign(N, Position{});
    s->ballPhysics.assign(N, BallPhysics{});
    s->ballEntity.assign(N, Entity{});
    s->grabbed.assign(N, Grabbed{});
    s->ballVel.assign(N, Velocity{});
    s->hoopPos.assign(N * 2, Position{});
    s->agentVel.assign(N * 2, Velocity{});
    s->inbounding.assign(N * 2, Inbounding{});
    s->cooldown.assign(N * 2, GrabCooldown{});
    s->attributes.assign(N * 2, Attributes{});
    s->curStep.assign(N * 2, CurStep{});
    s->worldClock.assign(N, WorldClock{0});
    s->rngCount.assign(N, 0u);
    s->scoringZone.assign(N * 2, ScoringZone{});

    for (int64_t w = 0; w < N; w++) {
        Ctx ctx = makeCtx(s, w);
        generateWorld(ctx); /* sim.cpp:95 */
        s->rngCount[w] = ctx.rng.count;
    }
    return s;
}

void orc_destroy(orc_sim *s) { delete s; }

void orc_step(orc_sim *s, int num_threads)
{
    parallelFor(s->N, num_threads, [s](int64_t lo, int64_t hi) {
        for (int64_t w = lo; w < hi; w++) stepWorld(s, w);
    });
}

void orc_fill_actions(orc_sim *s, uint32_t seed, int64_t step, int64_t world_offset)
{
    for (int64_t w = 0; w < s->N; w++)
        for (int i = 0; i < NUM_AGENTS; i++)
            s->action[w * 2 + i] = syntheticAction(seed, (uint32_t)(w + world_offset), (uint32_t)step, (uint32_t)i);
}

void orc_run_synthetic(orc_sim *s, uint32_t seed, int64_t first_step, int64_t nsteps, int64_t world_offset, int num_threads)
{
    parallelFor(s->N, num_threads, [=](int64_t lo, int64_t hi) {
        for (int64_t st = first_step; st < first_step + nsteps; st++)
            for (int64_t w = lo; w < hi; w++) {
                for (int i = 0; i < NUM_AGENTS; i++)
                    s->action[w * 2 + i] = syntheticAction(seed, (uint32_t)(w + world_offset), (uint32_t)st, (uint32_t)i);
                stepWorld(s, w);
            }
    });
}

void *orc_tensor(orc_sim *s, int slot, int *dtype, int64_t dims[4], int *ndim)
{
    int64_t N = s->N;
    auto set = [&](void *p, int dt, int nd, int64_t d1, int64_t d2) {
        *dtype = dt;
        *ndim = nd;
        dims[0] = N;
        dims[1] = d1;
        dims[2] = d2;
        dims[3] = 0;
        return p;
    };
    /* shapes and dtypes: mgr.cpp:317-445 */
    switch (slot) {
    case ORC_SLOT_RESET: return set(s->reset.data(), ORC_DTYPE_I32, 3, 2, 1);
    case ORC_SLOT_GAME_STATE: return set(s->gameState.data(), ORC_DTYPE_F32, 2, 14, 0);
    case ORC_SLOT_ACTION: return set(s->action.data(), ORC_DTYPE_I32, 3, 2, 6);
    case ORC_SLOT_ACTION_MASK: return set(s->actionMask.data(), ORC_DTYPE_I32, 3, 2, 4);
    case ORC_SLOT_AGENT_POS: return set(s->agentPos.data(), ORC_DTYPE_F32, 3, 2, 3);
    case ORC_SLOT_OBSERVATIONS: return set(s->obs.data(), ORC_DTYPE_F32, 3, 2, 128);
    case ORC_SLOT_REWARD: return set(s->reward.data(), ORC_DTYPE_F32, 2, 2, 0);
    case ORC_SLOT_DONE: return set(s->done.data(), ORC_DTYPE_F32, 2, 2, 0);
    case ORC_SLOT_AGENT_ENTITY_ID: return set(s->agentEntity.data(), ORC_DTYPE_I32, 2, 2, 0); /* quirk B10 */
    case ORC_SLOT_AGENT_POSSESSION: return set(s->possession.data(), ORC_DTYPE_I32, 3, 2, 3);
    case ORC_SLOT_ORIENTATION: return set(s->orientation.data(), ORC_DTYPE_F32, 3, 2, 4);
    case ORC_SLOT_TEAM_DATA: return set(s->team.data(), ORC_DTYPE_I32, 3, 2, 5);
    case ORC_SLOT_AGENT_STATS: return set(s->stats.data(), ORC_DTYPE_I32, 3, 2, 2);
    case ORC_SLOT_BASKETBALL_POS: return set(s->ballPos.data(), ORC_DTYPE_F32, 3, 1, 3);
    case ORC_SLOT_BALL_PHYSICS: return set(s->ballPhysics.data(), ORC_DTYPE_I32, 3, 1, 7);
    case ORC_SLOT_BALL_ENTITY_ID: return set(s->ballEntity.data(), ORC_DTYPE_I32, 2, 1, 0); /* quirk B10 */
    case ORC_SLOT_BALL_GRABBED: return set(s->grabbed.data(), ORC_DTYPE_I32, 3, 1, 2);
    case ORC_SLOT_BALL_VELOCITY: return set(s->ballVel.data(), ORC_DTYPE_F32, 3, 1, 3);
    case ORC_SLOT_HOOP_POS: return set(s->hoopPos.data(), ORC_DTYPE_F32, 3, 2, 3);
    case ORC_SLOT_AGENT_VELOCITY: return set(s->agentVel.data(), ORC_DTYPE_F32, 3, 2, 3);
    case ORC_SLOT_INBOUNDING: return set(s->inbounding.data(), ORC_DTYPE_I32, 3, 2, 2);
    case ORC_SLOT_GRAB_COOLDOWN: return set(s->cooldown.data(), ORC_DTYPE_F32, 2, 2, 0);
    case ORC_SLOT_ATTRIBUTES: return set(s->attributes.data(), ORC_DTYPE_F32, 3, 2, 9);
    case ORC_SLOT_CUR_STEP: return set(s->curStep.data(), ORC_DTYPE_I32, 2, 2, 0);
    case ORC_SLOT_WORLD_CLOCK: return set(s->worldClock.data(), ORC_DTYPE_I32, 1, 0, 0);
    case ORC_SLOT_RNG_COUNT: return set(s->rngCount.data(), ORC_DTYPE_I32, 1, 0, 0);
    default: return nullptr;
    }
}

void orc_checksum(orc_sim *s, uint64_t *out)
{
    /* canonical order: the true byte extent of each column for this world
     * (entity-id columns contribute their full 8-byte entities). */
    struct Col { const void *p; size_t bytes_per_world; };
    const Col cols[] = {
        {s->reset.data(), 8}, {s->gameState.data(), 56}, {s->action.data(), 48}, {s->actionMask.data(), 32},
        {s->agentPos.data(), 24}, {s->obs.data(), 1024}, {s->reward.data(), 8}, {s->done.data(), 8},
        {s->agentEntity.data(), 16}, {s->possession.data(), 24}, {s->orientation.data(), 32}, {s->team.data(), 40},
        {s->stats.data(), 16}, {s->ballPos.data(), 12}, {s->ballPhysics.data(), 28}, {s->ballEntity.data(), 8},
        {s->grabbed.data(), 8}, {s->ballVel.data(), 12}, {s->hoopPos.data(), 24},
        {s->agentVel.data(), 24}, {s->inbounding.data(), 16}, {s->cooldown.data(), 8}, {s->attributes.data(), 72},
        {s->curStep.data(), 8}, {s->worldClock.data(), 4}, {s->rngCount.data(), 4},
    };
    for (int64_t w = 0; w < s->N; w++) {
        uint64_t h = 0xcbf29ce484222325ull;
        for (const Col &c : cols) {
            const uint32_t *p = (const uint32_t *)((const char *)c.p + (size_t)w * c.bytes_per_world);
            for (size_t i = 0; i < c.bytes_per_world / 4; i++) h = (h ^ p[i]) * 0x100000001b3ull;
        }
        out[w] = h;
    }
}

void orc_threefry2x32(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t out[2])
{
    RandKey r = threefry2x32(RandKey{k0, k1}, c0, c1);
    out[0] = r.a;
    out[1] = r.b;
}

float orc_rng_uniform(uint32_t seed, uint32_t draw_index)
{
    RNG rng{split_i(initKey(seed), 0, 0), draw_index};
    return rng.sampleUniform();
}

/* Quat::rotateVec (madrona_prims.h) and the compile-time court constants (constants.hpp:67-93), for the
 * fixtures derived from the reference's own Python (scripts/viewer.py:59-65 rotate_vec, src/constants.py). */
void orc_rotate_vec(const float q[4], const float v[3], float out[3])
{
    Vector3 r = Quat{q[0], q[1], q[2], q[3]}.rotateVec(Vector3{v[0], v[1], v[2]});
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}

void orc_court_constants(float out[12])
{
    const float c[12] = {COURT_LENGTH_M, COURT_WIDTH_M, WORLD_WIDTH_M, WORLD_HEIGHT_M, COURT_MIN_X, COURT_MAX_X, COURT_MIN_Y,
                         COURT_MAX_Y, HOOP_FROM_BASELINE_M, ARC_RADIUS_M, CORNER_3_FROM_SIDELINE_M, CORNER_3_LENGTH_FROM_BASELINE_M};
    for (int i = 0; i < 12; i++) out[i] = c[i];
}

/* fn: 0 sinf, 1 cosf, 2 atanf, 3 atan2f(x, y), 4 acosf, 5 expf, 6 erff -- whichever arithmetic this
 * build uses (include/bb_math.h spec, or glibc with -DBB_USE_LIBM). */
void orc_math_eval(int fn, const float *x, const float *y, float *out, int64_t n)
{
    for (int64_t i = 0; i < n; i++) {
        float r = 0.f;
        switch (fn) {
        case 0: r = m_sinf(x[i]); break;
        case 1: r = m_cosf(x[i]); break;
        case 2: r = m_atanf(x[i]); break;
        case 3: r = m_atan2f(x[i], y[i]); break;
        case 4: r = m_acosf(x[i]); break;
        case 5: r = m_expf(x[i]); break;
        case 6: r = m_erff(x[i]); break;
        }
        out[i] = r;
    }
}

} // extern "C"
