/*
 * bb_oracle.h — C ABI of the CPU oracle.
 *
 * TEST INFRASTRUCTURE ONLY.  This is a CPU restatement of the reference's
 * per-world step (src/game.cpp, src/gen.cpp, src/helper.cpp of
 * davidj24/madrona_basketball @ 250dbe49) used as the checker for the CUDA
 * path.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may load it; the product (madrona_basketball_b200/)
 * never does.
 *
 * PARITY UNPINNED: the reference cannot be compiled here (its Madrona
 * submodule is absent) and ships no tests or golden vectors, so the Madrona
 * primitives restated in oracle/madrona_prims.h (RNG wrapping, Vector3/Quat
 * formulas, entity-id numbering) follow SURVEY.md App. C, not verified source.
 */
#ifndef BB_ORACLE_H
#define BB_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct orc_sim orc_sim;

/* Slot numbers 0..18 are the reference ExportID enum (src/types.hpp:10-42).
 * Slots >= 100 expose components the reference keeps internal, so that tests
 * can compare the complete world state. */
enum {
    ORC_SLOT_RESET = 0, ORC_SLOT_GAME_STATE, ORC_SLOT_ACTION, ORC_SLOT_ACTION_MASK,
    ORC_SLOT_AGENT_POS, ORC_SLOT_OBSERVATIONS, ORC_SLOT_REWARD, ORC_SLOT_DONE,
    ORC_SLOT_AGENT_ENTITY_ID, ORC_SLOT_AGENT_POSSESSION, ORC_SLOT_ORIENTATION,
    ORC_SLOT_TEAM_DATA, ORC_SLOT_AGENT_STATS, ORC_SLOT_BASKETBALL_POS,
    ORC_SLOT_BALL_PHYSICS, ORC_SLOT_BALL_ENTITY_ID, ORC_SLOT_BALL_GRABBED,
    ORC_SLOT_BALL_VELOCITY, ORC_SLOT_HOOP_POS, ORC_NUM_EXPORTS,

    ORC_SLOT_AGENT_VELOCITY = 100, /* f32 [N,2,3] */
    ORC_SLOT_INBOUNDING,           /* i32 [N,2,2] */
    ORC_SLOT_GRAB_COOLDOWN,        /* f32 [N,2]   */
    ORC_SLOT_ATTRIBUTES,           /* f32 [N,2,9] */
    ORC_SLOT_CUR_STEP,             /* i32 [N,2]   */
    ORC_SLOT_WORLD_CLOCK,          /* i32 [N]     */
    ORC_SLOT_RNG_COUNT,            /* i32 [N]     */
};

enum { ORC_DTYPE_I32 = 0, ORC_DTYPE_F32 = 1 };

/* flags */
enum {
    ORC_FLAG_TAG_MASK = 1u << 0, /* game.cpp:525-528 "FOR TAG" mask override (reference default) */
};

typedef struct orc_config {
    int64_t num_worlds;
    int64_t discrete_x, discrete_y; /* bindings.cpp:19-20 */
    float start_x, start_y;         /* bindings.cpp:21-22 */
    uint32_t rand_seed;             /* bindings.cpp:37 fixes this to 0 */
    uint32_t flags;                 /* ORC_FLAG_* ; reference behaviour = ORC_FLAG_TAG_MASK */
} orc_config;

orc_sim *orc_create(const orc_config *cfg);
void orc_destroy(orc_sim *sim);

/* One Manager::step() (mgr.cpp:243-246) over all worlds, on num_threads threads. */
void orc_step(orc_sim *sim, int num_threads);

/* Host pointer to a component column in the reference's export layout. */
void *orc_tensor(orc_sim *sim, int slot, int *dtype, int64_t dims[4], int *ndim);

/* Synthetic action stream of SURVEY.md section 8(d): overwrites Action for both agents
 * of every world with the (seed, global world, step, agent)-keyed draw. */
void orc_fill_actions(orc_sim *sim, uint32_t seed, int64_t step, int64_t world_offset);

/* nsteps x (fill_actions; step), each thread running its own world slice
 * without barriers.  Used for the timed CPU baseline. */
void orc_run_synthetic(orc_sim *sim, uint32_t seed, int64_t first_step, int64_t nsteps,
                       int64_t world_offset, int num_threads);

/* 64-bit FNV-1a over every 32-bit word of a world's state in canonical order
 * (slots 0..18 then 100..106; see bb_oracle.cpp:orc_checksum). */
void orc_checksum(orc_sim *sim, uint64_t *out_per_world);

/* Known-answer hooks for tests. */
void orc_threefry2x32(uint32_t k0, uint32_t k1, uint32_t c0, uint32_t c1, uint32_t out[2]);
float orc_rng_uniform(uint32_t seed, uint32_t draw_index);
void orc_rotate_vec(const float q[4], const float v[3], float out[3]);   /* Quat{w,x,y,z}.rotateVec(v) */
void orc_court_constants(float out[12]); /* COURT_LENGTH, COURT_WIDTH, WORLD_W, WORLD_H, MIN_X, MAX_X, MIN_Y, MAX_Y,
                                            HOOP_FROM_BASELINE, ARC_RADIUS, CORNER_3_FROM_SIDELINE, CORNER_3_LENGTH */
/* Elementwise libm call sites of SURVEY App. E in this build's arithmetic.
 * fn: 0 sinf, 1 cosf, 2 atanf, 3 atan2f(x, y), 4 acosf, 5 expf, 6 erff. */
void orc_math_eval(int fn, const float *x, const float *y, float *out, int64_t n);

#ifdef __cplusplus
}
#endif
#endif
