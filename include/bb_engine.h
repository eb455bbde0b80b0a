/*
 * bb_engine.h — C ABI of the B200 basketball step engine (libbball.so).
 *
 * This is the drop-in boundary for the reference's Manager
 * (/root/reference/src/mgr.hpp:17-81, mgr.cpp) as seen through its nanobind
 * module (/root/reference/src/bindings.cpp:14-101).  Plain pointers and sizes
 * only; no torch or C++ types.  Every entry point names the reference
 * interface it replaces.  The CUDA path is the only path: there is no CPU
 * executor behind this ABI and every call fails loudly (non-zero status +
 * bb_last_error()) when no sm_100 device is usable.
 */
#ifndef BB_ENGINE_H
#define BB_ENGINE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bb_engine bb_engine;

/* Export slots 0..18 are the reference's ExportID enum, in its order
 * (/root/reference/src/types.hpp:10-42); shapes/dtypes are those declared by the
 * 19 Manager::*Tensor() accessors (mgr.cpp:317-445).  Slots >= 100 expose the
 * components the reference keeps internal (debug / parity tests). */
enum bb_slot {
    BB_SLOT_RESET = 0,            /* i32 [N,2,1]   mgr.cpp:317 */
    BB_SLOT_GAME_STATE = 1,       /* f32 [N,14]    mgr.cpp:323 */
    BB_SLOT_ACTION = 2,           /* i32 [N,2,6]   mgr.cpp:332 */
    BB_SLOT_ACTION_MASK = 3,      /* i32 [N,2,4]   mgr.cpp:338 */
    BB_SLOT_AGENT_POS = 4,        /* f32 [N,2,3]   mgr.cpp:344 */
    BB_SLOT_OBSERVATIONS = 5,     /* f32 [N,2,128] mgr.cpp:350 */
    BB_SLOT_REWARD = 6,           /* f32 [N,2]     mgr.cpp:364 */
    BB_SLOT_DONE = 7,             /* f32 [N,2]     mgr.cpp:371 */
    BB_SLOT_AGENT_ENTITY_ID = 8,  /* i32 [N,2]     mgr.cpp:385 (8-byte entities underneath) */
    BB_SLOT_AGENT_POSSESSION = 9, /* i32 [N,2,3]   mgr.cpp:378 */
    BB_SLOT_ORIENTATION = 10,     /* f32 [N,2,4]   mgr.cpp:357 */
    BB_SLOT_TEAM_DATA = 11,       /* i32 [N,2,5]   mgr.cpp:392 */
    BB_SLOT_AGENT_STATS = 12,     /* i32 [N,2,2]   mgr.cpp:399 */
    BB_SLOT_BASKETBALL_POS = 13,  /* f32 [N,1,3]   mgr.cpp:406 */
    BB_SLOT_BALL_PHYSICS = 14,    /* i32 [N,1,7]   mgr.cpp:413 */
    BB_SLOT_BALL_ENTITY_ID = 15,  /* i32 [N,1]     mgr.cpp:427 */
    BB_SLOT_BALL_GRABBED = 16,    /* i32 [N,1,2]   mgr.cpp:420 */
    BB_SLOT_BALL_VELOCITY = 17,   /* f32 [N,1,3]   mgr.cpp:433 */
    BB_SLOT_HOOP_POS = 18,        /* f32 [N,2,3]   mgr.cpp:441 */
    BB_NUM_EXPORTS = 19,

    BB_SLOT_AGENT_VELOCITY = 100, /* f32 [N,2,3]  Velocity      types.hpp:86 */
    BB_SLOT_INBOUNDING = 101,     /* i32 [N,2,2]  Inbounding    types.hpp:144 */
    BB_SLOT_GRAB_COOLDOWN = 102,  /* f32 [N,2]    GrabCooldown  types.hpp:157 */
    BB_SLOT_ATTRIBUTES = 103,     /* f32 [N,2,9]  Attributes    types.hpp:176 */
    BB_SLOT_CUR_STEP = 104,       /* i32 [N,2]    CurStep       types.hpp:219 */
    BB_SLOT_WORLD_CLOCK = 105,    /* i32 [N]      WorldClock    types.hpp:69 */
    BB_SLOT_RNG_COUNT = 106       /* i32 [N]      RNG draw counter (sim.hpp:32) */
};

enum bb_dtype { BB_DTYPE_I32 = 0, BB_DTYPE_F32 = 1 };

enum bb_flags {
    /* game.cpp:525-528: the "FOR TAG" override that forces can_pass = can_grab = 0.
     * Set = reference behaviour.  Clearing it is a test hook that makes the
     * grab / pass / steal / inbound systems reachable. */
    BB_FLAG_TAG_MASK = 1u << 0
};

/* Manager::Config (mgr.hpp:19-25) + the GridState fields the constructor in
 * bindings.cpp:18-51 derives from its Python arguments. */
typedef struct bb_config {
    int64_t num_worlds;         /* Manager::Config::numWorlds */
    int64_t discrete_x;         /* bindings.cpp:19  -> GridState::width  = discrete_x / 1 */
    int64_t discrete_y;         /* bindings.cpp:20  -> GridState::height = discrete_y / 1 */
    float start_x;              /* bindings.cpp:21  -> GridState::startX */
    float start_y;              /* bindings.cpp:22  -> GridState::startY */
    int64_t max_episode_length; /* stored, never read by the step (sim.hpp:36) */
    uint32_t rand_seed;         /* Manager::Config::randSeed; bindings.cpp:37 passes 0 */
    int32_t gpu_id;             /* Manager::Config::gpuID */
    uint32_t flags;             /* bb_flags; reference behaviour = BB_FLAG_TAG_MASK */
} bb_config;

/* Manager::Manager (mgr.cpp:236-239, Impl::init 147-234): allocates every column on
 * cfg->gpu_id, runs world generation (Sim::Sim -> generateWorld, sim.cpp:86-96) on
 * the device.  Returns NULL on failure (see bb_last_error). */
bb_engine *bb_create(const bb_config *cfg);

/* Manager::~Manager (mgr.cpp:241). */
void bb_destroy(bb_engine *e);

/* Manager::step (mgr.cpp:243-246): one pass of the 19-system task graph
 * (game.cpp:1463-1526) over every world, as one fused kernel enqueued on
 * `cuda_stream` (a cudaStream_t; 0 = legacy default stream).  Asynchronous. */
int bb_step(bb_engine *e, void *cuda_stream);

/* n consecutive Manager::step() calls with no host work in between (the loop of scripts/run.py:16-20, or the
 * reference's own `for _ in range(n): sim.step()`), as ONE kernel launch on `cuda_stream`: worlds never interact, so each
 * warp takes its 16 worlds through all n steps in a row (bb_step_loop_kernel: no launch, drain or grid dependency between
 * steps; every step still writes its full state, observation rows, reward and done).  Results are bit-identical to n bb_step
 * calls.  The Action column is whatever the caller and hardCodeDefenseSystem left in it, exactly as with n bb_step calls.
 * BB_STEP_N_GRAPH=1 (or a tuning launch shape without a looping instantiation) selects round 1's form instead: the n step
 * kernels captured once per n into a CUDA graph with programmatic-dependent-launch edges, replayed with one graph launch. */
int bb_step_n(bb_engine *e, int32_t n, void *cuda_stream);

/* bb_step_n with a recorded action stream: step s takes its Action rows from actions_dev + s * N*2*6 (device, i32
 * [n,N,2,6]) -- n x (`action_tensor.copy_(actions[s]); step()`, scripts/env.py:143,155) as one launch of the looping
 * kernel.  Extension; bit-identical to n bb_step_with_actions calls. */
int bb_step_n_with_actions(bb_engine *e, int32_t n, const int32_t *actions_dev, void *cuda_stream);

/* bb_step, but Action is first taken from `actions_dev` (device, i32 [N,2,6]) —
 * the fused equivalent of `action_tensor.copy_(actions); step()`
 * (scripts/env.py:143,155).  actions_dev == NULL behaves like bb_step. */
int bb_step_with_actions(bb_engine *e, const int32_t *actions_dev, void *cuda_stream);

/* Host-buffer step: H2D copy of `actions_host` (i32 [N,2,6], may be NULL to keep
 * the device Action column), one step, D2H copy of whichever outputs are non-NULL
 * (reward f32 [N,2], done f32 [N,2], observations f32 [N,2,128]); returns after
 * the copies completed.  Buffers should be pinned for full PCIe rate. */
int bb_step_host(bb_engine *e, const int32_t *actions_host, float *reward_host, float *done_host,
                 float *obs_host);

/* bb_step_host / bb_serve_step with PACKED action rows: u8 [N,2,6], one byte per Action field (every field the
 * policy can emit is < 8: scripts/env.py:96-102; `rotate = -1`, which only hardCodeDefenseSystem writes, has no packed
 * form).  A declared extension for host-driven loops: the step reads a quarter of the bytes over PCIe (12 B instead of
 * 48 B per world); the Action column still ends up as the int32 rows the reference exports.  Pinned buffers only. */
int bb_step_host_u8(bb_engine *e, const uint8_t *actions_host, float *reward_host, float *done_host,
                    float *obs_host);

/* Persistent-server variant of bb_step_host for latency-bound host-driven loops (no reference
 * counterpart).  bb_serve_begin leaves one resident kernel on the device; each bb_serve_step rings
 * a doorbell in mapped pinned memory, the kernel runs one Manager::step() (action rows read from
 * `actions_host`, reward / done / observations mirrored into the non-NULL host buffers, device
 * columns updated as in bb_step) and publishes completion back; the call returns when the results
 * are in host memory.  No kernel launch and no stream wait per step.  All buffers must be pinned
 * (cudaHostAlloc / torch pin_memory) and must STAY pinned until bb_serve_end: their device aliases are looked up once per
 * serve session (a combination {actions, reward, done, obs} travels to a 64-slot device-side table on its first use and is
 * named by its slot afterwards; beyond 64 combinations slots are re-used in turn).  Between begin and end do not call bb_step* / setters, and
 * do not write the exported columns from other streams while a step is in flight.  The server
 * leaves the GPU by itself after 2 s without a step (BB_SERVE_IDLE_MS) and is restarted
 * transparently by the next bb_serve_step; device-wide synchronisation (cudaDeviceSynchronize,
 * cudaFree) stalls until then, so call bb_serve_end first. */
int bb_serve_begin(bb_engine *e);
int bb_serve_step(bb_engine *e, const int32_t *actions_host, float *reward_host, float *done_host,
                  float *obs_host);
int bb_serve_step_u8(bb_engine *e, const uint8_t *actions_host, float *reward_host, float *done_host,
                     float *obs_host);
int bb_serve_end(bb_engine *e);
/* Tuning aid: GPU-side nanoseconds between "doorbell seen" and "last block done" of the last served step. */
long long bb_serve_last_gpu_ns(bb_engine *e);
/* Tuning aid: eight GPU-side time marks (ns, relative to "doorbell seen") of the last served step: after the acquire
 * fence, after the release to the other blocks, block 0's tile start / end, the last block's tile end, its counter
 * increment, and just before the completion word leaves. */
int bb_serve_last_marks(bb_engine *e, long long out[8]);

/* Manager::*Tensor() (mgr.cpp:317-445) / Impl::exportTensor (mgr.cpp:121-126):
 * device pointer + declared dtype/shape of a column.  The engine owns the memory
 * for its lifetime; callers may read and write through the pointer between steps
 * (scripts/env.py:143,147,179,182 do).  Returns 0 on success. */
int bb_export(bb_engine *e, int slot, void **dev_ptr, int *dtype, int64_t dims[4], int *ndim,
              int *device);

/* Manager::setAction (mgr.cpp:270-293).  The reference dereferences the column on
 * the host, which is only valid for its CPU executor; here it is a 24-byte H2D
 * write ordered after previously enqueued work.  Out-of-range indices print the
 * reference's "ERROR: Invalid indices!" line and are ignored (mgr.cpp:288-292). */
int bb_set_action(bb_engine *e, int32_t world_idx, int32_t agent_idx, int32_t move_speed,
                  int32_t move_angle, int32_t rotate, int32_t grab, int32_t pass, int32_t shoot);

/* Manager::triggerReset (mgr.cpp:297-311): Reset.resetNow = 1 for both agents of
 * the world; out-of-range world_idx is silently ignored (mgr.cpp:303). */
int bb_trigger_reset(bb_engine *e, int32_t world_idx);

/* Synthetic benchmark input: fills `actions_dev` (or the Action
 * column when NULL) for both agents of every world with the
 * (seed, world_offset + world, step, agent)-keyed Threefry draw. */
int bb_fill_actions(bb_engine *e, int32_t *actions_dev, uint32_t seed, int64_t step,
                    int64_t world_offset, void *cuda_stream);

/* 64-bit FNV-1a per world over the complete state in canonical column order
 * (slots 0..18, then 100..106); out_dev is device memory, u64 [N]. */
int bb_checksum(bb_engine *e, uint64_t *out_dev, void *cuda_stream);

/* Test hook: evaluates one libm call site of the step path ON THE DEVICE with
 * the arithmetic spec include/bb_math.h; fn: 0 sinf, 1 cosf, 2 atanf, 3 atan2f(x, y), 4 acosf,
 * 5 expf, 6 erff.  x / y / out are device pointers, f32 [n] (y only read for atan2f). */
int bb_math_eval(int fn, const float *x_dev, const float *y_dev, float *out_dev, int64_t n, void *cuda_stream);

/* Number of kernels this engine has launched since creation (bench.py's
 * gpu_launches claim). */
int64_t bb_launch_count(const bb_engine *e);

int64_t bb_num_worlds(const bb_engine *e);

/* Last error message of the calling thread ("" if none). */
const char *bb_last_error(void);

#ifdef __cplusplus
}
#endif
#endif /* BB_ENGINE_H */
