/*
 * mgb200.h — C ABI of the B200-native batched Minigrid engine (libmgb200.so).
 *
 * The reference (Farama Minigrid 2.3.1 snapshot) is pure Python and has no FFI of its own; these
 * entry points are what a binding for its hot path binds instead of calling the Python methods
 * cited beside each function (file:line relative to the reference root).  Plain pointers and sizes
 * only: no torch types, no C++ types, no exceptions cross this boundary.  Every function returns
 * MG_OK (0) or a negative MG_E* code; mg_last_error() gives the text of the last failure on the
 * calling thread.
 *
 * Conventions
 *   "dev"  pointer = device memory on the engine's device, owned by the caller (e.g. a torch tensor)
 *   "host" pointer = ordinary (preferably pinned) host memory, owned by the caller
 *   stream         = a cudaStream_t passed as void* (NULL = the legacy default stream)
 *   Nothing synchronises unless its comment says so.  A handle is not thread-safe; distinct handles are independent.
 *
 * Layout of one env in HBM (struct-of-arrays uint8):
 *   cells[4][plane_stride]  planes 0 type, 1 color, 2 state, 3 aux; cell (x,y) of a W-wide env at y*W+x
 *                           (minigrid/core/grid.py:72).  None is (1,0,0,0) as Grid.encode writes it (grid.py:260-263).
 *                           aux: bits 0-3 type / bits 4-6 colour of Box.contains (core/world_object.py:274-276),
 *                           bit 7 = "is the mission target" (identity test, envs/keycorridor.py:133).
 *   plane_stride            = mg_plane_stride(): max over classes of W*H rounded up to 16.
 */
#ifndef MGB200_H
#define MGB200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MG_OK            0
#define MG_EINVAL      (-1)   /* bad argument */
#define MG_ECUDA       (-2)   /* CUDA runtime error (text in mg_last_error) */
#define MG_ENOMEM      (-3)
#define MG_ESTATE      (-4)   /* call order (e.g. step before layouts were uploaded) */
#define MG_EACTION     (-5)   /* an action outside 0..6 was seen (ValueError, minigrid_env.py:579-580) */

#define MG_VIEW          7    /* agent_view_size of every registered env (minigrid_env.py:42) */
#define MG_OBS_BYTES   147    /* 7*7*3 (minigrid_env.py:72-77) */
#define MG_MAX_CLASSES  16

#define MG_RULE_NONE           0
#define MG_RULE_PICKUP_TARGET  1   /* KeyCorridorEnv.step, envs/keycorridor.py:129-137; UnlockPickupEnv.step, envs/unlockpickup.py:98-106;
                                      BlockedUnlockPickupEnv.step, envs/blockedunlockpickup.py:108-116: picking up the object whose
                                      aux byte has bit 7 set pays and terminates */
#define MG_RULE_UNLOCK         2   /* UnlockEnv.step, envs/unlock.py:89-97: a toggle while the DOOR whose aux byte has bit 7 set is
                                      open pays and terminates */

#define MG_RULE_FETCH          3   /* FetchEnv.step, envs/fetch.py:162-176: carrying anything terminates; it pays iff the carried object's aux
                                      byte has bit 7 set (set it on every object of the target's type and colour) */
#define MG_RULE_GOTO           4   /* GoToDoorEnv.step, envs/gotodoor.py:133-150; GoToObjectEnv.step, envs/gotoobject.py:145-162: toggle
                                      terminates; done terminates and pays iff the agent is 4-adjacent to target_pos = the first cell whose
                                      aux byte has bit 7 set in the UPLOADED (initial) layout */
#define MG_RULE_REDBLUE        5   /* RedBlueDoorEnv.step, envs/redbluedoors.py:105-133; both doors carry aux bit 7 */
#define MG_RULE_MEMORY         6   /* MemoryEnv.step, envs/memory.py:153-165: pickup acts as toggle; the agent standing on success_pos (empty
                                      cell with aux 0x80) ends the episode with the reward, on failure_pos (aux 0x81) without */
#define MG_RULE_PUTNEAR        7   /* PutNearEnv.step, envs/putnear.py:176-201: picking up anything but the object to move (aux 0x80) ends the
                                      episode; a drop while carrying ends it and pays iff it happened within one cell of target_pos = the
                                      cell whose aux byte is 0x81 in the UPLOADED layout */

#define MG_RULE_DYNOBS         8   /* DynamicObstaclesEnv.step, envs/dynamicobstacles.py:136-167: actions >= 3 act as 0; before the base step every
                                      obstacle is re-placed in the 3x3 box around it with draws from the env's own PCG64 generator
                                      (mg_dyn_record); moving forward into a cell that was neither empty nor the goal BEFORE the
                                      obstacles moved ends the episode with reward -1 */

#define MG_AUTORESET_OFF        0  /* like the bare reference env: the caller resets (tests/test_envs.py:96-100) */
#define MG_AUTORESET_SAME_STEP  1  /* done env restarts inside the same step from pool slot (layout_id + R) % P, R = mg_reset_stride();
                                      obs/direction returned are the new episode's first ones */

#define MG_AUTORESET_NEXT_STEP  2  /* gymnasium >= 1.0's default: the step that ends an episode returns its last observation; the
                                      NEXT step of that env ignores its action and resets it (reward 0, flags clear, first
                                      observation of the new episode from pool slot (layout_id + R) % P) */

typedef struct mg_engine mg_engine;

/* One env class = the constructor constants step()/gen_obs() read (minigrid_env.py:28-111). */
typedef struct {
    int32_t width, height;        /* cells */
    int32_t max_steps;            /* truncation, minigrid_env.py:582-583 */
    int32_t see_through_walls;    /* skips process_vis, minigrid_env.py:611-616 */
    int32_t rule;                 /* MG_RULE_* */
    int32_t reserved[3];
} mg_class_desc;

typedef struct {
    int32_t device;               /* CUDA device ordinal */
    int32_t num_envs;             /* N on this device */
    int32_t num_classes;          /* <= MG_MAX_CLASSES */
    int32_t autoreset;            /* MG_AUTORESET_* */
    const mg_class_desc* classes; /* host, num_classes entries */
} mg_config;

/* Start pose of one pool layout. */
typedef struct { uint8_t x, y, dir, class_id; } mg_layout_agent;

/* Per-env dynamic state as mg_get_state/mg_set_state exchange it (16 bytes). */
typedef struct {
    uint8_t  x, y, dir, class_id;               /* agent_pos, agent_dir (minigrid_env.py:106-107) */
    uint8_t  carry_type, carry_color, carry_state, carry_aux;   /* carrying (:111); carry_type 0 = None */
    uint16_t step_count; uint16_t pad;          /* :146; pad bit 0 = next-step autoreset owed (MG_AUTORESET_NEXT_STEP) */
    uint32_t layout_id;                          /* pool slot this episode started from */
} mg_agent_state;

/* DynamicObstacles only: the env's numpy Generator(PCG64) (gymnasium seeding.np_random, minigrid_env.py:125) as
 * `bit_generator.state` spells it, and `self.obstacles` in creation order (envs/dynamicobstacles.py:129-132).  64 bytes. */
typedef struct {
    uint64_t state_lo, state_hi, inc_lo, inc_hi;   /* 128-bit LCG state and increment */
    uint32_t has_uint32, uinteger;                 /* buffered upper half of the last 64-bit draw */
    uint8_t  n_obstacles, pad[7];
    uint8_t  ox[8], oy[8];                         /* obstacle cells; unused slots 0xFF */
} mg_dyn_record;

const char* mg_last_error(void);
int  mg_version(void);

/* ---- lifetime ---- */
int  mg_create(mg_engine** out, const mg_config* cfg);      /* replaces gym.make x N (minigrid/__init__.py:23) */
void mg_destroy(mg_engine* h);
int  mg_num_envs(const mg_engine* h);
int  mg_plane_stride(const mg_engine* h);

/* ---- layouts: output of the host generators (_gen_grid, minigrid_env.py:132) ----
 * cells: host uint8[P][4][plane_stride]; agents: host [P].  Replaces any earlier pool.  Synchronous. */
int  mg_upload_layouts(mg_engine* h, const uint8_t* cells, const mg_layout_agent* agents, int32_t P);

/* DynamicObstacles (MG_RULE_DYNOBS): generator + obstacle record of every pool layout, host [P], P as uploaded.  Call after
 * mg_upload_layouts (which drops earlier records).  reset / autoreset copy pool record -> env record; the step kernel advances
 * the env's generator exactly as the reference's step does (same draws, same order).  Synchronous.
 * mg_get_dyn / mg_set_dyn exchange the N per-env records (generator state after the steps so far).  Synchronous. */
int  mg_upload_dyn(mg_engine* h, const mg_dyn_record* dyn_host, int32_t P);
int  mg_get_dyn(mg_engine* h, mg_dyn_record* dyn_host);
int  mg_set_dyn(mg_engine* h, const mg_dyn_record* dyn_host);

/* ---- NoDeath (minigrid/wrappers.py:799-872) as a mode of the step kernels ----
 * type_mask bit t set = OBJECT_TO_IDX type t is a "no death" type: a termination while moving forward into such a cell, or
 * while standing in one, is cancelled (so no autoreset either) and death_cost is added to the reward.  0 = off (default).
 * Goal (8) cannot be in the mask (wrappers.py:831).  Applies to every env of the handle from the next step on. */
int  mg_set_no_death(mg_engine* h, uint32_t type_mask, double death_cost);

/* ---- reset (MiniGridEnv.reset minus _gen_grid, minigrid_env.py:119-157) ----
 * env e <- pool[layout_ids[e]] where mask == NULL or mask[e] != 0; step_count = 0, carrying = None.
 * Writes the first observation of every env (reset or not): image dev uint8[N][7][7][3], direction dev uint8[N].
 * image/direction may be NULL. */
int  mg_reset(mg_engine* h, const int32_t* layout_ids_dev, const uint8_t* mask_dev,
              uint8_t* image_dev, uint8_t* direction_dev, void* stream);

/* ---- step (MiniGridEnv.step + gen_obs, minigrid_env.py:520-645; KeyCorridorEnv.step) ----
 * T consecutive steps in one launch; actions dev uint8[T][N]; every output is [T][N](...), any may be NULL.
 * reward is the fp64 value of _reward() (minigrid_env.py:235-240), 0.0 otherwise.  T = 1 is a plain step(). */
int  mg_step_many(mg_engine* h, const uint8_t* actions_dev, int32_t T,
                  uint8_t* image_dev, uint8_t* direction_dev, double* reward_dev,
                  uint8_t* terminated_dev, uint8_t* truncated_dev, void* stream);
/* mg_step_many + gymnasium's info["final_observation"] for MG_AUTORESET_SAME_STEP: where an episode ended in step t (and the
 * env was reset inside that step) final_image[t][e] / final_direction[t][e] receive the finished episode's LAST observation;
 * other entries are left untouched (terminated | truncated is the mask).  final_* may be NULL. */
int  mg_step_many_final(mg_engine* h, const uint8_t* actions_dev, int32_t T,
                        uint8_t* image_dev, uint8_t* direction_dev, double* reward_dev,
                        uint8_t* terminated_dev, uint8_t* truncated_dev,
                        uint8_t* final_image_dev, uint8_t* final_direction_dev, void* stream);
int  mg_step(mg_engine* h, const uint8_t* actions_dev,
             uint8_t* image_dev, uint8_t* direction_dev, double* reward_dev,
             uint8_t* terminated_dev, uint8_t* truncated_dev, void* stream);

/* Host-buffer convenience used by bindings without device memory of their own: copies actions H2D,
 * steps T times, copies all outputs D2H into the host buffers, and synchronises.  Same layouts as above. */
int  mg_step_many_host(mg_engine* h, const uint8_t* actions_host, int32_t T,
                       uint8_t* image_host, uint8_t* direction_host, double* reward_host,
                       uint8_t* terminated_host, uint8_t* truncated_host);

/* ---- errors raised by the kernels ----
 * An action outside 0..6 leaves that env untouched apart from step_count += 1 (minigrid_env.py:523 runs
 * before :579-580) and records (env, action).  Synchronises the device.  Returns MG_EACTION if one was
 * recorded since the last poll (and clears it), else MG_OK.  Also reported here: MG_EINVAL for a layout id outside the
 * pool (mg_reset), and MG_ECUDA if a step CTA waited ~2 s for the previous step launch of its envs — step launches of one
 * engine must be ordered by the caller (one stream, or events between streams); consecutive ones overlap at their edges
 * (a CTA waits for the same CTA of the previous launch, not for the whole grid), everything else follows stream order. */
int  mg_poll_error(mg_engine* h, int64_t* env_index, int32_t* action);

/* ---- state exchange (MiniGridEnv.grid / agent_pos / agent_dir / carrying / step_count) ----  Synchronous.
 * cells host uint8[N][4][plane_stride], agents host [N]; either may be NULL. */
int  mg_get_state(mg_engine* h, uint8_t* cells_host, mg_agent_state* agents_host);
int  mg_set_state(mg_engine* h, const uint8_t* cells_host, const mg_agent_state* agents_host);

/* ---- RGB rendering of the current state ----
 * What RGBImgPartialObsWrapper / RGBImgObsWrapper (minigrid/wrappers.py:337-381, 284-335) put into obs["image"] and
 * what env.render() returns (minigrid_env.py:711-737), for all N envs in one launch.  Tiles are rasterised on the device
 * once per tile_size (Grid.render_tile, core/grid.py:145-198) and cached in the handle; frames are bit-identical to
 * the reference's numpy rasteriser.  tile_size 1..64 (wrappers default 8, env.render() 32).
 *   mg_render_pov:  get_frame(agent_pov=True)  (minigrid_env.py:647-661)  image dev uint8[N][7*ts][7*ts][3]
 *   mg_render_full: get_frame(highlight, ts)   (minigrid_env.py:663-709)  image dev uint8[N][stride], env e's frame
 *                   uint8[H*ts][W*ts][3] of its class at e*stride, stride = mg_render_full_stride() (largest class,
 *                   rounded up to 16; equals H*W*ts*ts*3 for a single-class batch whose frame size is a multiple of 16)
 *   mg_render_atlas: copies the tile atlas to the host, uint8[1280][ts][ts][3]; tile index =
 *                   slot*256 + highlight*128 + code, slot 0 = no agent, 1..4 = agent facing 0..3,
 *                   code = type | color<<4 with closed / locked doors as type 13 / 14.  Synchronous. */
int  mg_render_pov(mg_engine* h, int32_t tile_size, uint8_t* image_dev, void* stream);
int  mg_render_full(mg_engine* h, int32_t tile_size, int32_t highlight, uint8_t* image_dev, void* stream);
int64_t mg_render_full_stride(const mg_engine* h, int32_t tile_size);
int  mg_render_atlas(mg_engine* h, int32_t tile_size, uint8_t* tiles_host);

/* ---- observation / reward wrappers as kernels over the batch (minigrid/wrappers.py) ----
 * Each call produces, for every env, what the wrapper's observation() / step() returns for the CURRENT state.
 *   mg_obs_fully     FullyObsWrapper (wrappers.py:384-423): image dev uint8[N][plane_stride*3]; env e's uint8[W][H][3]
 *                    (Grid.encode's [i][j] order) starts at e*plane_stride*3; the agent's cell is (10, 0, agent_dir).
 *   mg_obs_symbolic  SymbolicObsWrapper (:728-766): image dev int64[N][plane_stride*3], env e's int64[W][H][3] =
 *                    (x, y, type index, -1 for an empty slot, 10 at the agent).
 *   mg_obs_onehot    OneHotPartialObsWrapper (:220-287): image dev uint8[num_images][7][7][3] (e.g. mg_step_many's
 *                    output, num_images = T*N) -> out dev uint8[num_images][7][7][20]; out 16-byte aligned.
 *   mg_obs_flat      FlatObsWrapper (:553-623): out dev float32[N][147 + mission_len] = image bytes as floats, then
 *                    mission_table[mission_of_layout[layout_id of env]] (the one-hot string, float32[mission_len],
 *                    built by the host per distinct mission; mission_len = 28*maxStrLen).
 *   mg_bonus         ActionBonus (kind MG_BONUS_ACTION, :70-121) / PositionBonus (MG_BONUS_POSITION, :124-187), called
 *                    after a step: counts[key of the env's state after the step] += 1; reward[e] += 1/sqrt(count).
 *                    counts dev uint32, zero-initialised by the caller, [N][plane_stride*28] (key ((y*W+x)*4+dir)*7+action)
 *                    or [N][plane_stride] (key y*W+x); actions dev uint8[N] (ignored for MG_BONUS_POSITION). */
/* Pool slot each env's current episode started from (its mission is the slot's: minigrid_env.py:132,141), dev int32[N]. */
int  mg_layout_ids(mg_engine* h, int32_t* ids_dev, void* stream);
/* ViewSizeWrapper (wrappers.py:626-670): the partial observation at another agent_view_size (odd, 3..15) for the CURRENT
 * state: image dev uint8[N][V][V][3].  (The step kernels emit the registered envs' 7x7 view; this is the wrapper's path.) */
int  mg_obs_view(mg_engine* h, int32_t view_size, uint8_t* image_dev, void* stream);
/* DirectionObsWrapper (wrappers.py:673-725): out dev float64[N] = obs["goal_direction"] for the CURRENT agent positions.
 * goal_position dev int32[N][2] is the tuple the wrapper caches at its first reset (wrappers.py:699-709) — NOTE it holds
 * (int(idx / height), idx % width) of the first Goal in grid order, i.e. (row, column); the slope is
 * (goal_position[1] - agent_y) / (goal_position[0] - agent_x) exactly as the reference computes it.  angle != 0: np.arctan of
 * the slope (CUDA atan, within 1 ulp of numpy's); slope itself is bit-exact (one IEEE division; inf / nan as numpy). */
int  mg_obs_goal_direction(mg_engine* h, const int32_t* goal_position_dev, int32_t angle, double* out_dev, void* stream);
/* StochasticActionWrapper.action (wrappers.py:769-796) for a batch: out[e] = uniform[e] < prob ? actions[e] :
 * (random_action >= 0 ? random_action : replacement[e]).  The reference draws `uniform` from the process-global numpy state
 * and `replacement` from the env's generator (np_random.integers(0, 6)); the caller supplies the draws (dev float64[N],
 * dev uint8[N], replacement may be NULL when random_action >= 0). */
int  mg_stochastic_actions(mg_engine* h, const uint8_t* actions_dev, const double* uniform_dev, const uint8_t* replacement_dev,
                           double prob, int32_t random_action, uint8_t* out_dev, void* stream);
#define MG_BONUS_ACTION    0
#define MG_BONUS_POSITION  1
int  mg_obs_fully(mg_engine* h, uint8_t* image_dev, void* stream);
int  mg_obs_symbolic(mg_engine* h, int64_t* image_dev, void* stream);
int  mg_obs_onehot(mg_engine* h, const uint8_t* image_dev, int64_t num_images, uint8_t* out_dev, void* stream);
int  mg_obs_flat(mg_engine* h, const uint8_t* image_dev, const float* mission_table_dev, const int32_t* mission_of_layout_dev,
                 int32_t mission_len, float* out_dev, void* stream);
int  mg_bonus(mg_engine* h, int32_t kind, const uint8_t* actions_dev, uint32_t* counts_dev, double* reward_dev, void* stream);

/* Launch shape of the step kernel for this handle: resident CTAs per SM, threads per CTA, dynamic shared memory. */
int  mg_step_occupancy(mg_engine* h, int32_t* blocks_per_sm, int32_t* threads_per_block, int64_t* smem_bytes);

/* Number of kernels this handle has launched so far (bench.py's gpu_launches). */
int64_t mg_launch_count(const mg_engine* h);

/* Autoreset stride R: a finished env restarts from pool slot (layout_id + R) % P.  R = num_envs mod P moved up to the next
 * value coprime to P (1 if P divides num_envs), so every env walks through all P layouts before one repeats.  The reference
 * (MiniGridEnv.reset, minigrid_env.py:119-157) draws a fresh layout from the env's running generator instead; a pool has to
 * be at least as large as the number of distinct layouts the caller wants to see.  DynamicObstacles: the generator record
 * restarts from the pool slot's, it is not carried across episodes. */
int  mg_reset_stride(const mg_engine* h);

#ifdef __cplusplus
}
#endif
#endif /* MGB200_H */
