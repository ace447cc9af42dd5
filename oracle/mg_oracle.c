/*
 * mg_oracle.c — CPU restatement of the reference hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This file is the parity oracle for the CUDA engine.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may load it.  The product package
 * (minigrid_dynamicprogramming_b200/) never imports, links or executes anything in oracle/.
 *
 * It follows the reference (Farama Minigrid 2.3.1 snapshot, /root/reference) LITERALLY —
 * slice, then rotate_left (agent_dir+1) times, then the row-sequential process_vis, then
 * encode — and deliberately does NOT use the closed forms the CUDA kernels use, so that
 * a mistake in the closed form cannot hide in both sides.
 *
 * Parity pin: tests/test_oracle_golden.py replays tests/golden/ fixtures, which were recorded by
 * oracle/gen_golden.py from the UNMODIFIED reference imported in the build container
 * (known answers of the reference's own tests are re-checked by that script first).
 *
 * World representation: one cell = (type, color, state, aux) bytes.
 *   None (empty list slot, core/grid.py:35)     -> (1,0,0,0)   same bytes Grid.encode gives (grid.py:260-263)
 *   aux bits 0-3 : type of Box.contains (0 = None)   (core/world_object.py:274-276)
 *   aux bits 4-6 : colour of Box.contains
 *   aux bit  7   : "is the mission target object" (identity test, envs/keycorridor.py:133)
 * Planes are stored struct-of-arrays: cells[p*W*H + j*W + i], p = 0 type, 1 color, 2 state, 3 aux,
 * row-major j*W+i as core/grid.py:72.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { T_UNSEEN = 0, T_EMPTY = 1, T_WALL = 2, T_FLOOR = 3, T_DOOR = 4, T_KEY = 5,
       T_BALL = 6, T_BOX = 7, T_GOAL = 8, T_LAVA = 9, T_AGENT = 10 };     /* core/constants.py:25-37 */
enum { A_LEFT = 0, A_RIGHT = 1, A_FORWARD = 2, A_PICKUP = 3, A_DROP = 4, A_TOGGLE = 5, A_DONE = 6 }; /* core/actions.py:7-20 */
enum { S_OPEN = 0, S_CLOSED = 1, S_LOCKED = 2 };                          /* core/constants.py:42-46 */
enum { RULE_NONE = 0, RULE_PICKUP_TARGET = 1,                             /* envs/keycorridor.py:129-137 */
       RULE_UNLOCK = 2,                                                   /* envs/unlock.py:89-97 */
       RULE_FETCH = 3,                                                    /* envs/fetch.py:162-176 */
       RULE_GOTO = 4,                                                     /* envs/gotodoor.py:133-150, envs/gotoobject.py:145-162 */
       RULE_REDBLUE = 5,                                                  /* envs/redbluedoors.py:105-133 */
       RULE_MEMORY = 6,                                                   /* envs/memory.py:153-165 */
       RULE_PUTNEAR = 7,                                                  /* envs/putnear.py:176-201 */
       RULE_DYNOBS = 8 };                                                 /* envs/dynamicobstacles.py:136-167 */

#define VMAX 15

typedef struct { uint8_t type, color, state, aux; } cell_t;

/* cfg[] layout shared with the python wrapper */
enum { CFG_W = 0, CFG_H = 1, CFG_MAX_STEPS = 2, CFG_SEE_THROUGH = 3, CFG_RULE = 4, CFG_VIEW = 5, CFG_N = 6 };
/* st[] layout */
enum { ST_X = 0, ST_Y = 1, ST_DIR = 2, ST_STEPS = 3, ST_CT = 4, ST_CC = 5, ST_CS = 6, ST_CA = 7, ST_N = 8 };

static const int DIR_TO_VEC[4][2] = { {1, 0}, {0, 1}, {-1, 0}, {0, -1} };  /* core/constants.py:49-58 */

static inline cell_t cell_none(void) { cell_t c = { T_EMPTY, 0, 0, 0 }; return c; }
static inline cell_t cell_wall(void) { cell_t c = { T_WALL, 5, 0, 0 }; return c; } /* Wall() default grey, world_object.py:161-163 */
static inline int is_none(cell_t c) { return c.type == T_EMPTY; }

static inline cell_t grid_get(const uint8_t* cells, int W, int H, int i, int j)
{
    const int n = W * H, k = j * W + i;
    cell_t c = { cells[k], cells[n + k], cells[2 * n + k], cells[3 * n + k] };
    return c;
}
static inline void grid_set(uint8_t* cells, int W, int H, int i, int j, cell_t c)
{
    const int n = W * H, k = j * W + i;
    cells[k] = c.type; cells[n + k] = c.color; cells[2 * n + k] = c.state; cells[3 * n + k] = c.aux;
}

/* WorldObj capability flags: core/world_object.py:46-60 and the per-class overrides */
static inline int can_overlap(cell_t c)
{   /* Goal :114, Floor :129, Lava :142, Door iff open :178-180 */
    return c.type == T_GOAL || c.type == T_FLOOR || c.type == T_LAVA || (c.type == T_DOOR && c.state == S_OPEN);
}
static inline int can_pickup(cell_t c)
{   /* Key :244, Ball :266, Box :278 */
    return c.type == T_KEY || c.type == T_BALL || c.type == T_BOX;
}
static inline int see_behind(cell_t c)
{   /* base True :58-60; Wall False :165-166; Door iff open :182-183 */
    if (c.type == T_WALL) return 0;
    if (c.type == T_DOOR) return c.state == S_OPEN;
    return 1;
}

/* ---------------------------------------------------------------- observation */

/* MiniGridEnv._reward, minigrid_env.py:235-240: three separately rounded fp64 operations. */
static double reward_fn(int step_count, int max_steps)
{
    volatile double q = (double)step_count / (double)max_steps;
    volatile double m = 0.9 * q;
    volatile double r = 1.0 - m;
    return r;
}

/* gen_obs_grid, minigrid_env.py:592-627: g[j*V+i] = view cell (i,j), mask[i][j] = vis_mask. */
static void gen_obs_grid(const uint8_t* cells, const int32_t* cfg, const int32_t* st, cell_t* g, uint8_t mask[VMAX][VMAX])
{
    const int W = cfg[CFG_W], H = cfg[CFG_H], V = cfg[CFG_VIEW];
    const int ax = st[ST_X], ay = st[ST_Y], dir = st[ST_DIR];
    int topX, topY;
    /* get_view_exts, minigrid_env.py:448-479 */
    if (dir == 0)      { topX = ax;             topY = ay - V / 2; }
    else if (dir == 1) { topX = ax - V / 2;     topY = ay; }
    else if (dir == 2) { topX = ax - V + 1;     topY = ay - V / 2; }
    else               { topX = ax - V / 2;     topY = ay - V + 1; }

    /* Grid.slice, grid.py:124-143 (sub-grid is V x V, row-major j*V+i) */
    cell_t r[VMAX * VMAX];
    for (int j = 0; j < V; j++)
        for (int i = 0; i < V; i++) {
            int x = topX + i, y = topY + j;
            g[j * V + i] = (0 <= x && x < W && 0 <= y && y < H) ? grid_get(cells, W, H, x, y) : cell_wall();
        }

    /* rotate_left applied agent_dir+1 times, minigrid_env.py:606-607 / grid.py:110-122 */
    for (int k = 0; k < dir + 1; k++) {
        /* new grid has width=old height, height=old width (both V); set(j, newH-1-i, get(i,j)) */
        for (int i = 0; i < V; i++)
            for (int j = 0; j < V; j++)
                r[(V - 1 - i) * V + j] = g[j * V + i];
        memcpy(g, r, sizeof(cell_t) * V * V);
    }

    /* process_vis, grid.py:291-328; mask[i][j] */
    if (!cfg[CFG_SEE_THROUGH]) {
        memset(mask, 0, VMAX * VMAX);
        mask[V / 2][V - 1] = 1;
        for (int j = V - 1; j >= 0; j--) {
            for (int i = 0; i < V - 1; i++) {
                if (!mask[i][j]) continue;
                cell_t c = g[j * V + i];
                if (!is_none(c) && !see_behind(c)) continue;
                mask[i + 1][j] = 1;
                if (j > 0) { mask[i + 1][j - 1] = 1; mask[i][j - 1] = 1; }
            }
            for (int i = V - 1; i >= 1; i--) {
                if (!mask[i][j]) continue;
                cell_t c = g[j * V + i];
                if (!is_none(c) && !see_behind(c)) continue;
                mask[i - 1][j] = 1;
                if (j > 0) { mask[i - 1][j - 1] = 1; mask[i][j - 1] = 1; }
            }
        }
        for (int j = 0; j < V; j++)
            for (int i = 0; i < V; i++)
                if (!mask[i][j]) g[j * V + i] = cell_none();
    } else {
        memset(mask, 1, VMAX * VMAX);
    }

    /* carried object shown at the agent's own view cell, minigrid_env.py:618-625 */
    if (st[ST_CT] != 0) {
        cell_t c = { (uint8_t)st[ST_CT], (uint8_t)st[ST_CC], (uint8_t)st[ST_CS], (uint8_t)st[ST_CA] };
        g[(V - 1) * V + V / 2] = c;
    } else {
        g[(V - 1) * V + V / 2] = cell_none();
    }

}

/* gen_obs (minigrid_env.py:629-645) = gen_obs_grid + Grid.encode(vis_mask) (grid.py:244-268): array[i][j][:] */
static void gen_obs(const uint8_t* cells, const int32_t* cfg, const int32_t* st, uint8_t* image)
{
    const int V = cfg[CFG_VIEW];
    cell_t g[VMAX * VMAX];
    uint8_t mask[VMAX][VMAX];
    gen_obs_grid(cells, cfg, st, g, mask);
    for (int i = 0; i < V; i++)
        for (int j = 0; j < V; j++) {
            uint8_t* o = image + (i * V + j) * 3;
            if (mask[i][j]) {
                cell_t c = g[j * V + i];
                o[0] = c.type; o[1] = c.color; o[2] = c.state;
            } else {
                o[0] = 0; o[1] = 0; o[2] = 0;
            }
        }
}

/* ---------------------------------------------------------------- step */

/* MiniGridEnv.step, minigrid_env.py:520-590, plus KeyCorridorEnv.step, envs/keycorridor.py:129-137.
 * Returns 0, or -1 for an unknown action (ValueError at :579-580 — raised AFTER step_count += 1,
 * and before truncation / obs are computed). */
/* `target`: cell index (j*W+i) the mission's `target_pos` refers to (gotodoor.py:125, gotoobject.py:139) — the cell that
 * carried the 0x80 aux bit in the INITIAL layout (the object may have been carried away since); -1 if none. */
/* Packed as A | B << 16 (0xFFFF = none, so -1 = neither).  B = the second mission cell of Memory (failure_pos,
 * memory.py:149) and PutNear (target_pos, putnear.py:167): flagged with aux 0x81 (target bit + a low nibble no other rule
 * uses), everything else flagged 0x80 is A (first in scan order). */
static int find_target(const uint8_t* cells, int n, int rule)
{
    unsigned a = 0xFFFFu, b = 0xFFFFu;
    for (int k = 0; k < n; k++) {
        const uint8_t x = cells[3 * n + k];
        if (!(x & 0x80)) continue;
        if ((rule == RULE_MEMORY || rule == RULE_PUTNEAR) && (x & 0x0f) == 1) { if (b == 0xFFFFu) b = (unsigned)k; }
        else if (a == 0xFFFFu) a = (unsigned)k;
    }
    return (int)(a | (b << 16));
}
static int door_open(const uint8_t* cells, int n, int color)   /* is the flagged door of that colour open?  (redbluedoors.py:106-112) */
{
    for (int k = 0; k < n; k++)
        if (cells[k] == T_DOOR && (cells[3 * n + k] & 0x80) && cells[n + k] == color) return cells[2 * n + k] == S_OPEN;
    return 0;
}

/* ---- DynamicObstaclesEnv draws random numbers INSIDE step (dynamicobstacles.py:146-157): the env's generator is numpy's
 * Generator(PCG64) (gymnasium seeding.np_random; third-party, numpy >= 1.18, here 2.3.5 — SURVEY §8.3).  Restated from the
 * published algorithms:
 *   PCG64 = pcg_setseq_128_xsl_rr_64: state = state * 0x2360ED051FC65DA44385DF649FCCF645 + inc (mod 2^128), output =
 *   rotr64(hi ^ lo, state >> 122) of the NEW state; next_uint32 returns the low half of a 64-bit draw and buffers the
 *   high half (has_uint32 / uinteger).
 *   Generator.integers(low, high) (int64, endpoint=False, scalar): rng = high - 1 - low; rng == 0 -> low without a draw;
 *   rng < 2^32 - 1 -> Lemire's nearly-divisionless rejection on 32-bit draws (numpy/random/src/distributions.c,
 *   buffered_bounded_lemire_uint32).
 * Pinned by the reference's own rollouts (tests/golden/dynobs.npz).
 * dyn record (64 bytes, little endian): u64 state_lo, state_hi, inc_lo, inc_hi; u32 has_uint32, uinteger; u8 n_obstacles, pad[7];
 * u8 ox[8], oy[8] — the obstacles in creation order (`self.obstacles`, dynamicobstacles.py:129-132). */
typedef struct {
    uint64_t state_lo, state_hi, inc_lo, inc_hi;
    uint32_t has_uint32, uinteger;
    uint8_t n_obst, pad[7];
    uint8_t ox[8], oy[8];
} dyn_t;

static uint64_t pcg64_next64(dyn_t* d)
{
    const unsigned __int128 mult = ((unsigned __int128)0x2360ED051FC65DA4ull << 64) | 0x4385DF649FCCF645ull;
    unsigned __int128 st = ((unsigned __int128)d->state_hi << 64) | d->state_lo;
    const unsigned __int128 inc = ((unsigned __int128)d->inc_hi << 64) | d->inc_lo;
    st = st * mult + inc;
    d->state_hi = (uint64_t)(st >> 64); d->state_lo = (uint64_t)st;
    const uint64_t x = d->state_hi ^ d->state_lo;
    const unsigned rot = (unsigned)(d->state_hi >> 58);
    return (x >> rot) | (x << ((-rot) & 63));
}
static uint32_t pcg64_next32(dyn_t* d)
{
    if (d->has_uint32) { d->has_uint32 = 0; return d->uinteger; }
    const uint64_t next = pcg64_next64(d);
    d->has_uint32 = 1;
    d->uinteger = (uint32_t)(next >> 32);
    return (uint32_t)next;
}
static int rand_int(dyn_t* d, int low, int high)                          /* minigrid_env.py:244-249 */
{
    const uint32_t rng = (uint32_t)(high - 1 - low);
    if (rng == 0) return low;
    const uint32_t rng_excl = rng + 1;
    uint64_t m = (uint64_t)pcg64_next32(d) * rng_excl;
    uint32_t leftover = (uint32_t)m;
    if (leftover < rng_excl) {
        const uint32_t threshold = (0xFFFFFFFFu - rng) % rng_excl;
        while (leftover < threshold) { m = (uint64_t)pcg64_next32(d) * rng_excl; leftover = (uint32_t)m; }
    }
    return low + (int)(m >> 32);
}
/* the obstacle loop of DynamicObstaclesEnv.step (dynamicobstacles.py:146-157) with place_obj (minigrid_env.py:308-373) inlined */
static void move_obstacles(uint8_t* cells, int W, int H, const int32_t* st, dyn_t* d)
{
    for (int i = 0; i < d->n_obst; i++) {
        const int ox = d->ox[i], oy = d->oy[i];
        int tx = ox - 1, ty = oy - 1;
        if (tx < 0) tx = 0;                                                /* top = (max(top[0], 0), max(top[1], 0)) :331 */
        if (ty < 0) ty = 0;
        const int hx = tx + 3 < W ? tx + 3 : W, hy = ty + 3 < H ? ty + 3 : H;
        int tries = 0, px = 0, py = 0, found = 0;
        for (;;) {
            if (tries > 100) break;                                        /* RecursionError -> `except Exception: pass` */
            tries++;
            px = rand_int(d, tx, hx);
            py = rand_int(d, ty, hy);
            if (!is_none(grid_get(cells, W, H, px, py))) continue;         /* :355-356 */
            if (px == st[ST_X] && py == st[ST_Y]) continue;                /* :359-360 */
            found = 1;
            break;
        }
        if (found) {
            const cell_t ball = grid_get(cells, W, H, ox, oy);
            grid_set(cells, W, H, px, py, ball);
            grid_set(cells, W, H, ox, oy, cell_none());
            d->ox[i] = (uint8_t)px; d->oy[i] = (uint8_t)py;
        }
    }
}

static int step_env(uint8_t* cells, const int32_t* cfg, int32_t* st, int action, int target, uint8_t* dyn,
                    double* reward, uint8_t* terminated, uint8_t* truncated)
{
    int not_clear = 0;
    if (cfg[CFG_RULE] == RULE_DYNOBS) {                                    /* dynamicobstacles.py:137-157 */
        if (action >= 3) action = A_LEFT;                                  /* `if action >= self.action_space.n: action = 0` :138-139 */
        const int qx = st[ST_X] + DIR_TO_VEC[st[ST_DIR]][0], qy = st[ST_Y] + DIR_TO_VEC[st[ST_DIR]][1];
        const cell_t f0 = grid_get(cells, cfg[CFG_W], cfg[CFG_H], qx, qy);
        not_clear = !is_none(f0) && f0.type != T_GOAL;
        if (dyn) move_obstacles(cells, cfg[CFG_W], cfg[CFG_H], st, (dyn_t*)dyn);
    }
    const int W = cfg[CFG_W], H = cfg[CFG_H];
    st[ST_STEPS] += 1;                                                     /* :523 */
    *reward = 0.0; *terminated = 0; *truncated = 0;                        /* :525-527 */

    const int red_before = cfg[CFG_RULE] == RULE_REDBLUE ? door_open(cells, W * H, 0) : 0;      /* redbluedoors.py:106-107 */
    const int blue_before = cfg[CFG_RULE] == RULE_REDBLUE ? door_open(cells, W * H, 2) : 0;
    const int tgt_a = (target & 0xFFFF) == 0xFFFF ? -1 : (target & 0xFFFF);
    const int tgt_b = ((target >> 16) & 0xFFFF) == 0xFFFF ? -1 : ((target >> 16) & 0xFFFF);
    if (cfg[CFG_RULE] == RULE_MEMORY && action == A_PICKUP) action = A_TOGGLE;                  /* memory.py:154-155 */
    const int pre_carrying = st[ST_CT] != 0;                                                    /* putnear.py:177 */
    const int fx = st[ST_X] + DIR_TO_VEC[st[ST_DIR]][0];                   /* front_pos :413-419 */
    const int fy = st[ST_Y] + DIR_TO_VEC[st[ST_DIR]][1];
    /* :533.  Outside the grid the reference's Grid.get asserts (core/grid.py:75-76) — reachable only in envs whose room is
     * smaller than the grid (GoToDoor: through an opened door to the unwalled rim).  The oracle, like the engine, then sees the
     * Wall() that Grid.slice substitutes out of bounds, so nothing happens instead of an AssertionError. */
    cell_t fwd = (fx >= 0 && fy >= 0 && fx < W && fy < H) ? grid_get(cells, W, H, fx, fy) : cell_wall();

    if (action == A_LEFT) {                                                /* :536-539 */
        st[ST_DIR] -= 1;
        if (st[ST_DIR] < 0) st[ST_DIR] += 4;
    } else if (action == A_RIGHT) {                                        /* :542-543 */
        st[ST_DIR] = (st[ST_DIR] + 1) % 4;
    } else if (action == A_FORWARD) {                                      /* :546-553 */
        if (is_none(fwd) || can_overlap(fwd)) { st[ST_X] = fx; st[ST_Y] = fy; }
        if (!is_none(fwd) && fwd.type == T_GOAL) { *terminated = 1; *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]); }
        if (!is_none(fwd) && fwd.type == T_LAVA) { *terminated = 1; }
    } else if (action == A_PICKUP) {                                       /* :556-561 */
        if (!is_none(fwd) && can_pickup(fwd)) {
            if (st[ST_CT] == 0) {
                st[ST_CT] = fwd.type; st[ST_CC] = fwd.color; st[ST_CS] = fwd.state; st[ST_CA] = fwd.aux;
                grid_set(cells, W, H, fx, fy, cell_none());
            }
        }
    } else if (action == A_DROP) {                                         /* :564-568 */
        if (is_none(fwd) && st[ST_CT] != 0) {
            cell_t c = { (uint8_t)st[ST_CT], (uint8_t)st[ST_CC], (uint8_t)st[ST_CS], (uint8_t)st[ST_CA] };
            grid_set(cells, W, H, fx, fy, c);
            st[ST_CT] = st[ST_CC] = st[ST_CS] = st[ST_CA] = 0;
        }
    } else if (action == A_TOGGLE) {                                       /* :571-573 */
        if (!is_none(fwd)) {
            if (fwd.type == T_DOOR) {                                      /* Door.toggle, world_object.py:185-195 */
                if (fwd.state == S_LOCKED) {
                    if (st[ST_CT] == T_KEY && st[ST_CC] == fwd.color) { fwd.state = S_OPEN; grid_set(cells, W, H, fx, fy, fwd); }
                } else {
                    fwd.state = (fwd.state == S_OPEN) ? S_CLOSED : S_OPEN;
                    grid_set(cells, W, H, fx, fy, fwd);
                }
            } else if (fwd.type == T_BOX) {                                /* Box.toggle, world_object.py:291-294 */
                cell_t c = cell_none();
                if ((fwd.aux & 0x0f) != 0) { c.type = fwd.aux & 0x0f; c.color = (fwd.aux >> 4) & 7; c.state = 0; c.aux = 0; }
                grid_set(cells, W, H, fx, fy, c);
            }                                                              /* base WorldObj.toggle: no-op :62-64 */
        }
    } else if (action == A_DONE) {                                         /* :576-577 */
    } else {
        return -1;                                                         /* :579-580 */
    }

    if (st[ST_STEPS] >= cfg[CFG_MAX_STEPS]) *truncated = 1;                /* :582-583 */

    if (cfg[CFG_RULE] == RULE_PICKUP_TARGET && action == A_PICKUP) {       /* envs/keycorridor.py:132-135 */
        if (st[ST_CT] != 0 && (st[ST_CA] & 0x80)) {
            *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]);
            *terminated = 1;
        }
    }
    if (cfg[CFG_RULE] == RULE_FETCH && st[ST_CT] != 0) {                   /* envs/fetch.py:165-174: `if self.carrying:` */
        *terminated = 1;
        *reward = (st[ST_CA] & 0x80) ? reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]) : 0.0;   /* same type and colour as the target */
    }
    if (cfg[CFG_RULE] == RULE_GOTO) {                                      /* gotodoor.py:136-148, gotoobject.py:148-160 */
        if (action == A_TOGGLE) *terminated = 1;
        if (action == A_DONE) {
            const int tx = tgt_a >= 0 ? tgt_a % W : -9, ty = tgt_a >= 0 ? tgt_a / W : -9;
            const int ax = st[ST_X], ay = st[ST_Y];
            if ((ax == tx && abs(ay - ty) == 1) || (ay == ty && abs(ax - tx) == 1))
                *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]);
            *terminated = 1;
        }
    }
    if (cfg[CFG_RULE] == RULE_DYNOBS && action == A_FORWARD && not_clear) {   /* dynamicobstacles.py:163-166 */
        *reward = -1.0; *terminated = 1;
    }
    if (cfg[CFG_RULE] == RULE_MEMORY) {                                    /* memory.py:158-163 */
        const int here = st[ST_Y] * W + st[ST_X];
        if (here == tgt_a) { *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]); *terminated = 1; }
        if (here == tgt_b) { *reward = 0.0; *terminated = 1; }
    }
    if (cfg[CFG_RULE] == RULE_PUTNEAR) {                                   /* putnear.py:181-199 */
        if (action == A_PICKUP && st[ST_CT] != 0) {                        /* wrong object: type or colour differs from the one to move */
            if (!((st[ST_CA] & 0x80) && (st[ST_CA] & 0x0f) != 1)) *terminated = 1;
        }
        if (action == A_DROP && pre_carrying) {
            if (st[ST_CT] == 0) {                                          /* `grid.get(ox, oy) is preCarrying`: the drop happened */
                const int tx = tgt_b >= 0 ? tgt_b % W : -9, ty = tgt_b >= 0 ? tgt_b / W : -9;
                if (abs(fx - tx) <= 1 && abs(fy - ty) <= 1) *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]);
            }
            *terminated = 1;
        }
    }
    if (cfg[CFG_RULE] == RULE_REDBLUE) {                                   /* redbluedoors.py:114-131 */
        const int red_after = door_open(cells, W * H, 0), blue_after = door_open(cells, W * H, 2);
        if (blue_after) {
            *reward = red_before ? reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]) : 0.0;
            *terminated = 1;
        } else if (red_after) {
            if (blue_before) { *reward = 0.0; *terminated = 1; }
        }
    }
    if (cfg[CFG_RULE] == RULE_UNLOCK && action == A_TOGGLE) {              /* envs/unlock.py:92-95: self.door.is_open */
        for (int k = 0; k < W * H; k++)
            if (cells[k] == T_DOOR && (cells[3 * W * H + k] & 0x80) && cells[2 * W * H + k] == S_OPEN) {
                *reward = reward_fn(st[ST_STEPS], cfg[CFG_MAX_STEPS]);
                *terminated = 1;
            }
    }
    return 0;
}

/* ---------------------------------------------------------------- exported entry points (ctypes) */

/* view grid for the renderer (mg_oracle_render.c): g4[(j*V+i)*4 + {type,color,state,aux}], mask[i*V+j] */
void mgo_gen_obs_grid(const uint8_t* cells, const int32_t* cfg, const int32_t* st, uint8_t* g4, uint8_t* vis)
{
    const int V = cfg[CFG_VIEW];
    cell_t g[VMAX * VMAX];
    uint8_t mask[VMAX][VMAX];
    gen_obs_grid(cells, cfg, st, g, mask);
    for (int j = 0; j < V; j++)
        for (int i = 0; i < V; i++) {
            const cell_t c = g[j * V + i];
            uint8_t* o = g4 + (j * V + i) * 4;
            o[0] = c.type; o[1] = c.color; o[2] = c.state; o[3] = c.aux;
            vis[i * V + j] = mask[i][j];
        }
}

void mgo_gen_obs(const uint8_t* cells, const int32_t* cfg, const int32_t* st, uint8_t* image)
{
    gen_obs(cells, cfg, st, image);
}

int mgo_find_target(const uint8_t* cells, int n, int rule) { return find_target(cells, n, rule); }

int mgo_step_dyn(uint8_t* cells, const int32_t* cfg, int32_t* st, int action, int target, uint8_t* dyn,
                 uint8_t* image, double* reward, uint8_t* terminated, uint8_t* truncated)
{
    int rc = step_env(cells, cfg, st, action, target, dyn, reward, terminated, truncated);
    if (rc) return rc;
    gen_obs(cells, cfg, st, image);
    return 0;
}

int mgo_step_t(uint8_t* cells, const int32_t* cfg, int32_t* st, int action, int target,
               uint8_t* image, double* reward, uint8_t* terminated, uint8_t* truncated)
{
    int rc = step_env(cells, cfg, st, action, target, NULL, reward, terminated, truncated);
    if (rc) return rc;
    gen_obs(cells, cfg, st, image);
    return 0;
}

int mgo_step(uint8_t* cells, const int32_t* cfg, int32_t* st, int action,
             uint8_t* image, double* reward, uint8_t* terminated, uint8_t* truncated)
{
    int rc = step_env(cells, cfg, st, action, find_target(cells, cfg[CFG_W] * cfg[CFG_H], cfg[CFG_RULE]), NULL, reward, terminated, truncated);
    if (rc) return rc;
    gen_obs(cells, cfg, st, image);
    return 0;
}

double mgo_reward(int step_count, int max_steps) { return reward_fn(step_count, max_steps); }

/*
 * Batched rollout used (a) as the CPU baseline and (b) as the checker at full batch sizes.
 *   n envs, each with its own cell record (stride `cell_stride` bytes, 4 planes of W*H inside),
 *   class id -> cfg row, T steps of actions[t*n + e].
 * Autoreset (the build's own convention — the reference env has none):
 *   autoreset=1: when terminated|truncated, the env is re-initialised in the same step from
 *   layout pool slot (layout_id + reset_stride) % P (the engine's mg_reset_stride(): n mod P moved up to the next value
 *   coprime to P; the caller passes it) and the returned obs/direction are the first ones of
 *   the new episode; reward/terminated/truncated are those of the finished episode.
 *   autoreset=2 (gymnasium >= 1.0 "next step"): the step that ends an episode returns its last obs; the env's NEXT step
 *   ignores its action and only resets (reward 0, flags clear, first obs of the new episode).
 * Outputs are [T][n] (image [T][n][V*V*3]).  Any output pointer may be NULL to skip storing.
 * Returns 0, or 1+index of the first (t*n+e) with an invalid action.
 */
long mgo_rollout(int n, int T, int nthreads,
                 uint8_t* cells, long cell_stride, int32_t* st /*[n][ST_N]*/,
                 const uint8_t* class_id /*[n]*/, const int32_t* cfgs /*[C][CFG_N]*/,
                 const uint8_t* actions /*[T][n]*/,
                 int autoreset, int P, const uint8_t* pool_cells /*[P][cell_stride]*/,
                 const int32_t* pool_st /*[P][ST_N]*/, const uint8_t* pool_class /*[P]*/,
                 int32_t* layout_id /*[n]*/, uint8_t* class_id_mut /*[n] or NULL (== class_id if mixed pools)*/,
                 uint8_t* image, uint8_t* direction, double* reward, uint8_t* terminated, uint8_t* truncated,
                 uint8_t* dyn /*[n][64] or NULL*/, const uint8_t* pool_dyn /*[P][64] or NULL*/,
                 uint8_t* final_image /*[T][n][V*V*3] or NULL: last obs of an episode that ended in that step (autoreset=1)*/,
                 int reset_stride)
{
    long bad = 0;
    (void)nthreads;
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (int e = 0; e < n; e++) {
        uint8_t* c = cells + (long)e * cell_stride;
        int32_t* s = st + (long)e * ST_N;
        int cls = class_id[e];
        int target = find_target(c, cfgs[cls * CFG_N + CFG_W] * cfgs[cls * CFG_N + CFG_H], cfgs[cls * CFG_N + CFG_RULE]);   /* rollouts start from initial layouts */
        int pending = 0;
        for (int t = 0; t < T; t++) {
            const int32_t* cfg = cfgs + cls * CFG_N;
            const int V = cfg[CFG_VIEW];
            const long k = (long)t * n + e;
            double r = 0.0; uint8_t term = 0, trunc = 0;
            int rc = 0;
            if (!pending) rc = step_env(c, cfg, s, actions[k], target, dyn ? dyn + (long)e * 64 : NULL, &r, &term, &trunc);
            if (rc) {
#ifdef _OPENMP
#pragma omp critical
#endif
                { if (bad == 0 || k + 1 < bad) bad = k + 1; }
                break;
            }
            const int reset_now = (autoreset == 1 && (term || trunc)) || (autoreset == 2 && pending);
            pending = autoreset == 2 && !pending && (term || trunc);
            if (reset_now && final_image && autoreset == 1) gen_obs(c, cfg, s, final_image + k * (long)(V * V * 3));
            if (reset_now) {
                int lid = (int)(((long)layout_id[e] + reset_stride) % P);
                layout_id[e] = lid;
                memcpy(c, pool_cells + (long)lid * cell_stride, (size_t)cell_stride);
                memcpy(s, pool_st + (long)lid * ST_N, sizeof(int32_t) * ST_N);
                if (dyn && pool_dyn) memcpy(dyn + (long)e * 64, pool_dyn + (long)lid * 64, 64);
                cls = pool_class[lid];
                if (class_id_mut) class_id_mut[e] = (uint8_t)cls;
                cfg = cfgs + cls * CFG_N;
                target = find_target(c, cfg[CFG_W] * cfg[CFG_H], cfg[CFG_RULE]);
            }
            if (image) gen_obs(c, cfg, s, image + k * (long)(V * V * 3));
            else { uint8_t tmp[VMAX * VMAX * 3]; gen_obs(c, cfg, s, tmp); }
            if (direction) direction[k] = (uint8_t)s[ST_DIR];
            if (reward) reward[k] = r;
            if (terminated) terminated[k] = term;
            if (truncated) truncated[k] = trunc;
        }
    }
    return bad;
}
