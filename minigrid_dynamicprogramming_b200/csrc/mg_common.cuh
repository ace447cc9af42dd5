// mg_common.cuh — device-side world model shared by the step kernels and the DP frontier expansion.
//
// Restates, on (type, color, state, aux) bytes, the object behaviour the reference implements with
// virtual methods (minigrid/core/world_object.py:46-68 and per-class overrides) and the 7-way action
// switch of MiniGridEnv.step (minigrid/minigrid_env.py:520-583).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

// -DMG_BOUNDS_CHECK: device-side asserts on every index the kernels derive from env / table contents (agent pose, front cell,
// pool slot, successor id, hash slot).  compute-sanitizer is closed on this GPU pool, so the GPU test-suite is run once against a
// library built with this flag instead (tools/gpu_suite.sh bounds); a violated bound traps and fails the test loudly.
#ifdef MG_BOUNDS_CHECK
#include <assert.h>
#define MG_ASSERT(cond) assert(cond)
#else
#define MG_ASSERT(cond) ((void)0)
#endif

namespace mg {

// minigrid/core/constants.py:25-37
enum : int { T_UNSEEN = 0, T_EMPTY = 1, T_WALL = 2, T_FLOOR = 3, T_DOOR = 4, T_KEY = 5, T_BALL = 6,
             T_BOX = 7, T_GOAL = 8, T_LAVA = 9, T_AGENT = 10 };
// minigrid/core/actions.py:7-20
enum : int { A_LEFT = 0, A_RIGHT = 1, A_FORWARD = 2, A_PICKUP = 3, A_DROP = 4, A_TOGGLE = 5, A_DONE = 6 };
// minigrid/core/constants.py:42-46
enum : int { S_OPEN = 0, S_CLOSED = 1, S_LOCKED = 2 };
// per-class step rules on top of MiniGridEnv.step (MG_RULE_* in include/mgb200.h)
enum : int { RULE_NONE = 0, RULE_PICKUP_TARGET = 1 /* envs/keycorridor.py:129-137 */, RULE_UNLOCK = 2 /* envs/unlock.py:89-97 */,
             RULE_FETCH = 3 /* envs/fetch.py:162-176 */, RULE_GOTO = 4 /* envs/gotodoor.py:133-150, gotoobject.py:145-162 */,
             RULE_REDBLUE = 5 /* envs/redbluedoors.py:105-133 */, RULE_MEMORY = 6 /* envs/memory.py:153-165 */,
             RULE_PUTNEAR = 7 /* envs/putnear.py:176-201 */, RULE_DYNOBS = 8 /* envs/dynamicobstacles.py:136-167 */ };

constexpr int VIEW = 7;
constexpr int OBS_BYTES = VIEW * VIEW * 3;
constexpr int MAX_CLASSES = 16;

// One cell / one carried object as four bytes: type | color<<8 | state<<16 | aux<<24.
// None is (1,0,0,0): the bytes Grid.encode writes for an empty slot (core/grid.py:260-263).
typedef uint32_t cell_t;
constexpr cell_t CELL_NONE = 0x00000001u;
constexpr cell_t CELL_WALL = 0x00000502u;     // Wall() is grey (world_object.py:161-163)
__host__ __device__ __forceinline__ int c_type(cell_t c) { return c & 0xff; }
__host__ __device__ __forceinline__ int c_color(cell_t c) { return (c >> 8) & 0xff; }
__host__ __device__ __forceinline__ int c_state(cell_t c) { return (c >> 16) & 0xff; }
__host__ __device__ __forceinline__ int c_aux(cell_t c) { return (c >> 24) & 0xff; }
__host__ __device__ __forceinline__ cell_t make_cell(int t, int c, int s, int a)
{
    return (cell_t)t | ((cell_t)c << 8) | ((cell_t)s << 16) | ((cell_t)a << 24);
}

// can_overlap: Goal world_object.py:114, Floor :129, Lava :142, Door iff open :178-180; None is walkable (minigrid_env.py:547)
__host__ __device__ __forceinline__ bool can_overlap(cell_t c)
{
    const int t = c_type(c);
    return t == T_EMPTY || t == T_FLOOR || t == T_GOAL || t == T_LAVA || (t == T_DOOR && c_state(c) == S_OPEN);
}
// can_pickup: Key :244, Ball :266, Box :278
__host__ __device__ __forceinline__ bool can_pickup(cell_t c)
{
    const int t = c_type(c);
    return t == T_KEY || t == T_BALL || t == T_BOX;
}
// see_behind: base True :58-60, Wall False :165-166, Door iff open :182-183
__host__ __device__ __forceinline__ bool is_opaque(int type, int state)
{
    return type == T_WALL || (type == T_DOOR && state != S_OPEN);
}

struct ClassDesc {            // one env class (constructor constants, minigrid_env.py:28-111)
    int32_t W, H, max_steps, see_through, rule, pad0, pad1, pad2;
};
struct ClassTable { ClassDesc c[MAX_CLASSES]; };

// Per-env dynamic state, 16 bytes (mirrors mg_agent_state in include/mgb200.h)
struct __align__(16) AgentState {
    uint8_t x, y, dir, cls;
    uint8_t ct, cc, cs, ca;     // carrying; ct == 0 means None
    uint16_t steps, pad;        // pad bit 0: a next-step autoreset is owed (MG_AUTORESET_NEXT_STEP)
    uint32_t layout_id;
};

// _reward(), minigrid_env.py:235-240: 1 - 0.9 * (step_count / max_steps) as three separately rounded
// fp64 operations (no FMA contraction, or the last bit differs from CPython's).
__device__ __forceinline__ double reward_fn(int steps, int max_steps)
{
    return __dsub_rn(1.0, __dmul_rn(0.9, __ddiv_rn((double)steps, (double)max_steps)));
}

// Result of applying one action given the cell in front of the agent.
struct Transition {
    int x, y, dir;          // new pose
    cell_t carry;           // new carried object (0 = None)
    cell_t front;           // new content of the front cell
    bool wrote_front;       // front != old front
    bool terminated;        // goal / lava / target pickup
    bool success;           // the termination pays _reward()
    bool bad_action;
};

// MiniGridEnv.step's action switch (minigrid_env.py:536-580) + KeyCorridorEnv.step's rule
// (envs/keycorridor.py:132-135) on encoded cells.  (fx,fy) is the front cell position, `front` its content.
// Written as straight-line selects: the 32 envs of a warp take 7 different actions on different cells, so a
// switch would execute every arm anyway and pay for the divergence on top.
__host__ __device__ __forceinline__ Transition apply_action(int action, int x, int y, int dir, cell_t carry,
                                                            int fx, int fy, cell_t front, int rule)
{
    const int ft = c_type(front), fs = c_state(front);
    // capability flags as bit sets over the type index (world_object.py:46-60 and overrides)
    constexpr unsigned OVERLAP = (1u << T_EMPTY) | (1u << T_FLOOR) | (1u << T_GOAL) | (1u << T_LAVA);   // + open door
    constexpr unsigned PICKUP = (1u << T_KEY) | (1u << T_BALL) | (1u << T_BOX);
    const bool is_door = ft == T_DOOR, is_box = ft == T_BOX;
    const bool overlap = ((OVERLAP >> ft) & 1u) | (is_door & (fs == S_OPEN));
    const bool pickable = (PICKUP >> ft) & 1u;
    const bool a_fwd = action == A_FORWARD, a_pick = action == A_PICKUP, a_drop = action == A_DROP, a_tog = action == A_TOGGLE;

    Transition r;
    r.bad_action = (unsigned)action > (unsigned)A_DONE;                                   // :579-580
    r.dir = action == A_LEFT ? ((dir + 3) & 3) : (action == A_RIGHT ? ((dir + 1) & 3) : dir);   // :536-543
    const bool move = a_fwd & overlap;                                                    // :546-548
    r.x = move ? fx : x; r.y = move ? fy : y;
    r.success = a_fwd & (ft == T_GOAL);                                                   // :549-551
    r.terminated = a_fwd & ((ft == T_GOAL) | (ft == T_LAVA));                             // :552-553

    const bool do_pick = a_pick & pickable & (carry == 0);                                // :556-561
    const bool do_drop = a_drop & (ft == T_EMPTY) & (carry != 0);                         // :564-568 (only onto None)
    // Door.toggle (world_object.py:185-195): locked opens only with a key of the door's colour, else open <-> closed
    const bool key_fits = (c_type(carry) == T_KEY) & (c_color(carry) == c_color(front));
    const int ns = fs == S_LOCKED ? (key_fits ? S_OPEN : S_LOCKED) : (fs == S_OPEN ? S_CLOSED : S_OPEN);
    const bool door_changed = a_tog & is_door & (ns != fs);
    const bool box_opened = a_tog & is_box;                                               // Box.toggle, world_object.py:291-294
    const int aux = c_aux(front);
    const cell_t contents = (aux & 0x0f) ? make_cell(aux & 0x0f, (aux >> 4) & 7, 0, 0) : CELL_NONE;
    const cell_t door_next = (front & 0xff00ffffu) | ((cell_t)ns << 16);

    r.front = do_pick ? CELL_NONE : (do_drop ? carry : (door_changed ? door_next : (box_opened ? contents : front)));
    r.carry = do_pick ? front : (do_drop ? (cell_t)0 : carry);
    r.wrote_front = do_pick | do_drop | door_changed | box_opened;
    const bool got_target = (rule == RULE_PICKUP_TARGET) & a_pick & (r.carry != 0) & ((c_aux(r.carry) & 0x80) != 0);   // keycorridor.py:132-135
    r.terminated |= got_target; r.success |= got_target;
    return r;
}

__host__ __device__ __forceinline__ int dir_dx(int d) { return d == 0 ? 1 : (d == 2 ? -1 : 0); }   // constants.py:49-58
__host__ __device__ __forceinline__ int dir_dy(int d) { return d == 1 ? 1 : (d == 3 ? -1 : 0); }

}  // namespace mg
