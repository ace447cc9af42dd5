// mg_common.cuh — device-side world model shared by the step kernels and the DP frontier expansion.
//
// Restates, on (type, color, state, aux) bytes, the object behaviour the reference implements with
// virtual methods (minigrid/core/world_object.py:46-68 and per-class overrides) and the 7-way action
// switch of MiniGridEnv.step (minigrid/minigrid_env.py:520-583).
#pragma once
#include <stdint.h>
#include <cuda_runtime.h>

namespace mg {

// minigrid/core/constants.py:25-37
enum : int { T_UNSEEN = 0, T_EMPTY = 1, T_WALL = 2, T_FLOOR = 3, T_DOOR = 4, T_KEY = 5, T_BALL = 6,
             T_BOX = 7, T_GOAL = 8, T_LAVA = 9, T_AGENT = 10 };
// minigrid/core/actions.py:7-20
enum : int { A_LEFT = 0, A_RIGHT = 1, A_FORWARD = 2, A_PICKUP = 3, A_DROP = 4, A_TOGGLE = 5, A_DONE = 6 };
// minigrid/core/constants.py:42-46
enum : int { S_OPEN = 0, S_CLOSED = 1, S_LOCKED = 2 };

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
    uint16_t steps, pad;
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
__host__ __device__ __forceinline__ Transition apply_action(int action, int x, int y, int dir, cell_t carry,
                                                            int fx, int fy, cell_t front, int rule)
{
    Transition r;
    r.x = x; r.y = y; r.dir = dir; r.carry = carry; r.front = front;
    r.wrote_front = false; r.terminated = false; r.success = false; r.bad_action = false;
    const int ft = c_type(front);
    switch (action) {
    case A_LEFT:  r.dir = (dir + 3) & 3; break;                               // :536-539
    case A_RIGHT: r.dir = (dir + 1) & 3; break;                               // :542-543
    case A_FORWARD:                                                           // :546-553
        if (can_overlap(front)) { r.x = fx; r.y = fy; }
        if (ft == T_GOAL) { r.terminated = true; r.success = true; }
        if (ft == T_LAVA) { r.terminated = true; }
        break;
    case A_PICKUP:                                                            // :556-561
        if (can_pickup(front) && carry == 0) { r.carry = front; r.front = CELL_NONE; r.wrote_front = true; }
        break;
    case A_DROP:                                                              // :564-568 (only onto None)
        if (ft == T_EMPTY && carry != 0) { r.front = carry; r.carry = 0; r.wrote_front = true; }
        break;
    case A_TOGGLE:                                                            // :571-573
        if (ft == T_DOOR) {                                                   // Door.toggle, world_object.py:185-195
            const int s = c_state(front);
            int ns = s;
            if (s == S_LOCKED) { if (c_type(carry) == T_KEY && c_color(carry) == c_color(front)) ns = S_OPEN; }
            else ns = (s == S_OPEN) ? S_CLOSED : S_OPEN;
            if (ns != s) { r.front = (front & 0xff00ffffu) | ((cell_t)ns << 16); r.wrote_front = true; }
        } else if (ft == T_BOX) {                                             // Box.toggle, world_object.py:291-294
            const int a = c_aux(front);
            r.front = (a & 0x0f) ? make_cell(a & 0x0f, (a >> 4) & 7, 0, 0) : CELL_NONE;
            r.wrote_front = true;
        }
        break;
    case A_DONE: break;                                                       // :576-577
    default: r.bad_action = true; break;                                      // :579-580
    }
    if (rule == 1 && action == A_PICKUP && r.carry != 0 && (c_aux(r.carry) & 0x80)) {   // keycorridor.py:132-135
        r.terminated = true; r.success = true;
    }
    return r;
}

__host__ __device__ __forceinline__ int dir_dx(int d) { return d == 0 ? 1 : (d == 2 ? -1 : 0); }   // constants.py:49-58
__host__ __device__ __forceinline__ int dir_dy(int d) { return d == 1 ? 1 : (d == 3 ? -1 : 0); }

}  // namespace mg
