"""Host-side level generators: seed -> layout, for the env families the engine batches.

The reference builds a level in `MiniGridEnv.reset(seed)` (minigrid/minigrid_env.py:119-157) by calling
the task's `_gen_grid` with `self.np_random = Generator(PCG64(SeedSequence(seed)))`.  Generation is
rejection sampling with a data-dependent number of draws, runs once per episode and stays on the host;
only the finished layout (four uint8 planes + agent start) is uploaded to HBM.

Each generator below consumes the PCG64 stream in exactly the reference's order, so a layout is a pure
function of (env id, seed) and equals `gym.make(id).reset(seed=seed)` cell for cell
(tests/golden/layouts_*.npz, recorded from the reference by oracle/gen_golden.py, pin this).

Layout format (shared with the C-ABI, include/mgb200.h):
    cells  uint8[4, H*W]   planes type / color / state / aux, row-major j*W+i (core/grid.py:72);
                           None -> (1,0,0,0);  aux = box contents (type | color<<4) | 0x80 target bit
    agent  (x, y, dir)
"""
from __future__ import annotations

import dataclasses
import itertools as itt

import numpy as np

from .constants import (AUX_TARGET, AUX_TARGET_B, RULE_DYNOBS, RULE_MEMORY, RULE_PUTNEAR, COLOR_NAMES, COLOR_TO_IDX, RULE_NONE, RULE_PICKUP_TARGET, RULE_UNLOCK, RULE_FETCH, RULE_GOTO, RULE_REDBLUE, T_BALL, T_BOX,
                        T_DOOR, T_EMPTY, T_GOAL, T_KEY, T_LAVA, T_WALL)


@dataclasses.dataclass
class Layout:
    env_id: str
    seed: int
    width: int
    height: int
    max_steps: int
    see_through_walls: bool
    rule: int
    cells: np.ndarray          # uint8[4, H*W]
    agent_pos: tuple
    agent_dir: int
    mission: str
    agent_view_size: int = 7
    dyn: np.ndarray | None = None      # uint8[64], DynamicObstacles only: PCG64 state after _gen_grid + obstacle list (dyn_record)


_GREY = COLOR_TO_IDX["grey"]


class _Gen:
    """RNG helpers and placement logic of MiniGridEnv (minigrid_env.py:242-390) on a flat uint8 grid."""

    see_through_walls = False
    rule = RULE_NONE
    mission = ""

    def __init__(self, width, height, max_steps):
        self.width, self.height, self.max_steps = width, height, max_steps

    # -- Grid (core/grid.py:65-108) ------------------------------------------------------------
    def _new_grid(self):
        self.t = [T_EMPTY] * (self.width * self.height)
        self.c = [0] * (self.width * self.height)
        self.s = [0] * (self.width * self.height)
        self.a = [0] * (self.width * self.height)

    def _set(self, i, j, t=T_EMPTY, c=0, s=0, a=0):
        assert 0 <= i < self.width and 0 <= j < self.height            # grid.py:66-71
        k = j * self.width + i
        self.t[k], self.c[k], self.s[k], self.a[k] = t, c, s, a

    def _type(self, i, j):
        assert 0 <= i < self.width and 0 <= j < self.height            # grid.py:75-76
        return self.t[j * self.width + i]

    def _horz_wall(self, x, y, length=None):                            # grid.py:80-84
        if length is None:
            length = self.width - x
        for i in range(length):
            self._set(x + i, y, T_WALL, _GREY)

    def _vert_wall(self, x, y, length=None):                            # grid.py:86-90
        if length is None:
            length = self.height - y
        for j in range(length):
            self._set(x, y + j, T_WALL, _GREY)

    def _wall_rect(self, x, y, w, h):                                   # grid.py:92-96 (+ :104-108)
        self._horz_wall(x, y, w)
        self._horz_wall(x, y + h - 1, w)
        self._vert_wall(x, y, h)
        self._vert_wall(x + w - 1, y, h)

    # -- RNG helpers (minigrid_env.py:242-306) -------------------------------------------------
    def _rand_int(self, low, high):
        return int(self.rng.integers(low, high))

    def _rand_bool(self):
        return self._rand_int(0, 2) == 0

    def _rand_elem(self, lst):
        lst = list(lst)
        return lst[self._rand_int(0, len(lst))]

    def _rand_color(self):
        return self._rand_elem(COLOR_NAMES)

    # -- placement (minigrid_env.py:308-390) ---------------------------------------------------
    def _place_obj(self, obj, top=None, size=None, reject_fn=None, max_tries=float("inf")):
        top = (0, 0) if top is None else (max(top[0], 0), max(top[1], 0))
        if size is None:
            size = (self.width, self.height)
        num_tries = 0
        while True:
            if num_tries > max_tries:
                raise RecursionError("rejection sampling failed in place_obj")
            num_tries += 1
            pos = (self._rand_int(top[0], min(top[0] + size[0], self.width)),
                   self._rand_int(top[1], min(top[1] + size[1], self.height)))
            if self._type(*pos) != T_EMPTY:
                continue
            if pos == tuple(self.agent_pos):
                continue
            if reject_fn and reject_fn(self, pos):
                continue
            break
        if obj is not None:
            self._set(pos[0], pos[1], *obj)
        else:
            self._set(pos[0], pos[1])
        return pos

    def _place_agent(self, top=None, size=None, rand_dir=True, max_tries=float("inf")):
        self.agent_pos = (-1, -1)
        pos = self._place_obj(None, top, size, max_tries=max_tries)
        self.agent_pos = pos
        if rand_dir:
            self.agent_dir = self._rand_int(0, 4)
        return pos

    # -- reset (minigrid_env.py:119-157) -------------------------------------------------------
    def generate(self, env_id, seed):
        self.rng = np.random.Generator(np.random.PCG64(np.random.SeedSequence(seed)))   # gymnasium seeding.np_random
        self.agent_pos, self.agent_dir = (-1, -1), -1
        self._gen_grid(self.width, self.height)
        ax, ay = int(self.agent_pos[0]), int(self.agent_pos[1])
        assert ax >= 0 and ay >= 0 and self.agent_dir >= 0
        start = self._type(ax, ay)
        assert start in (T_EMPTY, T_GOAL, T_LAVA, 3) or (start == T_DOOR and self.s[ay * self.width + ax] == 0)
        cells = np.array([self.t, self.c, self.s, self.a], dtype=np.uint8)
        dyn = dyn_record(self.rng, self.obstacles) if getattr(self, "obstacles", None) is not None else None
        return Layout(env_id, seed, self.width, self.height, self.max_steps, self.see_through_walls, self.rule,
                      cells, (ax, ay), int(self.agent_dir), self.mission, dyn=dyn)


def dyn_record(rng, obstacles):
    """64-byte record the engine / oracle keep per DynamicObstacles env: the env's numpy Generator(PCG64) state
    (u64 state_lo, state_hi, inc_lo, inc_hi; u32 has_uint32, uinteger), u8 n_obstacles + 7 pad, u8 ox[8], oy[8] =
    `self.obstacles` in creation order (dynamicobstacles.py:129-132)."""
    st = rng.bit_generator.state
    assert st["bit_generator"] == "PCG64" and len(obstacles) <= 8
    m = (1 << 64) - 1
    s, inc = int(st["state"]["state"]), int(st["state"]["inc"])
    rec = np.zeros(64, np.uint8)
    rec[:32] = np.array([s & m, s >> 64, inc & m, inc >> 64], dtype="<u8").view(np.uint8)
    rec[32:40] = np.array([int(st["has_uint32"]), int(st["uinteger"])], dtype="<u4").view(np.uint8)
    rec[40] = len(obstacles)
    rec[48:64] = 0xFF
    for i, (x, y) in enumerate(obstacles):
        rec[48 + i], rec[56 + i] = x, y
    return rec


class EmptyGen(_Gen):
    """EmptyEnv, minigrid/envs/empty.py:68-114."""
    see_through_walls = True                                             # empty.py:88

    def __init__(self, size=8, agent_start_pos=(1, 1), agent_start_dir=0, max_steps=None):
        self.agent_start_pos, self.agent_start_dir = agent_start_pos, agent_start_dir
        super().__init__(size, size, 4 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self._set(width - 2, height - 2, T_GOAL, COLOR_TO_IDX["green"])
        if self.agent_start_pos is not None:
            self.agent_pos, self.agent_dir = self.agent_start_pos, self.agent_start_dir
        else:
            self._place_agent()
        self.mission = "get to the green goal square"


class DoorKeyGen(_Gen):
    """DoorKeyEnv, minigrid/envs/doorkey.py:62-100."""

    def __init__(self, size=8, max_steps=None):
        super().__init__(size, size, 10 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self._set(width - 2, height - 2, T_GOAL, COLOR_TO_IDX["green"])
        split = self._rand_int(2, width - 2)
        self._vert_wall(split, 0)
        self._place_agent(size=(split, height))
        door = self._rand_int(1, width - 2)
        self._set(split, door, T_DOOR, COLOR_TO_IDX["yellow"], 2)
        self._place_obj((T_KEY, COLOR_TO_IDX["yellow"], 0, 0), top=(0, 0), size=(split, height))
        self.mission = "use the key to open the door and then get to the goal"


class CrossingGen(_Gen):
    """CrossingEnv, minigrid/envs/crossing.py:87-184 (obstacle_type 'lava' or 'wall')."""

    def __init__(self, size=9, num_crossings=1, obstacle_type="lava", max_steps=None):
        self.num_crossings, self.obstacle_type = num_crossings, obstacle_type
        super().__init__(size, size, 4 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        assert width % 2 == 1 and height % 2 == 1
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self.agent_pos, self.agent_dir = (1, 1), 0
        self._set(width - 2, height - 2, T_GOAL, COLOR_TO_IDX["green"])
        v, h = 0, 1
        rivers = [(v, i) for i in range(2, height - 2, 2)]
        rivers += [(h, j) for j in range(2, width - 2, 2)]
        self.rng.shuffle(rivers)                     # Generator.shuffle on a python list: depends on len only
        rivers = rivers[: self.num_crossings]
        rivers_v = sorted(pos for d, pos in rivers if d == v)
        rivers_h = sorted(pos for d, pos in rivers if d == h)
        obstacle = (T_LAVA, COLOR_TO_IDX["red"]) if self.obstacle_type == "lava" else (T_WALL, _GREY)
        for i, j in itt.chain(itt.product(range(1, width - 1), rivers_h), itt.product(rivers_v, range(1, height - 1))):
            self._set(i, j, *obstacle)
        path = [h] * len(rivers_v) + [v] * len(rivers_h)
        self.rng.shuffle(path)
        limits_v = [0] + rivers_v + [height - 1]
        limits_h = [0] + rivers_h + [width - 1]
        room_i, room_j = 0, 0
        for d in path:
            if d == h:
                i = limits_v[room_i + 1]
                j = int(self.rng.choice(range(limits_h[room_j] + 1, limits_h[room_j + 1])))
                room_i += 1
            else:
                i = int(self.rng.choice(range(limits_v[room_i] + 1, limits_v[room_i + 1])))
                j = limits_h[room_j + 1]
                room_j += 1
            self._set(i, j)
        self.mission = ("avoid the lava and get to the green goal square" if self.obstacle_type == "lava"
                        else "find the opening and get to the green goal square")


class FourRoomsGen(_Gen):
    """FourRoomsEnv, minigrid/envs/fourrooms.py:59-128."""
    mission = "reach the goal"

    def __init__(self, agent_pos=None, goal_pos=None, max_steps=100):
        self._agent_default_pos, self._goal_default_pos = agent_pos, goal_pos
        super().__init__(19, 19, max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._horz_wall(0, 0)
        self._horz_wall(0, height - 1)
        self._vert_wall(0, 0)
        self._vert_wall(width - 1, 0)
        room_w, room_h = width // 2, height // 2
        for j in range(2):
            for i in range(2):
                xL, yT = i * room_w, j * room_h
                xR, yB = xL + room_w, yT + room_h
                if i + 1 < 2:
                    self._vert_wall(xR, yT, room_h)
                    self._set(xR, self._rand_int(yT + 1, yB))
                if j + 1 < 2:
                    self._horz_wall(xL, yB, room_w)
                    self._set(self._rand_int(xL + 1, xR), yB)
        if self._agent_default_pos is not None:
            self.agent_pos = tuple(self._agent_default_pos)
            self._set(*self.agent_pos)
            self.agent_dir = self._rand_int(0, 4)
        else:
            self._place_agent()
        goal = (T_GOAL, COLOR_TO_IDX["green"], 0, 0)
        if self._goal_default_pos is not None:
            self._set(*self._goal_default_pos, *goal)
        else:
            self._place_obj(goal)


class _Room:
    """Room, minigrid/core/roomgrid.py:23-66."""

    def __init__(self, top, size):
        self.top, self.size = top, size
        self.doors = [None] * 4          # right, down, left, up: None | True (wall removed) | door tuple
        self.door_pos = [None] * 4
        self.neighbors = [None] * 4
        self.locked = False


def _reject_next_to(gen, pos):            # roomgrid.py:11-20
    sx, sy = gen.agent_pos
    return abs(sx - pos[0]) + abs(sy - pos[1]) < 2


class RoomGridGen(_Gen):
    """RoomGrid, minigrid/core/roomgrid.py:69-394."""

    def __init__(self, room_size=7, num_rows=3, num_cols=3, max_steps=100):
        assert room_size >= 3 and num_rows > 0 and num_cols > 0
        self.room_size, self.num_rows, self.num_cols = room_size, num_rows, num_cols
        super().__init__((room_size - 1) * num_cols + 1, (room_size - 1) * num_rows + 1, max_steps)

    def _get_room(self, i, j):
        assert i < self.num_cols and j < self.num_rows
        return self.room_grid[j][i]

    def _gen_grid(self, width, height):                                   # roomgrid.py:123-179
        self._new_grid()
        rs = self.room_size
        self.room_grid = []
        for j in range(self.num_rows):
            row = []
            for i in range(self.num_cols):
                room = _Room((i * (rs - 1), j * (rs - 1)), (rs, rs))
                row.append(room)
                self._wall_rect(*room.top, *room.size)
            self.room_grid.append(row)
        for j in range(self.num_rows):
            for i in range(self.num_cols):
                room = self.room_grid[j][i]
                x_l, y_l = room.top[0] + 1, room.top[1] + 1
                x_m, y_m = room.top[0] + room.size[0] - 1, room.top[1] + room.size[1] - 1
                if i < self.num_cols - 1:
                    room.neighbors[0] = self.room_grid[j][i + 1]
                    room.door_pos[0] = (x_m, self._rand_int(y_l, y_m))
                if j < self.num_rows - 1:
                    room.neighbors[1] = self.room_grid[j + 1][i]
                    room.door_pos[1] = (self._rand_int(x_l, x_m), y_m)
                if i > 0:
                    room.neighbors[2] = self.room_grid[j][i - 1]
                    room.door_pos[2] = room.neighbors[2].door_pos[0]
                if j > 0:
                    room.neighbors[3] = self.room_grid[j - 1][i]
                    room.door_pos[3] = room.neighbors[3].door_pos[1]
        self.agent_pos = ((self.num_cols // 2) * (rs - 1) + rs // 2, (self.num_rows // 2) * (rs - 1) + rs // 2)
        self.agent_dir = 0

    def _add_object(self, i, j, kind=None, color=None):                   # roomgrid.py:197-229, 181-195
        if kind is None:
            kind = self._rand_elem(["key", "ball", "box"])
        if color is None:
            color = self._rand_color()
        t = {"key": T_KEY, "ball": T_BALL, "box": T_BOX}[kind]
        room = self._get_room(i, j)
        obj = (t, COLOR_TO_IDX[color], 0, 0)
        pos = self._place_obj(obj, room.top, room.size, reject_fn=_reject_next_to, max_tries=1000)
        return (kind, color), pos

    def _add_door(self, i, j, door_idx=None, color=None, locked=None):    # roomgrid.py:231-278
        room = self._get_room(i, j)
        if door_idx is None:
            while True:
                door_idx = self._rand_int(0, 4)
                if room.neighbors[door_idx] and room.doors[door_idx] is None:
                    break
        if color is None:
            color = self._rand_color()
        if locked is None:
            locked = self._rand_bool()
        assert room.doors[door_idx] is None, "door already exists"
        room.locked = locked
        pos = room.door_pos[door_idx]
        assert pos is not None
        self._set(pos[0], pos[1], T_DOOR, COLOR_TO_IDX[color], 2 if locked else 1)
        door = ("door", color)
        room.doors[door_idx] = door
        room.neighbors[door_idx].doors[(door_idx + 2) % 4] = door
        return door, pos

    def _remove_wall(self, i, j, wall_idx):                               # roomgrid.py:280-317
        room = self._get_room(i, j)
        assert 0 <= wall_idx < 4 and room.doors[wall_idx] is None and room.neighbors[wall_idx]
        tx, ty = room.top
        w, h = room.size
        if wall_idx == 0:
            for k in range(1, h - 1):
                self._set(tx + w - 1, ty + k)
        elif wall_idx == 1:
            for k in range(1, w - 1):
                self._set(tx + k, ty + h - 1)
        elif wall_idx == 2:
            for k in range(1, h - 1):
                self._set(tx, ty + k)
        else:
            for k in range(1, w - 1):
                self._set(tx + k, ty)
        room.doors[wall_idx] = True
        room.neighbors[wall_idx].doors[(wall_idx + 2) % 4] = True

    def _place_agent_in_room(self, i=None, j=None, rand_dir=True):        # roomgrid.py:319-340
        if i is None:
            i = self._rand_int(0, self.num_cols)
        if j is None:
            j = self._rand_int(0, self.num_rows)
        room = self.room_grid[j][i]
        while True:
            self._place_agent(room.top, room.size, rand_dir, max_tries=1000)
            dx, dy = ((1, 0), (0, 1), (-1, 0), (0, -1))[self.agent_dir]
            front = self._type(self.agent_pos[0] + dx, self.agent_pos[1] + dy)
            if front == T_EMPTY or front == T_WALL:
                break
        return self.agent_pos

    def _connect_all(self, door_colors=COLOR_NAMES, max_itrs=5000):       # roomgrid.py:342-394
        rs = self.room_size
        start_room = self.room_grid[self.agent_pos[1] // (rs - 1)][self.agent_pos[0] // (rs - 1)]

        def find_reach():
            reach, stack = set(), [start_room]
            while stack:
                room = stack.pop()
                if room in reach:
                    continue
                reach.add(room)
                for k in range(4):
                    if room.doors[k]:
                        stack.append(room.neighbors[k])
            return reach

        num_itrs = 0
        while True:
            if num_itrs > max_itrs:
                raise RecursionError("connect_all failed")
            num_itrs += 1
            if len(find_reach()) == self.num_rows * self.num_cols:
                break
            i = self._rand_int(0, self.num_cols)
            j = self._rand_int(0, self.num_rows)
            k = self._rand_int(0, 4)
            room = self._get_room(i, j)
            if not room.door_pos[k] or room.doors[k]:
                continue
            if room.locked or room.neighbors[k].locked:
                continue
            color = self._rand_elem(door_colors)
            self._add_door(i, j, k, color, False)


class KeyCorridorGen(RoomGridGen):
    """KeyCorridorEnv, minigrid/envs/keycorridor.py:76-137 (pickup-target terminal rule :129-137)."""
    rule = RULE_PICKUP_TARGET

    def __init__(self, num_rows=3, obj_type="ball", room_size=6, max_steps=None):
        self.obj_type = obj_type
        super().__init__(room_size=room_size, num_rows=num_rows,
                         max_steps=30 * room_size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        super()._gen_grid(width, height)
        for j in range(1, self.num_rows):
            self._remove_wall(1, j, 3)
        room_idx = self._rand_int(0, self.num_rows)
        door, _ = self._add_door(2, room_idx, 2, locked=True)
        obj, pos = self._add_object(2, room_idx, kind=self.obj_type)
        self._add_object(0, self._rand_int(0, self.num_rows), "key", door[1])
        self._place_agent_in_room(1, self.num_rows // 2)
        self._connect_all()
        self.a[pos[1] * self.width + pos[0]] |= AUX_TARGET              # `self.obj = obj`, keycorridor.py:126
        self.mission = f"pick up the {obj[1]} {obj[0]}"


class UnlockGen(RoomGridGen):
    """UnlockEnv, minigrid/envs/unlock.py:56-97 (rule: a toggle while the door is open pays and terminates, :89-97)."""
    rule = RULE_UNLOCK

    def __init__(self, max_steps=None):
        room_size = 6
        super().__init__(room_size=room_size, num_rows=1, num_cols=2,
                         max_steps=8 * room_size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        super()._gen_grid(width, height)
        door, pos = self._add_door(0, 0, 0, locked=True)
        self._add_object(0, 0, "key", door[1])
        self._place_agent_in_room(0, 0)
        self.a[pos[1] * self.width + pos[0]] |= AUX_TARGET              # `self.door = door`, unlock.py:86
        self.mission = "open the door"


class UnlockPickupGen(RoomGridGen):
    """UnlockPickupEnv, minigrid/envs/unlockpickup.py:63-106 (pickup-target rule :98-106, same as KeyCorridor's)."""
    rule = RULE_PICKUP_TARGET
    blocked = False

    def __init__(self, max_steps=None):
        room_size = 6
        k = 16 if self.blocked else 8                                    # blockedunlockpickup.py:75, unlockpickup.py:67
        super().__init__(room_size=room_size, num_rows=1, num_cols=2,
                         max_steps=k * room_size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        super()._gen_grid(width, height)
        obj, opos = self._add_object(1, 0, kind="box")
        door, pos = self._add_door(0, 0, 0, locked=True)
        if self.blocked:                                                  # blockedunlockpickup.py:99-101
            color = self._rand_color()
            self._set(pos[0] - 1, pos[1], T_BALL, COLOR_TO_IDX[color])
        self._add_object(0, 0, "key", door[1])
        self._place_agent_in_room(0, 0)
        self.a[opos[1] * self.width + opos[0]] |= AUX_TARGET            # `self.obj = obj`
        self.mission = f"pick up the {obj[1]} {obj[0]}"


class BlockedUnlockPickupGen(UnlockPickupGen):
    """BlockedUnlockPickupEnv, minigrid/envs/blockedunlockpickup.py:66-116."""
    blocked = True


class DistShiftGen(_Gen):
    """DistShiftEnv, minigrid/envs/distshift.py:65-121."""
    see_through_walls = True

    def __init__(self, width=9, height=7, agent_start_pos=(1, 1), agent_start_dir=0, strip2_row=2, max_steps=None):
        self.agent_start_pos, self.agent_start_dir, self.strip2_row = agent_start_pos, agent_start_dir, strip2_row
        super().__init__(width, height, 4 * width * height if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self._set(width - 2, 1, T_GOAL, COLOR_TO_IDX["green"])
        for i in range(width - 6):
            self._set(3 + i, 1, T_LAVA, COLOR_TO_IDX["red"])
            self._set(3 + i, self.strip2_row, T_LAVA, COLOR_TO_IDX["red"])
        if self.agent_start_pos is not None:
            self.agent_pos, self.agent_dir = self.agent_start_pos, self.agent_start_dir
        else:
            self._place_agent()
        self.mission = "get to the green goal square"


class LavaGapGen(_Gen):
    """LavaGapEnv, minigrid/envs/lavagap.py:69-133."""

    def __init__(self, size=5, obstacle_type="lava", max_steps=None):
        self.obstacle_type = obstacle_type
        super().__init__(size, size, 4 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        assert width >= 5 and height >= 5
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self.agent_pos, self.agent_dir = (1, 1), 0
        self._set(width - 2, height - 2, T_GOAL, COLOR_TO_IDX["green"])
        gap = (self._rand_int(2, width - 2), self._rand_int(1, height - 1))
        for j in range(height - 2):                                       # vert_wall(gap_x, 1, height - 2, obstacle_type)
            if self.obstacle_type == "lava":
                self._set(gap[0], 1 + j, T_LAVA, COLOR_TO_IDX["red"])
            else:
                self._set(gap[0], 1 + j, T_WALL, _GREY)
        self._set(gap[0], gap[1])
        self.mission = ("avoid the lava and get to the green goal square" if self.obstacle_type == "lava"
                        else "find the opening and get to the green goal square")


class FetchGen(_Gen):
    """FetchEnv, minigrid/envs/fetch.py:70-176.  Every object of the target's type AND colour carries the target bit
    (the rule compares type and colour, not identity: fetch.py:166-169)."""
    see_through_walls = True
    rule = RULE_FETCH

    def __init__(self, size=8, numObjs=3, max_steps=None):
        self.numObjs = numObjs
        super().__init__(size, size, 5 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._horz_wall(0, 0)
        self._horz_wall(0, height - 1)
        self._vert_wall(0, 0)
        self._vert_wall(width - 1, 0)
        objs = []
        while len(objs) < self.numObjs:
            kind = self._rand_elem(["key", "ball"])
            color = self._rand_elem(COLOR_NAMES)
            pos = self._place_obj(({"key": T_KEY, "ball": T_BALL}[kind], COLOR_TO_IDX[color], 0, 0))
            objs.append((kind, color, pos))
        self._place_agent()
        kind, color, _ = objs[self._rand_int(0, len(objs))]
        for k2, c2, pos in objs:
            if (k2, c2) == (kind, color):
                self.a[pos[1] * self.width + pos[0]] |= AUX_TARGET
        syntax = ["get a", "go get a", "fetch a", "go fetch a", "you must fetch a"][self._rand_int(0, 5)]
        self.mission = f"{syntax} {color} {kind}"


class GoToDoorGen(_Gen):
    """GoToDoorEnv, minigrid/envs/gotodoor.py:66-150."""
    see_through_walls = True
    rule = RULE_GOTO

    def __init__(self, size=5, max_steps=None):
        assert size >= 5
        super().__init__(size, size, 4 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        width = self._rand_int(5, width + 1)
        height = self._rand_int(5, height + 1)
        self._wall_rect(0, 0, width, height)
        pos = [(self._rand_int(2, width - 2), 0), (self._rand_int(2, width - 2), height - 1),
               (0, self._rand_int(2, height - 2)), (width - 1, self._rand_int(2, height - 2))]
        colors = []
        while len(colors) < len(pos):
            color = self._rand_elem(COLOR_NAMES)
            if color in colors:
                continue
            colors.append(color)
        for p, color in zip(pos, colors):
            self._set(p[0], p[1], T_DOOR, COLOR_TO_IDX[color], 1)
        self._place_agent(size=(width, height))
        k = self._rand_int(0, len(pos))
        self.a[pos[k][1] * self.width + pos[k][0]] |= AUX_TARGET        # `self.target_pos`, gotodoor.py:125
        self.mission = f"go to the {colors[k]} door"


class GoToObjectGen(_Gen):
    """GoToObjectEnv, minigrid/envs/gotoobject.py:64-162.  The target bit marks the cell `target_pos` names in the initial
    layout; the rule is about that POSITION (gotoobject.py:139,149), wherever the object is carried later."""
    see_through_walls = True
    rule = RULE_GOTO

    def __init__(self, size=6, numObjs=2, max_steps=None):
        self.numObjs = numObjs
        super().__init__(size, size, 5 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        objs, poss = [], []
        while len(objs) < self.numObjs:
            kind = self._rand_elem(["key", "ball", "box"])
            color = self._rand_elem(COLOR_NAMES)
            if (kind, color) in objs:
                continue
            pos = self._place_obj(({"key": T_KEY, "ball": T_BALL, "box": T_BOX}[kind], COLOR_TO_IDX[color], 0, 0))
            objs.append((kind, color))
            poss.append(pos)
        self._place_agent()
        k = self._rand_int(0, len(objs))
        self.a[poss[k][1] * self.width + poss[k][0]] |= AUX_TARGET
        self.mission = f"go to the {objs[k][1]} {objs[k][0]}"


class RedBlueDoorsGen(_Gen):
    """RedBlueDoorsEnv, minigrid/envs/redbluedoors.py:58-133; both doors carry the target bit (the rule tells them apart by colour)."""
    rule = RULE_REDBLUE

    def __init__(self, size=8, max_steps=None):
        self.size = size
        super().__init__(2 * size, size, 20 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        sz = self.size
        self._wall_rect(0, 0, 2 * sz, sz)
        self._wall_rect(sz // 2, 0, sz, sz)
        self._place_agent(top=(sz // 2, 0), size=(sz, sz))
        y = self._rand_int(1, sz - 1)
        self._set(sz // 2, y, T_DOOR, COLOR_TO_IDX["red"], 1, AUX_TARGET)
        y = self._rand_int(1, sz - 1)
        self._set(sz // 2 + sz - 1, y, T_DOOR, COLOR_TO_IDX["blue"], 1, AUX_TARGET)
        self.mission = "open the red door then the blue door"



_OBJ_T = {"key": T_KEY, "ball": T_BALL, "box": T_BOX}
_DIR_VEC = ((1, 0), (0, 1), (-1, 0), (0, -1))                           # constants.py:49-58


class ObstructedMazeGen(RoomGridGen):
    """ObstructedMazeEnv and its registered shapes, minigrid/envs/obstructedmaze.py:9-271 and obstructedmaze_v1.py:9-99
    (`v1=True`: doors and blocking balls first, keys afterwards).  Terminal rule = pickup of `self.obj`
    (obstructedmaze.py:126-134), i.e. KeyCorridor's.  A key hidden in a box is the box's aux byte."""
    rule = RULE_PICKUP_TARGET

    def __init__(self, shape="full", agent_room=(1, 1), key_in_box=True, blocked=True, num_quarters=4,
                 num_rooms_visited=25, v1=False, max_steps=None):
        self.shape, self.agent_room, self.key_in_box, self.blocked = shape, tuple(agent_room), key_in_box, blocked
        self.num_quarters, self.v1 = num_quarters, v1
        rows, cols = (1, 2) if shape == "1d" else (3, 3)
        if shape == "1d":
            num_rooms_visited = 2                                         # obstructedmaze.py:176
        room_size = 6
        super().__init__(room_size=room_size, num_rows=rows, num_cols=cols,
                         max_steps=4 * num_rooms_visited * room_size ** 2 if max_steps is None else max_steps)

    def _rand_subset(self, items, n):                                     # minigrid_env.py:272-287
        lst, out = list(items), []
        while len(out) < n:
            e = self._rand_elem(lst)
            lst.remove(e)
            out.append(e)
        return out

    def _place_in_room(self, i, j, obj):                                  # roomgrid.py:181-196
        room = self._get_room(i, j)
        return self._place_obj(obj, room.top, room.size, reject_fn=_reject_next_to, max_tries=1000)

    def _locked_door(self, i, j, door_idx, color, blocked):               # obstructedmaze.py:136-158 / _v1.py:75-84
        door, pos = self._add_door(i, j, door_idx, color, locked=True)
        if blocked:
            dx, dy = _DIR_VEC[door_idx]
            self._set(pos[0] - dx, pos[1] - dy, T_BALL, COLOR_TO_IDX[self.blocking_ball_color])
        return door, pos

    def _key(self, i, j, color, key_in_box):                              # obstructedmaze.py:160-166 / _v1.py:86-99
        if key_in_box:
            obj = (T_BOX, COLOR_TO_IDX[self.box_color], 0, T_KEY | (COLOR_TO_IDX[color] << 4))
        else:
            obj = (T_KEY, COLOR_TO_IDX[color], 0, 0)
        self._place_in_room(i, j, obj)

    def _gen_grid(self, width, height):
        super()._gen_grid(width, height)
        self.door_colors = self._rand_subset(COLOR_NAMES, len(COLOR_NAMES))       # obstructedmaze.py:114-122
        self.ball_to_find_color, self.blocking_ball_color, self.box_color = COLOR_NAMES[0], COLOR_NAMES[1], COLOR_NAMES[2]
        if self.shape == "1d":                                            # ObstructedMaze_1Dlhb, :181-195
            self._locked_door(0, 0, 0, self.door_colors[0], self.blocked)
            self._key(0, 0, self.door_colors[0], self.key_in_box)
            _, pos = self._add_object(1, 0, "ball", self.ball_to_find_color)
            self._place_agent_in_room(0, 0)
        else:                                                             # ObstructedMaze_Full, :222-256 / _v1.py:38-73
            side_rooms = [(2, 1), (1, 2), (0, 1), (1, 0)][: self.num_quarters]
            for i, side in enumerate(side_rooms):
                self._add_door(1, 1, i, self.door_colors[i], locked=False)
                for k in (-1, 1):
                    c = self.door_colors[(i + k) % len(self.door_colors)]
                    self._locked_door(side[0], side[1], (i + k) % 4, c, self.blocked)
                    if not self.v1:
                        self._key(side[0], side[1], c, self.key_in_box)
                if self.v1:
                    for k in (-1, 1):
                        self._key(side[0], side[1], self.door_colors[(i + k) % len(self.door_colors)], self.key_in_box)
            corners = [(2, 0), (2, 2), (0, 2), (0, 0)][: self.num_quarters]
            room = self._rand_elem(corners)
            _, pos = self._add_object(room[0], room[1], "ball", self.ball_to_find_color)
            self._place_agent_in_room(*self.agent_room)
        self.a[pos[1] * self.width + pos[0]] |= AUX_TARGET                # `self.obj`
        self.mission = f"pick up the {self.ball_to_find_color} ball"


class LockedRoomGen(_Gen):
    """LockedRoomEnv, minigrid/envs/lockedroom.py:24-174 (base step: reach the goal behind the locked door)."""

    def __init__(self, size=19, max_steps=None):
        super().__init__(size, size, 10 * size if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        lw, rw = width // 2 - 2, width // 2 + 2
        self._vert_wall(lw, 0)
        self._vert_wall(rw, 0)
        rooms = []                                                        # [top, size, door_pos, locked, color]
        for n in range(3):
            j = n * (height // 3)
            self._horz_wall(0, j, lw)
            self._horz_wall(rw, j, width - rw)
            size = (lw + 1, height // 3 + 1)
            rooms.append([(0, j), size, (lw, j + 3), False, None])
            rooms.append([(rw, j), size, (rw, j + 3), False, None])

        def rand_pos(room):                                               # lockedroom.py:18-21, minigrid_env.py:296-306
            (tx, ty), (sx, sy) = room[0], room[1]
            return (self._rand_int(tx + 1, tx + sx - 1), self._rand_int(ty + 1, ty + sy - 1))

        locked = self._rand_elem(rooms)
        locked[3] = True
        self._set(*rand_pos(locked), T_GOAL, COLOR_TO_IDX["green"])
        colors = set(COLOR_NAMES)
        for room in rooms:
            color = self._rand_elem(sorted(colors))
            colors.remove(color)
            room[4] = color
            self._set(room[2][0], room[2][1], T_DOOR, COLOR_TO_IDX[color], 2 if room[3] else 1)
        while True:
            key_room = self._rand_elem(rooms)
            if key_room is not locked:
                break
        self._set(*rand_pos(key_room), T_KEY, COLOR_TO_IDX[locked[4]])
        self._place_agent(top=(lw, 0), size=(rw - lw, height))
        self.mission = f"get the {locked[4]} key from the {key_room[4]} room, unlock the {locked[4]} door and go to the goal"


class MultiRoomGen(_Gen):
    """MultiRoomEnv, minigrid/envs/multiroom.py:17-281 (base step; 25x25 grid, recursive room placement)."""

    def __init__(self, minNumRooms, maxNumRooms, maxRoomSize=10, max_steps=None):
        assert minNumRooms > 0 and maxNumRooms >= minNumRooms and maxRoomSize >= 4
        self.minNumRooms, self.maxNumRooms, self.maxRoomSize = minNumRooms, maxNumRooms, maxRoomSize
        super().__init__(25, 25, maxNumRooms * 20 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        room_list = []
        num_rooms = self._rand_int(self.minNumRooms, self.maxNumRooms + 1)
        while len(room_list) < num_rooms:
            cur = []
            entry = (self._rand_int(0, width - 2), self._rand_int(0, width - 2))
            self._place_room(num_rooms, cur, 4, self.maxRoomSize, 2, entry)
            if len(cur) > len(room_list):
                room_list = cur
        self._new_grid()
        prev_color = None
        for idx, (top, size, entry) in enumerate(room_list):
            self._wall_rect(top[0], top[1], size[0], size[1])
            if idx > 0:
                colors = set(COLOR_NAMES)
                if prev_color:
                    colors.remove(prev_color)
                color = self._rand_elem(sorted(colors))
                self._set(entry[0], entry[1], T_DOOR, COLOR_TO_IDX[color], 1)
                prev_color = color
        self._place_agent(room_list[0][0], room_list[0][1])
        self._place_obj((T_GOAL, COLOR_TO_IDX["green"], 0, 0), room_list[-1][0], room_list[-1][1])
        self.mission = "traverse the rooms to get to the goal"

    def _place_room(self, num_left, room_list, min_sz, max_sz, entry_wall, entry_pos):       # multiroom.py:178-281
        sx, sy = self._rand_int(min_sz, max_sz + 1), self._rand_int(min_sz, max_sz + 1)
        if len(room_list) == 0:
            tx, ty = entry_pos
        elif entry_wall == 0:
            tx = entry_pos[0] - sx + 1
            ty = self._rand_int(entry_pos[1] - sy + 2, entry_pos[1])
        elif entry_wall == 1:
            tx = self._rand_int(entry_pos[0] - sx + 2, entry_pos[0])
            ty = entry_pos[1] - sy + 1
        elif entry_wall == 2:
            tx = entry_pos[0]
            ty = self._rand_int(entry_pos[1] - sy + 2, entry_pos[1])
        else:
            tx = self._rand_int(entry_pos[0] - sx + 2, entry_pos[0])
            ty = entry_pos[1]
        if tx < 0 or ty < 0:
            return False
        if tx + sx > self.width or ty + sy >= self.height:
            return False
        for (rt, rs, _) in room_list[:-1]:
            non_overlap = tx + sx < rt[0] or rt[0] + rs[0] <= tx or ty + sy < rt[1] or rt[1] + rs[1] <= ty
            if not non_overlap:
                return False
        room_list.append(((tx, ty), (sx, sy), entry_pos))
        if num_left == 1:
            return True
        for _ in range(8):
            walls = {0, 1, 2, 3}
            walls.remove(entry_wall)
            exit_wall = self._rand_elem(sorted(walls))
            if exit_wall == 0:
                exit_pos = (tx + sx - 1, ty + self._rand_int(1, sy - 1))
            elif exit_wall == 1:
                exit_pos = (tx + self._rand_int(1, sx - 1), ty + sy - 1)
            elif exit_wall == 2:
                exit_pos = (tx, ty + self._rand_int(1, sy - 1))
            else:
                exit_pos = (tx + self._rand_int(1, sx - 1), ty)
            if self._place_room(num_left - 1, room_list, min_sz, max_sz, (exit_wall + 2) % 4, exit_pos):
                break
        return True


class PlaygroundGen(_Gen):
    """PlaygroundEnv, minigrid/envs/playground.py:8-91 (no goal: only truncation ends an episode)."""

    def __init__(self, max_steps=100):
        super().__init__(19, 19, max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        rw, rh = width // 3, height // 3
        for j in range(3):
            for i in range(3):
                xl, yt = i * rw, j * rh
                xr, yb = xl + rw, yt + rh
                if i + 1 < 3:
                    self._vert_wall(xr, yt, rh)
                    pos = (xr, self._rand_int(yt + 1, yb - 1))
                    self._set(pos[0], pos[1], T_DOOR, COLOR_TO_IDX[self._rand_elem(COLOR_NAMES)], 1)
                if j + 1 < 3:
                    self._horz_wall(xl, yb, rw)
                    pos = (self._rand_int(xl + 1, xr - 1), yb)
                    self._set(pos[0], pos[1], T_DOOR, COLOR_TO_IDX[self._rand_elem(COLOR_NAMES)], 1)
        self._place_agent()
        for _ in range(12):
            kind = self._rand_elem(["key", "ball", "box"])
            color = self._rand_elem(COLOR_NAMES)
            self._place_obj((_OBJ_T[kind], COLOR_TO_IDX[color], 0, 0))
        self.mission = ""


class MemoryGen(_Gen):
    """MemoryEnv, minigrid/envs/memory.py:12-165.  `success_pos` / `failure_pos` (:144-149) are empty cells: they carry the
    target bit (AUX_TARGET) resp. AUX_TARGET_B in the aux plane; the rule (:153-165) compares the agent's position."""
    rule = RULE_MEMORY

    def __init__(self, size=8, random_length=False, max_steps=None):
        self.random_length = random_length
        super().__init__(size, size, 5 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        assert height % 2 == 1
        up, lo = height // 2 - 2, height // 2 + 2
        hallway_end = self._rand_int(4, width - 2) if self.random_length else width - 3
        for i in range(1, 5):
            self._set(i, up, T_WALL, _GREY)
            self._set(i, lo, T_WALL, _GREY)
        self._set(4, up + 1, T_WALL, _GREY)
        self._set(4, lo - 1, T_WALL, _GREY)
        for i in range(5, hallway_end):
            self._set(i, up + 1, T_WALL, _GREY)
            self._set(i, lo - 1, T_WALL, _GREY)
        for j in range(height):
            if j != height // 2:
                self._set(hallway_end, j, T_WALL, _GREY)
            self._set(hallway_end + 2, j, T_WALL, _GREY)
        self.agent_pos = (self._rand_int(1, hallway_end + 1), height // 2)
        self.agent_dir = 0
        green = COLOR_TO_IDX["green"]
        start = self._rand_elem([T_KEY, T_BALL])
        self._set(1, height // 2 - 1, start, green)
        other = self._rand_elem([[T_BALL, T_KEY], [T_KEY, T_BALL]])
        pos0, pos1 = (hallway_end + 1, height // 2 - 2), (hallway_end + 1, height // 2 + 2)
        self._set(pos0[0], pos0[1], other[0], green)
        self._set(pos1[0], pos1[1], other[1], green)
        if start == other[0]:
            success, failure = (pos0[0], pos0[1] + 1), (pos1[0], pos1[1] - 1)
        else:
            success, failure = (pos1[0], pos1[1] - 1), (pos0[0], pos0[1] + 1)
        self.a[success[1] * width + success[0]] = AUX_TARGET
        self.a[failure[1] * width + failure[0]] = AUX_TARGET_B
        self.mission = "go to the matching object at the end of the hallway"


class PutNearGen(_Gen):
    """PutNearEnv, minigrid/envs/putnear.py:13-201.  The object to move carries AUX_TARGET (the rule compares type and colour,
    and objects are unique in both, :127-131,186-192); the target object's initial cell (`target_pos`, :167) AUX_TARGET_B."""
    see_through_walls = True
    rule = RULE_PUTNEAR

    def __init__(self, size=6, numObjs=2, max_steps=None):
        self.numObjs = numObjs
        super().__init__(size, size, 5 * size if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        objs, poss = [], []

        def near_obj(gen, p1):                                            # putnear.py:115-121
            return any(abs(p1[0] - p2[0]) <= 1 and abs(p1[1] - p2[1]) <= 1 for p2 in poss)

        while len(objs) < self.numObjs:
            kind = self._rand_elem(["key", "ball", "box"])
            color = self._rand_elem(COLOR_NAMES)
            if (kind, color) in objs:
                continue
            pos = self._place_obj((_OBJ_T[kind], COLOR_TO_IDX[color], 0, 0), reject_fn=near_obj)
            objs.append((kind, color))
            poss.append(pos)
        self._place_agent()
        k = self._rand_int(0, len(objs))
        while True:
            t = self._rand_int(0, len(objs))
            if t != k:
                break
        self.a[poss[k][1] * width + poss[k][0]] = AUX_TARGET
        self.a[poss[t][1] * width + poss[t][0]] = AUX_TARGET_B
        self.mission = f"put the {objs[k][1]} {objs[k][0]} near the {objs[t][1]} {objs[t][0]}"



class DynamicObstaclesGen(_Gen):
    """DynamicObstaclesEnv, minigrid/envs/dynamicobstacles.py:14-167.  The obstacles (blue balls) move inside `step` with
    draws from the env's own generator: the layout carries the generator state after `_gen_grid` and the obstacle list."""
    see_through_walls = True
    rule = RULE_DYNOBS

    def __init__(self, size=8, agent_start_pos=(1, 1), agent_start_dir=0, n_obstacles=4, max_steps=None):
        self.agent_start_pos, self.agent_start_dir = agent_start_pos, agent_start_dir
        self.n_obstacles = int(n_obstacles) if n_obstacles <= size / 2 + 1 else int(size / 2)   # :84-88
        super().__init__(size, size, 4 * size ** 2 if max_steps is None else max_steps)

    def _gen_grid(self, width, height):
        self._new_grid()
        self._wall_rect(0, 0, width, height)
        self._set(width - 2, height - 2, T_GOAL, COLOR_TO_IDX["green"])
        if self.agent_start_pos is not None:
            self.agent_pos, self.agent_dir = tuple(self.agent_start_pos), self.agent_start_dir
        else:
            self._place_agent()
        self.obstacles = []
        for _ in range(self.n_obstacles):
            self.obstacles.append(self._place_obj((T_BALL, COLOR_TO_IDX["blue"], 0, 0), max_tries=100))
        self.mission = "get to the green goal square"


def _kc(rs, rows):
    return (KeyCorridorGen, {"room_size": rs, "num_rows": rows})


# id -> (generator class, kwargs): the rows of minigrid/__init__.py:23-1130 for these families
REGISTRY = {
    "MiniGrid-Empty-5x5-v0": (EmptyGen, {"size": 5}),
    "MiniGrid-Empty-Random-5x5-v0": (EmptyGen, {"size": 5, "agent_start_pos": None}),
    "MiniGrid-Empty-6x6-v0": (EmptyGen, {"size": 6}),
    "MiniGrid-Empty-Random-6x6-v0": (EmptyGen, {"size": 6, "agent_start_pos": None}),
    "MiniGrid-Empty-8x8-v0": (EmptyGen, {}),
    "MiniGrid-Empty-16x16-v0": (EmptyGen, {"size": 16}),
    "MiniGrid-DoorKey-5x5-v0": (DoorKeyGen, {"size": 5}),
    "MiniGrid-DoorKey-6x6-v0": (DoorKeyGen, {"size": 6}),
    "MiniGrid-DoorKey-8x8-v0": (DoorKeyGen, {"size": 8}),
    "MiniGrid-DoorKey-16x16-v0": (DoorKeyGen, {"size": 16}),
    "MiniGrid-FourRooms-v0": (FourRoomsGen, {}),
    "MiniGrid-KeyCorridorS3R1-v0": _kc(3, 1), "MiniGrid-KeyCorridorS3R2-v0": _kc(3, 2),
    "MiniGrid-KeyCorridorS3R3-v0": _kc(3, 3), "MiniGrid-KeyCorridorS4R3-v0": _kc(4, 3),
    "MiniGrid-KeyCorridorS5R3-v0": _kc(5, 3), "MiniGrid-KeyCorridorS6R3-v0": _kc(6, 3),
}
REGISTRY.update({
    "MiniGrid-Unlock-v0": (UnlockGen, {}),
    "MiniGrid-UnlockPickup-v0": (UnlockPickupGen, {}),
    "MiniGrid-BlockedUnlockPickup-v0": (BlockedUnlockPickupGen, {}),
    "MiniGrid-Fetch-5x5-N2-v0": (FetchGen, {"size": 5, "numObjs": 2}),
    "MiniGrid-Fetch-6x6-N2-v0": (FetchGen, {"size": 6, "numObjs": 2}),
    "MiniGrid-Fetch-8x8-N3-v0": (FetchGen, {}),
    "MiniGrid-GoToDoor-5x5-v0": (GoToDoorGen, {}),
    "MiniGrid-GoToDoor-6x6-v0": (GoToDoorGen, {"size": 6}),
    "MiniGrid-GoToDoor-8x8-v0": (GoToDoorGen, {"size": 8}),
    "MiniGrid-GoToObject-6x6-N2-v0": (GoToObjectGen, {}),
    "MiniGrid-GoToObject-8x8-N2-v0": (GoToObjectGen, {"size": 8, "numObjs": 2}),
    "MiniGrid-RedBlueDoors-6x6-v0": (RedBlueDoorsGen, {"size": 6}),
    "MiniGrid-RedBlueDoors-8x8-v0": (RedBlueDoorsGen, {}),
    "MiniGrid-DistShift1-v0": (DistShiftGen, {"strip2_row": 2}),
    "MiniGrid-DistShift2-v0": (DistShiftGen, {"strip2_row": 5}),
    "MiniGrid-LavaGapS5-v0": (LavaGapGen, {"size": 5}),
    "MiniGrid-LavaGapS6-v0": (LavaGapGen, {"size": 6}),
    "MiniGrid-LavaGapS7-v0": (LavaGapGen, {"size": 7}),
})
for _s, _n in ((9, 1), (9, 2), (9, 3), (11, 5)):
    REGISTRY[f"MiniGrid-LavaCrossingS{_s}N{_n}-v0"] = (CrossingGen, {"size": _s, "num_crossings": _n})
    REGISTRY[f"MiniGrid-SimpleCrossingS{_s}N{_n}-v0"] = (CrossingGen, {"size": _s, "num_crossings": _n,
                                                                       "obstacle_type": "wall"})


def _om(**kw):
    return (ObstructedMazeGen, kw)


REGISTRY.update({
    "MiniGrid-ObstructedMaze-1Dl-v0": _om(shape="1d", key_in_box=False, blocked=False),
    "MiniGrid-ObstructedMaze-1Dlh-v0": _om(shape="1d", key_in_box=True, blocked=False),
    "MiniGrid-ObstructedMaze-1Dlhb-v0": _om(shape="1d"),
    "MiniGrid-ObstructedMaze-2Dl-v0": _om(agent_room=(2, 1), key_in_box=False, blocked=False, num_quarters=1, num_rooms_visited=4),
    "MiniGrid-ObstructedMaze-2Dlh-v0": _om(agent_room=(2, 1), key_in_box=True, blocked=False, num_quarters=1, num_rooms_visited=4),
    "MiniGrid-ObstructedMaze-2Dlhb-v0": _om(agent_room=(2, 1), num_quarters=1, num_rooms_visited=4),
    "MiniGrid-ObstructedMaze-1Q-v0": _om(agent_room=(1, 1), num_quarters=1, num_rooms_visited=5),
    "MiniGrid-ObstructedMaze-2Q-v0": _om(agent_room=(2, 1), num_quarters=2, num_rooms_visited=11),
    "MiniGrid-ObstructedMaze-Full-v0": _om(),
    "MiniGrid-ObstructedMaze-2Dlhb-v1": _om(agent_room=(2, 1), num_quarters=1, num_rooms_visited=4, v1=True),
    "MiniGrid-ObstructedMaze-1Q-v1": _om(agent_room=(1, 1), num_quarters=1, num_rooms_visited=5, v1=True),
    "MiniGrid-ObstructedMaze-2Q-v1": _om(agent_room=(2, 1), num_quarters=2, num_rooms_visited=11, v1=True),
    "MiniGrid-ObstructedMaze-Full-v1": _om(v1=True),
    "MiniGrid-LockedRoom-v0": (LockedRoomGen, {}),
    "MiniGrid-MultiRoom-N2-S4-v0": (MultiRoomGen, {"minNumRooms": 2, "maxNumRooms": 2, "maxRoomSize": 4}),
    "MiniGrid-MultiRoom-N4-S5-v0": (MultiRoomGen, {"minNumRooms": 6, "maxNumRooms": 6, "maxRoomSize": 5}),
    "MiniGrid-MultiRoom-N6-v0": (MultiRoomGen, {"minNumRooms": 6, "maxNumRooms": 6}),
    "MiniGrid-Playground-v0": (PlaygroundGen, {}),
    "MiniGrid-MemoryS17Random-v0": (MemoryGen, {"size": 17, "random_length": True}),
    "MiniGrid-MemoryS13Random-v0": (MemoryGen, {"size": 13, "random_length": True}),
    "MiniGrid-MemoryS13-v0": (MemoryGen, {"size": 13}),
    "MiniGrid-MemoryS11-v0": (MemoryGen, {"size": 11}),
    "MiniGrid-MemoryS9-v0": (MemoryGen, {"size": 9}),
    "MiniGrid-MemoryS7-v0": (MemoryGen, {"size": 7}),
    "MiniGrid-PutNear-6x6-N2-v0": (PutNearGen, {}),
    "MiniGrid-PutNear-8x8-N3-v0": (PutNearGen, {"size": 8, "numObjs": 3}),
    "MiniGrid-Dynamic-Obstacles-5x5-v0": (DynamicObstaclesGen, {"size": 5, "n_obstacles": 2}),
    "MiniGrid-Dynamic-Obstacles-Random-5x5-v0": (DynamicObstaclesGen, {"size": 5, "agent_start_pos": None, "n_obstacles": 2}),
    "MiniGrid-Dynamic-Obstacles-6x6-v0": (DynamicObstaclesGen, {"size": 6, "n_obstacles": 3}),
    "MiniGrid-Dynamic-Obstacles-Random-6x6-v0": (DynamicObstaclesGen, {"size": 6, "agent_start_pos": None, "n_obstacles": 3}),
    "MiniGrid-Dynamic-Obstacles-8x8-v0": (DynamicObstaclesGen, {}),
    "MiniGrid-Dynamic-Obstacles-16x16-v0": (DynamicObstaclesGen, {"size": 16, "n_obstacles": 8}),
})


def make_generator(env_id, **overrides):
    """`gym.make(env_id, **overrides)` for the layout source only."""
    if env_id not in REGISTRY:
        raise KeyError(f"{env_id}: no built-in generator (supported: {sorted(REGISTRY)}); "
                       "pass layouts extracted from your own env instead")
    cls, kw = REGISTRY[env_id]
    kw = dict(kw)
    kw.update(overrides)
    return cls(**kw)


def generate(env_id, seed, **overrides) -> Layout:
    """Layout of `gym.make(env_id, **overrides).reset(seed=seed)`."""
    return make_generator(env_id, **overrides).generate(env_id, int(seed))


def generate_many(env_id, seeds, **overrides):
    gen = make_generator(env_id, **overrides)
    return [gen.generate(env_id, int(s)) for s in seeds]
