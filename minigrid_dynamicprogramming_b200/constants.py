"""Encodings of the reference world model (minigrid/core/constants.py:17-58, core/actions.py:7-20)."""
from __future__ import annotations

from enum import IntEnum

COLOR_NAMES = ["blue", "green", "grey", "purple", "red", "yellow"]        # sorted, constants.py:17
COLOR_TO_IDX = {"red": 0, "green": 1, "blue": 2, "purple": 3, "yellow": 4, "grey": 5}   # constants.py:20
IDX_TO_COLOR = {v: k for k, v in COLOR_TO_IDX.items()}
OBJECT_TO_IDX = {"unseen": 0, "empty": 1, "wall": 2, "floor": 3, "door": 4, "key": 5, "ball": 6,
                 "box": 7, "goal": 8, "lava": 9, "agent": 10}             # constants.py:25-37
IDX_TO_OBJECT = {v: k for k, v in OBJECT_TO_IDX.items()}
STATE_TO_IDX = {"open": 0, "closed": 1, "locked": 2}                      # constants.py:42-46
DIR_TO_VEC = ((1, 0), (0, 1), (-1, 0), (0, -1))                           # constants.py:49-58

T_UNSEEN, T_EMPTY, T_WALL, T_FLOOR, T_DOOR, T_KEY, T_BALL, T_BOX, T_GOAL, T_LAVA, T_AGENT = range(11)

AUX_TARGET = 0x80        # "is the mission target object" (identity test envs/keycorridor.py:133)

RULE_NONE = 0
RULE_PICKUP_TARGET = 1   # KeyCorridorEnv.step, envs/keycorridor.py:129-137 (also UnlockPickup, BlockedUnlockPickup)
RULE_UNLOCK = 2          # UnlockEnv.step, envs/unlock.py:89-97: a toggle while the target door is open pays and terminates
RULE_FETCH = 3           # FetchEnv.step, envs/fetch.py:162-176: carrying anything terminates; it pays iff it matches the target
RULE_GOTO = 4            # GoToDoorEnv / GoToObjectEnv.step: toggle terminates; done terminates and pays next to target_pos
RULE_REDBLUE = 5         # RedBlueDoorsEnv.step, envs/redbluedoors.py:105-133
RULE_MEMORY = 6          # MemoryEnv.step, envs/memory.py:153-165: pickup acts as toggle; standing on success_pos / failure_pos terminates
RULE_PUTNEAR = 7         # PutNearEnv.step, envs/putnear.py:176-201: wrong pickup terminates; a drop terminates and pays near target_pos
RULE_DYNOBS = 8          # DynamicObstaclesEnv.step, envs/dynamicobstacles.py:136-167: obstacles move by in-step RNG draws; bumping costs -1
AUX_TARGET_B = 0x81      # second mission cell (failure_pos of Memory, target_pos of PutNear): target bit + "contains None" nibble


class Actions(IntEnum):   # core/actions.py:7-20
    left = 0
    right = 1
    forward = 2
    pickup = 3
    drop = 4
    toggle = 5
    done = 6
