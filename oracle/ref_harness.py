"""Import the UNMODIFIED reference (/root/reference, Farama Minigrid 2.3.1) in the build container.

TEST INFRASTRUCTURE ONLY.  Used by oracle/gen_golden.py to record golden vectors and by the
container-only tests (skipped when /root/reference is absent, e.g. on the GPU box).
gymnasium / pygame are not installed here, so the stubs under oracle/refstub/ stand in for them
(only seeding affects results: Generator(PCG64(SeedSequence(seed))), pinned by the doctest
minigrid/wrappers.py:28-43 which gen_golden.py re-checks).
"""
from __future__ import annotations

import os
import sys

import numpy as np

REFERENCE_ROOT = "/root/reference"
_STUB = os.path.join(os.path.dirname(os.path.abspath(__file__)), "refstub")


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "minigrid"))


_gym = None


def gym():
    """Returns the (stub) gymnasium module with the reference's envs registered."""
    global _gym
    if _gym is None:
        if not available():
            raise RuntimeError("reference not present")
        for p in (REFERENCE_ROOT, _STUB):
            if p not in sys.path:
                sys.path.insert(0, p)
        import gymnasium
        import minigrid
        minigrid.register_minigrid_envs()          # plugin entry point is not installed
        _gym = gymnasium
    return _gym


def make(env_id, **kw):
    return gym().make(env_id, **kw)


def encode_obj(obj):
    """(type, color, state, aux) bytes of a reference WorldObj (or None) in the oracle/engine convention."""
    if obj is None:
        return (1, 0, 0, 0)
    t, c, s = obj.encode()
    aux = 0
    if obj.type == "box" and obj.contains is not None:
        ct, cc, cs = obj.contains.encode()
        assert cs == 0 and ct < 16 and obj.contains.type != "box", "nested / stateful box contents unsupported"
        aux = ct | (cc << 4)
    return (t, c, s, aux)


def extract_state(env):
    """Flatten a reference env into (cells uint8[4, H*W], st int32[8], cfg int32[6], mission str)."""
    env = env.unwrapped
    W, H = env.width, env.height
    cells = np.zeros((4, H * W), dtype=np.uint8)
    name = type(env).__name__
    # identity-tested mission objects: `self.obj` (keycorridor.py:133, unlockpickup.py:102, blockedunlockpickup.py:112),
    # `self.door` (unlock.py:93); position targets: `self.target_pos` (gotodoor.py:125, gotoobject.py:139 — valid at reset time
    # only, the flag marks the cell of the initial layout); both doors of RedBlueDoors; Fetch: every object of the target's
    # type and colour (fetch.py:166-169 compares type and colour)
    # ObstructedMaze*: `self.obj` (obstructedmaze.py:129); Memory: success_pos -> 0x80, failure_pos -> 0x81 (memory.py:144-149,
    # empty cells that never change); PutNear: the object to move (type and colour, putnear.py:186-192) -> 0x80, the cell
    # `target_pos` names in the initial layout (:167) -> 0x81
    targets, target_pos, fetch, target_b, flags = [], None, None, None, {}
    if name in ("KeyCorridorEnv", "UnlockPickupEnv", "BlockedUnlockPickupEnv") or name.startswith("ObstructedMaze"):
        targets = [env.obj]
    elif name == "MemoryEnv":
        flags = {tuple(env.success_pos): 0x80, tuple(env.failure_pos): 0x81}
    elif name == "PutNearEnv":
        fetch = (env.move_type, env.moveColor)
        target_b = tuple(int(v) for v in env.target_pos)
    elif name == "UnlockEnv":
        targets = [env.door]
    elif name == "RedBlueDoorEnv":
        targets = [env.red_door, env.blue_door]
    elif name in ("GoToDoorEnv", "GoToObjectEnv"):
        target_pos = tuple(int(v) for v in env.target_pos)
    elif name == "FetchEnv":
        fetch = (env.targetType, env.targetColor)
    target = None
    for j in range(H):
        for i in range(W):
            o = env.grid.get(i, j)
            t, c, s, a = encode_obj(o)
            if o is not None and (any(o is x for x in targets) or (fetch and (o.type, o.color) == fetch)):
                a |= 0x80
            if target_pos == (i, j) and env.step_count == 0:
                a |= 0x80
            if target_b == (i, j) and env.step_count == 0:
                a = 0x81
            a |= flags.get((i, j), 0)
            cells[:, j * W + i] = (t, c, s, a)
    st = np.zeros(8, dtype=np.int32)
    st[0], st[1], st[2] = int(env.agent_pos[0]), int(env.agent_pos[1]), int(env.agent_dir)
    st[3] = env.step_count
    if env.carrying is not None:
        t, c, s, a = encode_obj(env.carrying)
        if any(env.carrying is x for x in targets) or (fetch and (env.carrying.type, env.carrying.color) == fetch):
            a |= 0x80
        st[4:8] = (t, c, s, a)
    rule = {"KeyCorridorEnv": 1, "UnlockPickupEnv": 1, "BlockedUnlockPickupEnv": 1, "UnlockEnv": 2, "FetchEnv": 3,
            "GoToDoorEnv": 4, "GoToObjectEnv": 4, "RedBlueDoorEnv": 5, "MemoryEnv": 6, "PutNearEnv": 7, "DynamicObstaclesEnv": 8}.get(name, 0)
    if name.startswith("ObstructedMaze"):
        rule = 1
    cfg = np.array([W, H, env.max_steps, int(env.see_through_walls), rule, env.agent_view_size], dtype=np.int32)
    return cells, st, cfg, env.mission


def extract_dyn(env):
    """64-byte RNG + obstacle record of a reference DynamicObstaclesEnv (levelgen.dyn_record's layout)."""
    from minigrid_dynamicprogramming_b200.levelgen import dyn_record
    env = env.unwrapped
    return dyn_record(env.np_random, [tuple(int(v) for v in o.cur_pos) for o in env.obstacles])
