"""Record the RGB-render golden vectors from the UNMODIFIED reference (run in the build container only).

TEST INFRASTRUCTURE ONLY.  Writes
  tests/golden/atlas.npz   every tile Grid.render_tile (core/grid.py:145-198) can produce for the objects the
                           engine represents: (type, color, state) x agent_dir {None,0..3} x highlight, at
                           tile sizes 8 (wrappers.py:346), 32 (TILE_PIXELS, constants.py:5), 5 and 12
  tests/golden/frames.npz  get_frame(agent_pov=True) and get_frame(highlight=True/False) of real episodes
                           (minigrid_env.py:647-734) together with the env state they were rendered from

    python oracle/gen_golden_render.py
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness as R  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
TILE_SIZES = (8, 32, 5, 12)
COLORS6 = ["red", "green", "blue", "purple", "yellow", "grey"]
CONFIG_ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-DoorKey-16x16-v0",
               "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0", "MiniGrid-KeyCorridorS6R3-v0"]


def short(eid):
    return eid.replace("MiniGrid-", "").replace("-v0", "")


def object_keys():
    """(type, color, state) of every object the engine can hold, with a constructor for the reference object."""
    R.gym()
    from minigrid.core import world_object as wo
    keys = [((1, 0, 0), lambda: None)]
    for ci, c in enumerate(COLORS6):
        keys.append(((2, ci, 0), lambda c=c: wo.Wall(c)))
        keys.append(((3, ci, 0), lambda c=c: wo.Floor(c)))
        keys.append(((4, ci, 0), lambda c=c: wo.Door(c, is_open=True)))
        keys.append(((4, ci, 1), lambda c=c: wo.Door(c)))
        keys.append(((4, ci, 2), lambda c=c: wo.Door(c, is_locked=True)))
        keys.append(((5, ci, 0), lambda c=c: wo.Key(c)))
        keys.append(((6, ci, 0), lambda c=c: wo.Ball(c)))
        keys.append(((7, ci, 0), lambda c=c: wo.Box(c)))
    keys.append(((8, 1, 0), lambda: wo.Goal()))
    keys.append(((9, 0, 0), lambda: wo.Lava()))
    return keys


def record_atlas():
    R.gym()
    from minigrid.core.grid import Grid
    out = {}
    for ts in TILE_SIZES:
        K, T = [], []
        for (t, c, s), ctor in object_keys():
            obj = ctor()
            if obj is not None:
                assert tuple(obj.encode()) == (t, c, s), (obj.encode(), t, c, s)
            for ad in (None, 0, 1, 2, 3):
                for hl in (False, True):
                    Grid.tile_cache.clear()
                    img = Grid.render_tile(obj, agent_dir=ad, highlight=hl, tile_size=ts)
                    # Grid.render assigns the float64 tile into a uint8 frame (grid.py:240): same cast here
                    frame = np.zeros((ts, ts, 3), np.uint8)
                    frame[:, :, :] = img
                    K.append((t, c, s, -1 if ad is None else ad, int(hl)))
                    T.append(frame)
        out[f"ts{ts}/keys"] = np.array(K, np.int16)
        out[f"ts{ts}/tiles"] = np.stack(T)
        print("atlas", ts, len(K), flush=True)
    np.savez_compressed(os.path.join(GOLD, "atlas.npz"), **out)


def _frames_for(env, tag, out, actions, every):
    """Step `actions`; store state + frames every `every` steps and a digest over all frames."""
    h = hashlib.sha256()
    kept = []
    n = len(actions)
    for t in range(n + 1):
        u = env.unwrapped
        pov = u.get_frame(tile_size=8, agent_pov=True)
        full = u.get_frame(highlight=True, tile_size=8, agent_pov=False)
        plain = u.get_frame(highlight=False, tile_size=8, agent_pov=False)
        for a in (pov, full, plain):
            assert a.dtype == np.uint8
            h.update(a.tobytes())
        if t % every == 0 or t == n:
            cells, st, cfg, _ = R.extract_state(env)
            kept.append((t, cells, st, cfg, pov, full, plain))
        if t == n:
            break
        _, _, te, tr, _ = env.step(int(actions[t]))
        if te or tr:
            n = t + 1
    out[tag + "t"] = np.array([k[0] for k in kept], np.int32)
    out[tag + "cells"] = np.stack([k[1] for k in kept])
    out[tag + "st"] = np.stack([k[2] for k in kept])
    out[tag + "cfg"] = kept[0][3]
    out[tag + "pov"] = np.stack([k[4] for k in kept])
    out[tag + "full"] = np.stack([k[5] for k in kept])
    out[tag + "plain"] = np.stack([k[6] for k in kept])
    out[tag + "digest"] = np.frombuffer(h.digest(), np.uint8)


def record_frames():
    out = {}
    for eid in CONFIG_ENVS:
        tr = np.load(os.path.join(GOLD, f"traj_{short(eid)}.npz"))
        env = R.make(eid)
        env.reset(seed=0)
        first_done = np.flatnonzero(tr["terminated"] | tr["truncated"])
        n = min(40, int(first_done[0]) + 1 if len(first_done) else 40)
        _frames_for(env, f"{short(eid)}/random/", out, tr["actions"][:n], every=5)
        print("frames", eid, "random", flush=True)
    # greedy solutions: the agent carries keys / balls, doors open, boxes (KeyCorridor) are on screen
    for name, seeds in (("DoorKey-8x8", (0, 1)), ("KeyCorridorS3R3", (0, 1)), ("DoorKey-16x16", (0,))):
        g = np.load(os.path.join(GOLD, f"greedy_{name}.npz"))
        for s in seeds:
            env = R.make(f"MiniGrid-{name}-v0")
            env.reset(seed=s)
            _frames_for(env, f"{name}/greedy{s}/", out, g[f"s{s}/actions"], every=3)
            print("frames", name, "greedy", s, flush=True)
    # one big-tile frame pair (TILE_PIXELS = 32, what env.render() uses: minigrid_env.py:46,737)
    env = R.make("MiniGrid-DoorKey-8x8-v0")
    env.reset(seed=3)
    for a in (2, 1, 2, 3):
        env.step(a)
    cells, st, cfg, _ = R.extract_state(env)
    out["ts32/cells"], out["ts32/st"], out["ts32/cfg"] = cells, st, cfg
    out["ts32/pov"] = env.unwrapped.get_frame(tile_size=32, agent_pov=True)
    out["ts32/full"] = env.unwrapped.get_frame(highlight=True, tile_size=32, agent_pov=False)
    np.savez_compressed(os.path.join(GOLD, "frames.npz"), **out)


if __name__ == "__main__":
    if "--frames-only" not in sys.argv:
        record_atlas()
    record_frames()
