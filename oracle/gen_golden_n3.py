"""Goldens for the env ids added in the N3 row: Unlock, UnlockPickup, BlockedUnlockPickup (RoomGrid envs
with their own terminal rules, envs/unlock.py:89-97, unlockpickup.py:98-106, blockedunlockpickup.py:108-116), DistShift,
LavaGap — recorded from the UNMODIFIED reference in the build container.  TEST INFRASTRUCTURE ONLY.

Part "b" (n3b.npz): ObstructedMaze (13 ids, v0 and v1; obstructedmaze.py:126-134 = pickup of `self.obj`), LockedRoom,
MultiRoom, Playground (base step), Memory (memory.py:153-165), PutNear (putnear.py:176-201).

    python oracle/gen_golden_n3.py     ->  tests/golden/n3.npz
    python oracle/gen_golden_n3.py b   ->  tests/golden/n3b.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import gen_golden as G          # noqa: E402
from oracle import ref_harness as R         # noqa: E402

IDS = ["MiniGrid-Unlock-v0", "MiniGrid-UnlockPickup-v0", "MiniGrid-BlockedUnlockPickup-v0", "MiniGrid-DistShift1-v0",
       "MiniGrid-DistShift2-v0", "MiniGrid-LavaGapS5-v0", "MiniGrid-LavaGapS6-v0", "MiniGrid-LavaGapS7-v0",
       "MiniGrid-Fetch-5x5-N2-v0", "MiniGrid-Fetch-6x6-N2-v0", "MiniGrid-Fetch-8x8-N3-v0", "MiniGrid-GoToDoor-5x5-v0",
       "MiniGrid-GoToDoor-6x6-v0", "MiniGrid-GoToDoor-8x8-v0", "MiniGrid-GoToObject-6x6-N2-v0", "MiniGrid-GoToObject-8x8-N2-v0",
       "MiniGrid-RedBlueDoors-6x6-v0", "MiniGrid-RedBlueDoors-8x8-v0"]

IDS_B = ["MiniGrid-ObstructedMaze-1Dl-v0", "MiniGrid-ObstructedMaze-1Dlh-v0", "MiniGrid-ObstructedMaze-1Dlhb-v0",
         "MiniGrid-ObstructedMaze-2Dl-v0", "MiniGrid-ObstructedMaze-2Dlh-v0", "MiniGrid-ObstructedMaze-2Dlhb-v0",
         "MiniGrid-ObstructedMaze-1Q-v0", "MiniGrid-ObstructedMaze-2Q-v0", "MiniGrid-ObstructedMaze-Full-v0",
         "MiniGrid-ObstructedMaze-2Dlhb-v1", "MiniGrid-ObstructedMaze-1Q-v1", "MiniGrid-ObstructedMaze-2Q-v1",
         "MiniGrid-ObstructedMaze-Full-v1", "MiniGrid-LockedRoom-v0", "MiniGrid-MultiRoom-N2-S4-v0", "MiniGrid-MultiRoom-N4-S5-v0",
         "MiniGrid-MultiRoom-N6-v0", "MiniGrid-Playground-v0", "MiniGrid-MemoryS17Random-v0", "MiniGrid-MemoryS13Random-v0",
         "MiniGrid-MemoryS13-v0", "MiniGrid-MemoryS11-v0", "MiniGrid-MemoryS9-v0", "MiniGrid-MemoryS7-v0",
         "MiniGrid-PutNear-6x6-N2-v0", "MiniGrid-PutNear-8x8-N3-v0"]
ALL7 = tuple(range(7))
GREEDY_A = (("MiniGrid-Unlock-v0", ALL7, range(4)), ("MiniGrid-UnlockPickup-v0", ALL7, range(3)),
            ("MiniGrid-BlockedUnlockPickup-v0", ALL7, range(2)), ("MiniGrid-DistShift1-v0", (0, 1, 2), range(1)),
            ("MiniGrid-LavaGapS7-v0", (0, 1, 2), range(4)), ("MiniGrid-Fetch-8x8-N3-v0", ALL7, range(4)),
            ("MiniGrid-GoToDoor-8x8-v0", ALL7, range(4)), ("MiniGrid-GoToObject-8x8-N2-v0", ALL7, range(4)),
            ("MiniGrid-RedBlueDoors-8x8-v0", ALL7, range(3)))
# no drop in the big mazes: keeps the reachable set small (objects only leave the grid), the optimal episodes are unchanged
GREEDY_B = (("MiniGrid-ObstructedMaze-1Dl-v0", ALL7, range(3)), ("MiniGrid-ObstructedMaze-1Dlh-v0", ALL7, range(2)),
            ("MiniGrid-ObstructedMaze-1Dlhb-v0", (0, 1, 2, 3, 4, 5), range(2)), ("MiniGrid-LockedRoom-v0", (0, 1, 2, 3, 5), range(2)),
            ("MiniGrid-MultiRoom-N2-S4-v0", (0, 1, 2, 5), range(4)), ("MiniGrid-MultiRoom-N4-S5-v0", (0, 1, 2, 5), range(2)),
            ("MiniGrid-MemoryS13Random-v0", (0, 1, 2), range(4)), ("MiniGrid-MemoryS7-v0", (0, 1, 2), range(2)),
            ("MiniGrid-PutNear-6x6-N2-v0", ALL7, range(4)), ("MiniGrid-PutNear-8x8-N3-v0", ALL7, range(4)))


def record_rollout(eid, T=600):
    """reset(seed=0); T actions (forward / pickup / toggle weighted); on done reset(seed=k), k = 1, 2, ..."""
    env = R.make(eid)
    rng = np.random.default_rng(7)
    obs, _ = env.reset(seed=0)
    out = {k: [] for k in ("actions", "image", "direction", "reward", "terminated", "truncated", "reset_seed")}
    first = obs["image"].copy()
    k = 0
    for t in range(T):
        a = int(rng.choice([0, 1, 2, 2, 2, 3, 4, 5, 5, 6]))
        obs, r, te, tr, _ = env.step(a)
        rs = -1
        if te or tr:
            k += 1
            env.reset(seed=k)
            rs = k
        for key, v in zip(out, (a, obs["image"].copy(), obs["direction"], r, te, tr, rs)):
            out[key].append(v)
    d = {"actions": np.array(out["actions"], np.uint8), "image": np.stack(out["image"]),
         "direction": np.array(out["direction"], np.uint8), "reward": np.array(out["reward"], np.float64),
         "terminated": np.array(out["terminated"], np.uint8), "truncated": np.array(out["truncated"], np.uint8),
         "reset_seed": np.array(out["reset_seed"], np.int32), "first_image": first}
    return d


def main(ids=IDS, greedy=GREEDY_A, out="n3.npz"):
    from minigrid_dynamicprogramming_b200 import levelgen as L
    data = {}
    for eid in ids:
        env = R.make(eid)
        cells_all, st_all, missions = [], [], []
        for seed in range(32):
            env.reset(seed=seed)
            cells, st, cfg, mission = R.extract_state(env)
            lay = L.generate(eid, seed)
            assert (lay.cells == cells).all() and (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(st[:3]), (eid, seed)
            assert lay.mission == mission and lay.rule == cfg[4]
            cells_all.append(cells), st_all.append(st[:3]), missions.append(mission)
        p = G.short(eid) + "/"
        data[p + "cells"], data[p + "agent"], data[p + "cfg"] = np.stack(cells_all), np.stack(st_all).astype(np.int16), cfg
        data[p + "missions"] = np.array(missions)
        for k, v in record_rollout(eid).items():
            data[p + "traj/" + k] = v
        print("layouts + rollout", eid, int(data[p + "traj/terminated"].sum()), "terminations", flush=True)
    # successful episodes: the oracle's VI policy stepped in the reference (non-zero rewards through the new rules)
    for eid, acts, seeds in greedy:
        for k, v in G.record_greedy_rollouts(eid, acts, seeds, max_states=1 << 22).items():
            data[G.short(eid) + "/greedy/" + k] = v
        print("greedy", eid, flush=True)
    np.savez_compressed(os.path.join(G.GOLD, out), **data)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "b":
        main(IDS_B, GREEDY_B, "n3b.npz")
    else:
        main()
