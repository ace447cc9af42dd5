"""Record MiniGridEnv.hash / pprint_grid / agent_sees / Grid.decode goldens from the unmodified reference
(minigrid_env.py:159-229, 500-518; core/grid.py:270-289).  TEST INFRASTRUCTURE ONLY; build container only.

    python oracle/gen_golden_debug.py   ->  tests/golden/debug.json
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness as R  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-DoorKey-16x16-v0", "MiniGrid-LavaCrossingS11N5-v0",
        "MiniGrid-FourRooms-v0", "MiniGrid-KeyCorridorS6R3-v0", "MiniGrid-KeyCorridorS3R3-v0"]


def snapshot(env):
    u = env.unwrapped
    cells, st, cfg, _ = R.extract_state(env)
    sees = []
    for j in range(u.height):
        for i in range(u.width):
            if u.grid.get(i, j) is not None:
                sees.append([i, j, bool(u.agent_sees(i, j))])
    return {"cells": cells.tolist(), "st": st.tolist(), "cfg": cfg.tolist(), "hash": u.hash(), "hash8": u.hash(8),
            "pos_repr": str(u.agent_pos), "pprint": u.pprint_grid(), "sees": sees}


def main():
    out = {}
    rng = np.random.default_rng(1)
    for eid in ENVS:
        env = R.make(eid)
        env.reset(seed=0)
        snaps = [snapshot(env)]
        for k in range(3):
            for _ in range(12):
                env.step(int(rng.choice([0, 1, 2, 2, 2, 3, 5])))
            snaps.append(snapshot(env))
        out[eid] = snaps
        print(eid, [s["hash"] for s in snaps], flush=True)
    with open(os.path.join(GOLD, "debug.json"), "w") as f:
        json.dump(out, f)


if __name__ == "__main__":
    main()
