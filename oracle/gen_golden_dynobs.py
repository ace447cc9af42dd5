"""Goldens for DynamicObstaclesEnv (minigrid/envs/dynamicobstacles.py), the one env whose `step` draws from the env's random
generator — recorded from the UNMODIFIED reference in the build container.  TEST INFRASTRUCTURE ONLY.

Per env id: layouts of 16 seeds (cells, start pose, the 64-byte generator + obstacle record after `_gen_grid`), and a 500-step
rollout (reset(seed=0); actions mostly left / right / forward, some >= 3 which the env maps to `left`, :138-139; on done
reset(seed=k), k = 1, 2, ...) with, per step, the observation, reward (-1 when bumping, :163-166), flags, the FULL grid and
the generator + obstacle record — so a restatement is pinned draw by draw, not only through what the agent sees.

    python oracle/gen_golden_dynobs.py   ->  tests/golden/dynobs.npz
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import gen_golden as G          # noqa: E402
from oracle import ref_harness as R         # noqa: E402

IDS = ["MiniGrid-Dynamic-Obstacles-5x5-v0", "MiniGrid-Dynamic-Obstacles-Random-5x5-v0", "MiniGrid-Dynamic-Obstacles-6x6-v0",
       "MiniGrid-Dynamic-Obstacles-Random-6x6-v0", "MiniGrid-Dynamic-Obstacles-8x8-v0", "MiniGrid-Dynamic-Obstacles-16x16-v0"]


def main():
    from minigrid_dynamicprogramming_b200 import levelgen as L
    data = {}
    for eid in IDS:
        env = R.make(eid)
        p = G.short(eid) + "/"
        cells_all, st_all, dyn_all = [], [], []
        for seed in range(16):
            env.reset(seed=seed)
            cells, st, cfg, mission = R.extract_state(env)
            lay = L.generate(eid, seed)
            assert (lay.cells == cells).all() and (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(st[:3]), (eid, seed)
            assert lay.mission == mission and lay.rule == cfg[4] and (lay.dyn == R.extract_dyn(env)).all()
            cells_all.append(cells), st_all.append(st[:3]), dyn_all.append(lay.dyn)
        data[p + "cells"], data[p + "agent"], data[p + "cfg"] = np.stack(cells_all), np.stack(st_all).astype(np.int16), cfg
        data[p + "dyn"], data[p + "mission"] = np.stack(dyn_all), np.array(mission)
        rng = np.random.default_rng(5)
        obs, _ = env.reset(seed=0)
        out = {k: [] for k in ("actions", "image", "direction", "reward", "terminated", "truncated", "reset_seed", "grid", "dynrec")}
        data[p + "traj/first_image"] = obs["image"].copy()
        k = 0
        for t in range(500):
            a = int(rng.choice([0, 1, 2, 2, 2, 2, 3, 5, 6]))
            obs, r, te, tr, _ = env.step(a)
            grid, dyn = R.extract_state(env)[0], R.extract_dyn(env)          # state right after the step, before any reset
            rs = -1
            if te or tr:
                k += 1
                env.reset(seed=k)
                rs = k
            for key, v in zip(out, (a, obs["image"].copy(), obs["direction"], r, te, tr, rs, grid, dyn)):
                out[key].append(v)
        for key, dt in (("actions", np.uint8), ("direction", np.uint8), ("reward", np.float64), ("terminated", np.uint8),
                        ("truncated", np.uint8), ("reset_seed", np.int32)):
            data[p + "traj/" + key] = np.array(out[key], dt)
        for key in ("image", "grid", "dynrec"):
            data[p + "traj/" + key] = np.stack(out[key])
        print(eid, "bumps:", int((data[p + "traj/reward"] < 0).sum()), "goals:", int((data[p + "traj/reward"] > 0).sum()),
              "truncations:", int(data[p + "traj/truncated"].sum()), flush=True)
    np.savez_compressed(os.path.join(G.GOLD, "dynobs.npz"), **data)


if __name__ == "__main__":
    main()
