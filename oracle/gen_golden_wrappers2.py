"""Golden vectors of the last two wrappers of minigrid/wrappers.py, recorded from the UNMODIFIED reference in the build
container.  TEST INFRASTRUCTURE ONLY.

  DirectionObsWrapper      wrappers.py:673-725   obs["goal_direction"] along an episode, type "slope" and "angle"; the
                                                 wrapper's cached goal_position tuple is stored as the reference holds it
                                                 (it is (row, column) — int(idx / height), idx % width — not (x, y))
  StochasticActionWrapper  wrappers.py:769-796   action() for 400 calls under np.random.seed(k): the uniforms come from the
                                                 process-global numpy state, the replacement actions from the env's own
                                                 generator (np_random.integers(0, 6)); both draw sequences are stored

    python oracle/gen_golden_wrappers2.py
"""
from __future__ import annotations

import os
import sys
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness as R  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
DIR_ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0", "MiniGrid-DoorKey-8x8-v0"]


def short(eid):
    return eid.replace("MiniGrid-", "").replace("-v0", "")


def main():
    R.gym()
    from minigrid import wrappers as W
    out = {}
    warnings.simplefilter("ignore", RuntimeWarning)          # np.divide by zero when the agent is in the goal's row
    for eid in DIR_ENVS:
        for typ in ("slope", "angle"):
            env = W.DirectionObsWrapper(R.make(eid), type=typ)
            # the wrapper's reset() drops the seed (wrappers.py:697): seed the env first, as a caller has to
            env.unwrapped.reset(seed=3)
            env.env.unwrapped.np_random = np.random.Generator(np.random.PCG64(np.random.SeedSequence(3)))
            obs, _ = env.reset()
            u = env.unwrapped
            cells, st, cfg, _ = R.extract_state(u)
            p = f"dir/{short(eid)}/{typ}/"
            out[p + "cells"], out[p + "agent"], out[p + "cfg"] = cells, st[:3].copy(), cfg
            out[p + "goal_position"] = np.array(env.goal_position, dtype=np.int32)
            acts = np.random.default_rng(5).choice([0, 1, 2, 2, 2], size=60).astype(np.uint8)
            vals, used = [float(obs["goal_direction"])], []
            for a in acts:
                obs, r, te, tr, _ = env.step(int(a))
                used.append(a)
                vals.append(float(obs["goal_direction"]))
                if te or tr:
                    break
            out[p + "actions"] = np.array(used, dtype=np.uint8)
            out[p + "goal_direction"] = np.array(vals, dtype=np.float64)
    for k, (prob, ra) in enumerate([(0.9, None), (0.5, None), (0.7, 3)]):
        env = W.StochasticActionWrapper(R.make("MiniGrid-Empty-8x8-v0"), prob=prob, random_action=ra)
        env.reset(seed=11 + k)
        st0 = env.unwrapped.np_random.bit_generator.state
        np.random.seed(100 + k)
        given = np.random.default_rng(k).integers(0, 7, 400).astype(np.uint8)
        got = np.array([int(env.action(int(a))) for a in given], dtype=np.uint8)
        # the draw sequences behind it, replayed from the same seeds
        np.random.seed(100 + k)
        u = np.array([np.random.uniform() for _ in range(400)], dtype=np.float64)
        g = np.random.Generator(np.random.PCG64())
        g.bit_generator.state = st0
        nrep = int((u >= prob).sum())
        rep = g.integers(0, 6, size=0)
        rep = np.array([int(g.integers(0, high=6)) for _ in range(nrep)], dtype=np.uint8) if ra is None else np.zeros(0, np.uint8)
        p = f"stoch/{k}/"
        out[p + "prob"], out[p + "random_action"] = np.float64(prob), np.int32(-1 if ra is None else ra)
        out[p + "given"], out[p + "got"], out[p + "uniform"], out[p + "replacement"] = given, got, u, rep
    np.savez_compressed(os.path.join(GOLD, "wrappers2.npz"), **out)
    print("wrote wrappers2.npz", len(out), "arrays")


if __name__ == "__main__":
    main()
