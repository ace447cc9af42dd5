"""Record golden vectors of the reference's observation / reward wrappers (minigrid/wrappers.py), run in the
build container only.  TEST INFRASTRUCTURE ONLY.

For each env: reset(seed=0), then the first actions of the golden random rollout (until the episode ends, at most
40).  At every step the UNWRAPPED env state is stored together with what each wrapper returns for it:
  fully     FullyObsWrapper.observation         wrappers.py:384-423
  onehot    OneHotPartialObsWrapper.observation :220-287
  symbolic  SymbolicObsWrapper.observation      :728-766
  flat      FlatObsWrapper.observation          :553-623
  words     DictObservationSpaceWrapper         :426-550
  view3/5/9/11  ViewSizeWrapper.observation    :626-670
  abonus / pbonus   reward returned by ActionBonus / PositionBonus (:70-187) stepping the same actions

    python oracle/gen_golden_wrappers.py
"""
from __future__ import annotations

import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness as R  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0",
        "MiniGrid-KeyCorridorS6R3-v0"]


def short(eid):
    return eid.replace("MiniGrid-", "").replace("-v0", "")


def main():
    R.gym()
    from minigrid import wrappers as W
    out = {}
    for eid in ENVS:
        tr = np.load(os.path.join(GOLD, f"traj_{short(eid)}.npz"))
        done = np.flatnonzero(tr["terminated"] | tr["truncated"])
        n = min(40, int(done[0]) + 1 if len(done) else 40)
        acts = tr["actions"][:n]
        env = R.make(eid)
        obs, _ = env.reset(seed=0)
        fully, onehot, symbolic, flat, words = (W.FullyObsWrapper(env), W.OneHotPartialObsWrapper(env),
                                                W.SymbolicObsWrapper(env), W.FlatObsWrapper(env),
                                                W.DictObservationSpaceWrapper(env))
        views = {v: W.ViewSizeWrapper(env, agent_view_size=v) for v in (3, 5, 9, 11)}
        rec = {k: [] for k in ("cells", "st", "image", "fully", "onehot", "symbolic", "flat", "words", "view3", "view5", "view9", "view11")}
        for t in range(n + 1):
            cells, st, cfg, mission = R.extract_state(env)
            rec["cells"].append(cells), rec["st"].append(st), rec["image"].append(obs["image"].copy())
            rec["fully"].append(fully.observation(dict(obs))["image"].copy())
            rec["onehot"].append(onehot.observation(dict(obs))["image"].copy())
            rec["symbolic"].append(symbolic.observation(dict(obs))["image"].copy())
            rec["flat"].append(flat.observation(dict(obs)).copy())
            rec["words"].append(np.array(words.observation(dict(obs))["mission"], np.int64))
            for v, w in views.items():                      # ViewSizeWrapper.observation, wrappers.py:660-670
                rec[f"view{v}"].append(w.observation(dict(obs))["image"].copy())
            if t < n:
                obs, _, _, _, _ = env.step(int(acts[t]))
        p = short(eid) + "/"
        for k, v in rec.items():
            out[p + k] = np.stack(v)
        out[p + "cfg"] = cfg
        out[p + "actions"] = acts
        out[p + "mission"] = np.array(mission)
        assert out[p + "flat"].dtype == np.float32 and out[p + "symbolic"].dtype == np.int64, (out[p + "flat"].dtype, out[p + "symbolic"].dtype)
        # reward wrappers: fresh envs, same seed and actions
        for name, cls in (("abonus", W.ActionBonus), ("pbonus", W.PositionBonus)):
            e2 = cls(R.make(eid))
            e2.reset(seed=0)
            out[p + name] = np.array([e2.step(int(a))[1] for a in acts], np.float64)
        print(eid, n, {k: out[p + k].shape for k in ("fully", "onehot", "symbolic", "flat", "words")}, flush=True)
    # NoDeath (wrappers.py:799-872): lava does not kill, it costs; 300 forward-heavy steps, reset(seed=k) when done
    for eid, cost in (("MiniGrid-LavaCrossingS9N1-v0", -1.0), ("MiniGrid-DistShift1-v0", -0.3), ("MiniGrid-LavaGapS6-v0", -2.5)):
        env = W.NoDeath(R.make(eid), no_death_types=("lava",), death_cost=cost)
        rng = np.random.default_rng(3)
        obs, _ = env.reset(seed=0)
        A, I, Rw, Te, Tr, Rs = [], [obs["image"].copy()], [], [], [], []
        k = 0
        for t in range(300):
            a = int(rng.choice([0, 1, 2, 2, 2, 2]))
            obs, r, te, tr, _ = env.step(a)
            A.append(a), I.append(obs["image"].copy()), Rw.append(float(r)), Te.append(te), Tr.append(tr)
            if te or tr:
                k += 1
                env.reset(seed=k)
                Rs.append(k)
            else:
                Rs.append(-1)
        p = "nodeath/" + short(eid) + "/"
        out[p + "actions"], out[p + "image"] = np.array(A, np.uint8), np.stack(I)
        out[p + "reward"], out[p + "terminated"], out[p + "truncated"] = np.array(Rw, np.float64), np.array(Te, np.uint8), np.array(Tr, np.uint8)
        out[p + "reset_seed"], out[p + "cost"] = np.array(Rs, np.int32), np.array(cost)
        print("nodeath", eid, "costs paid:", int((np.array(Rw) < 0).sum()), "terminations:", int(np.sum(Te)), flush=True)
    np.savez_compressed(os.path.join(GOLD, "wrappers.npz"), **out)


if __name__ == "__main__":
    main()
