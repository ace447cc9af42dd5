"""Round-2 goldens at scale (checks (i) and (ii)), recorded from the UNMODIFIED reference in the build
container.  TEST INFRASTRUCTURE ONLY.

  mdp_KeyCorridorS6R3_rows.npz   100,000 uniformly sampled rows of the all-7-action KeyCorridorS6R3 (seed 0) successor table
                                 (16.8 M states): every sampled state is rebuilt inside the reference env, each of the 7
                                 actions is applied with the reference's own step() (minigrid_env.py:520-590 +
                                 envs/keycorridor.py:129-137), and the successor — a full grid + agent record — is mapped to
                                 its serial-BFS id.  ids int32[n], succ int32[n, 7] (-1 success, -2 failure), info uint8[n, 8].
                                 The oracle's own table must agree on every row (asserted here: that pins mg_oracle_dp.c at
                                 this size) and the GPU enumeration is compared with the file in tests/test_gpu_dp.py.
  greedy_KeyCorridorS3R3.npz     greedy goal-reaching rollouts stepped in the reference, seeds 0..15
  greedy_KeyCorridorS6R3.npz     the same for seeds 0..3 (each one a 16.8 M-state oracle solve)
  nodeath_dynobs.npz             NoDeath(("ball",)) on Dynamic-Obstacles-5x5 / 6x6 (wrappers.py:842-865 samples the front cell
                                 BEFORE the obstacles move; the reference's own test_no_death_wrapper covers this pairing)

    python oracle/gen_golden_big.py [rows] [greedy] [nodeath]          (default: all three)
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import oracle as O  # noqa: E402
from oracle import ref_harness as R  # noqa: E402

GOLD = os.path.join(os.path.dirname(HERE), "tests", "golden")
N_ROWS = 100_000


def _make_obj(rec, target_obj):
    """A reference WorldObj for the oracle's (type, color, state, aux) bytes; the mission's target keeps its identity."""
    from minigrid.core.constants import IDX_TO_COLOR
    from minigrid.core.world_object import Ball, Box, Door, Key
    t, c, s, a = (int(v) for v in rec)
    if t == 1:
        return None
    if (a & 0x80) and target_obj is not None:
        return target_obj
    col = IDX_TO_COLOR[c]
    if t == 4:
        return Door(col, is_open=(s == 0), is_locked=(s == 2))
    if t == 5:
        return Key(col)
    if t == 6:
        return Ball(col)
    if t == 7:
        inner = None if not (a & 0x0f) else _make_obj((a & 0x0f, (a >> 4) & 7, 0, 0), None)
        return Box(col, contains=inner)
    raise ValueError(f"cell type {t} cannot differ from the base layout")


def apply_state(env, shadow, cells, st, target_obj):
    """Turn the reference env into the state (cells, st): only cells that differ from `shadow` (what the env holds) are set."""
    W = env.width
    for k in np.flatnonzero((shadow != cells).any(axis=0)):
        env.grid.set(int(k % W), int(k // W), _make_obj(cells[:, k], target_obj))
    env.agent_pos = (int(st[0]), int(st[1]))
    env.agent_dir = int(st[2])
    env.step_count = 0
    env.carrying = None if st[4] == 0 else _make_obj(st[4:8], target_obj)


def record_rows():
    from minigrid_dynamicprogramming_b200 import levelgen as L
    eid = "MiniGrid-KeyCorridorS6R3-v0"
    acts = tuple(range(7))
    lay = L.generate(eid, 0)
    cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
    st0 = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
    t0 = time.time()
    sp = O.StateSpace(lay.cells, cfg, st0, acts, 1 << 25)
    print(f"oracle BFS: S = {sp.S} in {time.time() - t0:.0f} s", flush=True)
    ids = np.sort(np.random.default_rng(20261105).choice(sp.S, N_ROWS, replace=False)).astype(np.int32)
    env = R.make(eid, max_steps=10 ** 9).unwrapped
    env.reset(seed=0)
    base, st_ref, _, _ = R.extract_state(env)
    assert (base == lay.cells).all() and tuple(st_ref[:3]) == tuple(st0[:3])
    target = env.obj
    shadow = base.copy()
    succ = np.zeros((N_ROWS, 7), np.int32)
    info = np.zeros((N_ROWS, 8), np.uint8)
    t0 = time.time()
    for i, s in enumerate(ids):
        cells, st = sp.state(int(s))
        info[i] = (st[0], st[1], st[2], st[4], st[5], st[6], st[7], int((cells != base).any(axis=0).sum()))
        for a in acts:
            apply_state(env, shadow, cells, st, target)
            _, r, te, tr, _ = env.step(a)
            shadow, st2, _, _ = R.extract_state(env)
            assert not tr
            if te:
                succ[i, a] = -1 if r > 0 else -2
            else:
                st2[3] = 0
                succ[i, a] = sp.lookup(shadow, st2)
                assert succ[i, a] >= 0, (s, a, "the reference reached a state the oracle BFS never found")
        if i % 10000 == 0:
            print(f"  row {i} / {N_ROWS}  ({time.time() - t0:.0f} s)", flush=True)
    same = (succ == sp.succ[ids]).all()
    print("oracle table == reference step() on all sampled rows:", bool(same), flush=True)
    assert same
    np.savez_compressed(os.path.join(GOLD, "mdp_KeyCorridorS6R3_rows.npz"), ids=ids, succ=succ, info=info, S=np.array(sp.S))
    sp.close()


def record_greedy():
    from oracle.gen_golden import record_greedy_rollouts
    np.savez_compressed(os.path.join(GOLD, "greedy_KeyCorridorS3R3.npz"),
                        **record_greedy_rollouts("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7)), range(16)))
    print("greedy S3R3 seeds 0..15", flush=True)
    out = {}
    for seed in range(4):
        t0 = time.time()
        out.update(record_greedy_rollouts("MiniGrid-KeyCorridorS6R3-v0", tuple(range(7)), [seed], max_states=1 << 25))
        print(f"greedy S6R3 seed {seed}: V0, sweeps, S = {out[f's{seed}/V0_sweeps_S']} ({time.time() - t0:.0f} s)", flush=True)
    np.savez_compressed(os.path.join(GOLD, "greedy_KeyCorridorS6R3.npz"), **out)


def record_nodeath_dynobs():
    R.gym()
    from minigrid import wrappers as W
    out = {}
    for eid, cost in (("MiniGrid-Dynamic-Obstacles-5x5-v0", -0.25), ("MiniGrid-Dynamic-Obstacles-6x6-v0", -1.5)):
        env = W.NoDeath(R.make(eid), no_death_types=("ball",), death_cost=cost)
        rng = np.random.default_rng(7)
        obs, _ = env.reset(seed=0)
        A, I, Rw, Te, Tr, Rs, Dy = [], [obs["image"].copy()], [], [], [], [], [R.extract_dyn(env)]
        k = 0
        for t in range(400):
            a = int(rng.choice([0, 1, 2, 2, 2]))
            obs, r, te, tr, _ = env.step(a)
            A.append(a), I.append(obs["image"].copy()), Rw.append(float(r)), Te.append(te), Tr.append(tr)
            Dy.append(R.extract_dyn(env))               # generator state + obstacle list after the step (before any reset)
            if te or tr:
                k += 1
                env.reset(seed=k)
                Rs.append(k)
            else:
                Rs.append(-1)
        p = eid.replace("MiniGrid-", "").replace("-v0", "") + "/"
        out[p + "actions"], out[p + "image"] = np.array(A, np.uint8), np.stack(I)
        out[p + "reward"], out[p + "terminated"], out[p + "truncated"] = np.array(Rw, np.float64), np.array(Te, np.uint8), np.array(Tr, np.uint8)
        out[p + "reset_seed"], out[p + "cost"], out[p + "dyn"] = np.array(Rs, np.int32), np.array(cost), np.stack(Dy)
        cancelled = int(((np.array(Rw) == -1 + cost) & ~np.array(Te)).sum())
        print("nodeath", eid, "cancelled bumps:", cancelled, "terminations:", int(np.sum(Te)), "truncations:", int(np.sum(Tr)), flush=True)
        assert cancelled > 0
    np.savez_compressed(os.path.join(GOLD, "nodeath_dynobs.npz"), **out)


if __name__ == "__main__":
    what = sys.argv[1:] or ["nodeath", "greedy", "rows"]
    if "nodeath" in what:
        record_nodeath_dynobs()
    if "greedy" in what:
        record_greedy()
    if "rows" in what:
        record_rows()
