"""Record golden vectors from the UNMODIFIED reference (run in the build container only).

    python oracle/gen_golden.py            # writes tests/golden/*.npz + tests/golden/MANIFEST.json

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so everything the parity
tests need from the reference is recorded here as small fixtures:
  known_answers   the reference's own known-answer tests for this path, re-checked before recording
  layouts.npz     reset(seed) layouts for the config envs (pins levelgen.py)
  traj_<env>.npz  2,000-step random-action rollouts (SURVEY §8.3 protocol + digests) and greedy
                  goal-reaching rollouts (non-zero rewards), all stepped IN THE REFERENCE env
  mdp_<env>.npz   successor tables enumerated by FIFO BFS through the reference's own step()
"""
from __future__ import annotations

import copy
import hashlib
import json
import os
import struct
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import oracle as O          # noqa: E402
from oracle import ref_harness as R     # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
CONFIG_ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-DoorKey-16x16-v0",
               "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0", "MiniGrid-KeyCorridorS6R3-v0"]
# survey-time digests of the 2,000-step protocol, first 16 hex of sha256
SURVEY_DIGESTS = {"MiniGrid-Empty-8x8-v0": "6032d68c603be09b", "MiniGrid-DoorKey-8x8-v0": "6a7bf1df2078962e",
                  "MiniGrid-DoorKey-16x16-v0": "0570268516882fe6", "MiniGrid-LavaCrossingS11N5-v0": "c357a0e425777f4c",
                  "MiniGrid-FourRooms-v0": "fa096ba23d5cdaa8", "MiniGrid-KeyCorridorS6R3-v0": "7b8e82112d7f2d7b"}


def short(eid):
    return eid.replace("MiniGrid-", "").replace("-v0", "")


def check_known_answers():
    """The reference's own known-answer tests for this path."""
    gym = R.gym()
    out = {}
    # PCG64 stream doctest, minigrid/wrappers.py:28-43
    for seed, want in ((123, [0, 6, 5, 0, 9, 2, 2, 1, 3, 1]), (0, [8, 6, 5, 2, 3, 0, 0, 0, 1, 8]), (1, [4, 5, 7, 9, 0, 1, 8, 9, 2, 3])):
        env = gym.make("MiniGrid-Empty-5x5-v0")
        env.reset(seed=seed)
        got = [int(env.np_random.integers(10)) for _ in range(10)]
        assert got == want, (seed, got)
    out["pcg64_doctest"] = "ok"
    # Empty-5x5 first obs column, minigrid/wrappers.py:230-237
    env = gym.make("MiniGrid-Empty-5x5-v0")
    obs, _ = env.reset()
    assert (obs["image"][0] == np.array([[2, 5, 0]] * 7)).all()
    out["empty5x5_column"] = "ok"
    # lava death, tests/test_wrappers.py:362-376
    env = gym.make("MiniGrid-LavaCrossingS9N1-v0")
    env.reset(seed=2)
    _, _, term, *_ = env.step(1)
    _, rew, term, *_ = env.step(2)
    assert term is True and rew == 0
    out["lava_death_seed2"] = "ok"
    # exact truncation step, tests/test_envs.py:146-166
    env = gym.make("MiniGrid-Empty-8x8-v0", max_steps=50)
    env.reset()
    step_count = 0
    while True:
        _, _, terminated, truncated, _ = env.step(4)
        step_count += 1
        if truncated:
            assert step_count == 50
            step_count = 0
            break
    out["truncation_at_max_steps"] = "ok"
    return out


def record_layouts():
    from minigrid_dynamicprogramming_b200 import levelgen as L
    data, meta = {}, {}
    extra = ["MiniGrid-KeyCorridorS3R3-v0", "MiniGrid-Empty-Random-6x6-v0", "MiniGrid-SimpleCrossingS9N2-v0",
             "MiniGrid-DoorKey-5x5-v0", "MiniGrid-LavaCrossingS9N1-v0"]
    for eid in CONFIG_ENVS + extra:
        env = R.make(eid)
        n = 64
        cells_all, st_all, missions = [], [], []
        for seed in range(n):
            env.reset(seed=seed)
            cells, st, cfg, mission = R.extract_state(env)
            lay = L.generate(eid, seed)
            assert (lay.cells == cells).all() and (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(st[:3])
            cells_all.append(cells), st_all.append(st[:3]), missions.append(mission)
        data[short(eid) + "/cells"] = np.stack(cells_all)
        data[short(eid) + "/agent"] = np.stack(st_all).astype(np.int16)
        data[short(eid) + "/cfg"] = cfg
        meta[eid] = {"missions": missions, "hash_seed0": None}
        env.reset(seed=0)
        meta[eid]["hash_seed0"] = env.hash()
    np.savez_compressed(os.path.join(GOLD, "layouts.npz"), **data)
    return meta


def _digest_update(h, obs, rew=None, term=None, trunc=None):
    h.update(obs["image"].tobytes() + bytes([int(obs["direction"])]) + obs["mission"].encode())
    if rew is not None:
        h.update(struct.pack("<d??", float(rew), bool(term), bool(trunc)))


def record_random_rollout(eid, T=2000):
    """SURVEY §8.3 protocol: reset(seed=0); default_rng(0).integers(0,7) x T; on done reset(seed=k), k=1,2,…"""
    env = R.make(eid)
    rng = np.random.default_rng(0)
    obs, _ = env.reset(seed=0)
    h = hashlib.sha256()
    _digest_update(h, obs)
    actions = np.zeros(T, np.uint8)
    image = np.zeros((T, 7, 7, 3), np.uint8)
    direction = np.zeros(T, np.uint8)
    reward = np.zeros(T, np.float64)
    term = np.zeros(T, np.uint8)
    trunc = np.zeros(T, np.uint8)
    reset_seed = np.full(T, -1, np.int32)       # seed the env was reset with AFTER step t (or -1)
    first_image = obs["image"].copy()
    k = 0
    for t in range(T):
        a = int(rng.integers(0, 7))
        obs, r, te, tr, _ = env.step(a)
        _digest_update(h, obs, r, te, tr)
        actions[t], image[t], direction[t], reward[t], term[t], trunc[t] = a, obs["image"], obs["direction"], r, te, tr
        if te or tr:
            k += 1
            obs, _ = env.reset(seed=k)
            _digest_update(h, obs)
            reset_seed[t] = k
    digest = h.hexdigest()[:16]
    assert digest == SURVEY_DIGESTS[eid], (eid, digest)
    return dict(actions=actions, image=image, direction=direction, reward=reward, terminated=term,
                truncated=trunc, reset_seed=reset_seed, first_image=first_image), digest


def record_greedy_rollouts(eid, acts, seeds, max_states=1 << 24):
    """Greedy rollouts of the oracle's VI policy, stepped in the REFERENCE env: the non-zero-reward goldens."""
    from minigrid_dynamicprogramming_b200 import levelgen as L
    out = {}
    for seed in seeds:
        lay = L.generate(eid, seed)
        cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
        st = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
        succ, info = O.enumerate_mdp(lay.cells, cfg, st, acts, max_states)
        V, pi, sweeps, hist = O.value_iteration(succ)
        assert V[0] > 0, (eid, seed, "unsolvable")
        env = R.make(eid)
        obs, _ = env.reset(seed=seed)
        s, A, I, D, Rw, Te, Tr = 0, [], [obs["image"].copy()], [obs["direction"]], [], [], []
        while True:
            ai = int(pi[s])
            a = int(acts[ai])
            obs, r, te, tr, _ = env.step(a)
            A.append(a), I.append(obs["image"].copy()), D.append(obs["direction"]), Rw.append(r), Te.append(te), Tr.append(tr)
            s = int(succ[s, ai])
            if s < 0:
                break
        d = len(A)
        assert Te[-1] and Rw[-1] == 1 - 0.9 * (d / lay.max_steps), (eid, seed)
        # V(s0) is the repeated fp32 product gamma^(d-1)  
        v = np.float32(1.0)
        for _ in range(d - 1):
            v = np.float32(v * np.float32(0.99))
        assert V[0] == v, (eid, seed, V[0], v)
        p = f"s{seed}/"
        out[p + "actions"] = np.array(A, np.uint8)
        out[p + "image"] = np.stack(I)               # [d+1]: first obs + one per step
        out[p + "direction"] = np.array(D, np.uint8)
        out[p + "reward"] = np.array(Rw, np.float64)
        out[p + "terminated"] = np.array(Te, np.uint8)
        out[p + "truncated"] = np.array(Tr, np.uint8)
        out[p + "V0_sweeps_S"] = np.array([float(V[0]), sweeps, len(succ)], np.float64)
    return out


def _ref_key(env):
    cells, st, cfg, _ = R.extract_state(env)
    return cells.tobytes() + st[[0, 1, 2, 4, 5, 6, 7]].astype(np.uint8).tobytes()


def record_reference_mdp(eid, acts, seed=0):
    """FIFO BFS through the reference's own step(); ids in discovery order; sentinels -1 win / -2 lose."""
    env = R.make(eid, max_steps=10 ** 9)
    env.reset(seed=seed)
    target = getattr(env, "obj", None)
    snaps = [copy.deepcopy((env.grid, env.agent_pos, env.agent_dir, env.carrying, target))]
    ids = {_ref_key(env): 0}
    succ = []
    s = 0
    while s < len(snaps):
        row = []
        for a in acts:
            grid, pos, d, carrying, tgt = copy.deepcopy(snaps[s])
            env.grid, env.agent_pos, env.agent_dir, env.carrying, env.step_count = grid, pos, d, carrying, 0
            if target is not None:
                env.obj = tgt
            _, r, te, tr, _ = env.step(a)
            if te:
                row.append(-1 if r > 0 else -2)
                continue
            k = _ref_key(env)
            if k not in ids:
                ids[k] = len(snaps)
                snaps.append(copy.deepcopy((env.grid, env.agent_pos, env.agent_dir, env.carrying, getattr(env, "obj", None))))
            row.append(ids[k])
        succ.append(row)
        s += 1
    return np.array(succ, np.int32)


def main():
    os.makedirs(GOLD, exist_ok=True)
    manifest = {"reference": "Farama Minigrid 2.3.1 snapshot at /root/reference", "numpy": np.__version__}
    manifest["known_answers"] = check_known_answers()
    manifest["layouts"] = record_layouts()
    manifest["digests"] = {}
    greedy_acts = {"MiniGrid-Empty-8x8-v0": (0, 1, 2), "MiniGrid-DoorKey-8x8-v0": (0, 1, 2, 3, 5),
                   "MiniGrid-DoorKey-16x16-v0": (0, 1, 2, 3, 5), "MiniGrid-LavaCrossingS11N5-v0": (0, 1, 2),
                   "MiniGrid-FourRooms-v0": (0, 1, 2)}
    for eid in CONFIG_ENVS:
        data, digest = record_random_rollout(eid)
        manifest["digests"][eid] = digest
        np.savez_compressed(os.path.join(GOLD, f"traj_{short(eid)}.npz"), **data)
        print("random rollout", eid, digest, flush=True)
    for eid, acts in greedy_acts.items():
        np.savez_compressed(os.path.join(GOLD, f"greedy_{short(eid)}.npz"), **record_greedy_rollouts(eid, acts, range(8)))
        print("greedy", eid, flush=True)
    # KeyCorridor needs all 7 actions (the key must be dropped before the ball can be picked up)
    np.savez_compressed(os.path.join(GOLD, "greedy_KeyCorridorS3R3.npz"),
                        **record_greedy_rollouts("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7)), range(4)))
    if "--big" in sys.argv:
        np.savez_compressed(os.path.join(GOLD, "greedy_KeyCorridorS6R3.npz"),
                            **record_greedy_rollouts("MiniGrid-KeyCorridorS6R3-v0", tuple(range(7)), range(1)))
    mdps = {}
    for eid, acts in (("MiniGrid-Empty-8x8-v0", (0, 1, 2)), ("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5)),
                      ("MiniGrid-DoorKey-8x8-v0", tuple(range(7))), ("MiniGrid-DoorKey-16x16-v0", (0, 1, 2, 3, 5)),
                      ("MiniGrid-KeyCorridorS3R1-v0", tuple(range(7)))):
        succ = record_reference_mdp(eid, acts)
        mdps[f"{short(eid)}/a{len(acts)}/succ"] = succ
        mdps[f"{short(eid)}/a{len(acts)}/acts"] = np.array(acts, np.int32)
        print("reference BFS", eid, len(acts), succ.shape, flush=True)
    np.savez_compressed(os.path.join(GOLD, "mdp_reference.npz"), **mdps)
    with open(os.path.join(GOLD, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)


if __name__ == "__main__":
    main()
