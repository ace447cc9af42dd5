"""CPU: pin the oracle (oracle/mg_oracle*.c) and the host level generators against the golden
vectors recorded from the unmodified reference (oracle/gen_golden.py)."""
import json
import os

import numpy as np
import pytest

from oracle import oracle as O
from minigrid_dynamicprogramming_b200 import levelgen as L
from helpers import CONFIG_ENVS, GOLDEN, N3_IDS, N3B_IDS, golden, n3_golden, oracle_env_from_seed, short


def test_manifest_known_answers():
    m = json.load(open(os.path.join(GOLDEN, "MANIFEST.json")))
    assert set(m["known_answers"].values()) == {"ok"}
    assert len(m["digests"]) == 6


def test_layouts_match_reference():
    z = golden("layouts.npz")
    names = sorted({k.split("/")[0] for k in z.files})
    assert len(names) >= 11
    for name in names:
        eid = f"MiniGrid-{name}-v0"
        cells, agent, cfg = z[name + "/cells"], z[name + "/agent"], z[name + "/cfg"]
        for seed in range(cells.shape[0]):
            lay = L.generate(eid, seed)
            assert (lay.cells == cells[seed]).all(), (eid, seed)
            assert (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(agent[seed]), (eid, seed)
            assert [lay.width, lay.height, lay.max_steps, int(lay.see_through_walls), lay.rule, 7] == list(cfg)


@pytest.mark.parametrize("eid", CONFIG_ENVS)
def test_oracle_random_rollout(eid):
    z = golden(f"traj_{short(eid)}.npz")
    env = oracle_env_from_seed(eid, 0)
    img, _ = env.gen_obs()
    assert (img == z["first_image"]).all()
    for t in range(len(z["actions"])):
        img, d, r, te, tr = env.step(int(z["actions"][t]))
        assert (img == z["image"][t]).all(), t
        assert (d, r, te, tr) == (z["direction"][t], z["reward"][t], bool(z["terminated"][t]), bool(z["truncated"][t])), t
        if z["reset_seed"][t] >= 0:
            assert te or tr
            env = oracle_env_from_seed(eid, int(z["reset_seed"][t]))


@pytest.mark.parametrize("name", ["Empty-8x8", "DoorKey-8x8", "DoorKey-16x16", "LavaCrossingS11N5", "FourRooms",
                                  "KeyCorridorS3R3"])
def test_oracle_greedy_rollouts(name):
    z = golden(f"greedy_{name}.npz")
    eid = f"MiniGrid-{name}-v0"
    seeds = sorted({int(k.split("/")[0][1:]) for k in z.files})
    for seed in seeds:
        p = f"s{seed}/"
        env = oracle_env_from_seed(eid, seed)
        img, d = env.gen_obs()
        assert (img == z[p + "image"][0]).all() and d == z[p + "direction"][0]
        for t, a in enumerate(z[p + "actions"]):
            img, d, r, te, tr = env.step(int(a))
            assert (img == z[p + "image"][t + 1]).all()
            assert (d, r, te, tr) == (z[p + "direction"][t + 1], z[p + "reward"][t], bool(z[p + "terminated"][t]),
                                      bool(z[p + "truncated"][t]))
        assert te and r > 0


def test_oracle_mdp_matches_reference_bfs():
    z = golden("mdp_reference.npz")
    for key in [k for k in z.files if k.endswith("/succ")]:
        name, a, _ = key.split("/")
        acts = tuple(int(x) for x in z[f"{name}/{a}/acts"])
        lay = L.generate(f"MiniGrid-{name}-v0", 0)
        cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
        st = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
        succ, info = O.enumerate_mdp(lay.cells, cfg, st, acts, 1 << 16)
        assert succ.shape == z[key].shape and (succ == z[key]).all(), key


def test_oracle_vi_anchors():
    # sweeps to stop, exact V(s0) as repeated fp32 products
    for name, acts, S, sweeps, d in (("Empty-8x8", (0, 1, 2), 140, 13, 11), ("DoorKey-8x8", (0, 1, 2, 3, 5), 424, 24, 17),
                                     ("DoorKey-16x16", (0, 1, 2, 3, 5), 2680, 44, 29)):
        succ = golden("mdp_reference.npz")[f"{name}/a{len(acts)}/succ"]
        V, pi, k, hist = O.value_iteration(succ)
        v = np.float32(1.0)
        for _ in range(d - 1):
            v = np.float32(v * np.float32(0.99))
        assert (len(succ), k, hist[-1]) == (S, sweeps, 0.0) and V[0] == v
        V2, pi2, k2 = O.value_iteration_numpy(succ)
        assert k2 == k and (V2 == V).all() and (pi2 == pi).all()
        # policy evaluation of the greedy policy reproduces V
        V3, _, _, _ = O.value_iteration(succ, pi_fixed=pi)
        assert (V3 == V).all()


def test_reward_fp64():
    assert O.reward(11, 256) == 0.961328125
    assert O.reward(17, 640) == 1 - 0.9 * (17 / 640)


# ---------------------------------------------------------------- RGB rendering

@pytest.mark.parametrize("ts", [8, 5, 12, 32])
def test_oracle_render_tile_atlas(ts):
    """mg_oracle_render.c against every tile Grid.render_tile produced in the reference (atlas.npz)."""
    d = golden("atlas.npz")
    for k, want in zip(d[f"ts{ts}/keys"], d[f"ts{ts}/tiles"]):
        got = O.render_tile(k[0], k[1], k[2], None if k[3] < 0 else k[3], k[4], ts)
        assert np.array_equal(got, want), k


def frame_tags(f):
    return sorted({k.rsplit("/", 1)[0] for k in f.files if k.endswith("/pov") and not k.startswith("ts32")})


def test_oracle_render_frames():
    """get_frame(agent_pov=True) / get_frame(highlight=True|False) of reference episodes (frames.npz)."""
    f = golden("frames.npz")
    tags = frame_tags(f)
    assert len(tags) == 11
    for tag in tags:
        cfg = f[tag + "/cfg"]
        for i in range(len(f[tag + "/t"])):
            c, st = f[tag + "/cells"][i], f[tag + "/st"][i]
            assert np.array_equal(O.render_pov(c, cfg, st, 8), f[tag + "/pov"][i]), (tag, i)
            assert np.array_equal(O.render_full(c, cfg, st, True, 8), f[tag + "/full"][i]), (tag, i)
            assert np.array_equal(O.render_full(c, cfg, st, False, 8), f[tag + "/plain"][i]), (tag, i)
    assert np.array_equal(O.render_pov(f["ts32/cells"], f["ts32/cfg"], f["ts32/st"], 32), f["ts32/pov"])
    assert np.array_equal(O.render_full(f["ts32/cells"], f["ts32/cfg"], f["ts32/st"], True, 32), f["ts32/full"])


# ---------------------------------------------------------------- wrappers

def test_oracle_wrappers_vs_reference():
    """oracle/wrappers_np.py against what the reference's wrappers returned (wrappers.npz)."""
    from oracle import wrappers_np as WN
    z = golden("wrappers.npz")
    names = sorted({k.split("/")[0] for k in z.files} - {"nodeath"})
    assert len(names) == 5
    for name in names:
        cfg, mission = z[name + "/cfg"], str(z[name + "/mission"])
        acts = z[name + "/actions"]
        ab, pb = WN.BonusCounter("action"), WN.BonusCounter("position")
        for t in range(len(z[name + "/st"])):
            cells, st, img = z[name + "/cells"][t], z[name + "/st"][t], z[name + "/image"][t]
            assert np.array_equal(WN.fully_obs(cells, cfg, st), z[name + "/fully"][t]), (name, t)
            assert np.array_equal(WN.onehot_obs(img), z[name + "/onehot"][t]), (name, t)
            assert np.array_equal(WN.symbolic_obs(cells, cfg, st), z[name + "/symbolic"][t]), (name, t)
            got = WN.flat_obs(img, mission)
            assert got.dtype == np.float32 and np.array_equal(got, z[name + "/flat"][t]), (name, t)
            assert np.array_equal(WN.mission_words(mission), z[name + "/words"][t]), (name, t)
            if t > 0:
                # reward of the wrapped step = env reward (0 on these prefixes unless terminal) + bonus
                base = golden(f"traj_{name}.npz")["reward"][t - 1]
                assert z[name + "/abonus"][t - 1] == base + ab.bonus(st, acts[t - 1]), (name, t)
                assert z[name + "/pbonus"][t - 1] == base + pb.bonus(st, acts[t - 1]), (name, t)


# ---------------------------------------------------------------- more env rules



@pytest.mark.parametrize("eid", N3_IDS + N3B_IDS)
def test_n3_layouts_and_oracle_rollout(eid):
    """Generators of the N3 ids bit-equal to the reference on 32 seeds; the oracle replays the reference's 600-step rollout
    (with resets) and, where recorded, the successful greedy episodes that exercise the env's own terminal rule."""
    z, p = n3_golden(eid), short(eid) + "/"
    cells, agent, cfg = z[p + "cells"], z[p + "agent"], z[p + "cfg"]
    for seed in range(cells.shape[0]):
        lay = L.generate(eid, seed)
        assert (lay.cells == cells[seed]).all() and (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(agent[seed])
        assert [lay.width, lay.height, lay.max_steps, int(lay.see_through_walls), lay.rule, 7] == list(cfg)
        assert lay.mission == str(z[p + "missions"][seed])
    env = oracle_env_from_seed(eid, 0)
    assert (env.gen_obs()[0] == z[p + "traj/first_image"]).all()
    for t, a in enumerate(z[p + "traj/actions"]):
        img, d, r, te, tr = env.step(int(a))
        assert (img == z[p + "traj/image"][t]).all() and d == z[p + "traj/direction"][t], t
        assert (r, te, tr) == (z[p + "traj/reward"][t], bool(z[p + "traj/terminated"][t]), bool(z[p + "traj/truncated"][t])), t
        if z[p + "traj/reset_seed"][t] >= 0:
            assert te or tr
            env = oracle_env_from_seed(eid, int(z[p + "traj/reset_seed"][t]))
    seeds = sorted({int(k.split("/")[2][1:]) for k in z.files if k.startswith(p + "greedy/")})
    for s in seeds:
        env = oracle_env_from_seed(eid, s)
        g = p + f"greedy/s{s}/"
        assert (env.gen_obs()[0] == z[g + "image"][0]).all()
        for t, a in enumerate(z[g + "actions"]):
            img, d, r, te, tr = env.step(int(a))
            assert (img == z[g + "image"][t + 1]).all() and r == z[g + "reward"][t] and te == bool(z[g + "terminated"][t]), (s, t)
        assert te and r > 0


@pytest.mark.parametrize("eid", ["MiniGrid-LavaCrossingS9N1-v0", "MiniGrid-DistShift1-v0", "MiniGrid-LavaGapS6-v0"])
def test_oracle_no_death_vs_reference(eid):
    """NoDeath (wrappers.py:799-872) restated around the oracle's step, against the reference's wrapped episodes."""
    z, p = golden("wrappers.npz"), "nodeath/" + short(eid) + "/"
    cost = float(z[p + "cost"])
    env = oracle_env_from_seed(eid, 0)
    assert (env.gen_obs()[0] == z[p + "image"][0]).all()
    for t, a in enumerate(z[p + "actions"]):
        img, d, r, te, tr = O.step_no_death(env, int(a), (9,), cost)
        assert (img == z[p + "image"][t + 1]).all() and r == z[p + "reward"][t], t
        assert te == bool(z[p + "terminated"][t]) and tr == bool(z[p + "truncated"][t]), t
        if z[p + "reset_seed"][t] >= 0:
            env = oracle_env_from_seed(eid, int(z[p + "reset_seed"][t]))


@pytest.mark.parametrize("eid", ["MiniGrid-Dynamic-Obstacles-5x5-v0", "MiniGrid-Dynamic-Obstacles-6x6-v0"])
def test_oracle_no_death_on_dynamic_obstacles_vs_reference(eid):
    """NoDeath(("ball",)) around DynamicObstacles: the front cell is sampled before the obstacles move (wrappers.py:852-858)."""
    z, p = golden("nodeath_dynobs.npz"), short(eid) + "/"
    cost = float(z[p + "cost"])
    env = oracle_env_from_seed(eid, 0)
    assert (env.gen_obs()[0] == z[p + "image"][0]).all() and (env.dyn == z[p + "dyn"][0]).all()
    for t, a in enumerate(z[p + "actions"]):
        img, d, r, te, tr = O.step_no_death(env, int(a), (6,), cost)
        assert (img == z[p + "image"][t + 1]).all() and r == z[p + "reward"][t], t
        assert te == bool(z[p + "terminated"][t]) and tr == bool(z[p + "truncated"][t]), t
        assert (env.dyn == z[p + "dyn"][t + 1]).all(), t
        if z[p + "reset_seed"][t] >= 0:
            env = oracle_env_from_seed(eid, int(z[p + "reset_seed"][t]))


def test_oracle_autoreset_walks_the_pool_with_the_reset_stride():
    """mgo_rollout's same-step autoreset: every reset moves an env from pool slot l to (l + reset_stride) % P — checked on the
    slot ids the rollout leaves behind and on the first observation after a reset (it is the NEW layout's)."""
    eid, N, P, T = "MiniGrid-DoorKey-8x8-v0", 24, 40, 90
    pool = L.generate_many(eid, range(P), max_steps=20)
    lays = pool[:N]
    acts = np.random.default_rng(3).integers(0, 3, (T, N)).astype(np.uint8)
    from helpers import oracle_env
    _np = np
    cells = _np.stack([l.cells.reshape(-1) for l in lays]).astype(_np.uint8)
    pc = _np.stack([l.cells.reshape(-1) for l in pool]).astype(_np.uint8)
    st = _np.zeros((N, 8), _np.int32)
    pst = _np.zeros((P, 8), _np.int32)
    for i, l in enumerate(pool):
        pst[i, :3] = (*l.agent_pos, l.agent_dir)
    st[:] = pst[:N]
    lid = _np.arange(N, dtype=_np.int32)
    cfgs = O.make_cfg(8, 8, 20, False, 0, 7).reshape(1, -1)
    img, d, r, te, tr = O.rollout(cells, st, _np.zeros(N, _np.uint8), cfgs, acts, autoreset=True, pool_cells=pc, pool_st=pst,
                                  pool_class=_np.zeros(P, _np.uint8), layout_id=lid)
    s = O.reset_stride(N, P)
    assert s == 27 and (tr.sum(axis=0) >= 4).all()
    k = (te | tr).sum(axis=0)
    assert (lid == (_np.arange(N) + k * s) % P).all()
    t0 = int(_np.flatnonzero(tr[:, 5])[0])                  # env 5's first reset: the returned obs is slot (5 + s) % P's first
    assert d[t0, 5] == pool[(5 + s) % P].agent_dir
    assert (img[t0, 5] == oracle_env(pool[(5 + s) % P]).gen_obs()[0]).all()


def test_reset_stride_rule():
    """The autoreset stride (include/mgb200.h mg_reset_stride; the oracle's restatement): coprime to the pool size, N mod P when
    that already is, never 0 for P > 1 — so an env visits all P layouts before one repeats."""
    import math
    assert O.reset_stride(65536, 4096) == 1 and O.reset_stride(256, 256) == 1 and O.reset_stride(5, 1) == 0
    assert O.reset_stride(128, 192) == 131 and O.reset_stride(1, 7) == 1 and O.reset_stride(65536, 4099) == 65536 % 4099
    for n in (1, 7, 64, 100, 65536):
        for P in range(2, 70):
            s = O.reset_stride(n, P)
            assert 0 < s < P and math.gcd(s, P) == 1 and s >= min(n % P, P - 1) * (math.gcd(n % P or 1, P) == 1)
            assert len({(k * s) % P for k in range(P)}) == P


# ---------------------------------------------------------------- DynamicObstacles: RNG draws inside step (N3, last row)

from helpers import DYN_IDS  # noqa: E402


@pytest.mark.parametrize("eid", DYN_IDS)
def test_dynamic_obstacles_oracle_vs_reference(eid):
    """The oracle's PCG64 / Generator.integers / place_obj restatement against the reference draw by draw: after every step
    of the recorded rollout the full grid, the generator state and the obstacle list are identical (not only the obs)."""
    z, p = golden("dynobs.npz"), short(eid) + "/"
    for seed in range(z[p + "cells"].shape[0]):
        lay = L.generate(eid, seed)
        assert (lay.cells == z[p + "cells"][seed]).all() and (lay.dyn == z[p + "dyn"][seed]).all()
        assert (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir) == tuple(z[p + "agent"][seed])
        assert [lay.width, lay.height, lay.max_steps, int(lay.see_through_walls), lay.rule, 7] == list(z[p + "cfg"])
    env = oracle_env_from_seed(eid, 0)
    assert (env.gen_obs()[0] == z[p + "traj/first_image"]).all()
    for t, a in enumerate(z[p + "traj/actions"]):
        img, d, r, te, tr = env.step(int(a))
        assert (img == z[p + "traj/image"][t]).all() and d == z[p + "traj/direction"][t], t
        assert (r, te, tr) == (z[p + "traj/reward"][t], bool(z[p + "traj/terminated"][t]), bool(z[p + "traj/truncated"][t])), t
        assert (env.cells.reshape(4, -1) == z[p + "traj/grid"][t]).all() and (env.dyn == z[p + "traj/dynrec"][t]).all(), t
        if z[p + "traj/reset_seed"][t] >= 0:
            env = oracle_env_from_seed(eid, int(z[p + "traj/reset_seed"][t]))
    assert (z[p + "traj/reward"] == -1).sum() > 20


def test_direction_and_stochastic_wrappers_oracle_vs_reference():
    """oracle/wrappers_np.py goal_position / goal_direction / stochastic_action against the reference's DirectionObsWrapper
    and StochasticActionWrapper outputs (wrappers2.npz, oracle/gen_golden_wrappers2.py)."""
    from oracle import wrappers_np as WN
    z = golden("wrappers2.npz")
    for name in sorted({k.split("/")[1] for k in z.files if k.startswith("dir/")}):
        for typ in ("slope", "angle"):
            p = f"dir/{name}/{typ}/"
            cfg = z[p + "cfg"]
            gp = WN.goal_position(z[p + "cells"], cfg)
            assert gp == tuple(z[p + "goal_position"])
            st = np.array([*z[p + "agent"], 0, 0, 0, 0, 0], dtype=np.int32)
            env = O.OracleEnv(z[p + "cells"], st, cfg)
            want = z[p + "goal_direction"]
            got = [WN.goal_direction(gp, env.st, typ)]
            for a in z[p + "actions"]:
                env.step(int(a))
                got.append(WN.goal_direction(gp, env.st, typ))
            assert np.array_equal(np.array(got), want, equal_nan=True), p
    for k in range(3):
        p = f"stoch/{k}/"
        ra = None if int(z[p + "random_action"]) < 0 else int(z[p + "random_action"])
        rep = iter(z[p + "replacement"])
        for a, u, want in zip(z[p + "given"], z[p + "uniform"], z[p + "got"]):
            r = None
            if u >= float(z[p + "prob"]) and ra is None:
                r = next(rep)
            assert WN.stochastic_action(a, u, r, float(z[p + "prob"]), ra) == want
