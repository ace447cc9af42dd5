"""GPU parity of the wrapper kernels through the host mirror minigrid_dynamicprogramming_b200.wrappers:
against what the reference's own wrappers returned (tests/golden/wrappers.npz) and against oracle/wrappers_np.py on
larger random batches — bit-exact, including the float32 FlatObs vector and the fp64 bonus rewards."""
import numpy as np
import pytest

from helpers import CONFIG_ENVS, L, O, golden, oracle_rollout, short

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _mods():
    from minigrid_dynamicprogramming_b200 import wrappers as W
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    return BatchedMiniGrid, W


GOLDEN_ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0",
               "MiniGrid-KeyCorridorS6R3-v0"]


@pytest.mark.parametrize("eid", GOLDEN_ENVS)
def test_wrappers_match_reference_episode(eid):
    """One env stepped through the recorded actions; every wrapper's output compared with the reference's at each step."""
    Env, W = _mods()
    z, name = golden("wrappers.npz"), short(eid)
    acts = z[name + "/actions"]

    def fresh():
        return Env(eid, 1, seeds=[0], autoreset="off")

    wrapped = {"fully": W.FullyObsWrapper(fresh()), "onehot": W.OneHotPartialObsWrapper(fresh()),
               "symbolic": W.SymbolicObsWrapper(fresh()), "flat": W.FlatObsWrapper(fresh()),
               "words": W.DictObservationSpaceWrapper(fresh()), "img": W.ImgObsWrapper(fresh()),
               "view3": W.ViewSizeWrapper(fresh(), 3), "view5": W.ViewSizeWrapper(fresh(), 5),
               "view9": W.ViewSizeWrapper(fresh(), 9), "view11": W.ViewSizeWrapper(fresh(), 11)}
    bonus = {"abonus": W.ActionBonus(fresh()), "pbonus": W.PositionBonus(fresh())}
    obs = {k: w.reset()[0] for k, w in wrapped.items()}
    for w in bonus.values():
        w.reset()

    def check(t):
        assert np.array_equal(obs["fully"]["image"][0].cpu().numpy(), z[name + "/fully"][t]), t
        assert np.array_equal(obs["onehot"]["image"][0].cpu().numpy(), z[name + "/onehot"][t]), t
        sym = obs["symbolic"]["image"][0].cpu().numpy()
        assert sym.dtype == np.int64 and np.array_equal(sym, z[name + "/symbolic"][t]), t
        flat = obs["flat"][0].cpu().numpy()
        assert flat.dtype == np.float32 and np.array_equal(flat, z[name + "/flat"][t]), t
        assert np.array_equal(obs["words"]["mission"][0].cpu().numpy(), z[name + "/words"][t]), t
        assert np.array_equal(obs["img"][0].cpu().numpy(), z[name + "/image"][t]), t
        for v in (3, 5, 9, 11):
            assert np.array_equal(obs[f"view{v}"]["image"][0].cpu().numpy(), z[f"{name}/view{v}"][t]), (v, t)

    check(0)
    for t, a in enumerate(acts):
        act = torch.tensor([a], dtype=torch.uint8, device="cuda")
        for k, w in wrapped.items():
            obs[k] = w.step(act)[0]
        check(t + 1)
        for k, w in bonus.items():
            r = w.step(act)[1]
            assert float(r[0]) == z[name + "/" + k][t], (k, t)


def test_wrappers_batch_vs_oracle():
    """256 DoorKey-8x8 envs, 60 random steps one at a time (autoreset off): FullyObs / Symbolic / OneHot / Flat / bonuses
    of the whole batch against oracle/wrappers_np.py applied to the oracle's states."""
    from oracle import wrappers_np as WN
    Env, W = _mods()
    eid, N, T = "MiniGrid-DoorKey-8x8-v0", 256, 60
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(3).integers(0, 7, (T, N)).astype(np.uint8)
    env = Env(eid, N, layouts=lays, autoreset="off")
    ab = W.ActionBonus(env)
    env.reset()
    counters = [WN.BonusCounter("action") for _ in range(N)]
    cfg = O.make_cfg(8, 8, lays[0].max_steps, False, 0)
    for t in range(T):
        ref = oracle_rollout(lays, acts[: t + 1])            # oracle state after t+1 steps (small: replays from the start)
        obs, r, te, tr, _ = ab.step(torch.from_numpy(acts[t]).cuda())
        r = r.cpu().numpy()
        if t % 10 == 9 or t == 0:
            fully = W.FullyObsWrapper(env).observation(obs)["image"].cpu().numpy()
            sym = W.SymbolicObsWrapper(env).observation(obs)["image"].cpu().numpy()
            onehot = W.OneHotPartialObsWrapper(env).observation(obs)["image"].cpu().numpy()
            flat = W.FlatObsWrapper(env).observation(obs).cpu().numpy()
            img = obs["image"].cpu().numpy()
            for e in range(0, N, 7):
                c = ref["cells"][e, : 4 * 64].reshape(4, 64)
                assert np.array_equal(fully[e], WN.fully_obs(c, cfg, ref["st"][e])), (t, e)
                assert np.array_equal(sym[e], WN.symbolic_obs(c, cfg, ref["st"][e])), (t, e)
                assert np.array_equal(onehot[e], WN.onehot_obs(img[e])), (t, e)
                assert np.array_equal(flat[e], WN.flat_obs(img[e], lays[e].mission)), (t, e)
        for e in range(N):
            want = ref["reward"][t, e] + counters[e].bonus(ref["st"][e], acts[t, e])
            assert r[e] == want, (t, e)


def test_onehot_on_a_whole_rollout_and_ragged_sizes():
    """mg_obs_onehot over the [T, N] images of one step_many launch; N*T*49 not a multiple of the 128-cell CTA."""
    from oracle import wrappers_np as WN
    Env, W = _mods()
    env = Env("MiniGrid-KeyCorridorS3R3-v0", 37, autoreset="same_step")
    env.reset()
    acts = torch.randint(0, 7, (11, 37), dtype=torch.uint8, device="cuda")
    obs, *_ = env.step_many(acts)
    out = W.OneHotPartialObsWrapper(env).observation(obs)["image"]
    assert out.shape == (11, 37, 7, 7, 20)
    assert np.array_equal(out.cpu().numpy(), WN.onehot_obs(obs["image"].cpu().numpy()))


def test_rgb_wrappers_and_composition():
    Env, W = _mods()
    env = Env("MiniGrid-DoorKey-8x8-v0", 16, autoreset="off")
    w = W.ImgObsWrapper(W.RGBImgPartialObsWrapper(env, tile_size=8))
    img, _ = w.reset()
    assert img.shape == (16, 56, 56, 3) and img.dtype == torch.uint8
    assert torch.equal(img, env.get_frame(agent_pov=True, tile_size=8))
    full = W.RGBImgObsWrapper(env, tile_size=4).reset()[0]["image"]
    assert full.shape == (16, 32, 32, 3)
    rs = W.ReseedWrapper(env, seeds=[5, 9], seed_idx=0)
    a = rs.reset()[0]["image"].clone()
    b = rs.reset()[0]["image"].clone()
    c = rs.reset()[0]["image"].clone()
    assert torch.equal(a, c) and not torch.equal(a, b)
    ref = Env("MiniGrid-DoorKey-8x8-v0", 16, seeds=range(5, 21), autoreset="off").reset()[0]["image"]
    assert torch.equal(a, ref)


@pytest.mark.parametrize("V", [3, 5, 7, 9, 13, 15])
def test_view_size_batch_vs_oracle(V):
    """mg_obs_view on 160 playground grids (every object type, carried objects) against the oracle's gen_obs at view V;
    V = 7 must also equal the step kernel's own observation."""
    from test_gpu_parity import _random_playground
    Env, W = _mods()
    rng = np.random.default_rng(V)
    N, T, Wd, Hd = 160, 30, 10, 8
    lays = [_random_playground(rng, Wd, Hd, s) for s in range(N)]
    acts = rng.integers(0, 7, (T, N)).astype(np.uint8)
    ref = oracle_rollout(lays, acts, autoreset=False, pool=lays)
    env = Env("playground", N, layouts=lays, autoreset="off")
    env.reset()
    obs, *_ = env.step_many(torch.from_numpy(acts).cuda())
    got = W.ViewSizeWrapper(env, V).observation({"image": obs["image"][-1]})["image"].cpu().numpy()
    if V == 7:
        assert np.array_equal(got, obs["image"][-1].cpu().numpy())
    cfg = O.make_cfg(Wd, Hd, 200, False, 1, V)
    for e in range(N):
        c = ref["cells"][e, : 4 * Wd * Hd].reshape(4, Wd * Hd)
        want, _ = O.OracleEnv(c, ref["st"][e], cfg).gen_obs()
        assert np.array_equal(got[e], want), e


@pytest.mark.parametrize("eid", ["MiniGrid-LavaCrossingS9N1-v0", "MiniGrid-DistShift1-v0", "MiniGrid-LavaGapS6-v0"])
def test_no_death_matches_reference(eid):
    """NoDeath as a step-kernel mode against the reference's wrapped episodes (lava costs instead of killing)."""
    Env, W = _mods()
    z, p = golden("wrappers.npz"), "nodeath/" + short(eid) + "/"
    cost = float(z[p + "cost"])
    nres = int(z[p + "reset_seed"].max()) + 1
    base = Env(eid, 1, seeds=range(max(nres, 1)), autoreset="off")
    env = W.NoDeath(base, no_death_types=("lava",), death_cost=cost)
    obs, _ = env.reset()
    assert (obs["image"][0].cpu().numpy() == z[p + "image"][0]).all()
    for t, a in enumerate(z[p + "actions"]):
        obs, r, te, tr, _ = env.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))
        assert (obs["image"][0].cpu().numpy() == z[p + "image"][t + 1]).all(), t
        assert float(r[0]) == z[p + "reward"][t] and bool(te[0]) == bool(z[p + "terminated"][t]) and bool(tr[0]) == bool(z[p + "truncated"][t]), t
        if z[p + "reset_seed"][t] >= 0:
            base.reset_envs([int(z[p + "reset_seed"][t])])


@pytest.mark.parametrize("view", [3, 5, 9, 11])
def test_agent_view_size_constructor_knob(view):
    """MiniGridEnv(agent_view_size=V) (minigrid_env.py:42,66-68): stepping with another view size — against the oracle's literal
    gen_obs at that size, with same-step autoreset, and against the reference's ViewSizeWrapper goldens (the wrapper rebuilds
    exactly the observation the constructor knob would give, wrappers.py:660-670)."""
    import dataclasses
    Env, W = _mods()
    eid, N, T = "MiniGrid-DoorKey-8x8-v0", 96, 60
    lays = [dataclasses.replace(l, agent_view_size=view, max_steps=25) for l in L.generate_many(eid, range(N))]
    acts = np.random.default_rng(view).integers(0, 7, (T, N)).astype(np.uint8)
    ref = oracle_rollout(lays, acts, autoreset=True, pool=lays)
    env = Env(eid, N, layouts=lays, autoreset="same_step", agent_view_size=view)
    assert env.observation_space["image"].shape == (view, view, 3)
    obs0, _ = env.reset()
    assert obs0["image"].shape == (N, view, view, 3)
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    assert (obs["image"].cpu().numpy() == ref["image"]).all()
    assert (r.cpu().numpy() == ref["reward"]).all() and (tr.cpu().numpy().astype(np.uint8) == ref["truncated"]).all()
    assert ref["truncated"].sum() >= N
    z = golden("wrappers.npz")
    env1 = Env("MiniGrid-DoorKey-8x8-v0", 1, seeds=[0], autoreset="off", agent_view_size=view)
    o, _ = env1.reset()
    assert (o["image"][0].cpu().numpy() == z[f"DoorKey-8x8/view{view}"][0]).all()
    for t, a in enumerate(z["DoorKey-8x8/actions"]):
        o = env1.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))[0]
        assert (o["image"][0].cpu().numpy() == z[f"DoorKey-8x8/view{view}"][t + 1]).all(), t


@pytest.mark.parametrize("eid", ["MiniGrid-Dynamic-Obstacles-5x5-v0", "MiniGrid-Dynamic-Obstacles-6x6-v0"])
def test_no_death_on_dynamic_obstacles_matches_reference(eid):
    """NoDeath(("ball",)) on DynamicObstacles (the pairing of the reference's own test_no_death_wrapper): the wrapper looks at
    the front cell BEFORE env.step, i.e. before the obstacles move (wrappers.py:852-858) — a ball that was in front and then
    moved away still cancels the death.  Recorded from the reference (tests/golden/nodeath_dynobs.npz)."""
    Env, W = _mods()
    z, p = golden("nodeath_dynobs.npz"), short(eid) + "/"
    cost = float(z[p + "cost"])
    nres = int(z[p + "reset_seed"].max()) + 1
    base = Env(eid, 1, seeds=range(max(nres, 1)), autoreset="off")
    env = W.NoDeath(base, no_death_types=("ball",), death_cost=cost)
    obs, _ = env.reset()
    assert (obs["image"][0].cpu().numpy() == z[p + "image"][0]).all()
    assert (base.get_dyn()[0] == z[p + "dyn"][0]).all()
    cancelled = 0
    for t, a in enumerate(z[p + "actions"]):
        obs, r, te, tr, _ = env.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))
        assert (obs["image"][0].cpu().numpy() == z[p + "image"][t + 1]).all(), t
        assert float(r[0]) == z[p + "reward"][t] and bool(te[0]) == bool(z[p + "terminated"][t]) and bool(tr[0]) == bool(z[p + "truncated"][t]), t
        assert (base.get_dyn()[0] == z[p + "dyn"][t + 1]).all(), t
        cancelled += float(r[0]) == -1 + cost
        if z[p + "reset_seed"][t] >= 0:
            base.reset_envs([int(z[p + "reset_seed"][t])])
    assert cancelled > 20


def test_no_death_batch_with_autoreset():
    """512 LavaCrossing envs with same-step autoreset: a cancelled death must not reset the env; against the oracle."""
    Env, W = _mods()
    eid, N, T = "MiniGrid-LavaCrossingS9N2-v0", 512, 120
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(9).choice(np.array([0, 1, 2, 2, 2], np.uint8), size=(T, N))
    env = W.NoDeath(Env(eid, N, layouts=lays, autoreset="same_step"), ("lava",), -0.5)
    env.reset()
    from helpers import oracle_env
    oenvs = [oracle_env(l) for l in lays]
    lid = list(range(N))
    paid = 0
    for t in range(T):
        obs, r, te, tr, _ = env.step(torch.from_numpy(acts[t]).cuda())
        img, r, te, tr = obs["image"].cpu().numpy(), r.cpu().numpy(), te.cpu().numpy(), tr.cpu().numpy()
        for e in range(0, N, 5):
            wi, wd, wr, wte, wtr = O.step_no_death(oenvs[e], int(acts[t, e]), (9,), -0.5)
            assert r[e] == wr and te[e] == wte and tr[e] == wtr, (t, e)
            paid += wr < 0
            if wte or wtr:                                   # same-step autoreset: next pool slot, obs of the new episode
                lid[e] = (lid[e] + O.reset_stride(N, N)) % N
                oenvs[e] = oracle_env(lays[lid[e]])
                wi, _ = oenvs[e].gen_obs()
            assert (img[e] == wi).all(), (t, e)
    assert paid > 20


def test_direction_obs_wrapper_matches_reference():
    """DirectionObsWrapper: slope bit-exact (one fp64 division, inf / nan included), angle within 1 ulp of numpy's arctan;
    the cached goal_position is the reference's (row, column) tuple."""
    import dataclasses
    Env, W = _mods()
    z = golden("wrappers2.npz")
    for name in sorted({k.split("/")[1] for k in z.files if k.startswith("dir/")}):
        for typ in ("slope", "angle"):
            p = f"dir/{name}/{typ}/"
            cfg = z[p + "cfg"]
            lay = L.Layout("MiniGrid-" + name + "-v0", 3, int(cfg[0]), int(cfg[1]), int(cfg[2]), bool(cfg[3]), int(cfg[4]),
                           z[p + "cells"], (int(z[p + "agent"][0]), int(z[p + "agent"][1])), int(z[p + "agent"][2]), "")
            env = W.DirectionObsWrapper(Env(lay.env_id, 1, layouts=[lay], autoreset="off"), type=typ)
            obs, _ = env.reset()
            assert tuple(env.goal_position[0]) == tuple(z[p + "goal_position"])
            got = [float(obs["goal_direction"][0])]
            for a in z[p + "actions"]:
                obs = env.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))[0]
                got.append(float(obs["goal_direction"][0]))
            got, want = np.array(got), z[p + "goal_direction"]
            if typ == "slope":
                assert np.array_equal(got, want, equal_nan=True), p
            else:
                assert np.all((got == want) | (np.abs(got - want) <= np.spacing(np.abs(want))) | (np.isnan(got) & np.isnan(want))), p


def test_direction_obs_batch_vs_oracle():
    from oracle import wrappers_np as WN
    Env, W = _mods()
    eid, N, T = "MiniGrid-LavaCrossingS11N5-v0", 192, 40
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(9).choice(np.array([0, 1, 2, 2], np.uint8), size=(T, N))
    env = W.DirectionObsWrapper(Env(eid, N, layouts=lays, autoreset="off"))
    env.reset()
    cfg = O.make_cfg(11, 11, lays[0].max_steps, False, 0)
    for t in range(T):
        obs = env.step(torch.from_numpy(acts[t]).cuda())[0]
    ref = oracle_rollout(lays, acts)
    got = obs["goal_direction"].cpu().numpy()
    for e in range(N):
        gp = WN.goal_position(lays[e].cells, cfg)
        want = WN.goal_direction(gp, ref["st"][e])
        assert got[e] == want or (np.isnan(got[e]) and np.isnan(want)), e


def test_stochastic_action_wrapper_replays_reference_draws():
    """StochasticActionWrapper.action with the reference's own draw sequences: the 400 calls of the golden as one batch."""
    Env, W = _mods()
    z = golden("wrappers2.npz")
    for k in range(3):
        p = f"stoch/{k}/"
        prob, ra = float(z[p + "prob"]), int(z[p + "random_action"])
        given, u, want = z[p + "given"], z[p + "uniform"], z[p + "got"]
        rep = np.zeros(len(given), np.uint8)
        rep[u >= prob] = z[p + "replacement"] if ra < 0 else 0
        env = W.StochasticActionWrapper(Env("MiniGrid-Empty-8x8-v0", len(given), pool_size=4, autoreset="off"), prob=prob,
                                        random_action=None if ra < 0 else ra)
        env.reset()
        out = env.action(given, uniform=u, replacement=rep)
        assert np.array_equal(out.cpu().numpy(), want)
        out2 = env.action(given)                       # own draws: still only kept or replaced values, about prob kept
        kept = (out2.cpu().numpy() == given).mean()
        assert kept >= prob - 0.1
        env.step(given)
