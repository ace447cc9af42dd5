"""GPU parity: the CUDA engine (through the C ABI) against the oracle and the golden vectors — bit-exact."""
import numpy as np
import pytest

from helpers import CONFIG_ENVS, L, O, golden, oracle_rollout, short

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _engine():
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    return BatchedMiniGrid


def _np(x):
    return x.cpu().numpy()


def _compare(obs, r, te, tr, ref, what=""):
    assert (_np(obs["image"]) == ref["image"]).all(), f"{what}: image"
    assert (_np(obs["direction"]) == ref["direction"]).all(), f"{what}: direction"
    assert (_np(r) == ref["reward"]).all(), f"{what}: reward (bit-exact fp64)"
    assert (_np(te).astype(np.uint8) == ref["terminated"]).all(), f"{what}: terminated"
    assert (_np(tr).astype(np.uint8) == ref["truncated"]).all(), f"{what}: truncated"


@pytest.mark.parametrize("eid", CONFIG_ENVS)
def test_golden_random_rollout(eid):
    """The reference's own 2,000-step rollout (tests/golden/traj_*.npz), one step() per action."""
    z = golden(f"traj_{short(eid)}.npz")
    nres = int(z["reset_seed"].max()) + 1
    env = _engine()(eid, 1, seeds=range(max(nres, 1)), autoreset="off")
    obs, _ = env.reset()
    assert (_np(obs["image"])[0] == z["first_image"]).all()
    acts = z["actions"]
    for t in range(len(acts)):
        obs, r, te, tr, _ = env.step(torch.tensor([acts[t]], dtype=torch.uint8, device="cuda"))
        assert (_np(obs["image"])[0] == z["image"][t]).all(), t
        assert (int(obs["direction"][0]), float(r[0]), bool(te[0]), bool(tr[0])) == (
            z["direction"][t], z["reward"][t], bool(z["terminated"][t]), bool(z["truncated"][t])), t
        if z["reset_seed"][t] >= 0:
            env.reset_envs([int(z["reset_seed"][t])])


@pytest.mark.parametrize("name", ["Empty-8x8", "DoorKey-8x8", "DoorKey-16x16", "LavaCrossingS11N5", "FourRooms",
                                  "KeyCorridorS3R3", "KeyCorridorS6R3"])
def test_golden_greedy_rollouts(name):
    """Goal-reaching rollouts recorded in the reference: non-zero fp64 rewards, door/key/pickup/drop paths."""
    import os
    from helpers import GOLDEN
    if not os.path.exists(os.path.join(GOLDEN, f"greedy_{name}.npz")):
        pytest.skip("fixture not recorded")
    z = golden(f"greedy_{name}.npz")
    eid = f"MiniGrid-{name}-v0"
    seeds = sorted({int(k.split("/")[0][1:]) for k in z.files})
    env = _engine()(eid, len(seeds), seeds=seeds, autoreset="off")
    obs, _ = env.reset()
    T = max(len(z[f"s{s}/actions"]) for s in seeds)
    for i, s in enumerate(seeds):
        assert (_np(obs["image"])[i] == z[f"s{s}/image"][0]).all()
    acts = np.full((T, len(seeds)), 6, np.uint8)         # pad with `done` (no-op) after the episode ends
    for i, s in enumerate(seeds):
        a = z[f"s{s}/actions"]
        acts[: len(a), i] = a
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    for i, s in enumerate(seeds):
        d = len(z[f"s{s}/actions"])
        assert (_np(obs["image"])[:d, i] == z[f"s{s}/image"][1:]).all(), (name, s)
        assert (_np(obs["direction"])[:d, i] == z[f"s{s}/direction"][1:]).all()
        assert (_np(r)[:d, i] == z[f"s{s}/reward"]).all(), (name, s, _np(r)[:d, i][-1], z[f"s{s}/reward"][-1])
        assert (_np(te)[:d, i] == z[f"s{s}/terminated"].astype(bool)).all()
        assert (_np(tr)[:d, i] == z[f"s{s}/truncated"].astype(bool)).all()
        assert _np(r)[d - 1, i] > 0


@pytest.mark.parametrize("eid", CONFIG_ENVS)
@pytest.mark.parametrize("autoreset", ["off", "same_step"])
def test_batch_vs_oracle(eid, autoreset):
    N, T = 256, 400
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(1).integers(0, 7, (T, N)).astype(np.uint8)
    # bias towards forward so that goals / lava are reached and doors get used
    acts[np.random.default_rng(2).random((T, N)) < 0.3] = 2
    ref = oracle_rollout(lays, acts, autoreset=autoreset == "same_step", pool=lays)
    env = _engine()(eid, N, layouts=lays, autoreset=autoreset)
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    _compare(obs, r, te, tr, ref, eid)
    st = env.get_state(cells=False)
    assert (st["x"] == ref["st"][:, 0]).all() and (st["y"] == ref["st"][:, 1]).all()
    assert (st["step_count"] == ref["st"][:, 3]).all() and (st["carry_type"] == ref["st"][:, 4]).all()


def test_step_many_equals_steps():
    eid, N, T = "MiniGrid-DoorKey-8x8-v0", 512, 64
    acts = torch.from_numpy(np.random.default_rng(3).integers(0, 7, (T, N)).astype(np.uint8)).cuda()
    a, b = _engine()(eid, N), _engine()(eid, N)
    a.reset(), b.reset()
    oa, ra, tea, tra, _ = a.step_many(acts)
    for t in range(T):
        ob, rb, teb, trb, _ = b.step(acts[t])
        assert torch.equal(ob["image"], oa["image"][t]) and torch.equal(rb, ra[t])
        assert torch.equal(teb, tea[t]) and torch.equal(trb, tra[t]) and torch.equal(ob["direction"], oa["direction"][t])


@pytest.mark.parametrize("eid,N", [("MiniGrid-DoorKey-8x8-v0", 65536), ("MiniGrid-LavaCrossingS11N5-v0", 20000), ("MiniGrid-DoorKey-5x5-v0", 40)])
def test_back_to_back_launches_vs_oracle(eid, N):
    """Per-CTA launch chaining (mg_step.cu: a step CTA waits for the same CTA of the previous launch, not for the whole grid):
    hundreds of short launches enqueued back to back WITHOUT a host sync in between — more than one wave, one wave and a
    single CTA (where dozens of launches are resident at once, each waiting for the one before) — must leave what one long
    launch leaves: every output of every step, the final planes and agents equal to the oracle's."""
    T, P = 192, 257
    lays = L.generate_many(eid, range(P))
    acts = np.random.default_rng(21).integers(0, 7 if "DoorKey" in eid else 3, (T, N)).astype(np.uint8)
    ref = oracle_rollout([lays[i % P] for i in range(N)], acts, autoreset=True, pool=lays)
    env = _engine()(eid, N, layouts=lays, autoreset="same_step")
    env.reset_envs(torch.arange(N, device="cuda") % P)
    a = torch.from_numpy(acts).cuda()
    imgs, rews, terms = [], [], []
    t = 0
    for chunk in [1] * 40 + [3] * 20 + [2] * 46:                     # 192 steps in 106 launches
        obs, r, te, tr, _ = env.step_many(a[t:t + chunk])
        imgs.append(obs["image"].clone()), rews.append(r.clone()), terms.append(te.clone())       # (device copies: no host sync)
        t += chunk
    assert t == T
    assert (_np(torch.cat(imgs)) == ref["image"]).all()
    assert (_np(torch.cat(rews)) == ref["reward"]).all() and (_np(torch.cat(terms)).astype(np.uint8) == ref["terminated"]).all()
    n = lays[0].width * lays[0].height
    assert (env.get_state()[1][:, :, :n].reshape(N, -1) == ref["cells"]).all()
    env.check_errors()


def test_graph_replay_of_chained_launches():
    """A CUDA graph of 16 single-step launches, replayed: the chaining counters live in the engine and count on their own, so a
    replay continues the chain (nothing of a launch's position in the chain is baked into its parameters).  Three replays ==
    48 eager steps with the same actions."""
    eid, N = "MiniGrid-DoorKey-8x8-v0", 4096
    acts = torch.from_numpy(np.random.default_rng(4).integers(0, 7, (16, N)).astype(np.uint8)).cuda()
    a, b = _engine()(eid, N), _engine()(eid, N)
    a.reset(), b.reset()
    outs = [a.step_many(acts[i:i + 1]) for i in range(1)]              # allocates the T = 1 outputs before the capture
    b.step_many(acts[0:1])
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    kept = []
    with torch.cuda.stream(s):
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for i in range(16):
                obs, r, te, tr, _ = a.step_many(acts[i:i + 1])
                kept.append((obs["image"].clone(), r.clone()))
    for rep in range(3):
        g.replay()
        torch.cuda.synchronize()
        for i in range(16):
            ob, rb, teb, trb, _ = b.step_many(acts[i:i + 1])
            assert torch.equal(kept[i][0], ob["image"]) and torch.equal(kept[i][1], rb), (rep, i)
    a.check_errors()


def test_chain_counters_wrap(monkeypatch):
    """The chaining counters count modulo 2^32 (a T = 1 loop at 100 k launches/s gets there in 12 hours): an engine whose counters
    start 20 launches before the wrap (test hook MG_CHAIN_START, read by mg_create) steps through it — 60 single-step launches
    back to back == one launch of 60 steps on an engine that starts at zero."""
    eid, N, T = "MiniGrid-DoorKey-8x8-v0", 3000, 60
    acts = torch.from_numpy(np.random.default_rng(8).integers(0, 7, (T, N)).astype(np.uint8)).cuda()
    b = _engine()(eid, N)
    monkeypatch.setenv("MG_CHAIN_START", str(0xFFFFFFFF - 20))
    a = _engine()(eid, N)
    monkeypatch.delenv("MG_CHAIN_START")
    a.reset(), b.reset()
    ob, rb, teb, trb, _ = b.step_many(acts)
    imgs, rews = [], []
    for t in range(T):
        oa, ra, tea, tra, _ = a.step_many(acts[t:t + 1])
        imgs.append(oa["image"].clone()), rews.append(ra.clone())
    assert torch.equal(torch.cat(imgs), ob["image"]) and torch.equal(torch.cat(rews), rb)
    a.check_errors()
    assert (a.get_state()[1] == b.get_state()[1]).all()


def test_mixed_layout_batch():
    """BASELINE config 4 in small: even envs LavaCrossingS11N5 (11x11), odd envs FourRooms (19x19, max_steps 100)."""
    ids = ["MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0"]
    N, T = 512, 300
    env = _engine()(ids, N, seeds=[e // 2 for e in range(N)], autoreset="same_step")
    lays = env._pool_layouts
    assert lays[0].width == 11 and lays[1].width == 19
    acts = np.random.default_rng(4).integers(0, 7, (T, N)).astype(np.uint8)
    acts[np.random.default_rng(5).random((T, N)) < 0.3] = 2
    ref = oracle_rollout(lays, acts, autoreset=True, pool=lays)
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    _compare(obs, r, te, tr, ref, "mixed")
    assert ref["terminated"].sum() > 0 and ref["truncated"].sum() > 0
    # missions follow the pool slot an env currently plays (autoreset moves envs between the two classes: stride 1)
    missions, lid = list(obs["mission"]), env.get_state(cells=False)["layout_id"]
    assert all(missions[e] == lays[lid[e]].mission for e in range(N))
    assert any(m.startswith("avoid the lava") for m in missions) and any(m == "reach the goal" for m in missions)
    assert (lid != np.arange(N)).any()


def test_truncation_known_answer():
    """tests/test_envs.py:146-166: max_steps=50, action drop repeated -> truncated first at step 50."""
    env = _engine()("MiniGrid-Empty-8x8-v0", 4, max_steps=50, autoreset="off")
    env.reset()
    acts = torch.full((60, 4), 4, dtype=torch.uint8, device="cuda")
    _, r, te, tr, _ = env.step_many(acts)
    tr = _np(tr)
    assert not tr[:49].any() and tr[49:].all() and not _np(te).any() and (_np(r) == 0).all()


def test_invalid_action_raises_after_counting():
    """minigrid_env.py:523 runs before :579-580: step_count is incremented, then ValueError."""
    env = _engine()("MiniGrid-Empty-8x8-v0", 8, autoreset="off")
    env.reset()
    with pytest.raises(ValueError):
        env.step(np.array([0, 1, 2, 7, 0, 0, 0, 0]))
    st = env.get_state(cells=False)
    assert (st["step_count"] == 1).all() and st["dir"][3] == 0
    env.step(np.zeros(8, np.int64))        # the engine keeps working afterwards
    env.check_errors()


def test_reset_seed_semantics_and_first_obs():
    """wrappers.py:230-237: Empty-5x5 first obs column is 7x[2,5,0]; reset(seed=list) regenerates layouts."""
    env = _engine()("MiniGrid-Empty-5x5-v0", 2, autoreset="off")
    obs, _ = env.reset()
    assert (_np(obs["image"])[0, 0] == np.array([[2, 5, 0]] * 7)).all()
    env = _engine()("MiniGrid-DoorKey-8x8-v0", 3, autoreset="off")
    obs, _ = env.reset(seed=[5, 0, 7])
    assert (env.agent_pos[1] == (3, 4)).all() and env.agent_dir[1] == 1        # seed 0 -> (3,4) dir 1
    obs2, _ = env.reset(seed=5)                                                  # seeds 5,6,7
    assert (_np(obs2["image"])[0] == _np(obs["image"])[0]).all()


def test_full_batch_autoreset_changes_layouts_vs_oracle():
    """The benchmarked configuration at full size, with resets that CHANGE layouts: 65,536 DoorKey-8x8 envs, T = 64 in one
    launch, same-step autoreset from a pool of 4,099 layouts (a prime: it does not divide N, and the reset stride is coprime
    to it anyway), max_steps = 24 so that every env is truncated and reloaded at least twice.  Every output of every env is
    compared with the C oracle, and so are the planes, agents and pool slots the rollout ends in."""
    eid, N, P, T = "MiniGrid-DoorKey-8x8-v0", 65536, 4099, 64
    pool = L.generate_many(eid, range(P), max_steps=24)
    lays = [pool[e % P] for e in range(N)]
    acts = np.random.default_rng(66).integers(0, 7, (T, N)).astype(np.uint8)
    ref = oracle_rollout(lays, acts, autoreset=True, pool=pool)
    env = _engine()(eid, N, layouts=pool, autoreset="same_step")
    assert env.reset_stride == O.reset_stride(N, P) and np.gcd(env.reset_stride, P) == 1
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    _compare(obs, r, te, tr, ref, "65536 autoreset P=4099")
    assert ref["truncated"].sum() >= 2 * N
    agents, cells = env.get_state()
    assert (cells[:, :, :64].reshape(N, -1) == ref["cells"]).all()
    assert (agents["x"] == ref["st"][:, 0]).all() and (agents["y"] == ref["st"][:, 1]).all() and (agents["dir"] == ref["st"][:, 2]).all()
    assert (agents["step_count"] == ref["st"][:, 3]).all()
    # every env has moved through the pool: slot after k resets = (e + k * stride) % P
    k = ref["truncated"].sum(axis=0) + ref["terminated"].sum(axis=0) - (ref["truncated"] & ref["terminated"]).sum(axis=0)
    assert (agents["layout_id"] == (np.arange(N) + k.astype(np.int64) * env.reset_stride) % P).all()
    assert len(np.unique(agents["layout_id"] - np.arange(N) % P)) > 1


def test_reset_rejects_layout_id_outside_pool():
    from minigrid_dynamicprogramming_b200 import _capi
    env = _engine()("MiniGrid-Empty-8x8-v0", 8, pool_size=4, autoreset="off")
    env.reset()
    before = env.get_state(cells=False).copy()
    env.reset_envs([0, 1, 2, 3, 4, 1, 2, 3])                  # slot 4 does not exist
    with pytest.raises(_capi.MgError):
        env.check_errors()
    assert env.get_state(cells=False)["layout_id"][4] == before["layout_id"][4]
    env.check_errors()


def test_full_batch_65536_vs_oracle():
    """BASELINE config 2 size: 65,536 DoorKey-8x8 envs (4,096-layout pool), autoreset, 8 steps, all envs checked."""
    eid, N, P, T = "MiniGrid-DoorKey-8x8-v0", 65536, 4096, 8
    pool = L.generate_many(eid, range(P))
    lays = [pool[e % P] for e in range(N)]
    acts = np.random.default_rng(6).integers(0, 7, (T, N)).astype(np.uint8)
    ref = oracle_rollout(lays, acts, autoreset=False)
    env = _engine()(eid, N, layouts=pool, autoreset="off")
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    _compare(obs, r, te, tr, ref, "65536")


def _random_playground(rng, W, H, seed):
    """A hand-made layout exercising every object type the engine supports, incl. boxes with hidden contents,
    floor tiles, open/closed/locked doors of several colours and lava (objects the six config envs never place)."""
    from minigrid_dynamicprogramming_b200.levelgen import Layout
    n = W * H
    cells = np.zeros((4, n), np.uint8)
    cells[0] = 1
    for x in range(W):
        for y in (0, H - 1):
            cells[:3, y * W + x] = (2, 5, 0)
    for y in range(H):
        for x in (0, W - 1):
            cells[:3, y * W + x] = (2, 5, 0)
    interior = [(x, y) for x in range(1, W - 1) for y in range(1, H - 1)]
    rng.shuffle(interior)
    ax, ay = interior.pop()
    k = 0
    for (x, y) in interior:
        u = rng.random()
        i = y * W + x
        col = int(rng.integers(0, 6))
        if u < 0.08:
            cells[:3, i] = (2, 5, 0)                                  # wall
        elif u < 0.14:
            cells[:3, i] = (4, col, int(rng.integers(0, 3)))          # door open / closed / locked
        elif u < 0.20:
            cells[:3, i] = (5, col, 0)                                # key
        elif u < 0.25:
            cells[:3, i] = (6, col, 0)                                # ball
            if k == 0:
                cells[3, i] = 0x80                                    # mission target
                k = 1
        elif u < 0.32:
            cells[:3, i] = (7, col, 0)                                # box ...
            c = rng.random()
            if c < 0.4:
                cells[3, i] = 5 | (int(rng.integers(0, 6)) << 4)      # ... hiding a key
            elif c < 0.7:
                cells[3, i] = 6 | (int(rng.integers(0, 6)) << 4)      # ... hiding a ball
        elif u < 0.36:
            cells[:3, i] = (3, col, 0)                                # floor
        elif u < 0.39:
            cells[:3, i] = (9, 0, 0)                                  # lava
        elif u < 0.41:
            cells[:3, i] = (8, 1, 0)                                  # goal
    return Layout("playground", seed, W, H, 200, False, 1, cells, (ax, ay), int(rng.integers(0, 4)), "playground")


@pytest.mark.parametrize("W,H", [(9, 9), (13, 7), (5, 5), (22, 22)])
def test_playground_layouts_all_object_types(W, H):
    """Boxes (toggle reveals contents, carried boxes keep them), floor, multi-colour doors/keys, target pickup rule."""
    rng = np.random.default_rng(W * 100 + H)
    N, T = 192, 600
    lays = [_random_playground(rng, W, H, s) for s in range(N)]
    acts = rng.integers(0, 7, (T, N)).astype(np.uint8)
    acts[rng.random((T, N)) < 0.25] = 2
    acts[rng.random((T, N)) < 0.10] = 5
    for autoreset in ("off", "same_step"):
        ref = oracle_rollout(lays, acts, autoreset=autoreset == "same_step", pool=lays)
        env = _engine()("playground", N, layouts=lays, autoreset=autoreset)
        env.reset()
        obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
        _compare(obs, r, te, tr, ref, f"playground {W}x{H} {autoreset}")
        st, cells = env.get_state()
        if autoreset == "off":
            for e in range(N):                       # the authoritative SoA planes followed every mutation
                n = W * H
                assert (cells[e, :, :n].reshape(-1) == ref["cells"][e, : 4 * n]).all(), e
            assert (st["carry_type"] == ref["st"][:, 4]).all() and (st["carry_aux"] == ref["st"][:, 7]).all()
    assert ref["reward"].max() > 0 and ref["terminated"].sum() > 10


@pytest.mark.parametrize("N", [1, 63, 65, 100, 1000])
def test_ragged_batch_sizes(N):
    """N not a multiple of the CTA size / of 16 (the image block of the last CTA is then not 16-byte aligned)."""
    eid, T = "MiniGrid-DoorKey-8x8-v0", 50
    lays = L.generate_many(eid, range(min(N, 64)))
    lays = [lays[i % len(lays)] for i in range(N)]
    acts = np.random.default_rng(N).integers(0, 7, (T, N)).astype(np.uint8)
    ref = oracle_rollout(lays, acts)
    env = _engine()(eid, N, layouts=lays, autoreset="off")
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    _compare(obs, r, te, tr, ref, f"N={N}")


def test_set_state_roundtrip_and_host_api():
    """mg_get_state / mg_set_state and the host-buffer entry point mg_step_many_host."""
    import ctypes as C
    from minigrid_dynamicprogramming_b200 import _capi
    eid, N, T = "MiniGrid-KeyCorridorS3R3-v0", 128, 40
    a, b = _engine()(eid, N, autoreset="off"), _engine()(eid, N, seeds=range(1000, 1000 + N), autoreset="off")
    a.reset(), b.reset()
    acts = np.random.default_rng(7).integers(0, 7, (T, N)).astype(np.uint8)
    a.step_many(torch.from_numpy(acts[:20]).cuda())
    st, cells = a.get_state()
    b.set_state(st, cells)                                   # b continues where a is
    oa = a.step_many(torch.from_numpy(acts[20:]).cuda())
    img = np.zeros((20, N, 147), np.uint8); d = np.zeros((20, N), np.uint8); r = np.zeros((20, N), np.float64)
    te = np.zeros((20, N), np.uint8); tr = np.zeros((20, N), np.uint8)
    h_acts = np.ascontiguousarray(acts[20:])
    _capi.check(_capi.lib.mg_step_many_host(b._handle, h_acts.ctypes.data, 20, img.ctypes.data, d.ctypes.data,
                                            r.ctypes.data, te.ctypes.data, tr.ctypes.data), "mg_step_many_host")
    assert (img.reshape(20, N, 7, 7, 3) == _np(oa[0]["image"])).all() and (r == _np(oa[1])).all()
    assert (d == _np(oa[0]["direction"])).all() and (te == _np(oa[2])).all() and (tr == _np(oa[3])).all()


def test_upload_rejects_unrepresentable_cells():
    from minigrid_dynamicprogramming_b200._capi import MgError
    lay = L.generate("MiniGrid-Empty-8x8-v0", 0)
    lay.cells = lay.cells.copy()
    lay.cells[2, 9] = 1                    # a non-door with a state byte: not representable, must be refused loudly
    with pytest.raises(MgError):
        _engine()("x", 4, layouts=[lay])


def test_debug_surface_hash_pprint_agent_sees():
    """BatchedMiniGrid.hash / pprint_grid / agent_sees on engine state against the reference's values (debug.json)."""
    import json
    import os
    from helpers import GOLDEN
    with open(os.path.join(GOLDEN, "debug.json")) as f:
        gold = json.load(f)
    for eid in ("MiniGrid-DoorKey-8x8-v0", "MiniGrid-KeyCorridorS3R3-v0"):
        s = gold[eid][0]
        env = _engine()(eid, 2, seeds=[0, 1], autoreset="off")
        env.reset()
        assert env.hash(0) == s["hash"] and env.pprint_grid(0) == s["pprint"]
        for x, y, want in s["sees"]:
            assert env.agent_sees(0, x, y) == want, (eid, x, y)


@pytest.mark.parametrize("eid", ["MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0", "MiniGrid-Dynamic-Obstacles-6x6-v0"])
def test_next_step_autoreset_vs_oracle(eid):
    """gymnasium >= 1.0's autoreset mode: the step after an episode ends ignores its action and resets (reward 0, flags
    clear).  Run as two launches so that the owed reset crosses a launch boundary (AgentState.pad)."""
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    N, T = 256, 300
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(21).choice(np.array([0, 1, 2, 2, 2], np.uint8), size=(T, N))
    ref = oracle_rollout(lays, acts, autoreset=2, pool=lays)
    env = BatchedMiniGrid(eid, N, layouts=lays, autoreset="next_step")
    env.reset()
    a = torch.from_numpy(acts).cuda()
    outs = []
    for lo, hi in ((0, 137), (137, T)):
        obs, r, te, tr, _ = env.step_many(a[lo:hi])
        outs.append((obs["image"].cpu().numpy(), r.cpu().numpy(), te.cpu().numpy(), tr.cpu().numpy()))
    img, r, te, tr = (np.concatenate([o[i] for o in outs]) for i in range(4))
    assert (img == ref["image"]).all() and (r == ref["reward"]).all()
    assert (te.astype(np.uint8) == ref["terminated"]).all() and (tr.astype(np.uint8) == ref["truncated"]).all()
    done = ref["terminated"] | ref["truncated"]
    assert done.sum() > 100 and not (done[1:] & done[:-1]).any()      # the step after a done step is the reset step


def test_final_observation_same_step_autoreset():
    """info["final_observation"]: the finished episode's last obs where an env was reset inside the step."""
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    eid, N, T = "MiniGrid-LavaCrossingS11N5-v0", 256, 200
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(22).choice(np.array([0, 1, 2, 2, 2], np.uint8), size=(T, N))
    ref = oracle_rollout(lays, acts, autoreset=True, pool=lays, final_obs=True)
    env = BatchedMiniGrid(eid, N, layouts=lays, autoreset="same_step", final_obs=True)
    env.reset()
    obs, r, te, tr, info = env.step_many(torch.from_numpy(acts).cuda())
    m = info["_final_observation"].cpu().numpy()
    assert (m == (ref["terminated"] | ref["truncated"]).astype(bool)).all() and m.sum() > 100
    assert (obs["image"].cpu().numpy() == ref["image"]).all()
    assert (info["final_observation"]["image"].cpu().numpy()[m] == ref["final_image"][m]).all()
    obs, r, te, tr, info = env.step(torch.from_numpy(acts[0]).cuda())     # T = 1 path
    assert info["final_observation"]["image"].shape == (N, 7, 7, 3)


@pytest.mark.parametrize("eid", ["MiniGrid-DoorKey-16x16-v0", "MiniGrid-FourRooms-v0"])
@pytest.mark.parametrize("N", [1, 15, 17, 33, 100, 1000])
def test_ragged_batch_sizes_large_grids(eid, N):
    """The rows-only tile layouts (16x16: half-size staging buffer, two copy-out phases per step; 19x19: full staging) with
    batch sizes that leave the last CTA with a partial first or second half and unaligned image blocks."""
    T = 60
    lays = L.generate_many(eid, range(min(N, 48)))
    lays = [lays[i % len(lays)] for i in range(N)]
    acts = np.random.default_rng(N).choice(np.array([0, 1, 2, 2, 2, 3, 5], np.uint8), size=(T, N))
    for autoreset in (False, True):
        ref = oracle_rollout(lays, acts, autoreset=autoreset, pool=lays)
        env = _engine()(eid, N, layouts=lays, autoreset="same_step" if autoreset else "off")
        env.reset()
        obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
        _compare(obs, r, te, tr, ref, f"{eid} N={N} autoreset={autoreset}")
        b, t, sm = env.launch_shape()
        assert b >= (14 if "16x16" in eid else 10)            # the layout the engine picked keeps that many CTAs per SM


def test_attribute_surface_of_the_reference_env():
    """The attributes callers of MiniGridEnv read: spaces, reward_range, render(), grid.encode()."""
    eid = "MiniGrid-DoorKey-8x8-v0"
    env = _engine()(eid, 4, seeds=range(4), autoreset="off")
    env.reset()
    assert env.action_space.n == 7 and env.reward_range == (0, 1) and env.width == 8 and env.max_steps == 640
    assert env.observation_space["image"].shape == (7, 7, 3) and env.observation_space["image"].dtype == np.uint8
    assert env.observation_space["direction"].n == 4
    lay = L.generate(eid, 2)
    enc = env.grid_encode(2)
    assert enc.shape == (8, 8, 3) and (enc == lay.cells[:3].reshape(3, 8, 8).transpose(2, 1, 0)).all()
    fr = env.render()
    assert tuple(fr.shape) == (4, 256, 256, 3) and fr.dtype == torch.uint8
    assert torch.equal(fr, env.get_frame(highlight=True, tile_size=32))
    assert env.carrying == [None] * 4 and list(env.step_count) == [0] * 4 and env.mission[0] == lay.mission
