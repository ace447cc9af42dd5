"""GPU parity for the env ids of the N3 row (44 ids: Unlock / UnlockPickup / BlockedUnlockPickup / Fetch / GoToDoor /
GoToObject / RedBlueDoors / Memory / PutNear / ObstructedMaze with their own terminal rules, DistShift, LavaGap, LockedRoom,
MultiRoom, Playground on the base step): the engine through the C ABI against the reference's goldens (tests/golden/n3.npz) and against
the oracle on batches with autoreset."""
import numpy as np
import pytest

from helpers import DYN_IDS, L, N3_IDS, N3B_IDS, O, golden, n3_golden, oracle_rollout, short

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")



def _engine():
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    return BatchedMiniGrid


@pytest.mark.parametrize("eid", N3_IDS + N3B_IDS)
def test_n3_golden_rollouts(eid):
    z, p = n3_golden(eid), short(eid) + "/"
    acts = z[p + "traj/actions"]
    nres = int(z[p + "traj/reset_seed"].max()) + 1
    env = _engine()(eid, 1, seeds=range(max(nres, 1)), autoreset="off")
    obs, _ = env.reset()
    assert (obs["image"][0].cpu().numpy() == z[p + "traj/first_image"]).all()
    for t, a in enumerate(acts):
        obs, r, te, tr, _ = env.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))
        assert (obs["image"][0].cpu().numpy() == z[p + "traj/image"][t]).all(), t
        assert int(obs["direction"][0]) == z[p + "traj/direction"][t] and float(r[0]) == z[p + "traj/reward"][t], t
        assert bool(te[0]) == bool(z[p + "traj/terminated"][t]) and bool(tr[0]) == bool(z[p + "traj/truncated"][t]), t
        if z[p + "traj/reset_seed"][t] >= 0:
            env.reset_envs([int(z[p + "traj/reset_seed"][t])])
    seeds = sorted({int(k.split("/")[2][1:]) for k in z.files if k.startswith(p + "greedy/")})
    for s in seeds:                                   # successful episodes: the env's own terminal rule pays the fp64 reward
        g = p + f"greedy/s{s}/"
        env = _engine()(eid, 1, seeds=[s], autoreset="off")
        obs, _ = env.reset()
        assert (obs["image"][0].cpu().numpy() == z[g + "image"][0]).all()
        a = torch.from_numpy(z[g + "actions"])[:, None].cuda()
        obs, r, te, tr, _ = env.step_many(a)
        assert (obs["image"][:, 0].cpu().numpy() == z[g + "image"][1:]).all()
        assert (r[:, 0].cpu().numpy() == z[g + "reward"]).all() and (te[:, 0].cpu().numpy() == z[g + "terminated"].astype(bool)).all()
        assert float(r[-1, 0]) > 0


@pytest.mark.parametrize("eid", N3_IDS[:3] + ["MiniGrid-LavaGapS6-v0", "MiniGrid-Fetch-8x8-N3-v0", "MiniGrid-GoToDoor-8x8-v0",
                                              "MiniGrid-GoToObject-8x8-N2-v0", "MiniGrid-RedBlueDoors-8x8-v0",
                                              "MiniGrid-ObstructedMaze-1Dlhb-v0", "MiniGrid-ObstructedMaze-Full-v1", "MiniGrid-LockedRoom-v0",
                                              "MiniGrid-MultiRoom-N6-v0", "MiniGrid-Playground-v0", "MiniGrid-MemoryS13Random-v0",
                                              "MiniGrid-MemoryS7-v0", "MiniGrid-PutNear-6x6-N2-v0", "MiniGrid-PutNear-8x8-N3-v0"])
@pytest.mark.parametrize("autoreset", ["off", "same_step"])
def test_n3_batch_vs_oracle(eid, autoreset):
    """192 envs x 500 steps biased towards forward / pickup / toggle so that doors get unlocked and boxes picked up."""
    N, T = 192, 500
    lays = L.generate_many(eid, range(N))
    rng = np.random.default_rng(11)
    acts = rng.choice(np.array([0, 1, 2, 2, 2, 3, 3, 4, 5, 5, 6], np.uint8), size=(T, N))
    ref = oracle_rollout(lays, acts, autoreset=autoreset == "same_step", pool=lays)
    env = _engine()(eid, N, layouts=lays, autoreset=autoreset)
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    assert (obs["image"].cpu().numpy() == ref["image"]).all()
    assert (r.cpu().numpy() == ref["reward"]).all()
    assert (te.cpu().numpy().astype(np.uint8) == ref["terminated"]).all() and (tr.cpu().numpy().astype(np.uint8) == ref["truncated"]).all()
    if eid in ("MiniGrid-Unlock-v0", "MiniGrid-UnlockPickup-v0"):          # max_steps 288 < T: truncation + autoreset exercised
        assert ref["truncated"].sum() > 0


def test_n3_unlock_mdp_and_finite_horizon():
    """Unlock's rule in the GPU enumeration (a toggle while the mission's door is open is terminal): successor table ==
    oracle BFS through step(); VI bit-exact; finite-horizon V_0(s0) == the reference's reward for the golden episode."""
    from minigrid_dynamicprogramming_b200 import dp
    eid, acts = "MiniGrid-Unlock-v0", tuple(range(7))
    z = golden("n3.npz")
    for seed in range(4):
        lay = L.generate(eid, seed)
        m = dp.enumerate_mdp(lay, acts)
        cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
        st = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
        succ, info = O.enumerate_mdp(lay.cells, cfg, st, acts, 1 << 20)
        assert m.S == len(succ) and (m.succ_sa() == succ).all(), seed
        V, pi, k = dp.value_iteration(m)
        Vo, pio, ko, _ = O.value_iteration(succ)
        assert k == ko and (V.cpu().numpy() == Vo).all() and (pi.cpu().numpy() == pio).all()
        assert float(dp.finite_horizon(m, lay.max_steps)[0]) == z[f"Unlock/greedy/s{seed}/reward"][-1]


# ---------------------------------------------------------------- DynamicObstacles: PCG64 draws inside the step kernel

@pytest.mark.parametrize("eid", DYN_IDS)
def test_dynamic_obstacles_golden(eid):
    """The reference's recorded rollout (tests/golden/dynobs.npz): obs, reward (-1 on a bump), flags AND, after every step,
    the full grid, the generator state and the obstacle list the reference held."""
    z, p = golden("dynobs.npz"), short(eid) + "/"
    acts = z[p + "traj/actions"]
    nres = int(z[p + "traj/reset_seed"].max()) + 1
    env = _engine()(eid, 1, seeds=range(max(nres, 1)), autoreset="off")
    obs, _ = env.reset()
    assert (obs["image"][0].cpu().numpy() == z[p + "traj/first_image"]).all()
    n = env._pool_layouts[0].width * env._pool_layouts[0].height
    for t, a in enumerate(acts):
        obs, r, te, tr, _ = env.step(torch.tensor([a], dtype=torch.uint8, device="cuda"))
        assert (obs["image"][0].cpu().numpy() == z[p + "traj/image"][t]).all(), t
        assert int(obs["direction"][0]) == z[p + "traj/direction"][t] and float(r[0]) == z[p + "traj/reward"][t], t
        assert bool(te[0]) == bool(z[p + "traj/terminated"][t]) and bool(tr[0]) == bool(z[p + "traj/truncated"][t]), t
        if t % 8 == 0 or z[p + "traj/reset_seed"][t] >= 0:
            assert (env.get_state()[1][0][:, :n] == z[p + "traj/grid"][t]).all(), t
            assert (env.get_dyn()[0] == z[p + "traj/dynrec"][t]).all(), t
        if z[p + "traj/reset_seed"][t] >= 0:
            env.reset_envs([int(z[p + "traj/reset_seed"][t])])


@pytest.mark.parametrize("eid", ["MiniGrid-Dynamic-Obstacles-Random-6x6-v0", "MiniGrid-Dynamic-Obstacles-8x8-v0",
                                 "MiniGrid-Dynamic-Obstacles-16x16-v0"])
@pytest.mark.parametrize("autoreset", ["off", "same_step"])
def test_dynamic_obstacles_batch_vs_oracle(eid, autoreset):
    """192 envs x 400 steps, all 7 action values (>= 3 act as `left`); same-step autoreset reloads generator + obstacles."""
    N, T = 192, 400
    lays = L.generate_many(eid, range(N))
    acts = np.random.default_rng(5).choice(np.array([0, 1, 2, 2, 2, 2, 3, 6], np.uint8), size=(T, N))
    ref = oracle_rollout(lays, acts, autoreset=autoreset == "same_step", pool=lays)
    env = _engine()(eid, N, layouts=lays, autoreset=autoreset)
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    assert (obs["image"].cpu().numpy() == ref["image"]).all()
    assert (r.cpu().numpy() == ref["reward"]).all() and (ref["reward"] == -1).sum() > 50
    assert (te.cpu().numpy().astype(np.uint8) == ref["terminated"]).all() and (tr.cpu().numpy().astype(np.uint8) == ref["truncated"]).all()
    assert (env.get_dyn() == ref["dyn"]).all()                       # generator states and obstacle lists after T steps
    n = lays[0].width * lays[0].height
    assert (env.get_state()[1][:, :, :n].reshape(N, -1) == ref["cells"]).all()


def test_dynamic_obstacles_mixed_with_other_classes():
    """A pool that mixes DynamicObstacles-8x8 with DoorKey-8x8 layouts (same grid size, different rules) under autoreset.
    300 steps: every env ends up on a DoorKey layout (its episodes outlast the run); 24 steps: both classes are still running
    when the launch ends — the state left in HBM (obstacle moves reach it at the end of a launch; envs change class on reset) is
    compared too: grids and generator records."""
    N, P = 128, 192                   # (layout_id + reset_stride) % P walks through slots of both classes
    pool = [L.generate("MiniGrid-Dynamic-Obstacles-8x8-v0" if i % 3 else "MiniGrid-DoorKey-8x8-v0", i) for i in range(P)]
    for T in (300, 24):
        acts = np.random.default_rng(6).integers(0, 7, (T, N)).astype(np.uint8)
        ref = oracle_rollout(pool[:N], acts, autoreset=True, pool=pool)
        env = _engine()(["MiniGrid-DoorKey-8x8-v0"], N, layouts=pool)
        env.reset()
        obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
        assert (obs["image"].cpu().numpy() == ref["image"]).all() and (r.cpu().numpy() == ref["reward"]).all()
        assert (te.cpu().numpy().astype(np.uint8) == ref["terminated"]).all()
        got, want = env.get_dyn(), ref["dyn"].copy()
        isdyn = want[:, 40] > 0                                      # n_obstacles: the env currently runs a DynamicObstacles layout
        if T == 24:
            assert isdyn.sum() > 20 and (~isdyn).sum() > 20
        want[~isdyn, 16] = 1         # (the engine gives the unused records of other classes an odd PCG64 increment, the oracle zeros)
        assert (got == want).all()
        assert (env.get_state()[1][:, :, :64].reshape(N, -1) == ref["cells"]).all()


def test_dynamic_obstacles_launch_lengths_agree():
    """The same 96 steps as one launch, as launches of 7 steps (+ a ragged last one) and as single steps: obstacle moves are kept
    in shared memory during a launch and written to HBM when it ends, so every split must leave the same grids, generator
    records and observations behind — also when render / get_state look at HBM between launches."""
    eid, N, T = "MiniGrid-Dynamic-Obstacles-8x8-v0", 100, 96
    lays = L.generate_many(eid, range(N))
    acts = torch.from_numpy(np.random.default_rng(11).choice(np.array([0, 1, 2, 2, 2], np.uint8), size=(T, N))).cuda()
    outs = []
    for chunk in (T, 7, 1):
        env = _engine()(eid, N, layouts=lays, autoreset="same_step")
        env.reset()
        imgs, states = [], []
        for t0 in range(0, T, chunk):
            obs, r, te, tr, _ = env.step_many(acts[t0:t0 + chunk])
            imgs.append(obs["image"].cpu().numpy().copy())
            if chunk == 7:
                states.append(env.get_state()[1].copy())          # HBM planes between launches
        outs.append((np.concatenate(imgs), env.get_dyn().copy(), env.get_state()[1].copy(), states))
    for o in outs[1:]:
        assert (o[0] == outs[0][0]).all() and (o[1] == outs[0][1]).all() and (o[2] == outs[0][2]).all()
    # planes between launches show exactly the obstacles the next observation is built from: 4 balls per env, always
    for st in outs[1][3]:
        assert ((st[:, 0, :64] == 6).sum(axis=1) == 4).all()


def test_dynamic_obstacles_needs_records():
    """A MG_RULE_DYNOBS class without mg_upload_dyn is a call-order error, not a silent no-move."""
    import dataclasses
    from minigrid_dynamicprogramming_b200 import _capi
    lay = dataclasses.replace(L.generate("MiniGrid-Dynamic-Obstacles-5x5-v0", 0), dyn=None)
    env = _engine()("MiniGrid-Dynamic-Obstacles-5x5-v0", 1, layouts=[lay], autoreset="off")
    env.reset()
    with pytest.raises(_capi.MgError):
        env.step(torch.zeros(1, dtype=torch.uint8, device="cuda"))


@pytest.mark.parametrize("eid", ["MiniGrid-Fetch-5x5-N2-v0", "MiniGrid-Fetch-8x8-N3-v0", "MiniGrid-GoToDoor-5x5-v0", "MiniGrid-GoToDoor-8x8-v0",
                                 "MiniGrid-RedBlueDoors-6x6-v0", "MiniGrid-RedBlueDoors-8x8-v0", "MiniGrid-MemoryS7-v0", "MiniGrid-MemoryS9-v0",
                                 "MiniGrid-ObstructedMaze-1Dl-v0", "MiniGrid-ObstructedMaze-1Dlh-v0", "MiniGrid-ObstructedMaze-1Dlhb-v0",
                                 "MiniGrid-PutNear-6x6-N2-v0", "MiniGrid-PutNear-8x8-N3-v0", "MiniGrid-GoToObject-6x6-N2-v0"])
def test_n3_rules_in_gpu_enumeration(eid):
    """Fetch / GoToDoor / GoToObject / RedBlueDoors / Memory / PutNear terminal rules and BOXES (hidden keys in ObstructedMaze,
    empty boxes in PutNear: an opened box becomes its contents or disappears) in the GPU enumeration (K3b): successor table == the oracle's
    BFS through step() with the same rule, VI bit-exact, and the finite-horizon V_0(s0) equals the reward the reference paid
    for the golden successful episode of that seed."""
    from minigrid_dynamicprogramming_b200 import dp
    acts = tuple(range(7))
    z, p = n3_golden(eid), short(eid) + "/"
    for seed in range(3):
        lay = L.generate(eid, seed)
        m = dp.enumerate_mdp(lay, acts, max_states=1 << 21)
        cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
        st = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
        succ, info = O.enumerate_mdp(lay.cells, cfg, st, acts, 1 << 21)
        assert m.S == len(succ) and (m.succ_sa() == succ).all(), (eid, seed)
        V, pi, k = dp.value_iteration(m)
        Vo, pio, ko, _ = O.value_iteration(succ)
        assert k == ko and (V.cpu().numpy() == Vo).all() and (pi.cpu().numpy() == pio).all()
        key = p + f"greedy/s{seed}/reward"
        if key in z.files:
            assert float(dp.finite_horizon(m, lay.max_steps)[0]) >= z[key][-1] > 0      # the golden episode is a lower bound...
            if float(Vo[0]) > 0:
                d = int(round(np.log(float(Vo[0])) / np.log(0.99))) + 1                 # ...and VI's shortest success path has its length
                assert len(z[p + f"greedy/s{seed}/actions"]) >= d
