"""GPU parity for the env ids of the N3 row (44 ids: Unlock / UnlockPickup / BlockedUnlockPickup / Fetch / GoToDoor /
GoToObject / RedBlueDoors / Memory / PutNear / ObstructedMaze with their own terminal rules, DistShift, LavaGap, LockedRoom,
MultiRoom, Playground on the base step): the engine through the C ABI against the reference's goldens (tests/golden/n3.npz) and against
the oracle on batches with autoreset."""
import numpy as np
import pytest

from helpers import L, N3_IDS, N3B_IDS, O, golden, n3_golden, oracle_rollout, short

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
