"""GPU parity for the DP path: K3b enumeration and K4 Bellman sweeps against the oracle and the reference BFS goldens."""
import numpy as np
import pytest

from helpers import L, O, golden, short

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")

CASES = [("MiniGrid-Empty-8x8-v0", (0, 1, 2)), ("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5)),
         ("MiniGrid-DoorKey-8x8-v0", tuple(range(7))), ("MiniGrid-DoorKey-16x16-v0", (0, 1, 2, 3, 5)),
         ("MiniGrid-DoorKey-16x16-v0", tuple(range(7))), ("MiniGrid-LavaCrossingS11N5-v0", (0, 1, 2)),
         ("MiniGrid-FourRooms-v0", (0, 1, 2)), ("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7))),
         ("MiniGrid-KeyCorridorS3R1-v0", tuple(range(7)))]


def _dp():
    from minigrid_dynamicprogramming_b200 import dp
    return dp


def _oracle_mdp(lay, acts, max_states=1 << 20):
    cfg = O.make_cfg(lay.width, lay.height, lay.max_steps, lay.see_through_walls, lay.rule)
    st = np.array([*lay.agent_pos, lay.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
    return O.enumerate_mdp(lay.cells, cfg, st, acts, max_states)


@pytest.mark.parametrize("eid,acts", CASES)
def test_enumerate_and_solve_vs_oracle(eid, acts):
    dp = _dp()
    for seed in (0, 3):
        lay = L.generate(eid, seed)
        succ, info = _oracle_mdp(lay, acts)
        m = dp.enumerate_mdp(lay, acts, max_states=1 << 20)
        assert m.S == len(succ), (eid, seed, m.S, len(succ))
        assert (m.succ_sa() == succ).all(), (eid, seed)
        assert (m.info.cpu().numpy()[:, :7] == info[:, :7]).all()
        Vo, pio, ko, hist = O.value_iteration(succ)
        for method in ("cooperative", "launch"):
            V, pi, k = dp.value_iteration(m, method=method)
            assert k == ko, (eid, method, k, ko)
            assert (V.cpu().numpy() == Vo).all() and (pi.cpu().numpy() == pio).all(), (eid, method)
        Vp, kp = dp.policy_evaluation(m, pi)
        Vpo, _, kpo, _ = O.value_iteration(succ, pi_fixed=pio)
        assert kp == kpo and (Vp.cpu().numpy() == Vpo).all()


def test_masked_sweep_equals_unmasked():
    """The row mask (mgdp_rowmask) only skips rows whose 32 successors are all self-loops: mask bytes == numpy's, and V, pi,
    sweep counts of the masked solve / sweeps are bit-identical to the unmasked ones (both methods), also on a concatenated
    table and on a table whose size is not a multiple of the tile."""
    dp = _dp()
    for eid, acts in (("MiniGrid-DoorKey-8x8-v0", tuple(range(7))), ("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7))),
                      ("MiniGrid-DoorKey-16x16-v0", (0, 1, 2, 3, 5)), ("MiniGrid-Empty-8x8-v0", (0, 1, 2))):
        m = dp.enumerate_mdp(eid, acts, seed=2)
        succ = m.succ_sa()
        S, A = succ.shape
        ns = np.zeros((m.S_pad, A), bool)
        ns[:S] = succ != np.arange(S)[:, None]
        want = (ns.reshape(-1, 32, A).any(1) * (1 << np.arange(A))).sum(1).astype(np.uint8)
        assert (m.rowmask.cpu().numpy() == want).all(), eid
        assert 8 < m.bytes_per_backup() <= (4 * A + 8) * m.S_pad / S + 1 / 32          # whole 32-state rows are read
        for method in ("cooperative", "launch"):
            V0, pi0, k0 = dp.value_iteration(m, method=method, masked=False)
            V1, pi1, k1 = dp.value_iteration(m, method=method, masked=True)
            assert k0 == k1 and torch.equal(V0, V1) and torch.equal(pi0, pi1), (eid, method)
    big = dp.concat_mdps([dp.enumerate_mdp("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5), seed=s) for s in range(5)])
    V0, pi0, k0 = dp.value_iteration(big, masked=False)
    V1, pi1, k1 = dp.value_iteration(big, masked=True)
    assert k0 == k1 and torch.equal(V0, V1) and torch.equal(pi0, pi1)


def test_enumerate_matches_reference_bfs_goldens():
    """Successor tables enumerated through the reference's own step() (tests/golden/mdp_reference.npz)."""
    dp = _dp()
    z = golden("mdp_reference.npz")
    for key in [k for k in z.files if k.endswith("/succ")]:
        name, a, _ = key.split("/")
        acts = tuple(int(x) for x in z[f"{name}/{a}/acts"])
        m = dp.enumerate_mdp(f"MiniGrid-{name}-v0", acts, seed=0)
        assert m.succ_sa().shape == z[key].shape and (m.succ_sa() == z[key]).all(), key


def test_vi_survey_anchors():
    """sweeps to stop and V(s0) = repeated fp32 product 0.99^(d-1)."""
    dp = _dp()
    for eid, acts, S, sweeps, d in (("MiniGrid-Empty-8x8-v0", (0, 1, 2), 140, 13, 11),
                                    ("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5), 424, 24, 17),
                                    ("MiniGrid-DoorKey-16x16-v0", (0, 1, 2, 3, 5), 2680, 44, 29)):
        m = dp.enumerate_mdp(eid, acts, seed=0)
        V, pi, k = dp.value_iteration(m)
        v = np.float32(1.0)
        for _ in range(d - 1):
            v = np.float32(v * np.float32(0.99))
        assert (m.S, k) == (S, sweeps) and V[0].item() == v


def test_greedy_policy_rollout_through_engine():
    """The extracted policy, rolled out by the step kernel, reaches the goal in d steps with reward 1-0.9*d/max_steps."""
    dp = _dp()
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    for eid, acts in (("MiniGrid-DoorKey-8x8-v0", tuple(range(7))), ("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7)))):
        lay = L.generate(eid, 1)
        m = dp.enumerate_mdp(lay, acts)
        V, pi, k = dp.value_iteration(m)
        succ = m.succ_sa()
        pi_h = pi.cpu().numpy()
        s, seq = 0, []
        while s >= 0:
            seq.append(acts[pi_h[s]])
            s = succ[s, pi_h[s]]
        assert s == -1
        env = BatchedMiniGrid(eid, 1, layouts=[lay], autoreset="off")
        env.reset()
        obs, r, te, tr, _ = env.step_many(torch.tensor(seq, dtype=torch.uint8, device="cuda")[:, None])
        d = len(seq)
        assert bool(te[d - 1, 0]) and not te[: d - 1].any()
        assert r[d - 1, 0].item() == 1 - 0.9 * (d / lay.max_steps)
        v = np.float32(1.0)
        for _ in range(d - 1):
            v = np.float32(v * np.float32(0.99))
        assert V[0].item() == v


def test_batched_layouts_equal_individual():
    dp = _dp()
    acts = (0, 1, 2, 3, 5)
    ms = [dp.enumerate_mdp("MiniGrid-DoorKey-8x8-v0", acts, seed=s) for s in range(12)]
    big = dp.concat_mdps(ms)
    V, pi, k = dp.value_iteration(big)
    ks = []
    for i, m in enumerate(ms):
        Vi, pii, ki = dp.value_iteration(m)
        ks.append(ki)
        a, b = big.offsets[i], big.offsets[i + 1]
        assert torch.equal(V[a:b], Vi) and torch.equal(pi[a:b], pii)
    assert k == max(ks)


def test_streamed_solve_equals_cooperative():
    """mgdp_solve_stream (one gated launch per sweep, stop flag raised by the last CTA of the converging sweep, host one chunk
    of 8 launches behind) == mgdp_solve / mgdp_solve_masked bit for bit: V, pi, sweep count — for every action-count
    instantiation, with and without the row mask, when max_sweeps cuts the solve short (inside a chunk, at a chunk boundary,
    at one sweep), and twice in a row on one device (the control words are cleared by the call)."""
    dp = _dp()
    for eid, acts in (("MiniGrid-DoorKey-8x8-v0", tuple(range(7))), ("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5)),
                      ("MiniGrid-Empty-8x8-v0", (0, 1, 2)), ("MiniGrid-KeyCorridorS3R3-v0", (0, 1, 2, 3, 4, 5)),
                      ("MiniGrid-DoorKey-16x16-v0", tuple(range(7)))):
        m = dp.enumerate_mdp(eid, acts, seed=3)
        for masked in (False, True):
            V0, pi0, k0 = dp.value_iteration(m, method="cooperative", masked=masked)
            for _ in range(2):
                V1, pi1, k1 = dp.value_iteration(m, method="stream", masked=masked)
                assert k0 == k1 and k0 > 8 and torch.equal(V0, V1) and torch.equal(pi0, pi1), (eid, masked, k0, k1)
        k_full = k0
        for cap in (1, 5, 8, 9, 16):
            V0, pi0, k0 = dp.value_iteration(m, max_sweeps=cap, method="cooperative")
            V1, pi1, k1 = dp.value_iteration(m, max_sweeps=cap, method="stream")
            assert k0 == k1 == min(cap, k_full) and torch.equal(V0, V1) and torch.equal(pi0, pi1), (eid, cap)
    with pytest.raises(ValueError):
        dp.value_iteration(m, method="stream", pi_fixed=pi0)


def test_keycorridor_s6r3_full_space():
    """BASELINE config 5: all-7-action KeyCorridorS6R3 (16.8 M states) enumerated and solved on the GPU, for every recorded
    seed; pinned by the oracle solves behind tests/golden/greedy_KeyCorridorS6R3.npz (S, sweep count, V(s0)) and by the
    greedy action sequences that were stepped in the reference env."""
    dp = _dp()
    z = golden("greedy_KeyCorridorS6R3.npz")
    acts = tuple(range(7))
    seeds = sorted({int(k.split("/")[0][1:]) for k in z.files})
    for seed in seeds:
        V0, sweeps, S = z[f"s{seed}/V0_sweeps_S"]
        m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", acts, seed=seed, max_states=1 << 25)
        assert m.S == int(S), (seed, m.S, S)
        V, pi, k = dp.value_iteration(m)
        assert k == int(sweeps) and V[0].item() == np.float32(V0), seed
        # greedy walk through the table reproduces the golden action sequence recorded in the reference env
        s, seq = 0, []
        while s >= 0:
            a = int(pi[s].item())
            seq.append(acts[a])
            s = int(m.succ[a, s].item())
        assert s == -1 and seq == [int(a) for a in z[f"s{seed}/actions"]], seed
        del m, V, pi


def test_keycorridor_s6r3_sampled_rows_vs_reference_step():
    """100,000 uniformly sampled rows of the 16.8 M-state table.  The file was recorded by
    rebuilding each sampled state inside the reference env and applying the reference's own step() for all 7 actions
    (oracle/gen_golden_big.py); ids are serial-BFS ids, which the GPU enumeration reproduces."""
    import os
    from helpers import GOLDEN
    if not os.path.exists(os.path.join(GOLDEN, "mdp_KeyCorridorS6R3_rows.npz")):
        pytest.skip("fixture not recorded")
    dp = _dp()
    z = golden("mdp_KeyCorridorS6R3_rows.npz")
    m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", tuple(range(7)), seed=0, max_states=1 << 25)
    assert m.S == int(z["S"])
    ids = torch.from_numpy(z["ids"].astype(np.int64)).cuda()
    assert len(ids) >= 100000
    rows = m.succ[:, ids].t().cpu().numpy()
    assert (rows == z["succ"]).all()
    assert (m.info[ids].cpu().numpy() == z["info"]).all()


def test_batched_enumeration_equals_individual():
    """mgdp_enumerate_batch: B layouts in one BFS; layout l's rows == its own enumeration shifted by offsets[l] (SURVEY §8.4
    config 3b), info records equal, and the solve over the concatenated table gives every layout its single-solve V and pi."""
    dp = _dp()
    for eid, acts, B in (("MiniGrid-DoorKey-8x8-v0", tuple(range(7)), 24), ("MiniGrid-DoorKey-16x16-v0", (0, 1, 2, 3, 5), 40),
                         ("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7)), 5), ("MiniGrid-Empty-8x8-v0", (0, 1, 2), 3)):
        lays = L.generate_many(eid, range(B))
        big = dp.enumerate_batch(lays, acts)
        assert len(big.offsets) == B + 1 and big.offsets[0] == 0 and big.offsets[-1] == big.S
        V, pi, k = dp.value_iteration(big)
        ks = []
        for i, lay in enumerate(lays):
            m = dp.enumerate_mdp(lay, acts)
            a, b = big.offsets[i], big.offsets[i + 1]
            assert b - a == m.S, (eid, i)
            want = m.succ[:, :m.S]
            want = torch.where(want >= 0, want + a, want)
            assert torch.equal(big.succ[:, a:b], want), (eid, i)
            assert torch.equal(big.info[a:b], m.info), (eid, i)
            Vi, pii, ki = dp.value_iteration(m)
            ks.append(ki)
            assert torch.equal(V[a:b], Vi) and torch.equal(pi[a:b], pii), (eid, i)
        assert k == max(ks)
    res = dp.value_iteration_batched("MiniGrid-DoorKey-8x8-v0", range(10, 20), (0, 1, 2, 3, 5))
    m = dp.enumerate_mdp("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5), seed=13)
    Vi, pii, _ = dp.value_iteration(m)
    assert torch.equal(res.layout(3)[0], Vi) and torch.equal(res.layout(3)[1], pii)


def _playground_sub_layout(keep_doors):
    """Playground seed 0 (12 movables, 12 doors, 19x19) with all but the `keep_doors` doors nearest to the agent walled up:
    the state space stays enumerable by the serial oracle while the key still carries 12 movable positions."""
    import dataclasses
    lay = L.generate("MiniGrid-Playground-v0", 0)
    cells = lay.cells.copy()
    doors = np.flatnonzero(cells[0] == 4)
    ax, ay = lay.agent_pos
    for d in sorted(doors, key=lambda d: abs(d % lay.width - ax) + abs(d // lay.width - ay))[keep_doors:]:
        cells[:, d] = (2, 5, 0, 0)
    return dataclasses.replace(lay, cells=cells)


@pytest.mark.parametrize("case", ["ObstructedMaze-2Dlhb", "ObstructedMaze-1Q", "ObstructedMaze-2Q", "Playground-sub", "Playground-pickup"])
def test_wide_key_enumeration_vs_oracle(case):
    """States that need more than 64 bits (2Dlhb / 1Q: 66, 2Q: 111, Playground: 156): multi-word keys, tables == the oracle's
    serial BFS through its own step, V / pi == the oracle's solve."""
    dp = _dp()
    from minigrid_dynamicprogramming_b200 import _capi
    acts = (0, 1, 2, 3, 5)
    if case == "Playground-sub":
        lay = _playground_sub_layout(4)
    elif case == "Playground-pickup":
        lay, acts = L.generate("MiniGrid-Playground-v0", 0), (0, 1, 2, 3)
    else:
        lay = L.generate(f"MiniGrid-{case}-v0", 0)
    succ, info = _oracle_mdp(lay, acts, 1 << 21)
    m = dp.enumerate_mdp(lay, acts, max_states=1 << 21)
    assert m.S == len(succ) and (m.succ_sa() == succ).all(), (case, m.S, len(succ))
    assert (m.info.cpu().numpy()[:, :7] == info[:, :7]).all()
    V, pi, k = dp.value_iteration(m)
    Vo, pio, ko, _ = O.value_iteration(succ)
    assert k == ko and (V.cpu().numpy() == Vo).all() and (pi.cpu().numpy() == pio).all()
    # the same layouts batched (layout id on top of a multi-word key)
    if case == "ObstructedMaze-2Dlhb":
        lays = L.generate_many("MiniGrid-ObstructedMaze-2Dlhb-v0", range(3))
        big = dp.enumerate_batch(lays, acts)
        a, b = big.offsets[0], big.offsets[1]
        assert b - a == m.S and torch.equal(big.succ[:, a:b], m.succ[:, :m.S])


def test_enumerate_reports_overflow_and_limits():
    """More reachable states than max_states -> MG_ENOMEM (also when the hash table fills up inside a level), never a hang."""
    dp = _dp()
    from minigrid_dynamicprogramming_b200 import _capi
    with pytest.raises(_capi.MgError, match="max_states"):
        dp.enumerate_mdp("MiniGrid-DoorKey-16x16-v0", tuple(range(7)), seed=0, max_states=1000)
    with pytest.raises(_capi.MgError, match="max_states"):
        dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", tuple(range(7)), seed=0, max_states=1 << 16)


@pytest.mark.parametrize("eid,acts", [("MiniGrid-DoorKey-8x8-v0", (0, 1, 2, 3, 5)), ("MiniGrid-Empty-8x8-v0", (0, 1, 2)),
                                      ("MiniGrid-LavaCrossingS11N5-v0", (0, 1, 2)), ("MiniGrid-KeyCorridorS3R3-v0", tuple(range(7)))])
def test_finite_horizon_dp(eid, acts):
    """Exact finite-horizon DP (time-dependent fp64 reward) against the numpy oracle — V_0 and the whole pi_t table
    bit-exact — and against the reference: V_0(s0) is the reward the reference paid at the end of the golden greedy
    (shortest-path) episode, 1 - 0.9 * d / max_steps."""
    from minigrid_dynamicprogramming_b200 import dp
    for seed in (0, 1):
        lay = L.generate(eid, seed)
        m = dp.enumerate_mdp(lay, acts)
        V0, pi = dp.finite_horizon(m, lay.max_steps, return_policy=True)
        Vo, pio = O.finite_horizon_numpy(m.succ_sa(), lay.max_steps, return_policy=True)
        assert np.array_equal(V0.cpu().numpy(), Vo) and np.array_equal(pi.cpu().numpy(), pio), (eid, seed)
        g = golden(f"greedy_{short(eid)}.npz")
        assert float(V0[0]) == g[f"s{seed}/reward"][-1], (eid, seed)
        # a horizon shorter than the solution length cannot succeed
        d = len(g[f"s{seed}/actions"])
        assert float(dp.finite_horizon(m, d - 1)[0]) == 0.0
        assert float(dp.finite_horizon(m, d)[0]) == 1 - 0.9 * (d / d)
        Vg = dp.finite_horizon(m, lay.max_steps, gamma=0.99)
        assert np.array_equal(Vg.cpu().numpy(), O.finite_horizon_numpy(m.succ_sa(), lay.max_steps, 0.99))


def test_sharded_vi_two_gpus_push_and_allgather():
    """2 ranks over NCCL: value_iteration_sharded with the fused push exchange (mgdp_sweep_push over symmetric memory) and
    with the all-gather baseline, both bit-exact against the single-GPU solve.  Skipped on a 1-GPU box."""
    import os
    import subprocess
    import sys
    import torch as T
    if T.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", "29617", os.path.join(root, "tools", "dp_sharded_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SHARDED_DP_OK" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
