import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np, torch
from helpers import L, O, oracle_rollout
from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
for eid, N, T in (("MiniGrid-Unlock-v0", 192, 500), ("MiniGrid-DoorKey-8x8-v0", 256, 700), ("MiniGrid-Empty-8x8-v0", 64, 300)):
    lays = L.generate_many(eid, range(N))
    rng = np.random.default_rng(11)
    acts = rng.choice(np.array([0, 1, 2, 2, 2, 3, 3, 4, 5, 5, 6], np.uint8), size=(T, N))
    ref = oracle_rollout(lays, acts, autoreset=True, pool=lays)
    env = BatchedMiniGrid(eid, N, layouts=lays, autoreset="same_step")
    print(eid, "stride", env.reset_stride, O.reset_stride(N, N))
    env.reset()
    obs, r, te, tr, _ = env.step_many(torch.from_numpy(acts).cuda())
    img = obs["image"].cpu().numpy(); r = r.cpu().numpy(); te = te.cpu().numpy(); tr = tr.cpu().numpy()
    bad = (img != ref["image"]).reshape(T, N, -1).any(-1)
    print(" image mismatches", bad.sum(), "reward", (r != ref["reward"]).sum(), "term", (te != ref["terminated"].astype(bool)).sum(), "trunc", (tr != ref["truncated"].astype(bool)).sum())
    if bad.any():
        t, e = np.argwhere(bad)[0]
        done = ref["terminated"] | ref["truncated"]
        print(" first mismatch t", t, "env", e, "resets of env before/at t:", np.flatnonzero(done[:t + 1, e]))
        print(" envs mismatching at t:", np.flatnonzero(bad[t])[:40])
        print(" got\n", img[t, e, :, :, 0].T, "\n want\n", ref["image"][t, e, :, :, 0].T)
        print(" dir got", obs["direction"][t, e].item(), "want", ref["direction"][t, e])
