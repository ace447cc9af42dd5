"""Small invocations of every hot kernel, for compute-sanitizer (tools/sanitize.sh):
    compute-sanitizer --tool memcheck  python tools/sanitize_targets.py step
    compute-sanitizer --tool racecheck python tools/sanitize_targets.py step
modes: step (all three shared-memory layouts x lean / EXT, ragged N = 100, T = 8, same-step autoreset), solve (sweep, cooperative
solve, policy evaluation, finite horizon, static solve), enum (single + batched + wide-key BFS), render (odd / even tile sizes),
wrappers."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from minigrid_dynamicprogramming_b200 import dp, levelgen as L
from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid

mode = sys.argv[1] if len(sys.argv) > 1 else "step"
rng = np.random.default_rng(0)

if mode == "step":
    for eid in ("MiniGrid-DoorKey-8x8-v0", "MiniGrid-DoorKey-16x16-v0", "MiniGrid-FourRooms-v0",            # lean: DUAL / ROWS_HALF / ROWS
                "MiniGrid-Unlock-v0", "MiniGrid-MemoryS17Random-v0", "MiniGrid-Dynamic-Obstacles-16x16-v0",   # EXT
                "MiniGrid-LockedRoom-v0", "MiniGrid-Dynamic-Obstacles-8x8-v0"):
        N, T = 100, 8
        env = BatchedMiniGrid(eid, N, pool_size=37, autoreset="same_step", max_steps=5)
        env.reset()
        for _ in range(2):
            env.step_many(torch.from_numpy(rng.integers(0, 7, (T, N)).astype(np.uint8)).cuda())
        env.step(torch.from_numpy(rng.integers(0, 7, N).astype(np.uint8)).cuda())
        torch.cuda.synchronize()
        print("step", eid, env.launch_shape(), flush=True)
        env.close()
    env = BatchedMiniGrid("MiniGrid-LavaCrossingS9N1-v0", 100, autoreset="next_step", max_steps=7)
    env.reset()
    env.step_many(torch.from_numpy(rng.integers(0, 3, (8, 100)).astype(np.uint8)).cuda())
    env = BatchedMiniGrid("MiniGrid-DoorKey-8x8-v0", 63, autoreset="same_step", final_obs=True, max_steps=6)
    env.reset()
    env.step_many(torch.from_numpy(rng.integers(0, 7, (8, 63)).astype(np.uint8)).cuda())
    torch.cuda.synchronize()
elif mode == "solve":
    import ctypes as C
    from minigrid_dynamicprogramming_b200 import _capi
    m = dp.enumerate_mdp("MiniGrid-DoorKey-8x8-v0", dp.ALL_ACTIONS, seed=0)
    for method in ("cooperative", "launch"):
        V, pi, k = dp.value_iteration(m, method=method)
    dp.policy_evaluation(m, pi)
    dp.finite_horizon(m, 40, return_policy=True)
    V2 = torch.zeros(m.S_pad, dtype=torch.float32, device="cuda")
    r2 = torch.zeros(2, dtype=torch.int32, device="cuda")
    _capi.check(_capi.lib.mgdp_solve_static(m.succ.data_ptr(), m.S_pad, m.A, m.S, V2.data_ptr(), None, None, 0.99, 1e-6, 1000,
                                            r2.data_ptr(), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "static")
    torch.cuda.synchronize()
    print("solve", m.S, k, int(r2[0]), flush=True)
elif mode == "enum":
    m = dp.enumerate_mdp("MiniGrid-KeyCorridorS3R3-v0", dp.ALL_ACTIONS, seed=0)
    b = dp.enumerate_batch(L.generate_many("MiniGrid-DoorKey-8x8-v0", range(9)), dp.ALL_ACTIONS)
    w = dp.enumerate_mdp("MiniGrid-ObstructedMaze-2Dlhb-v0", (0, 1, 2, 3, 5), seed=0)
    torch.cuda.synchronize()
    print("enum", m.S, b.S, w.S, flush=True)
elif mode == "render":
    env = BatchedMiniGrid("MiniGrid-KeyCorridorS3R3-v0", 37)
    env.reset()
    for ts in (5, 8, 16):
        env.get_frame(agent_pov=True, tile_size=ts)
        env.get_frame(agent_pov=False, tile_size=ts)
    torch.cuda.synchronize()
    print("render ok", flush=True)
elif mode == "wrappers":
    from minigrid_dynamicprogramming_b200 import wrappers as W
    base = BatchedMiniGrid("MiniGrid-DoorKey-8x8-v0", 37, autoreset="off")
    for wrap in (W.FullyObsWrapper, W.SymbolicObsWrapper, W.OneHotPartialObsWrapper, W.FlatObsWrapper, W.ActionBonus, W.PositionBonus):
        env = wrap(base)
        env.reset()
        env.step(torch.from_numpy(rng.integers(0, 7, 37).astype(np.uint8)).cuda())
    env = W.ViewSizeWrapper(base, agent_view_size=5)
    env.reset()
    torch.cuda.synchronize()
    print("wrappers ok", flush=True)
