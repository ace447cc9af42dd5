"""Two launches of the EXT-less step kernel on the mixed LavaCrossingS11N5 + FourRooms pool (config 4's per-GPU share), for ncu:
    ncu --set full --import-source on --clock-control none -k regex:^step_many_kernel -c 3 -o gpurun_out/mixed python tools/profile_mixed.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid

N, POOL = 131072, 4096
env = BatchedMiniGrid(["MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0"], N, seeds=[i // 2 for i in range(POOL)], autoreset="same_step")
env.reset_envs(torch.arange(N, device="cuda") % POOL)
acts = torch.randint(0, 7, (32, N), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
for _ in range(2):
    env.step_many(acts)
torch.cuda.synchronize()
print("done")
