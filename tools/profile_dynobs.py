"""Two launches of the EXT step kernel on DynamicObstacles-8x8 (for ncu):
    ncu --set full --import-source on --clock-control none -k regex:^step_many_kernel -c 3 -o gpurun_out/dyn python tools/profile_dynobs.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid

N = 65536
env = BatchedMiniGrid("MiniGrid-Dynamic-Obstacles-8x8-v0", N, pool_size=4096)
env.reset()
acts = torch.randint(0, 7, (32, N), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
for _ in range(2):
    env.step_many(acts)
torch.cuda.synchronize()
print("done")
