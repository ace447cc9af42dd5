import torch, time, sys
sys.path.insert(0, '.')
from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
for N in (16384, 32768, 49152, 65536, 131072):
    env = BatchedMiniGrid("MiniGrid-DoorKey-8x8-v0", N, seeds=range(4096), autoreset="same_step")
    env.reset()
    T = 32
    a = torch.randint(0, 7, (T, N), dtype=torch.uint8, device="cuda")
    for _ in range(3): env.step_many(a)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): env.step_many(a)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 20
    print(N, env.launch_shape(), round(ms * 1e3, 1), "us", round(N * T / ms / 1e6, 2), "G steps/s")
    env.close()
