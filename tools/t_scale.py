"""Throughput of mg_step_many vs T (env-steps per launch) at the headline config, incl. a CUDA-graph replay of T=1."""
import sys, torch
sys.path.insert(0, '.')
from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
N = 65536
env = BatchedMiniGrid("MiniGrid-DoorKey-8x8-v0", N, seeds=range(4096), autoreset="same_step")
env.reset()
for T in (1, 2, 4, 8, 32, 128):
    a = torch.randint(0, 7, (T, N), dtype=torch.uint8, device="cuda")
    for _ in range(3): env.step_many(a)
    torch.cuda.synchronize()
    reps = max(10, 256 // T)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): env.step_many(a)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"T={T:4d}  {ms*1e3:8.1f} us/launch  {N*T/ms/1e6:7.2f} G env-steps/s")
# T=1 via step() captured in a CUDA graph
a = torch.randint(0, 7, (N,), dtype=torch.uint8, device="cuda")
env.step(a); torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
s = torch.cuda.Stream()
with torch.cuda.stream(s):
    env.step(a)
    with torch.cuda.graph(g, stream=s):
        for _ in range(16): env.step(a)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20): g.replay()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / (20 * 16)
print(f"T=1 in a CUDA graph (16 steps per replay): {ms*1e3:.1f} us/step  {N/ms/1e6:.2f} G env-steps/s")
