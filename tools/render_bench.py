"""Frame-render throughput: mg_render_pov / mg_render_full at 65,536 DoorKey-8x8 envs.
    python tools/render_bench.py [N] [env_id]
"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid

N = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
eid = sys.argv[2] if len(sys.argv) > 2 else "MiniGrid-DoorKey-8x8-v0"
env = BatchedMiniGrid(eid, N, pool_size=4096)
env.reset()
acts = torch.randint(0, 7, (64, N), dtype=torch.uint8, device="cuda", generator=torch.Generator("cuda").manual_seed(0))
env.step_many(acts)


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        fn()
        ev[i + 1].record()
    torch.cuda.synchronize()
    return np.median([ev[i].elapsed_time(ev[i + 1]) for i in range(reps)]) * 1e-3


SIZES = tuple(int(x) for x in os.environ.get('RB_TS', '8,16,32').split(','))
for ts in SIZES:
    for pov in (True, False):
        try:
            out = env.get_frame(agent_pov=pov, tile_size=ts)
        except Exception as e:  # noqa: BLE001
            print(ts, pov, "skipped:", e)
            continue
        nbytes = out.numel()
        dt = timeit(lambda: env.get_frame(agent_pov=pov, tile_size=ts, out=out if pov else None))
        print(f"ts={ts:3d} {'pov ' if pov else 'full'}  {dt*1e6:9.1f} us  {nbytes/1e6:9.1f} MB  {nbytes/dt/1e9:8.1f} GB/s  "
              f"{N/dt/1e6:8.1f} M frames/s", flush=True)
        del out
        torch.cuda.empty_cache()
