"""Fused multi-GPU solve (mgdp_solve_push) of the KeyCorridorS6R3 space: unconditional push vs changed-only push.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29613 tools/dp_fused_bench.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from minigrid_dynamicprogramming_b200 import dp

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
os.environ["NCCL_DEBUG"] = "WARN"
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", dp.ALL_ACTIONS, seed=0, device=dev, max_states=1 << 25)
solver = dp.FusedShardedSolver(m)
MODES = (("1", "0"),) if os.environ.get("MGDP_BENCH_QUICK") else (("0", "0"), ("1", "0"), ("0", "1"), ("1", "1"))
for filt, mcast in MODES:
    os.environ["MGDP_PUSH_FILTER"] = filt
    os.environ["MGDP_PUSH_MULTICAST"] = mcast
    best = None
    for _ in range(3):
        V, pi, k, ms = solver.solve()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = t.item() if best is None else min(best, t.item())
    if rank == 0:
        print(f"world={world} filter={filt} multicast={int(solver.multicast)}: {k} sweeps, {best:.3f} ms, {best / k * 1e3:.1f} us/sweep, "
              f"{m.S * k / best / 1e6:.1f} G backups/s, V0={float(V[0]):.6f}", flush=True)
dist.destroy_process_group()
