"""Multi-GPU parity of the state-sharded value iteration (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tools/dp_sharded_check.py
Both exchange modes (NCCL all-gather; V' pushed by the sweep kernel over P2P) must reproduce the single-GPU solve bit for bit."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

from minigrid_dynamicprogramming_b200 import dp

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for eid, acts in (("MiniGrid-DoorKey-8x8-v0", dp.ALL_ACTIONS), ("MiniGrid-DoorKey-16x16-v0", dp.ALL_ACTIONS),
                  ("MiniGrid-KeyCorridorS3R3-v0", dp.ALL_ACTIONS)):
    m = dp.enumerate_mdp(eid, acts, seed=0, device=dev)
    V1, pi1, k1 = dp.value_iteration(m)
    for ex in ("allgather", "push", "fused"):
        V, pi, k = dp.value_iteration_sharded(m, exchange=ex)
        same = bool(k == k1 and torch.equal(V, V1) and torch.equal(pi, pi1))
        ok &= same
        if rank == 0:
            print(f"{eid} S={m.S} exchange={ex}: sweeps {k} (single {k1}) bit-exact={same}", flush=True)
# layout-sharded batched solve (no data-path collective): every rank's layouts equal their single-layout solves
from minigrid_dynamicprogramming_b200 import levelgen
lays = levelgen.generate_many("MiniGrid-DoorKey-8x8-v0", range(20, 20 + 4 * world + 1))
res = dp.value_iteration_batched(lays, None, dp.ALL_ACTIONS, device=dev, shard="layout")
lo, hi = res.layout_range
same, ks = True, []
for i in range(hi - lo):
    m1 = dp.enumerate_mdp(lays[lo + i], dp.ALL_ACTIONS, device=dev)
    V1, pi1, k1 = dp.value_iteration(m1)
    ks.append(k1)
    Vi, pii = res.layout(i)
    same &= bool(torch.equal(Vi, V1) and torch.equal(pii, pi1))
kmax = torch.tensor([max(ks) if ks else 0], dtype=torch.int32, device=dev)
dist.all_reduce(kmax, op=dist.ReduceOp.MAX)
same &= res.sweeps == int(kmax.item())
ok &= same
print(f"rank {rank}: layout-sharded batched VI, layouts [{lo}, {hi}) of {len(lays)}: sweeps {res.sweeps}, bit-exact={same}", flush=True)
t = torch.tensor([1.0 if ok else 0.0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
if rank == 0:
    print("SHARDED_DP_OK" if t.item() > 0 else "SHARDED_DP_MISMATCH", flush=True)
sys.exit(0 if t.item() > 0 else 1)
