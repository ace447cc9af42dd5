"""State-sharded fused solve (mgdp_solve_push) of the 16.8 M-state KeyCorridorS6R3 table on N GPUs: bit-exactness against the
single-GPU solve of the SAME table, time per sweep, and — with the instrumentation build — where a sweep's time goes.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29617 tools/dp_push_timing.py
    # per-sweep breakdown (%globaltimer stamps inside solve_push_kernel):
    python -m minigrid_dynamicprogramming_b200.build --force -DMGDP_TIMING=1 -o=variants/timing.so
    MGB200_LIB=$PWD/variants/timing.so python -m torch.distributed.run ... tools/dp_push_timing.py
    # round 1's synchronisation protocol for comparison: build with -DMGDP_PUSH_SYNC=0 as well
Stamps per sweep k: [0] first CTA arrives  [1] last CTA arrives  [2] flags + residual published  [3] all peers' flags seen
[4] generation released.  Reported per rank as medians over sweeps 5..k-1 (us):
    work  = [1](k) - [4](k-1)      last CTA done after the previous release (tile loop incl. the peer stores)
    skew  = [1] - [0]              first to last CTA of this GPU
    pub   = [2] - [1]              sys-scope release of this GPU's stores + flags
    wait  = [3] - [2]              waiting for the slowest peer (its work + its publication)
    rel   = [4] - [3]              residual reduce + local release
"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

from minigrid_dynamicprogramming_b200 import _capi, dp

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
timing = hasattr(_capi.lib, "mgdp_debug_push_timing")
m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", dp.ALL_ACTIONS, seed=0, device=dev, max_states=1 << 25)
V1, pi1, k1 = dp.value_iteration(m)
solver = dp.FusedShardedSolver(m)
ok = True
modes = [None] if os.environ.get("MGDP_BENCH_QUICK") else ([None, "0", "1"] if world > 2 else [None, "0"])
for mcast in modes:
    if mcast is None:
        os.environ.pop("MGDP_PUSH_MULTICAST", None)
    else:
        os.environ["MGDP_PUSH_MULTICAST"] = mcast
    solver.solve()
    best = None
    for rep in range(3):
        if timing:
            assert _capi.lib.mgdp_debug_push_timing(None, 1) == 0
        V, pi, k, ms = solver.solve()
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        best = t.item() if best is None else min(best, t.item())
    same = bool(k == k1 and torch.equal(V, V1) and torch.equal(pi, pi1))
    ok &= same
    if rank == 0:
        print(f"world={world} multicast={'auto' if mcast is None else mcast} (used {int(solver.multicast)}): {k} sweeps (single GPU {k1}), "
              f"{best:.3f} ms, {best / k * 1e3:.1f} us/sweep, {m.S * k / best / 1e6:.1f} G backups/s, "
              f"bit_exact_vs_single_gpu(V, pi over {m.S} states)={same}", flush=True)
    if timing:
        tt = np.zeros((5, 128), np.uint64)
        _capi.lib.mgdp_debug_push_timing.argtypes = [C.c_void_p, C.c_int32]
        assert _capi.lib.mgdp_debug_push_timing(tt.ctypes.data, 0) == 0
        n = min(k, 128)
        a = tt[:, :n].astype(np.int64)
        work = a[1, 1:] - a[4, :-1]
        row = [np.median(x[5:]) / 1e3 for x in (a[4, 1:] - a[4, :-1], work, (a[1] - a[0])[1:], (a[2] - a[1])[1:], (a[3] - a[2])[1:], (a[4] - a[3])[1:])]
        rows = [None] * world
        dist.all_gather_object(rows, row)
        if rank == 0:
            print("   rank   period     work     skew      pub     wait      rel   (us, medians)")
            for r, x in enumerate(rows):
                print(f"   {r:4d} " + " ".join(f"{v:8.1f}" for v in x), flush=True)
t = torch.tensor([1.0 if ok else 0.0], device=dev)
dist.all_reduce(t, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
if rank == 0:
    print("FUSED_DP_OK" if t.item() > 0 else "FUSED_DP_MISMATCH", flush=True)
sys.exit(0 if t.item() > 0 else 1)
