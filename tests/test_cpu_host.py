"""CPU: host-side logic — the C ABI exports, state sharding over gloo (world_size 2), layout generators."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    from minigrid_dynamicprogramming_b200 import build
    lib = ctypes.CDLL(build.build())
    n = 0
    for hdr in ("mgb200.h", "mgb200_dp.h"):
        src = open(os.path.join(ROOT, "include", hdr)).read()
        src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
        for name in sorted(set(re.findall(r"\b(mg(?:dp)?_[a-z0-9_]+)\s*\(", src))):
            assert hasattr(lib, name), f"{hdr}: {name} not exported"
            n += 1
    assert n >= 25
    lib.mg_last_error.restype = ctypes.c_char_p
    assert lib.mg_version() == 100


def test_shard_bounds_cover_the_state_range():
    import importlib.util
    spec = importlib.util.find_spec("minigrid_dynamicprogramming_b200.dp")
    assert spec is not None
    src = open(spec.origin).read()
    ns = {}
    exec(src[src.index("def shard_bounds"):src.index("def value_iteration_sharded")], ns)
    for S in (1, 31, 32, 33, 2680, 357680, 16786688):
        for world in (1, 2, 3, 4, 8):
            prev = 0
            for r in range(world):
                lo, hi, per = ns["shard_bounds"](S, r, world)
                assert lo == min(prev, S) and hi >= lo and per % 32 == 0
                prev = lo + per
            assert hi == S or prev >= S


def _worker(rank, world, port, succ, q, exchange="allgather"):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import types, sys
    # the product module imports the CUDA library at import time; for the host-logic test load only the pure-python driver
    src = open(os.path.join(ROOT, "minigrid_dynamicprogramming_b200", "dp.py")).read()
    body = src[src.index("def shard_bounds"):src.index("def greedy_actions")]
    ns = {"torch": torch}
    exec(body, ns)
    S, A = succ.shape
    S_pad = (S + 31) // 32 * 32
    t = torch.full((A, S_pad), -1, dtype=torch.int32)
    t[:, :S] = torch.from_numpy(succ).t()
    mdp = types.SimpleNamespace(succ=t, S=S, A=A, S_pad=S_pad)
    g = np.float32(0.99)

    def sweep_fn(lo, hi, V_in, V_out, pi, resid, k):   # numpy stand-in for the CUDA sweep kernel
        sp = succ[lo:hi]
        Vn = V_in.numpy()
        q = g * Vn[np.where(sp < 0, 0, sp)]
        q[sp == -1] = 1.0
        q[sp == -2] = 0.0
        best = q.max(axis=1)
        V_out[lo:hi] = torch.from_numpy(best)
        pi[lo:hi] = torch.from_numpy(q.argmax(axis=1).astype(np.uint8))
        resid[0] = float(np.max(np.abs(best - Vn[lo:hi]))) if hi > lo else 0.0
        if exchange == "push":
            # stand-in for mgdp_sweep_push's P2P stores: the "kernel" itself delivers its range to every rank
            per = V_out.numel() // world
            dist.all_gather_into_tensor(V_out, V_out[rank * per: (rank + 1) * per].clone())

    bufs = [torch.zeros(((S + world - 1) // world + 31) // 32 * 32 * world) for _ in range(2)] if exchange == "push" else None
    V, pi, k = ns["value_iteration_sharded"](mdp, sweep_fn=sweep_fn, exchange=exchange, buffers=bufs)
    if rank == 0:
        q.put((V.numpy().copy(), pi.numpy().copy(), k))
    dist.destroy_process_group()


@pytest.mark.parametrize("exchange", ["allgather", "push"])
def test_state_sharded_vi_gloo_world2(exchange):
    """value_iteration_sharded over gloo, world_size 2 — both exchange modes of the driver loop (all-gather after the
    sweep; or the sweep delivering V' itself, as mgdp_sweep_push does over NVLink) — reproduces the single-process
    oracle bit for bit."""
    import torch.multiprocessing as mp
    from helpers import golden
    from oracle import oracle as O
    succ = golden("mdp_reference.npz")["DoorKey-8x8/a7/succ"]
    Vo, pio, ko, _ = O.value_iteration(succ)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + (7 if exchange == "push" else 0)
    procs = [ctx.Process(target=_worker, args=(r, 2, port, succ, q, exchange)) for r in range(2)]
    for p in procs:
        p.start()
    V, pi, k = q.get(timeout=120)
    for p in procs:
        p.join(timeout=60)
    assert k == ko and (V == Vo).all() and (pi == pio).all()


def _layout_worker(rank, world, port, tables, q):
    import types
    import torch
    import torch.distributed as dist
    from oracle import oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    src = open(os.path.join(ROOT, "minigrid_dynamicprogramming_b200", "dp.py")).read()
    body = src[src.index("class BatchedSolve"):src.index("def concat_mdps")]

    def enumerate_batch(layouts, actions, device=None, max_states=None):      # stand-in for the GPU BFS: "layouts" are tables
        offs, parts, o = [0], [], 0
        for t in layouts:
            parts.append(np.where(t >= 0, t + o, t))
            o += len(t)
            offs.append(o)
        return types.SimpleNamespace(succ=np.concatenate(parts), S=o, offsets=offs)

    def value_iteration(m, gamma, tol, max_sweeps):                            # stand-in for mgdp_solve
        V, pi, k, _ = O.value_iteration(m.succ, gamma, tol, max_sweeps)
        return torch.from_numpy(V), torch.from_numpy(pi), k

    ns = {"torch": torch, "enumerate_batch": enumerate_batch, "value_iteration": value_iteration, "ALL_ACTIONS": tuple(range(7)),
          "levelgen": None}
    exec(body, ns)
    res = ns["value_iteration_batched"](tables, device="cpu", shard="layout")
    n = res.layout_range[1] - res.layout_range[0]
    q.put((rank, res.layout_range, res.sweeps, [tuple(x.numpy().copy() for x in res.layout(i)) for i in range(n)]))
    dist.destroy_process_group()


def test_layout_sharded_vi_gloo_world2():
    """value_iteration_batched(shard="layout") over gloo, world_size 2: every rank solves its own block of layouts (no data-path
    collective), the blocks cover all layouts, every layout's V / pi equal its single solve, and the reported sweep count is the
    maximum over ALL layouts (the one 4-byte all-reduce)."""
    import torch.multiprocessing as mp
    from helpers import golden
    from oracle import oracle as O
    z = golden("mdp_reference.npz")
    tables = [z["DoorKey-8x8/a7/succ"], z["Empty-8x8/a3/succ"], z["DoorKey-8x8/a5/succ"], z["DoorKey-16x16/a5/succ"], z["Empty-8x8/a3/succ"]]
    tables = [np.ascontiguousarray(np.pad(t, ((0, 0), (0, 7 - t.shape[1])), constant_values=-2)) for t in tables]   # same action count
    single = [O.value_iteration(t) for t in tables]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + 13
    procs = [ctx.Process(target=_layout_worker, args=(r, 2, port, tables, q)) for r in range(2)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    assert [o[1] for o in out] == [(0, 3), (3, 5)]
    assert out[0][2] == out[1][2] == max(s[2] for s in single)
    for rank, (lo, hi), _, sols in out:
        for i, (V, pi) in enumerate(sols):
            assert (V == single[lo + i][0]).all() and (pi == single[lo + i][1]).all(), (rank, i)


def test_launch_chaining_protocol_model():
    """A small-integer model of the step kernel's per-CTA launch chaining (csrc/mg_step.cu prologue / epilogue), run under random
    schedules: one word per CTA index holds `tickets taken` (high half) and `launches finished` (low half), here 3 bits each so
    that every run wraps several times.  The only ordering the model grants is the one the hardware gives — a CTA of launch k+1
    exists only after EVERY CTA of launch k has taken its ticket (griddepcontrol.launch_dependents) — and it checks what the
    kernel relies on: CTA i of launch k+1 never runs before CTA i of launch k has finished, in launch order, and the wrap of the
    finished half never leaks into the ticket half (the release add of the wrapping launch is 1 - 2^H)."""
    import random
    H = 3
    M, W = 1 << H, 1 << (2 * H)
    rng = random.Random(5)
    for trial in range(300):
        G, K = rng.randint(1, 4), rng.randint(10, 40)          # CTAs per launch, launches
        word = [((s << H) | s) % W for s in [rng.randrange(M)] * G]      # counters start anywhere (MG_CHAIN_START)
        start = word[0] & (M - 1)
        # per CTA (k, i): 0 = not launched, 1 = resident (no ticket yet), 2 = has ticket / waiting, 3 = running, 4 = done
        state = [[0] * G for _ in range(K)]
        ticket = [[None] * G for _ in range(K)]
        ran = [[] for _ in range(G)]                           # launch indices in the order CTA index i actually ran them
        steps = 0
        while any(s != 4 for row in state for s in row):
            steps += 1
            assert steps < 100000, "model deadlocked"
            moves = []
            oldest = min(k for k in range(K) if any(x != 4 for x in state[k]))
            for k in range(K):
                for i in range(G):
                    s = state[k][i]
                    if s == 0 and (k == 0 or all(x >= 2 for x in state[k - 1])) and k - oldest < M - 1:
                        moves.append((k, i))                   # may become resident (fewer than 2^H launches in flight: on the
                                                               # GPU the resident CTAs bound that far below 2^32)
                    elif s in (1, 3):
                        moves.append((k, i))                   # take the ticket / finish
                    elif s == 2 and (word[i] & (M - 1)) == ticket[k][i]:
                        moves.append((k, i))                   # predecessor through: run
            k, i = rng.choice(moves)
            s = state[k][i]
            if s == 0:
                state[k][i] = 1
            elif s == 1:                                       # old = atomicAdd(word, 1 << H)
                ticket[k][i] = word[i] >> H
                word[i] = (word[i] + (1 << H)) % W
                state[k][i] = 2
            elif s == 2:
                assert all(state[j][i] == 4 for j in range(k)), "ran before its predecessor finished"
                ran[i].append(k)
                state[k][i] = 3
            else:                                              # finished++ (carry taken back out on the wrap)
                inc = (1 - (1 << H)) % W if ticket[k][i] == M - 1 else 1
                word[i] = (word[i] + inc) % W
                state[k][i] = 4
        for i in range(G):
            assert ran[i] == list(range(K))
            assert word[i] == ((((start + K) % M) << H) | ((start + K) % M))
