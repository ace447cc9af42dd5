"""Dynamic programming over the MDP that `MiniGridEnv.step` defines — enumerate, value iteration, policy evaluation.

The reference contains no DP code; these are the "value-iteration/DP entry points" BASELINE.json
asks for, with the conventions of SURVEY §8.3:

    state   everything step() reads except step_count (agent x,y,dir, carrying, door states, object positions)
    ids     serial FIFO-BFS discovery order from the reset state (id 0), actions in the order given
    succ    int32[A, S_pad] ACTION-MAJOR on the device; -1 = terminal with reward 1 (goal / target pickup:
            the transitions on which the reference pays `_reward()`, minigrid_env.py:549-551, keycorridor.py:133-135),
            -2 = terminal with reward 0 (lava, :552-553)
    backup  fp32 Jacobi from V = 0:  V'[s] = max_a (1 | 0 | gamma * V[succ[a, s]])
    stop    after the first sweep with max|V' - V| < tol; pi = lowest action INDEX attaining the max

Everything runs in the sm_100a kernels of csrc/mg_dp.cu through the C ABI (include/mgb200_dp.h); no CPU path.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch

from . import _capi
from . import levelgen

WIN, LOSE = -1, -2
ALL_ACTIONS = (0, 1, 2, 3, 4, 5, 6)
NAV_ACTIONS = (0, 1, 2, 3, 5)          # drop / done are documented "Unused" for DoorKey & co (envs/doorkey.py:25-33)


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class MDP:
    """Successor table of one (or several concatenated) enumerated MDPs, resident on one GPU."""

    def __init__(self, succ, S, actions, info=None, levels=0, offsets=None):
        self.succ = succ                    # int32[A, S_pad] cuda
        self.S = int(S)
        self.actions = tuple(int(a) for a in actions)
        self.A = len(self.actions)
        self.info = info                    # uint8[S, 8] cuda: x, y, dir, carry type/color/state/aux, #modified cells
        self.levels = levels
        self.offsets = offsets if offsets is not None else [0, self.S]   # start id of every member MDP (+ end)
        self._rowmask = None

    @property
    def rowmask(self):
        """uint8[S_pad / 32] cuda: bit a of byte w = action row a has a non-self successor among states 32 w .. 32 w + 31
        (mgdp_rowmask; built on first use).  The masked sweeps skip the other rows: same values, fewer bytes."""
        if self._rowmask is None:
            dev = self.succ.device
            m = torch.empty(self.S_pad // 32, dtype=torch.uint8, device=dev)
            with torch.cuda.device(dev):
                _capi.check(_capi.lib.mgdp_rowmask(self.succ.data_ptr(), self.S_pad, self.A, self.S, m.data_ptr(), _stream(dev)),
                            "mgdp_rowmask")
            self._rowmask = m
        return self._rowmask

    def bytes_per_backup(self, masked=True):
        """HBM bytes a sweep moves per state: 4 per successor row read + 4 V read + 4 V write (+ the mask byte per 32 states)."""
        if not masked:
            return 4.0 * self.A + 8.0
        bits = torch.tensor([bin(i).count("1") for i in range(256)], dtype=torch.int64, device=self.succ.device)
        rows = int(bits[self.rowmask.long()].sum().item())
        return 4.0 * rows * 32 / self.S + 8.0 + 1.0 / 32

    @property
    def S_pad(self):
        return self.succ.shape[1]

    @property
    def device(self):
        return self.succ.device

    def succ_sa(self):
        """int32[S, A] numpy copy in the oracle's state-major layout."""
        return self.succ[:, : self.S].t().contiguous().cpu().numpy()

    @staticmethod
    def from_table(succ_sa, actions=None, device="cuda:0"):
        """Upload a state-major int32[S, A] table (numpy / tensor)."""
        t = torch.as_tensor(np.ascontiguousarray(succ_sa), dtype=torch.int32)
        S, A = t.shape
        S_pad = (S + 31) // 32 * 32
        dev = torch.device(device)
        out = torch.full((A, S_pad), WIN, dtype=torch.int32, device=dev)
        out[:, :S] = t.to(dev).t()
        return MDP(out, S, actions if actions is not None else range(A))


def enumerate_mdp(layout, actions=ALL_ACTIONS, *, seed=None, device="cuda:0", max_states=1 << 20, **env_kwargs):
    """Reachable state space of one layout, enumerated on the GPU (K3b) with the step kernel's transition function.

    layout: a `levelgen.Layout`, or an env id (then `seed` picks the layout: gym.make(id).reset(seed=seed))."""
    if isinstance(layout, str):
        layout = levelgen.generate(layout, 0 if seed is None else seed, **env_kwargs)
    dev = torch.device(device)
    acts = np.asarray(actions, dtype=np.int32)
    cells = np.ascontiguousarray(layout.cells, dtype=np.uint8)
    h = C.c_void_p()
    with torch.cuda.device(dev):
        _capi.check(_capi.lib.mgdp_enumerate(C.byref(h), dev.index or 0, cells.ctypes.data, layout.width, layout.height,
                                             layout.agent_pos[0], layout.agent_pos[1], layout.agent_dir, layout.rule,
                                             acts.ctypes.data, len(acts), int(max_states)), "mgdp_enumerate")
        try:
            S = _capi.lib.mgdp_num_states(h)
            S_pad = _capi.lib.mgdp_padded_states(h)
            succ = torch.empty((len(acts), S_pad), dtype=torch.int32, device=dev)
            info = torch.empty((S, 8), dtype=torch.uint8, device=dev)
            _capi.check(_capi.lib.mgdp_copy_out(h, succ.data_ptr(), info.data_ptr(), _stream(dev)), "mgdp_copy_out")
            levels = _capi.lib.mgdp_num_levels(h)
        finally:
            _capi.lib.mgdp_free(h)
    return MDP(succ, S, acts, info, levels)


def enumerate_batch(layouts, actions=ALL_ACTIONS, *, device="cuda:0", max_states=None):
    """Reachable state spaces of MANY layouts of one env class (same width, height and rule) in ONE batched BFS on the GPU
    (mgdp_enumerate_batch: five launches per BFS level whatever the batch size).  Returns one MDP whose `offsets` give every
    layout its contiguous id range; rows of layout l equal `enumerate_mdp(layouts[l])` shifted by offsets[l].
    max_states bounds the total (default: estimated from the first layout, doubled on MG_ENOMEM)."""
    dev = torch.device(device)
    B = len(layouts)
    l0 = layouts[0]
    if any((l.width, l.height, l.rule) != (l0.width, l0.height, l0.rule) for l in layouts):
        raise ValueError("enumerate_batch: all layouts must share width, height and rule (one env class)")
    acts = np.asarray(actions, dtype=np.int32)
    cells = np.ascontiguousarray(np.stack([l.cells for l in layouts]), dtype=np.uint8)
    agents = np.ascontiguousarray([[l.agent_pos[0], l.agent_pos[1], l.agent_dir] for l in layouts], dtype=np.int32)
    if max_states is None:
        probe = enumerate_mdp(l0, actions, device=dev, max_states=1 << 22)
        max_states = int(probe.S * B * 1.25) + 4096
        del probe
    with torch.cuda.device(dev):
        for _ in range(6):
            h = C.c_void_p()
            max_states = min(int(max_states), (1 << 31) - 17)
            rc = _capi.lib.mgdp_enumerate_batch(C.byref(h), dev.index or 0, B, cells.ctypes.data, l0.width, l0.height,
                                                agents.ctypes.data, l0.rule, acts.ctypes.data, len(acts), max_states)
            if rc != _capi.MG_ENOMEM or max_states >= (1 << 31) - 17:
                break
            max_states *= 2
        _capi.check(rc, "mgdp_enumerate_batch")
        try:
            S = _capi.lib.mgdp_num_states(h)
            S_pad = _capi.lib.mgdp_padded_states(h)
            succ = torch.empty((len(acts), S_pad), dtype=torch.int32, device=dev)
            info = torch.empty((S, 8), dtype=torch.uint8, device=dev)
            _capi.check(_capi.lib.mgdp_copy_out(h, succ.data_ptr(), info.data_ptr(), _stream(dev)), "mgdp_copy_out")
            levels = _capi.lib.mgdp_num_levels(h)
            offs = np.zeros(B + 1, dtype=np.int64)
            _capi.check(_capi.lib.mgdp_layout_offsets(h, offs.ctypes.data), "mgdp_layout_offsets")
        finally:
            _capi.lib.mgdp_free(h)
    return MDP(succ, S, acts, info, levels, [int(o) for o in offs])


class BatchedSolve:
    """Result of value_iteration_batched: the concatenated MDP of this rank's layouts, V / pi over its states, the sweep count
    (max over the layouts — and over the ranks when sharded), and the wall-clock split."""

    def __init__(self, mdp, V, pi, sweeps, enumerate_s, solve_s, layout_range):
        self.mdp, self.V, self.pi, self.sweeps = mdp, V, pi, sweeps
        self.enumerate_s, self.solve_s, self.layout_range = enumerate_s, solve_s, layout_range

    def layout(self, i):
        """(V, pi) of the i-th layout OF THIS RANK (global index layout_range[0] + i)."""
        a, b = self.mdp.offsets[i], self.mdp.offsets[i + 1]
        return self.V[a:b], self.pi[a:b]


def value_iteration_batched(layouts, seeds=None, actions=ALL_ACTIONS, *, device="cuda:0", gamma=0.99, tol=1e-6,
                            max_sweeps=100000, shard=None, group=None, max_states=None, **env_kwargs):
    """Value iteration over MANY layouts at once (BASELINE config 3 in the HBM regime):
    one batched enumeration + one solve over the concatenated table.

    layouts: a list of `levelgen.Layout`s, or an env id (then `seeds` picks them).  shard="layout": rank r of `group` takes
    the r-th contiguous block of the layouts — independent MDPs, so there is NO collective on the data path; one 4-byte
    all-reduce(MAX) of the sweep count at the end.  Every layout's V / pi equal its own single-layout solve bit for bit
    (a converged layout's values are a fixed point of further Jacobi sweeps)."""
    import time
    if isinstance(layouts, str):
        gen = levelgen.make_generator(layouts, **env_kwargs)
        eid = layouts
        layouts = None
    lo, hi = 0, len(seeds) if layouts is None else len(layouts)
    if shard == "layout":
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        per = (hi + world - 1) // world
        lo, hi = min(rank * per, hi), min((rank + 1) * per, hi)
    elif shard is not None:
        raise ValueError('shard must be None or "layout"')
    mine = [gen.generate(eid, int(s)) for s in list(seeds)[lo:hi]] if layouts is None else list(layouts[lo:hi])
    dev = torch.device(device)

    def sync():
        if dev.type == "cuda":
            torch.cuda.synchronize(dev)

    sync()
    t0 = time.perf_counter()
    if mine:
        m = enumerate_batch(mine, actions, device=dev, max_states=max_states)
        sync()
        t1 = time.perf_counter()
        V, pi, k = value_iteration(m, gamma, tol, max_sweeps)
        sync()
        t2 = time.perf_counter()
    else:                                                   # more ranks than layouts
        m, V, pi, k, t1, t2 = None, None, None, 0, t0, t0
    if shard == "layout":
        kk = torch.tensor([k], dtype=torch.int32, device=dev if dist.get_backend(group) == "nccl" else "cpu")
        dist.all_reduce(kk, op=dist.ReduceOp.MAX, group=group)
        k = int(kk.item())
    return BatchedSolve(m, V, pi, k, t1 - t0, t2 - t1, (lo, hi))


def concat_mdps(mdps):
    """One table holding several independent MDPs back to back (the batched-layouts regime of BASELINE config 3)."""
    A = mdps[0].A
    dev = mdps[0].device
    total = sum(m.S for m in mdps)
    S_pad = (total + 31) // 32 * 32
    out = torch.full((A, S_pad), WIN, dtype=torch.int32, device=dev)
    offsets, o = [0], 0
    for m in mdps:
        assert m.A == A
        blk = m.succ[:, : m.S]
        out[:, o: o + m.S] = torch.where(blk >= 0, blk + o, blk)
        o += m.S
        offsets.append(o)
    return MDP(out, total, mdps[0].actions, None, max(m.levels for m in mdps), offsets)


def _as_mdp(mdp, device="cuda:0"):
    return mdp if isinstance(mdp, MDP) else MDP.from_table(mdp, device=device)


# tables above this many successor-id bytes do not stay in the 126 MB L2 between sweeps: method="auto" streams them
STREAM_MIN_TABLE_BYTES = 96 << 20


def value_iteration(mdp, gamma=0.99, tol=1e-6, max_sweeps=100000, *, pi_fixed=None, method="auto", masked=True):
    """fp32 Jacobi value iteration (policy evaluation when `pi_fixed` is given).  masked=True (default) skips the action rows
    that are all self-loops (`MDP.rowmask`); the result is bit-identical to masked=False.

    Returns (V float32[S], pi uint8[S] (action INDEX into mdp.actions), sweeps).  method "cooperative" runs all
    sweeps inside one cooperative launch (no host round trip: right for tables that live in L2, where a sweep takes
    microseconds); "stream" issues one gated launch per sweep with the convergence test on the device
    (`mgdp_solve_stream`: ~8 % less time per sweep in the HBM regime); "auto" (default) picks by table size; "launch"
    issues one kernel per sweep and reads the residual back after each (the path the multi-GPU driver uses).
    All methods give bit-identical V, pi and sweep counts."""
    mdp = _as_mdp(mdp)
    dev, S, S_pad, A = mdp.device, mdp.S, mdp.S_pad, mdp.A
    V = torch.zeros(S_pad, dtype=torch.float32, device=dev)
    Vt = torch.zeros(S_pad, dtype=torch.float32, device=dev)
    pi = torch.zeros(S_pad, dtype=torch.uint8, device=dev)
    pf = None if pi_fixed is None else torch.as_tensor(pi_fixed, dtype=torch.uint8, device=dev).contiguous()
    pfp = None if pf is None else pf.data_ptr()
    rm = mdp.rowmask if (masked and pf is None and S_pad % 32 == 0) else None
    if method == "auto":
        method = "stream" if (pf is None and 4 * A * S_pad >= STREAM_MIN_TABLE_BYTES) else "cooperative"
    if method not in ("cooperative", "stream", "launch"):
        raise ValueError(f"unknown method {method!r}")
    if method == "stream" and pf is not None:
        raise ValueError('method="stream" is value iteration only (policy evaluation: "cooperative" or "launch")')
    with torch.cuda.device(dev):
        if method == "stream":
            res = torch.zeros(5, dtype=torch.int32, device=dev)
            _capi.check(_capi.lib.mgdp_solve_stream(mdp.succ.data_ptr(), None if rm is None else rm.data_ptr(), S_pad, A, S,
                                                    V.data_ptr(), Vt.data_ptr(), pi.data_ptr(), gamma, tol, int(max_sweeps),
                                                    res.data_ptr(), _stream(dev)), "mgdp_solve_stream")
            sweeps = int(res[0].item())
        elif method == "cooperative":
            res = torch.zeros(5, dtype=torch.int32, device=dev)
            if rm is not None:
                _capi.check(_capi.lib.mgdp_solve_masked(mdp.succ.data_ptr(), rm.data_ptr(), S_pad, A, S, V.data_ptr(), Vt.data_ptr(),
                                                        pi.data_ptr(), gamma, tol, int(max_sweeps), res.data_ptr(), _stream(dev)),
                            "mgdp_solve_masked")
            else:
                _capi.check(_capi.lib.mgdp_solve(mdp.succ.data_ptr(), S_pad, A, S, V.data_ptr(), Vt.data_ptr(), pi.data_ptr(), pfp,
                                                 gamma, tol, int(max_sweeps), res.data_ptr(), _stream(dev)), "mgdp_solve")
            sweeps = int(res[0].item())
        else:
            resid = torch.zeros(1, dtype=torch.float32, device=dev)
            sweeps = 0
            while sweeps < max_sweeps:
                resid.zero_()
                if rm is not None:
                    _capi.check(_capi.lib.mgdp_sweep_masked(mdp.succ.data_ptr(), rm.data_ptr(), S_pad, A, 0, S, V.data_ptr(), Vt.data_ptr(),
                                                            pi.data_ptr(), gamma, resid.data_ptr(), _stream(dev)), "mgdp_sweep_masked")
                else:
                    _capi.check(_capi.lib.mgdp_sweep(mdp.succ.data_ptr(), S_pad, A, 0, S, V.data_ptr(), Vt.data_ptr(), pi.data_ptr(),
                                                     pfp, gamma, resid.data_ptr(), _stream(dev)), "mgdp_sweep")
                V, Vt = Vt, V
                sweeps += 1
                if float(resid.item()) < tol:
                    break
    return V[:S], pi[:S], sweeps


def policy_evaluation(mdp, pi, gamma=0.99, tol=1e-6, max_sweeps=100000, **kw):
    """V^pi by fp32 Jacobi sweeps.  Returns (V, sweeps)."""
    V, _, k = value_iteration(mdp, gamma, tol, max_sweeps, pi_fixed=pi, **kw)
    return V, k


def finite_horizon(mdp, max_steps, gamma=1.0, *, return_policy=False):
    """Exact finite-horizon backward induction with the reference's time-dependent reward.

    V_0 float64[S]: the return the reference's episode pays under optimal play from every state at step_count 0
    (for gamma = 1: 1 - 0.9 * d / max_steps with d the shortest number of steps to success, 0 if d > max_steps).
    return_policy=True also returns pi uint8[max_steps, S], the greedy action index at every (step_count, state)."""
    mdp = _as_mdp(mdp)
    dev, S = mdp.device, mdp.S
    V0 = torch.zeros(S, dtype=torch.float64, device=dev)
    Vt = torch.zeros(S, dtype=torch.float64, device=dev)
    pi = torch.zeros((int(max_steps), S), dtype=torch.uint8, device=dev) if return_policy else None
    with torch.cuda.device(dev):
        _capi.check(_capi.lib.mgdp_finite_horizon(mdp.succ.data_ptr(), S, mdp.A, mdp.S_pad, int(max_steps), float(gamma),
                                                  V0.data_ptr(), Vt.data_ptr(), None if pi is None else pi.data_ptr(),
                                                  _stream(dev)), "mgdp_finite_horizon")
    return (V0, pi) if return_policy else V0


def shard_bounds(S, rank, world, align=32):
    """Contiguous state range [lo, hi) owned by `rank`; every shard has the same padded length `per`."""
    per = ((S + world - 1) // world + align - 1) // align * align
    lo = min(rank * per, S)
    hi = min(lo + per, S)
    return lo, hi, per


def _symmetric_buffers(n, dev, group, want_multicast=False):
    """Two float32[n] buffers every rank of `group` can store into: torch symmetric memory (CUDA VMM handles exchanged
    through the process group — plumbing); returns (tensors, per-buffer list of device pointers indexed by rank)
    [+ per-buffer NVLS multicast pointer, 0 where the fabric offers none, with want_multicast]."""
    import torch.distributed as dist
    import torch.distributed._symmetric_memory as symm
    g = group if group is not None else dist.group.WORLD
    bufs, ptrs, mcs = [], [], []
    for _ in range(2):
        t = symm.empty(n, dtype=torch.float32, device=dev)
        h = symm.rendezvous(t, g)
        t.zero_()
        bufs.append(t)
        ptrs.append([int(p) for p in h.buffer_ptrs])
        mcs.append(int(getattr(h, "multicast_ptr", 0) or 0))
    torch.cuda.synchronize(dev)
    dist.barrier(group=group)                                            # nobody pushes into a buffer a peer is still zeroing
    return (bufs, ptrs, mcs) if want_multicast else (bufs, ptrs)


class FusedShardedSolver:
    """exchange="fused": ALL sweeps in one cooperative launch per GPU (mgdp_solve_push) — V' pushed to the peers by the
    sweep itself, cross-GPU barrier from epoch flags in symmetric memory, convergence decided on the device.
    Holds the symmetric buffers so that repeated solves (policy evaluation loops, benchmarks) reuse them."""

    def __init__(self, mdp, group=None):
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm
        self.mdp, self.group = mdp, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.lo, self.hi, self.per = shard_bounds(mdp.S, self.rank, self.world)
        dev = self.dev = mdp.succ.device
        g = group if group is not None else dist.group.WORLD
        self.bufs, ptrs, self._mc = _symmetric_buffers(self.per * self.world, dev, group, want_multicast=True)
        self.ctrl = symm.empty(64, dtype=torch.int32, device=dev)
        hc = symm.rendezvous(self.ctrl, g)
        others = [r for r in range(self.world) if r != self.rank]

        def arr(ps):
            return (C.c_void_p * max(len(others), 1))(*[int(ps[r]) for r in others])

        self._peers = (arr(ptrs[0]), arr(ptrs[1]), arr(hc.buffer_ptrs))
        self.pi = torch.zeros(self.per * self.world, dtype=torch.uint8, device=dev)
        self.scratch = torch.zeros(8, dtype=torch.float32, device=dev)
        self.res = torch.zeros(4, dtype=torch.int32, device=dev)

    def solve(self, gamma=0.99, tol=1e-6, max_sweeps=100000):
        """Returns (V float32[S], pi uint8[S], sweeps, milliseconds of the launch measured with CUDA events)."""
        import torch.distributed as dist
        m, dev = self.mdp, self.dev
        for b in self.bufs:
            b.zero_()
        self.ctrl.zero_()
        torch.cuda.synchronize(dev)
        dist.barrier(group=self.group)                     # every rank's buffers and flags are clean before anyone pushes
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.device(dev):
            e0.record()
            # NVLS multicast (one multimem.st per changed value instead of world-1 unicast stores) when the fabric maps one.
            # Measured with the changed-only filter (us per sweep, unicast / multicast): 2 GPUs 86.7 / 86.4, 4 GPUs 56.5 / 58.9,
            # 8 GPUs 53.6 / 51.4 — the switch-side replication pays once there are more than four destinations.
            # MGDP_PUSH_MULTICAST=0 / 1 forces the unicast / multicast form (A/B runs).
            want = os.environ.get("MGDP_PUSH_MULTICAST")
            use_mc = all(self._mc) and (want == "1" or (want is None and self.world > 4))
            mc = self._mc if use_mc else (None, None)
            self.multicast = mc[0] is not None
            masked = os.environ.get("MGDP_PUSH_MASKED", "1") != "0" and m.S_pad % 32 == 0 and self.lo % 32 == 0
            if masked:
                _capi.check(_capi.lib.mgdp_solve_push_masked(m.succ.data_ptr(), m.rowmask.data_ptr(), m.S_pad, m.A, self.lo, self.hi,
                                                             self.bufs[0].data_ptr(), self.bufs[1].data_ptr(), self._peers[0],
                                                             self._peers[1], mc[0], mc[1], self.ctrl.data_ptr(), self._peers[2],
                                                             self.world - 1, self.rank, self.pi.data_ptr(), gamma, tol, int(max_sweeps),
                                                             self.scratch.data_ptr(), self.res.data_ptr(), _stream(dev)),
                            "mgdp_solve_push_masked")
            else:
                _capi.check(_capi.lib.mgdp_solve_push_mc(m.succ.data_ptr(), m.S_pad, m.A, self.lo, self.hi, self.bufs[0].data_ptr(),
                                                         self.bufs[1].data_ptr(), self._peers[0], self._peers[1], mc[0], mc[1],
                                                         self.ctrl.data_ptr(), self._peers[2], self.world - 1, self.rank,
                                                         self.pi.data_ptr(), gamma, tol, int(max_sweeps), self.scratch.data_ptr(),
                                                         self.res.data_ptr(), _stream(dev)), "mgdp_solve_push_mc")
            self.masked = masked
            e1.record()
        r = self.res.cpu().numpy()
        if r[3]:
            raise RuntimeError("mgdp_solve_push: a peer rank did not reach the barrier within the timeout")
        ms = e0.elapsed_time(e1)
        per, rank = self.per, self.rank
        dist.all_gather_into_tensor(self.pi, self.pi[rank * per: (rank + 1) * per].clone(), group=self.group)
        return self.bufs[int(r[2])][:m.S].clone(), self.pi[:m.S].clone(), int(r[0]), ms


def value_iteration_sharded(mdp, gamma=0.99, tol=1e-6, max_sweeps=100000, *, group=None, sweep_fn=None, exchange="allgather",
                            buffers=None):
    """State-sharded Jacobi VI: every rank sweeps ITS contiguous state range; bit-identical to the single-GPU result
    (Jacobi reads only the previous iterate).  Two ways to exchange V' between sweeps:

      exchange="push"       the sweep kernel itself stores every V'[s] into all peers' buffers over NVLink
                            (mgdp_sweep_push, symmetric memory); per sweep only a 4-byte all-reduce(MAX) of the residual
                            follows, which also orders the peers' stores before the next sweep.
      exchange="fused"      all sweeps in ONE cooperative launch per GPU (mgdp_solve_push): push as above, plus a cross-GPU
                            barrier from epoch flags in symmetric memory and the convergence test on the device — no
                            host round trip and no NCCL call per sweep.
      exchange="allgather"  sweep, then all_gather_into_tensor(V') + all-reduce(MAX): the NCCL baseline.

    `sweep_fn(lo, hi, V_in, V_out, pi, resid, k)` defaults to the CUDA kernel; tests on CPU/gloo inject a stand-in
    (and `buffers` = two plain tensors)."""
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    S, A = mdp.S, mdp.A
    lo, hi, per = shard_bounds(S, rank, world)
    dev = mdp.succ.device
    push = exchange == "push"
    if exchange == "fused":
        bufs, ptrs = None, None
    elif buffers is not None:
        bufs, ptrs = buffers, None
    elif push:
        bufs, ptrs = _symmetric_buffers(per * world, dev, group)
    else:
        bufs, ptrs = [torch.zeros(per * world, dtype=torch.float32, device=dev) for _ in range(2)], None
    pi = torch.zeros(per * world, dtype=torch.uint8, device=dev)
    resid = torch.zeros(1, dtype=torch.float32, device=dev)
    if sweep_fn is None:
        if push:
            peer_arr = []
            for k in range(2):
                others = [ptrs[k][r] for r in range(world) if r != rank]
                peer_arr.append((C.c_void_p * max(len(others), 1))(*others))

            def sweep_fn(lo, hi, V_in, V_out, pi, resid, k):
                with torch.cuda.device(dev):
                    _capi.check(_capi.lib.mgdp_sweep_push(mdp.succ.data_ptr(), mdp.S_pad, A, lo, hi, V_in.data_ptr(), V_out.data_ptr(),
                                                          peer_arr[k], world - 1, pi.data_ptr(), gamma, resid.data_ptr(),
                                                          _stream(dev)), "mgdp_sweep_push")
        else:
            def sweep_fn(lo, hi, V_in, V_out, pi, resid, k):
                with torch.cuda.device(dev):
                    _capi.check(_capi.lib.mgdp_sweep(mdp.succ.data_ptr(), mdp.S_pad, A, lo, hi, V_in.data_ptr(), V_out.data_ptr(),
                                                     pi.data_ptr(), None, gamma, resid.data_ptr(), _stream(dev)), "mgdp_sweep")
    if exchange == "fused":
        return FusedShardedSolver(mdp, group).solve(gamma, tol, max_sweeps)[:3]
    sweeps, cur = 0, 0
    while sweeps < max_sweeps:
        resid.zero_()
        V, Vn = bufs[cur], bufs[1 - cur]
        sweep_fn(lo, hi, V, Vn, pi, resid, 1 - cur)
        if not push:
            # the kernel wrote Vn[lo:hi] in place in the full-length buffer: gather the shards in place
            dist.all_gather_into_tensor(Vn, Vn[rank * per: (rank + 1) * per].clone(), group=group)
        dist.all_reduce(resid, op=dist.ReduceOp.MAX, group=group)
        cur = 1 - cur
        sweeps += 1
        if float(resid.item()) < tol:
            break
    dist.all_gather_into_tensor(pi, pi[rank * per: (rank + 1) * per].clone(), group=group)
    return bufs[cur][:S].clone(), pi[:S], sweeps


def greedy_actions(mdp, pi):
    """Map policy action indices to env action ids."""
    return torch.as_tensor(mdp.actions, device=pi.device, dtype=torch.uint8)[pi.long()]
