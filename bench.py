#!/usr/bin/env python
"""bench.py — headline benchmark: batched step + partial observation, env-steps/s.

Workload (BASELINE.json configs[1]): MiniGrid-DoorKey-8x8-v0, 65,536 envs per GPU, uniform random actions,
same-step autoreset from a 4,096-layout pool (seeds 0..4095 of the reference generator, levelgen.py).
One bench "step" = LAUNCHES_PER_STEP mg_step_many launches of ROLLOUT env-steps each (every obs written), so that
`--steps 20` times more than half a second of back-to-back kernels (a sustained figure, not a 2 ms burst).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # this engine
    python bench.py --impl reference ...                           # the CPU port (oracle/) on the host cores

Prints ONE JSON line (rank 0), as the LAST line of the run: NCCL's own log lines (NCCL_DEBUG, if the caller sets it)
are left alone and come before it.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ENV_ID = "MiniGrid-DoorKey-8x8-v0"
N_PER_GPU = 65536
POOL = 4096
# other workloads: BASELINE configs 3 / 4 / 5 grids (reported in the `workloads` block of the line; `--workload X` makes one
# of them the timed headline for A/B runs)
WORKLOADS = {
    "doorkey8": (ENV_ID, 65536),
    "doorkey16": ("MiniGrid-DoorKey-16x16-v0", 65536),
    "mixed": (["MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0"], 131072),     # config 4: 1,048,576 envs over 8 GPUs
    "keycorridor": ("MiniGrid-KeyCorridorS6R3-v0", 65536),
    "dynobs8": ("MiniGrid-Dynamic-Obstacles-8x8-v0", 65536),       # EXT kernel: PCG64 draws + obstacle moves inside every step
}
ROLLOUT = 32                 # env-steps per launch
LAUNCHES_PER_STEP = 384      # launches per bench step (0.65 s timed at --steps 20)
BYTES_PER_ENV_STEP = 322     # 147 window + 147 obs + 1 action + 8 reward + 2 flags + 1 dir + 16 agent state
D2H_BYTES_PER_ENV_STEP = 147 + 1 + 8 + 1 + 1
METRIC = "env_steps_per_sec_incl_obs"


def workload_config(n_gpus, workload="doorkey8"):
    """The `config` of the line — the SAME dict in the engine arm and in the reference arm."""
    env_id, n = WORKLOADS[workload]
    name = env_id if isinstance(env_id, str) else " + ".join(env_id) + " (even/odd envs)"
    return {"workload": f"{name} step+partial obs, {n} envs per GPU x {n_gpus} GPU(s), uniform random actions 0..6, "
                        f"same-step autoreset from a {POOL}-layout pool",
            "envs_per_gpu": n, "n_envs": n * n_gpus, "env_steps_per_launch_per_env": ROLLOUT,
            "launches_per_step": LAUNCHES_PER_STEP, "env_steps_per_step": n * n_gpus * ROLLOUT * LAUNCHES_PER_STEP,
            "l2": f"no flush: every launch writes {n * ROLLOUT * D2H_BYTES_PER_ENV_STEP / 1e6:.0f} MB of fresh outputs per GPU "
                  "(> 126 MB L2); the env state (planes + packed tiles) is the engine's resident working set"}


def hbm_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def committed_traffic(key):
    """DRAM bytes per launch from the committed ncu --set full capture (profiles/traffic.json) — a pointer to the profile,
    not something this run measured."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            d = json.load(f)
        return d.get(key), "profiles/traffic.json <- profiles/r2/ncu_summary.txt (ncu --set full; committed, not measured by this run)"
    except Exception:
        return None, None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "50"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.1)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1] or [r for _, r in self.rows]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        try:
            sm = float(np.median([float(r[0]) for r in rows]))
            mx = float(rows[0][1])
        except Exception:
            sm, mx = None, None
        return {"sm_mhz": sm, "sm_max_mhz": mx, "reasons": reasons, "samples": len(rows)}


def cpu_port_rollout(n_envs, T, nthreads, seed=0, repeat=1):
    """Times the C port of the reference path (oracle/) on the host cores.  Returns (best env-steps/s, times)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import L
    from oracle import oracle as O
    pool = L.generate_many(ENV_ID, range(POOL))
    n = n_envs
    pc = np.stack([l.cells.reshape(-1) for l in pool]).astype(np.uint8)
    pst = np.zeros((POOL, 8), np.int32)
    for i, l in enumerate(pool):
        pst[i, :3] = (*l.agent_pos, l.agent_dir)
    cells = np.ascontiguousarray(pc[np.arange(n) % POOL])
    st = np.ascontiguousarray(pst[np.arange(n) % POOL])
    cfgs = O.make_cfg(8, 8, 640, False, 0, 7).reshape(1, -1)
    cls = np.zeros(n, np.uint8)
    lid = (np.arange(n) % POOL).astype(np.int32)
    rng = np.random.default_rng(seed)
    best = 0.0
    times = []
    for _ in range(repeat):
        acts = rng.integers(0, 7, (T, n)).astype(np.uint8)
        t0 = time.perf_counter()
        O.rollout(cells, st, cls, cfgs, acts, autoreset=True, pool_cells=pc, pool_st=pst,
                  pool_class=np.zeros(POOL, np.uint8), layout_id=lid, nthreads=nthreads)
        dt = time.perf_counter() - t0
        times.append(dt)
        best = max(best, n * T / dt)
    return best, times


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path.  The reference is pure Python and
    /root/reference does not exist on the GPU box, so this is the oracle's C port of it (kind "port"),
    with all host threads, on the engine arm's config (args.gpus x 65,536 envs): each step is a bounded sample of that
    workload — ONE launch's worth (32 env-steps of all envs) instead of the engine's LAUNCHES_PER_STEP."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    K, W = args.steps, args.warmup
    n = N_PER_GPU * max(args.gpus, 1)
    _, times = cpu_port_rollout(n, ROLLOUT, cores, repeat=K + W)
    dt = sum(times[W:])
    value = n * ROLLOUT * K / dt
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": 1e3 * dt / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(max(args.gpus, 1)),
            "cpu_baseline": {"value": value, "unit": "env-steps/s", "cores": cores, "kind": "port",
                             "sample": f"each step = {n} envs x {ROLLOUT} env-steps ({n * ROLLOUT} env-steps: one launch's worth of the "
                                       f"engine's {LAUNCHES_PER_STEP}-launch step), C port oracle/mg_oracle.c, OpenMP over envs"},
            "e2e": {"value": value, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def bind_to_gpu_numa(index):
    """Pin this process to the CPUs next to its GPU (sysfs local_cpulist of the PCI device) BEFORE it allocates pinned
    buffers, so that they land on the GPU's NUMA node.  Returns a short description for the line."""
    try:
        bus = subprocess.check_output(["nvidia-smi", "-i", str(index), "--query-gpu=pci.bus_id", "--format=csv,noheader"],
                                      text=True, timeout=20).strip().lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        base = f"/sys/bus/pci/devices/{bus}"
        node = open(base + "/numa_node").read().strip()
        cpus = open(base + "/local_cpulist").read().strip()
        want = set()
        for part in cpus.split(","):
            if part:
                a, _, b = part.partition("-")
                want |= set(range(int(a), int(b or a) + 1))
        have = os.sched_getaffinity(0)
        use = (want & have) or have
        if use != have:
            os.sched_setaffinity(0, use)
        return f"gpu {index}: numa_node {node}, local cpus {cpus}, bound to {len(use)} of {len(have)} cpus"
    except Exception as e:  # noqa: BLE001
        return f"unbound ({type(e).__name__})"


def measure_step_workload(name, dev, rank, world, launches=5):
    """Device-resident step throughput of another BASELINE grid: value + §8.4 roofline fraction, max over ranks."""
    import torch
    import torch.distributed as dist
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    env_id, N = WORKLOADS[name]
    seeds = range(POOL) if isinstance(env_id, str) else [i // 2 for i in range(POOL)]
    env = BatchedMiniGrid(env_id, N, seeds=seeds, autoreset="same_step", device=dev)
    env.reset_envs((torch.arange(N, device=dev) + rank * N) % POOL)
    gen = torch.Generator(device=dev)
    gen.manual_seed(99 + rank)
    acts = [torch.randint(0, 7, (ROLLOUT, N), dtype=torch.uint8, device=dev, generator=gen) for _ in range(2)]
    for i in range(3):
        env.step_many(acts[i % 2])
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(launches):
        env.step_many(acts[i % 2])
    e1.record()
    torch.cuda.synchronize()
    env.check_errors()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    b_sm, thr, smem = env.launch_shape()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    v = world * N * ROLLOUT * launches / (float(ms.item()) * 1e-3)
    peak, _ = hbm_peak()
    env.close()
    del env, acts
    torch.cuda.empty_cache()
    return {"envs_per_gpu": N, "n_envs": N * world, "value": v, "unit": "env-steps/s", "launches": launches,
            "frac": v * BYTES_PER_ENV_STEP / 1e9 / (peak * world), "resident_ctas_per_sm": b_sm,
            "waves": round(((N + thr - 1) // thr) / (b_sm * sms), 3)}


def d2h_ceiling(dev, world, nbytes, reps=6):
    """What the box can move device->host when every rank copies at once: `reps` plain pinned cudaMemcpyAsync of `nbytes`
    per rank (one call per copy), wall clock between two barriers.  GB/s summed over ranks."""
    import torch
    import torch.distributed as dist
    src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    dst = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        dst.copy_(src, non_blocking=True)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    del src, dst
    return world * nbytes * reps / float(dt.item()) / 1e9


def run_engine(args):
    import torch
    import torch.distributed as dist
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    from minigrid_dynamicprogramming_b200 import _capi

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    binding = bind_to_gpu_numa(local)               # before any pinned allocation
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    env_id, N = WORKLOADS[args.workload]
    K, W, T, LPS = args.steps, args.warmup, ROLLOUT, LAUNCHES_PER_STEP
    headline = args.workload == "doorkey8"

    # env shard of this rank: global env g = rank*N + e uses pool slot g % POOL (no cross-rank data path)
    pool_seeds = range(POOL) if isinstance(env_id, str) else [i // 2 for i in range(POOL)]
    env = BatchedMiniGrid(env_id, N, seeds=pool_seeds, autoreset="same_step", device=dev)
    env.reset_envs((torch.arange(N, device=dev) + rank * N) % POOL)
    gen = torch.Generator(device=dev)
    gen.manual_seed(1234 + rank)
    nbuf = 8
    actions = [torch.randint(0, 7, (T, N), dtype=torch.uint8, device=dev, generator=gen) for _ in range(nbuf)]
    env.step_many(actions[0])
    torch.cuda.synchronize()
    b_sm, thr, smem = env.launch_shape()
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    launch_shape = {"threads_per_cta": thr, "ctas": (N + thr - 1) // thr, "resident_ctas_per_sm": b_sm, "smem_per_cta": smem,
                    "waves": round(((N + thr - 1) // thr) / (b_sm * sms), 3), "autoreset_stride": env.reset_stride}

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`): K steps of LPS back-to-back launches ----
    j = 0
    for _ in range(W):
        for _ in range(LPS):
            env.step_many(actions[j % nbuf])
            j += 1
    barrier()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = env.launch_count
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    t_wall0 = time.perf_counter()
    evs[0].record()
    for i in range(K):
        for _ in range(LPS):
            env.step_many(actions[j % nbuf])
            j += 1
        evs[i + 1].record()
    barrier()
    t_wall1 = time.perf_counter()
    launches = env.launch_count - l0
    ms_total = evs[0].elapsed_time(evs[K])
    step_ms = [evs[i].elapsed_time(evs[i + 1]) for i in range(K)]
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    env.check_errors()
    tmax = torch.tensor([ms_total], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms_total_max = float(tmax.item())
    value = world * N * T * LPS * K / (ms_total_max * 1e-3)

    # ---- end to end through the C ABI with HOST buffers (`e2e`): one e2e step = one mg_step_many_host call ----
    n = T * N
    h_act = torch.randint(0, 7, (T, N), dtype=torch.uint8).pin_memory()
    h_img = torch.empty((T, N, 147), dtype=torch.uint8).pin_memory()
    h_dir = torch.empty((T, N), dtype=torch.uint8).pin_memory()
    h_rew = torch.empty((T, N), dtype=torch.float64).pin_memory()
    h_te = torch.empty((T, N), dtype=torch.uint8).pin_memory()
    h_tr = torch.empty((T, N), dtype=torch.uint8).pin_memory()

    def host_step():
        _capi.check(_capi.lib.mg_step_many_host(env._handle, h_act.data_ptr(), T, h_img.data_ptr(), h_dir.data_ptr(),
                                                h_rew.data_ptr(), h_te.data_ptr(), h_tr.data_ptr()), "mg_step_many_host")

    Ke = 4 * K
    for _ in range(3):
        host_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(Ke):
        host_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    te = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())
    e2e_value = world * N * T * Ke / e2e_s
    ceiling = d2h_ceiling(dev, world, n * 147)
    del h_img, h_rew

    wl = None
    if headline and not args.no_workloads:
        wl = {name: measure_step_workload(name, dev, rank, world) for name in ("mixed", "keycorridor", "doorkey16")}
    extras = run_extras(env, dev, N) if (world == 1 and headline and not args.no_dp) else None
    env.close()
    del env
    torch.cuda.empty_cache()
    dp_result = None if (args.no_dp or not headline) else run_dp(args, dev, rank, world)

    ok = True
    line = None
    if rank == 0:
        peak, peak_src = hbm_peak()
        avg_ms = float(np.mean(step_ms)) / LPS
        achieved = BYTES_PER_ENV_STEP * N * T / (avg_ms * 1e-3) / 1e9
        traffic, traffic_src = committed_traffic("step_many_kernel_dram_bytes_per_launch") if headline else (None, None)
        cores = os.cpu_count() or 1
        cpu = None
        if world == 1 and not args.no_cpu_baseline and headline:
            v, _ = cpu_port_rollout(4096, 2048, cores, repeat=1)
            cpu = {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                   "sample": "4096 envs x 2048 env-steps (8.4M env-steps) of the same workload, C port (oracle/mg_oracle.c), OpenMP over envs"}
        line = {"metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms_total_max / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic",
                "config": workload_config(world, args.workload),
                "launch_shape": launch_shape,
                "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                             "traffic": traffic, "traffic_source": traffic_src,
                             "frac_dram": None if not traffic else traffic / (avg_ms * 1e-3) / 1e9 / peak,
                             "kernel": "mg::step_many_kernel", "peak_source": peak_src,
                             "bytes_per_env_step": BYTES_PER_ENV_STEP, "avg_launch_ms": avg_ms,
                             "note": "frac = SURVEY 8.4's algorithmic 322 B/env-step (the 147-B window and the 16-B agent record stay in "
                                     "shared memory / L2 across the 32 steps of a launch, so this is an accounting ceiling and frac can "
                                     "exceed 1); frac_dram = DRAM bytes per launch of the committed ncu capture over this run's time per "
                                     "launch: the physical HBM figure.  avg_launch_ms = timed region / launches: consecutive step launches "
                                     "overlap (a CTA waits for the same CTA of the previous launch, not for the whole grid), so this is the "
                                     "steady-state time per launch, shorter than one launch run alone (ncu: ~104 us)"},
                "cpu_baseline": cpu,
                "e2e": {"value": e2e_value, "unit": "env-steps/s", "h2d_bytes_per_step": n, "d2h_bytes_per_step": n * D2H_BYTES_PER_ENV_STEP,
                        "api": "mg_step_many_host (C ABI, pinned host buffers; 4 chunks alternating on 2 streams)",
                        "step": f"one call = {T} env-steps of all {N} envs per GPU", "steps": Ke, "seconds": e2e_s,
                        "d2h_gbs": e2e_value * D2H_BYTES_PER_ENV_STEP / 1e9, "d2h_ceiling_gbs": ceiling,
                        "frac_of_d2h_ceiling": e2e_value * D2H_BYTES_PER_ENV_STEP / 1e9 / ceiling,
                        "ceiling_how": f"{world} rank(s) x 6 plain pinned cudaMemcpyAsync D2H of {n * 147 / 1e6:.0f} MB each, concurrently",
                        "cpu_binding": binding},
                "gpu_launches": int(launches), "clocks": clocks, "timed_region_s": ms_total_max * 1e-3,
                "workloads": wl, "dp": dp_result, "extras": extras}
        if dp_result is not None:
            fs = dp_result.get("fused_solve") or {}
            line["dp_summary"] = {"n_gpus": world, "exchange": dp_result["exchange"], "sweep_us": round(dp_result["sweep_ms"] * 1e3, 2),
                                  "backups_per_s": round(dp_result["value"] / 1e9, 2), "unit": "G backups/s",
                                  "frac": round(dp_result["roofline"]["frac"], 4), "sweeps": fs.get("sweeps", (dp_result.get("solve") or {}).get("sweeps")),
                                  "bit_exact": fs.get("bit_exact_vs_single_gpu", None if world > 1 else True)}
            ok = not (world > 1 and fs and fs.get("bit_exact_vs_single_gpu") is False)
            ok = ok and (dp_result.get("solve") or {}).get("stream_bit_exact_vs_cooperative") is not False
            line["dp_summary"]["solve_ms"] = round(1e3 * (dp_result.get("solve") or {}).get("seconds", 0.0), 3) or None
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    finish(rank, world, line, ok)
    if not ok:
        sys.exit(3)


def finish(rank, world, line, ok=True):
    """Rank 0 prints the JSON line LAST: the other ranks have torn NCCL down and flushed before it writes."""
    sys.stdout.flush()
    sys.stderr.flush()
    tag = os.path.join("/tmp", f"mgb200_bench_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}")
    if world > 1 and rank != 0:
        open(f"{tag}_{rank}", "w").close()
        os._exit(0)                      # no static destructors: NCCL would print its "Closing env plugin" line after rank 0's JSON
    if world > 1:
        t0 = time.time()
        while time.time() - t0 < 10 and not all(os.path.exists(f"{tag}_{r}") for r in range(1, world)):
            time.sleep(0.05)
        time.sleep(0.3)
        for r in range(1, world):
            try:
                os.remove(f"{tag}_{r}")
            except OSError:
                pass
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        # NCCL (with NCCL_DEBUG=INFO) prints one more line from a static destructor at process exit ("Closing env plugin"),
        # i.e. AFTER anything Python can print: rank 0 leaves without running destructors so that the JSON line stays last
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0 if ok else 3)


def run_extras(env, dev, N):
    """Secondary measurements on the same batch (not the headline): one step() per launch, reset latency, the RGB renderers
    (the analogues of minigrid/benchmark.py:13-49's reset time / agent-view FPS / render FPS)."""
    import torch

    def timed(fn, reps):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e-3 / reps

    peak, _ = hbm_peak()
    out = {}
    a1 = torch.randint(0, 7, (1, N), dtype=torch.uint8, device=dev)
    dt = timed(lambda: env.step_many(a1), 200)
    out["single_step"] = {"what": "T=1 closed loop: one mg_step launch per env step (python call included)", "us_per_step": dt * 1e6,
                          "env_steps_per_s": N / dt}
    ids = torch.arange(N, dtype=torch.int32, device=dev) % POOL
    dt = timed(lambda: env.reset_envs(ids), 50)
    out["reset"] = {"what": f"mg_reset of all {N} envs from the layout pool + first observation (benchmark.py:22-29 'Env reset time'; "
                            "the layouts themselves are generated on the host once)", "ms_per_reset_of_all_envs": dt * 1e3,
                    "ns_per_env": dt * 1e9 / N}
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(s):
        env.step_many(a1)
        torch.cuda.synchronize()
        with torch.cuda.graph(g, stream=s):
            for _ in range(16):
                env.step_many(a1)
    dt = timed(g.replay, 50) / 16
    out["single_step_graph"] = {"what": "T=1 launches captured in a CUDA graph (16 per replay)", "us_per_step": dt * 1e6,
                                "env_steps_per_s": N / dt}
    a128 = torch.randint(0, 7, (128, N), dtype=torch.uint8, device=dev)
    dt = timed(lambda: env.step_many(a128), 5)
    out["rollout_128"] = {"what": "T=128 env-steps per launch (a PPO-length rollout; the headline uses T=32)", "us_per_launch": dt * 1e6,
                          "env_steps_per_s": 128 * N / dt,
                          "roofline_frac": 128 * N / dt * BYTES_PER_ENV_STEP / 1e9 / peak}
    del a128
    env.step_many(a1)                                    # shrink the engine's output buffers back to T = 1
    for name, pov in (("render_pov_ts8", True), ("render_full_ts8", False)):
        buf = env.get_frame(agent_pov=pov, tile_size=8)
        nbytes = buf.numel()
        dt = timed(lambda: env.get_frame(agent_pov=pov, tile_size=8, out=buf if pov else None), 20)
        out[name] = {"what": ("RGBImgPartialObsWrapper frame 56x56x3" if pov else "RGBImgObsWrapper frame 64x64x3") + f" of {N} envs per launch",
                     "frames_per_s": N / dt, "us_per_launch": dt * 1e6,
                     "roofline": {"bound": "hbm", "achieved": nbytes / dt / 1e9, "peak": peak, "unit": "GB/s",
                                  "frac": nbytes / dt / 1e9 / peak, "bytes_per_frame": nbytes // N}}
        del buf
    return out


def dp_cpu_baseline(m, V_gpu):
    """The CPU side of the DP metric on the host cores of the GPU box (SURVEY §8.4 "CPU DP baseline"; the reference has
    no DP code, F2): the same Jacobi iteration on the same DoorKey-16x16 successor table as (a) the oracle's C loop,
    (b) numpy gathers, (c) a plain Python double loop over (s, a) — the last on a bounded sample of one partial sweep.
    One core each.  The C solve must reproduce the GPU's V bit for bit."""
    from oracle import oracle as O
    succ = np.ascontiguousarray(m.succ_sa())
    S, A = succ.shape
    t0 = time.perf_counter()
    Vc, _, kc, _ = O.value_iteration(succ)
    tc = time.perf_counter() - t0
    t0 = time.perf_counter()
    Vn, _, kn = O.value_iteration_numpy(succ)
    tn = time.perf_counter() - t0
    n_py = min(S, 150000)
    V = Vc.tolist()
    rows = succ[:n_py].tolist()
    t0 = time.perf_counter()
    out = [0.0] * n_py
    for s in range(n_py):                              # the "Python DP loop": one Bellman backup per state, 7 actions each
        best = -1.0
        for sp in rows[s]:
            q = 1.0 if sp == -1 else (0.0 if sp == -2 else 0.99 * V[sp])
            if q > best:
                best = q
        out[s] = best
    tp = time.perf_counter() - t0
    return {"unit": "backups/s", "cores": 1, "kind": "port",
            "c_loop": S * kc / tc, "numpy": S * kn / tn, "python_loop": n_py / tp,
            "bit_exact_vs_gpu": bool(np.array_equal(Vc[:S], V_gpu[:S])),
            "sample": f"DoorKey-16x16 all-7-action table ({S} states): C loop {kc} sweeps to convergence, numpy {kn} sweeps, "
                      f"Python loop one pass over the first {n_py} states"}


def run_dp(args, dev, rank, world):
    """Second metric of BASELINE.json: Bellman state-backups/s and % of the HBM roofline (36 B per backup at A=7).

    HBM regime: the all-7-action KeyCorridorS6R3 space (16.8 M states, 470 MB successor table; BASELINE config 5),
    enumerated on the GPU, K4 sweeps timed with CUDA events.  With N > 1 the sweep is state-sharded; the headline is the
    fused solve (one cooperative launch per GPU, V' pushed to the peers by the sweep, epoch-flag barrier) and its V and
    policy are compared bit for bit with the single-GPU solve of the same table — a mismatch fails the run.
    Latency regime: DoorKey-16x16 seed 0, all 7 actions (357,680 states; BASELINE config 3): time to convergence of
    the cooperative on-device solve — reported as time, never divided by the HBM peak."""
    import torch
    import torch.distributed as dist
    from minigrid_dynamicprogramming_b200 import dp, _capi
    import ctypes as C
    out = {}
    t0 = time.perf_counter()
    m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", dp.ALL_ACTIONS, seed=0, device=dev, max_states=1 << 25)
    torch.cuda.synchronize()
    enum_s = time.perf_counter() - t0
    S, A, S_pad = m.S, m.A, m.S_pad
    lo, hi, per = dp.shard_bounds(S, rank, world)
    resid = torch.zeros(1, dtype=torch.float32, device=dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    n_v = per * world if world > 1 else S_pad

    def time_sweeps(bufs, one_sweep, K=30):
        cur = 0
        for _ in range(5):
            one_sweep(bufs[cur], bufs[1 - cur], 1 - cur)
            cur = 1 - cur
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(K):
            one_sweep(bufs[cur], bufs[1 - cur], 1 - cur)
            cur = 1 - cur
        e1.record()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item()) / K

    def sweep_allgather(V, Vn, k):
        _capi.check(_capi.lib.mgdp_sweep(m.succ.data_ptr(), S_pad, A, lo, hi, V.data_ptr(), Vn.data_ptr(), None, None,
                                         0.99, resid.data_ptr(), stream), "mgdp_sweep")
        if world > 1:
            dist.all_gather_into_tensor(Vn, Vn[rank * per:(rank + 1) * per].clone())
            dist.all_reduce(resid, op=dist.ReduceOp.MAX)

    rowmask = m.rowmask
    bytes_moved = m.bytes_per_backup(masked=True)

    def sweep_masked(V, Vn, k):
        _capi.check(_capi.lib.mgdp_sweep_masked(m.succ.data_ptr(), rowmask.data_ptr(), S_pad, A, lo, hi, V.data_ptr(), Vn.data_ptr(), None,
                                                0.99, resid.data_ptr(), stream), "mgdp_sweep_masked")
        if world > 1:
            dist.all_gather_into_tensor(Vn, Vn[rank * per:(rank + 1) * per].clone())
            dist.all_reduce(resid, op=dist.ReduceOp.MAX)

    plain = [torch.zeros(n_v, dtype=torch.float32, device=dev) for _ in range(2)]
    unmasked_ms = time_sweeps(plain, sweep_allgather)
    for b in plain:
        b.zero_()
    sweep_ms = time_sweeps(plain, sweep_masked)
    exchange, allgather_ms = ("none" if world == 1 else "allgather"), None
    del plain
    fused = None
    if world > 1:
        # the whole solve in ONE cooperative launch per GPU: push + cross-GPU flag barrier, no host / NCCL call per sweep
        try:
            solver = dp.FusedShardedSolver(m)
            solver.solve()
            best = None
            for _ in range(3):
                Vf, pif, kf, msf = solver.solve()
                tm = torch.tensor([msf], dtype=torch.float64, device=dev)
                dist.all_reduce(tm, op=dist.ReduceOp.MAX)
                best = float(tm.item()) if best is None else min(best, float(tm.item()))
            fused = {"masked": bool(getattr(solver, "masked", False)), "what": "mgdp_solve_push: all sweeps to convergence in one cooperative launch per GPU (V' pushed to the peers by the "
                             "sweep; one arrival counter + epoch-flag round per sweep)", "multicast": bool(solver.multicast),
                     "sweeps": kf, "solve_ms": best, "sweep_ms": best / kf, "backups_per_s": S * kf / (best * 1e-3), "V_s0": float(Vf[0])}
            del solver
        except Exception as e:  # noqa: BLE001
            print(f"[bench] rank {rank}: fused solve unavailable ({type(e).__name__}: {e})", file=sys.stderr)
        okf = torch.tensor([1.0 if fused is not None else 0.0], device=dev)
        dist.all_reduce(okf, op=dist.ReduceOp.MIN)
        if float(okf.item()) == 0:
            fused = None
    if fused is not None:
        # headline of the multi-GPU DP = the fused solve (per-sweep time averaged over the whole solve)
        allgather_ms, sweep_ms, exchange = sweep_ms, fused["sweep_ms"], "fused"
    peak, peak_src = hbm_peak()
    backups = S / (sweep_ms * 1e-3)
    bytes_per_backup = 4 * A + 8
    traffic, traffic_src = committed_traffic("sweep_kernel_7_masked_dram_bytes_per_launch")
    out.update({"metric": "bellman_state_backups_per_sec", "value": backups, "unit": "backups/s", "n_gpus": world,
                "row_mask": {"what": "the sweep skips action rows whose 32 successors are all self-loops (mgdp_rowmask): fewer bytes, "
                                     "bit-identical V / pi; `roofline.frac` keeps SURVEY 8.4's 36 B/backup accounting and can therefore "
                                     "exceed 1 — `frac_moved` divides the bytes the masked sweep actually reads and writes by the peak",
                             "bytes_moved_per_backup": bytes_moved,
                             "frac_moved": backups * bytes_moved / 1e9 / (peak * world),
                             "unmasked_sweep_ms": unmasked_ms if world == 1 else None,
                             "unmasked_backups_per_s": S / (unmasked_ms * 1e-3) if world == 1 else None,
                             "unmasked_frac": S / (unmasked_ms * 1e-3) * bytes_per_backup / 1e9 / peak if world == 1 else None},
                "workload": "MiniGrid-KeyCorridorS6R3-v0 seed 0, all 7 actions, gamma 0.99 fp32 Jacobi" +
                            ({"none": "", "allgather": ", state-sharded + NCCL all-gather(V') + all-reduce(max residual) per sweep",
                             "fused": ", state-sharded, whole solve in one cooperative launch per GPU: V' pushed to all peers by the sweep (P2P stores / NVLS multicast), cross-GPU epoch-flag barrier, no NCCL call per sweep"}[exchange]),
                "exchange": exchange, "allgather_sweep_ms": allgather_ms, "fused_solve": fused,
                "states": S, "actions": A, "sweep_ms": sweep_ms, "enumerate_s": enum_s, "bfs_levels": m.levels,
                "roofline": {"bound": "hbm", "achieved": backups * bytes_per_backup / 1e9, "peak": peak * world, "unit": "GB/s",
                             "frac": backups * bytes_per_backup / 1e9 / (peak * world), "bytes_per_backup": bytes_per_backup,
                             "traffic": traffic if world == 1 else None, "traffic_source": traffic_src if world == 1 else None,
                             "frac_dram": None if (world > 1 or not traffic) else traffic / (sweep_ms * 1e-3) / 1e9 / peak,
                             "kernel": "mgdp::sweep_kernel<7, masked>", "peak_source": peak_src,
                             "note": "V (67 MB) is L2-resident; the 470 MB successor table streams from HBM every sweep"}})
    # full solve (on-device convergence test) on ONE GPU — rank 0; with N > 1 it is also the parity reference.
    # Both drivers: one cooperative launch for all sweeps, and the stream of gated per-sweep launches (the default at this size).
    if rank == 0:
        timed = {}
        for method in ("cooperative", "stream"):
            dp.value_iteration(m, method=method)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = dp.value_iteration(m, method=method)
            torch.cuda.synchronize()
            timed[method] = (time.perf_counter() - t0, r)
        (tc, (Vc, pic, kc)), (ts, (Vs, pis, k)) = timed["cooperative"], timed["stream"]
        out["solve"] = {"sweeps": k, "seconds": ts, "V_s0": float(Vs[0]), "method": "stream (mgdp_solve_stream: one gated launch per sweep, "
                        "each launch gates on the previous sweep's residual on the device; host one chunk of 8 launches behind)", "us_per_sweep": 1e6 * ts / k,
                        "cooperative_seconds": tc, "cooperative_us_per_sweep": 1e6 * tc / kc,
                        "stream_bit_exact_vs_cooperative": bool(k == kc and torch.equal(Vs, Vc) and torch.equal(pis, pic))}
        if not out["solve"]["stream_bit_exact_vs_cooperative"]:
            print("bench.py: mgdp_solve_stream and mgdp_solve_masked disagree", file=sys.stderr)
        del Vc, pic, timed, r
        if fused is not None:
            fused["bit_exact_vs_single_gpu"] = bool(k == kf and torch.equal(Vs, Vf) and torch.equal(pis, pif))
            fused["compared"] = f"V float32[{S}] and pi uint8[{S}] of the fused {world}-GPU solve == rank 0's single-GPU mgdp_solve of the same table; sweeps {kf} == {k}"
        del Vs, pis
    del m
    torch.cuda.empty_cache()
    if rank == 0:
        t0 = time.perf_counter()
        m2 = dp.enumerate_mdp("MiniGrid-DoorKey-16x16-v0", dp.ALL_ACTIONS, seed=0, device=dev, max_states=1 << 19)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        dp.value_iteration(m2)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        V2, _, k2 = dp.value_iteration(m2)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        out["latency_regime"] = {"workload": "MiniGrid-DoorKey-16x16-v0 seed 0, all 7 actions (L2-resident: not a roofline case)",
                                 "states": m2.S, "enumerate_s": t1 - t0, "solve_s": t3 - t2, "sweeps": k2,
                                 "us_per_sweep": 1e6 * (t3 - t2) / k2, "V_s0": float(V2[0])}
        if world == 1 and not args.no_cpu_baseline:
            out["cpu_baseline"] = dp_cpu_baseline(m2, V2.cpu().numpy())
        del m2, V2
    if hasattr(dp, "value_iteration_batched"):
        out["batched_layouts"] = run_dp_batched(dev, rank, world, peak, bytes_per_backup)
    out["gpu_launches"] = int(_capi.lib.mgdp_launch_count())
    return out


def run_dp_batched(dev, rank, world, peak, bytes_per_backup):
    """BASELINE config 3 in the HBM regime: B DoorKey-16x16 layouts enumerated in ONE
    batched BFS and solved as one table; with N > 1 the layouts are sharded across the GPUs — no collective on the data
    path (the sweep counts are max-reduced at the end).  Two shapes: 480 all-7-action layouts per GPU (306 k states each)
    and 16,384 navigation-action layouts per GPU (actions 0,1,2,3,5: 2.7 k states each, 44 M states — the low end of SURVEY
    §8.4 config 3b's B = 16,384-65,536; the launch-latency regime when enumerated one at a time)."""
    import torch
    import torch.distributed as dist
    from minigrid_dynamicprogramming_b200 import dp, levelgen
    eid = "MiniGrid-DoorKey-16x16-v0"
    out = {}
    for name, B_per, acts in (("all7", 480, dp.ALL_ACTIONS), ("nav5", int(os.environ.get("MGDP_BENCH_NAV5", 16384)), dp.NAV_ACTIONS)):
        bpb = 4 * len(acts) + 8
        lays = levelgen.generate_many(eid, range(rank * B_per, (rank + 1) * B_per))
        res = dp.value_iteration_batched(lays, None, acts, device=dev)                          # warm-up
        cap = int(res.mdp.S * 1.02) + 4096
        del res
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        res = dp.value_iteration_batched(lays, None, acts, device=dev, max_states=cap)
        torch.cuda.synchronize()
        dt = torch.tensor([res.enumerate_s, res.solve_s], dtype=torch.float64, device=dev)
        tot = torch.tensor([float(res.mdp.S) * res.sweeps, float(res.mdp.S)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        # parity: two layouts of this rank against their own single-layout enumeration + solve
        same = 1.0
        t1 = time.perf_counter()
        for i in (0, B_per // 2):
            m1 = dp.enumerate_mdp(lays[i], acts, device=dev, max_states=1 << 19)
            V1, pi1, _ = dp.value_iteration(m1)
            a, b = res.mdp.offsets[i], res.mdp.offsets[i + 1]
            if not (b - a == m1.S and torch.equal(res.V[a:b], V1) and torch.equal(res.pi[a:b], pi1)):
                same = 0.0
        torch.cuda.synchronize()
        single_us = (time.perf_counter() - t1) * 1e6 / 2
        ok = torch.tensor([same], device=dev)
        if world > 1:
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        bps = float(tot[0].item()) / float(dt[1].item())
        out[name] = {"layouts": B_per * world, "actions": len(acts), "states": int(tot[1].item()),
                     "enumerate_s": float(dt[0].item()), "enumerate_us_per_layout": float(dt[0].item()) * 1e6 / B_per,
                     "single_layout_enumerate_plus_solve_us": single_us,
                     "solve_s": float(dt[1].item()), "sweeps_rank0": res.sweeps, "backups_per_s": bps,
                     "roofline_frac": bps * bpb / 1e9 / (peak * world), "bytes_per_backup": bpb,
                     "layouts_equal_single_solve": bool(ok.item() > 0)}
        del res, lays
        torch.cuda.empty_cache()
    out["workload"] = (f"{eid} layouts (seeds B*rank ..): one batched BFS (mgdp_enumerate_batch) + one cooperative solve per GPU, "
                       "layout-sharded, no data-path collective; max over ranks of the times, sum of the work")
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="engine", choices=["engine", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-dp", action="store_true", help="skip the Bellman-sweep section")
    ap.add_argument("--no-workloads", action="store_true", help="skip the configs 3/4/5 step workloads block")
    ap.add_argument("--workload", default="doorkey8", choices=sorted(WORKLOADS), help="default: the headline config")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "engine" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_engine(args)


if __name__ == "__main__":
    main()
