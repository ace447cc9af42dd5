"""Time the UNMODIFIED reference's CPU path in the build container (BASELINE config 1).  TEST INFRASTRUCTURE.

The reference is pure Python and does not travel to the GPU box, so bench.py's CPU arm there is the C port; this script
records what the real thing does on this container's cores (stub gymnasium / pygame, oracle/refstub):
  (a) minigrid/benchmark.py's own protocol: reset time, env.step(0) + render, agent-view (RGBImgPartialObs) FPS;
  (b) the random-action rollout BASELINE names: reset(seed=0), default_rng(0).integers(0, 7), reset on done.

    python oracle/run_reference_cpu.py > profiles/r1_reference_cpu_container.txt
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle import ref_harness as R  # noqa: E402


def rollout(eid, steps):
    env = R.make(eid)
    env.reset(seed=0)
    acts = np.random.default_rng(0).integers(0, 7, steps)
    k = 0
    t0 = time.perf_counter()
    for a in acts:
        _, _, te, tr, _ = env.step(int(a))
        if te or tr:
            k += 1
            env.reset(seed=k)
    return steps / (time.perf_counter() - t0)


def benchmark_protocol(eid, num_resets=50, num_frames=1000):
    """minigrid/benchmark.py:13-49 without its prints."""
    from minigrid.wrappers import ImgObsWrapper, RGBImgPartialObsWrapper
    env = R.make(eid, render_mode="rgb_array")
    t0 = time.perf_counter()
    for _ in range(num_resets):
        env.reset()
    reset_ms = 1000 * (time.perf_counter() - t0) / num_resets
    t0 = time.perf_counter()
    for _ in range(num_frames):
        env.render()
    render_fps = num_frames / (time.perf_counter() - t0)
    env = ImgObsWrapper(RGBImgPartialObsWrapper(R.make(eid)))
    env.reset()
    t0 = time.perf_counter()
    for _ in range(num_frames):
        env.step(0)
    view_fps = num_frames / (time.perf_counter() - t0)
    return reset_ms, render_fps, view_fps


if __name__ == "__main__":
    R.gym()
    print(f"reference = Farama Minigrid 2.3.1 snapshot at {R.REFERENCE_ROOT}, python {sys.version.split()[0]}, numpy {np.__version__}, 1 core of this container")
    for eid in ("MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-KeyCorridorS6R3-v0"):
        print(f"{eid}: random-action rollout {rollout(eid, 20000):,.0f} env-steps/s (incl. gen_obs, resets on done)")
    for eid in ("MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0"):
        r, f, v = benchmark_protocol(eid)
        print(f"{eid}: benchmark.py protocol: reset {r:.2f} ms, render (rgb_array, tile 32) {f:,.0f} FPS, agent view (RGBImgPartialObs, tile 8) {v:,.0f} FPS")
