"""Where does a sweep of the cooperative solve spend its time?  Needs the instrumentation build:
    python -m minigrid_dynamicprogramming_b200.build --force -DMGDP_TIMING=1 -o=variants/timing.so
    MGB200_LIB=$PWD/variants/timing.so python tools/solve_timing.py
Per sweep k (globaltimer, ns): first / last CTA arriving at the grid barrier, first CTA released."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from minigrid_dynamicprogramming_b200 import _capi, dp

m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", dp.ALL_ACTIONS, seed=0, max_states=1 << 25)
dp.value_iteration(m)
torch.cuda.synchronize()
_capi.lib.mgdp_debug_timing.argtypes = [C.c_void_p, C.c_int32]
assert _capi.lib.mgdp_debug_timing(None, 1) == 0
_, _, k = dp.value_iteration(m)
torch.cuda.synchronize()
t = np.zeros((3, 128), np.uint64)
assert _capi.lib.mgdp_debug_timing(t.ctypes.data, 0) == 0
first, last, rel = (t[i, :k].astype(np.int64) for i in range(3))
work_first = first[1:] - rel[:-1]          # release of barrier k-1 -> first arrival at barrier k
work_last = last[1:] - rel[:-1]            # ... -> last arrival
bar = rel - last                           # last arrival -> first release
period = rel[1:] - rel[:-1]
print(f"{k} sweeps; medians over sweeps 5..{k - 1} (us): sweep period {np.median(period[5:]) / 1e3:.1f}, "
      f"fastest CTA done after {np.median(work_first[5:]) / 1e3:.1f}, slowest after {np.median(work_last[5:]) / 1e3:.1f}, "
      f"barrier (last arrival -> first release) {np.median(bar[5:]) / 1e3:.1f}")
