"""Time of one Bellman sweep (mgdp_sweep) over the KeyCorridorS6R3 all-action space; A/B tool (MGB200_LIB=variant)."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from minigrid_dynamicprogramming_b200 import _capi, dp

m = dp.enumerate_mdp("MiniGrid-KeyCorridorS6R3-v0", dp.ALL_ACTIONS, seed=0, max_states=1 << 25)
V = torch.zeros(m.S_pad, dtype=torch.float32, device="cuda")
Vn = torch.zeros_like(V)
resid = torch.zeros(1, dtype=torch.float32, device="cuda")
st = C.c_void_p(torch.cuda.current_stream().cuda_stream)


def sweep():
    global V, Vn
    _capi.check(_capi.lib.mgdp_sweep(m.succ.data_ptr(), m.S_pad, m.A, 0, m.S, V.data_ptr(), Vn.data_ptr(), None, None, 0.99,
                                     resid.data_ptr(), st), "mgdp_sweep")
    V, Vn = Vn, V


for _ in range(10):
    sweep()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(40):
    sweep()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 40
chk = int(V[:m.S].view(torch.int32).to(torch.int64).sum())        # bit pattern checksum after 50 sweeps: equal across variants
print(f"{os.environ.get('MGB200_LIB', 'default')}: {ms*1e3:.1f} us/sweep  {m.S/ms/1e6:.1f} G backups/s  {m.S*36/ms/1e6:.0f} GB/s  chk {chk}")
torch.cuda.synchronize()
import time
dp.value_iteration(m)
torch.cuda.synchronize()
t0 = time.perf_counter()
_, _, k = dp.value_iteration(m)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
print(f"   cooperative solve: {k} sweeps in {dt*1e3:.2f} ms = {dt/k*1e6:.1f} us/sweep")
