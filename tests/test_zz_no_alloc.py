"""mgdp_solve_static — the no-allocation variant of the cooperative Bellman solve (include/mgb200_dp.h) — against
mgdp_solve, through the C ABI.  Collected last on purpose."""
import ctypes as C
import gc

import numpy as np
import pytest

from helpers import L

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")


def _solve_pair(dp, _capi, m, check_mem=True):
    """(V, pi, sweeps, residual bits) from mgdp_solve and from mgdp_solve_static on the same table."""
    dev, S, S_pad, A = m.device, m.S, m.S_pad, m.A
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    V1 = torch.zeros(S_pad, dtype=torch.float32, device=dev)
    Vt = torch.zeros(S_pad, dtype=torch.float32, device=dev)
    pi1 = torch.zeros(S_pad, dtype=torch.uint8, device=dev)
    r1 = torch.zeros(5, dtype=torch.int32, device=dev)
    _capi.check(_capi.lib.mgdp_solve(m.succ.data_ptr(), S_pad, A, S, V1.data_ptr(), Vt.data_ptr(), pi1.data_ptr(), None,
                                     0.99, 1e-6, 100000, r1.data_ptr(), st), "mgdp_solve")
    # inputs and outputs of the static call are allocated BEFORE it; the call itself must not allocate
    V2 = torch.zeros(S_pad, dtype=torch.float32, device=dev)
    pi2 = torch.zeros(S_pad, dtype=torch.uint8, device=dev)
    r2 = torch.full((2,), -7, dtype=torch.int32, device=dev)
    gc.collect()           # engines of earlier tests free their device memory in __del__: not in the middle of the measurement
    torch.cuda.synchronize(dev)
    free0 = torch.cuda.mem_get_info(dev)[0]
    _capi.check(_capi.lib.mgdp_solve_static(m.succ.data_ptr(), S_pad, A, S, V2.data_ptr(), pi2.data_ptr(), None,
                                            0.99, 1e-6, 100000, r2.data_ptr(), st), "mgdp_solve_static")
    torch.cuda.synchronize(dev)
    if check_mem:          # (not on the process's very first call: the driver may still be loading the module lazily)
        assert torch.cuda.mem_get_info(dev)[0] >= free0, "mgdp_solve_static took device memory"
    return (V1[:S], pi1[:S], r1[:2]), (V2[:S], pi2[:S], r2)


def test_solve_static_matches_solve_on_two_calls_in_a_row():
    """Two calls in a row on DIFFERENT tables (the second is smaller, so the shared scratch still holds the first call's
    values past its end): V, pi, sweep count and last residual bit-identical to mgdp_solve (the DP tests' bar is bit-exact);
    then the first table again."""
    from minigrid_dynamicprogramming_b200 import _capi, dp
    mdps = [dp.enumerate_mdp(L.generate("MiniGrid-DoorKey-8x8-v0", 0), dp.ALL_ACTIONS),
            dp.enumerate_mdp(L.generate("MiniGrid-Empty-8x8-v0", 0), (0, 1, 2)),
            dp.enumerate_mdp(L.generate("MiniGrid-DoorKey-8x8-v0", 0), dp.ALL_ACTIONS)]
    seen = []
    for i, m in enumerate(mdps):
        (V1, pi1, r1), (V2, pi2, r2) = _solve_pair(dp, _capi, m, check_mem=i > 0)
        assert int(r1[0]) > 5 and torch.equal(r1, r2), (r1, r2)
        assert torch.equal(V1, V2) and torch.equal(pi1, pi2)
        assert float(V2[0]) > 0
        seen.append((V2.clone(), int(r2[0])))
    assert seen[0][1] == seen[2][1] and torch.equal(seen[0][0], seen[2][0]) and seen[0][1] != seen[1][1]


def test_solve_static_policy_evaluation_and_limit():
    """pi_fixed path through the static scratch, and the size limit: more states than MGDP_STATIC_MAX_STATES is rejected
    with a message that names the limit (nothing is launched)."""
    from minigrid_dynamicprogramming_b200 import _capi, dp
    m = dp.enumerate_mdp(L.generate("MiniGrid-DoorKey-8x8-v0", 1), (0, 1, 2, 3, 5))
    V, pi, k = dp.value_iteration(m)
    Vp, kp = dp.policy_evaluation(m, pi)
    dev = m.device
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    V2 = torch.zeros(m.S_pad, dtype=torch.float32, device=dev)
    r2 = torch.zeros(2, dtype=torch.int32, device=dev)
    pf = torch.zeros(m.S_pad, dtype=torch.uint8, device=dev)
    pf[:m.S] = pi
    _capi.check(_capi.lib.mgdp_solve_static(m.succ.data_ptr(), m.S_pad, m.A, m.S, V2.data_ptr(), None, pf.data_ptr(),
                                            0.99, 1e-6, 100000, r2.data_ptr(), st), "mgdp_solve_static")
    torch.cuda.synchronize(dev)
    assert int(r2[0]) == kp and torch.equal(V2[:m.S], Vp)
    lim = int(_capi.lib.mgdp_static_max_states())
    assert lim == 1 << 20
    rc = _capi.lib.mgdp_solve_static(m.succ.data_ptr(), lim + 32, m.A, lim + 1, V2.data_ptr(), None, None,
                                     0.99, 1e-6, 10, r2.data_ptr(), st)
    assert rc == _capi.MG_EINVAL
    msg = _capi.lib.mg_last_error().decode()
    assert "MGDP_STATIC_MAX_STATES" in msg and str(lim) in msg, msg
    torch.cuda.synchronize(dev)
