"""ctypes front-end of the C oracle (oracle/mg_oracle.c, oracle/mg_oracle_dp.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  Never imported by the product package.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libmgoracle.so")

CFG_N = 6   # W, H, max_steps, see_through, rule, view
ST_N = 8    # x, y, dir, step_count, carry type/color/state/aux

_lib = None


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("mg_oracle.c", "mg_oracle_dp.c", "mg_oracle_render.c")]
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.check_call(["make", "-s", "-C", _HERE, "-B" if force else "-s"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.mgo_reward.restype = C.c_double
        _lib.mgo_reward.argtypes = [C.c_int, C.c_int]
        _lib.mgo_rollout.restype = C.c_long
        _lib.mgo_enumerate.restype = C.c_long
        _lib.mgo_space_build.restype = C.c_long
        _lib.mgo_space_lookup.restype = C.c_long
        _lib.mgo_space_lookup.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
        _lib.mgo_space_state.argtypes = [C.c_void_p, C.c_long, C.c_void_p, C.c_void_p]
        _lib.mgo_space_free.argtypes = [C.c_void_p]
        _lib.mgo_vi_sweeps.restype = C.c_int
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def make_cfg(W, H, max_steps, see_through=False, rule=0, view=7):
    return np.array([W, H, max_steps, int(see_through), rule, view], dtype=np.int32)


class OracleEnv:
    """One env in the oracle's flat representation.

    cells: uint8[4, H*W] (type, color, state, aux planes); st: int32[8]; cfg: int32[6].
    """

    def __init__(self, cells, st, cfg, dyn=None):
        self.dyn = None if dyn is None else np.ascontiguousarray(dyn, dtype=np.uint8).copy()    # DynamicObstacles: RNG + obstacle list
        self.cfg = np.ascontiguousarray(cfg, dtype=np.int32).copy()
        self.cells = np.ascontiguousarray(cells, dtype=np.uint8).copy()
        self.st = np.ascontiguousarray(st, dtype=np.int32).copy()
        self.view = int(self.cfg[5])
        # the cell the mission's target_pos refers to = the flagged cell of the INITIAL layout (gotodoor.py:125, gotoobject.py:139)
        n = int(self.cfg[0]) * int(self.cfg[1])
        assert self.cells.size == 4 * n
        self.target = int(lib().mgo_find_target(_p(self.cells), n, int(self.cfg[4])))     # A | B << 16, 0xFFFF = none

    def gen_obs(self):
        img = np.zeros((self.view, self.view, 3), dtype=np.uint8)
        lib().mgo_gen_obs(_p(self.cells), _p(self.cfg), _p(self.st), _p(img))
        return img, int(self.st[2])

    def step(self, action):
        img = np.zeros((self.view, self.view, 3), dtype=np.uint8)
        r = C.c_double(0.0)
        term = C.c_uint8(0)
        trunc = C.c_uint8(0)
        rc = lib().mgo_step_dyn(_p(self.cells), _p(self.cfg), _p(self.st), int(action), int(self.target), _p(self.dyn), _p(img),
                                C.byref(r), C.byref(term), C.byref(trunc))
        if rc != 0:
            raise ValueError(f"Unknown action: {action}")
        return img, int(self.st[2]), float(r.value), bool(term.value), bool(trunc.value)


def reward(step_count, max_steps):
    return float(lib().mgo_reward(int(step_count), int(max_steps)))


def reset_stride(n, P):
    """The engine's autoreset stride (include/mgb200.h mg_reset_stride): n mod P moved up to the next value coprime to P."""
    import math
    if P <= 1:
        return 0
    s = n % P or 1
    while math.gcd(s, P) != 1:
        s += 1
    return s % P


def rollout(cells, st, class_id, cfgs, actions, *, autoreset=False, pool_cells=None, pool_st=None,
            pool_class=None, layout_id=None, nthreads=1, store_image=True, dyn=None, pool_dyn=None, final_image=None,
            rstride=None):
    """Batched rollout; cells uint8[n, stride], st int32[n, 8], actions uint8[T, n].  In-place on cells/st (and on dyn
    uint8[n, 64], the RNG + obstacle records of DynamicObstacles envs)."""
    n, stride = cells.shape
    T = actions.shape[0]
    cfgs = np.ascontiguousarray(cfgs, dtype=np.int32).reshape(-1, CFG_N)
    V = int(cfgs[0, 5])
    image = np.zeros((T, n, V, V, 3), dtype=np.uint8) if store_image else None
    direction = np.zeros((T, n), dtype=np.uint8)
    rew = np.zeros((T, n), dtype=np.float64)
    term = np.zeros((T, n), dtype=np.uint8)
    trunc = np.zeros((T, n), dtype=np.uint8)
    P = 0 if pool_cells is None else pool_cells.shape[0]
    if layout_id is None:
        layout_id = np.arange(n, dtype=np.int32)
    class_id = np.ascontiguousarray(class_id, dtype=np.uint8)
    bad = lib().mgo_rollout(
        C.c_int(n), C.c_int(T), C.c_int(nthreads), _p(cells), C.c_long(stride), _p(st), _p(class_id), _p(cfgs),
        _p(np.ascontiguousarray(actions, dtype=np.uint8)), C.c_int(int(autoreset)), C.c_int(P), _p(pool_cells),
        _p(pool_st), _p(pool_class), _p(layout_id), _p(class_id), _p(image), _p(direction), _p(rew), _p(term),
        _p(trunc), _p(dyn), _p(pool_dyn), _p(final_image), C.c_int(reset_stride(n, P) if rstride is None else int(rstride)))
    if bad:
        raise ValueError(f"Unknown action at flat index {bad - 1}")
    return image, direction, rew, term, trunc


WIN, LOSE = -1, -2


def enumerate_mdp(cells0, cfg, st0, actions=(0, 1, 2, 3, 4, 5, 6), max_states=1 << 20):
    """Serial FIFO BFS through the oracle step.  Returns (succ int32[S, A], info uint8[S, 8])."""
    acts = np.asarray(actions, dtype=np.int32)
    A = len(acts)
    succ = np.empty((max_states, A), dtype=np.int32)
    info = np.empty((max_states, 8), dtype=np.uint8)
    cells0 = np.ascontiguousarray(cells0, dtype=np.uint8)
    cfg = np.ascontiguousarray(cfg, dtype=np.int32)
    st0 = np.ascontiguousarray(st0, dtype=np.int32)
    S = lib().mgo_enumerate(_p(cells0), _p(cfg), _p(st0), _p(acts), C.c_int(A), C.c_long(max_states),
                            _p(succ), _p(info))
    if S < 0:
        raise RuntimeError(f"mgo_enumerate failed: {S}")
    return succ[:S].copy(), info[:S].copy()


class StateSpace:
    """A finished serial BFS kept alive: maps full states (cells, st) <-> BFS ids (oracle/gen_golden_big.py)."""

    def __init__(self, cells0, cfg, st0, actions=(0, 1, 2, 3, 4, 5, 6), max_states=1 << 20):
        acts = np.asarray(actions, dtype=np.int32)
        self.A = len(acts)
        succ = np.empty((max_states, self.A), dtype=np.int32)
        self._h = C.c_void_p()
        cells0 = np.ascontiguousarray(cells0, dtype=np.uint8)
        self.ncell = cells0.size // 4
        S = lib().mgo_space_build(_p(cells0), _p(np.ascontiguousarray(cfg, dtype=np.int32)), _p(np.ascontiguousarray(st0, dtype=np.int32)),
                                  _p(acts), C.c_int(self.A), C.c_long(max_states), _p(succ), None, C.byref(self._h))
        if S < 0:
            raise RuntimeError(f"mgo_space_build failed: {S}")
        self.S = int(S)
        self.succ = succ[:S]

    def lookup(self, cells, st):
        return int(lib().mgo_space_lookup(self._h, _p(np.ascontiguousarray(cells, dtype=np.uint8)), _p(np.ascontiguousarray(st, dtype=np.int32))))

    def state(self, i):
        cells = np.zeros((4, self.ncell), np.uint8)
        st = np.zeros(ST_N, np.int32)
        if lib().mgo_space_state(self._h, C.c_long(int(i)), _p(cells), _p(st)):
            raise IndexError(i)
        return cells, st

    def close(self):
        if self._h:
            lib().mgo_space_free(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def value_iteration(succ, gamma=0.99, tol=1e-6, max_sweeps=100000, pi_fixed=None):
    """fp32 Jacobi VI (or policy evaluation when pi_fixed is given).  Returns (V, pi, sweeps, resid_hist)."""
    succ = np.ascontiguousarray(succ, dtype=np.int32)
    S, A = succ.shape
    V = np.zeros(S, dtype=np.float32)
    pi = np.zeros(S, dtype=np.uint8)
    hist = np.zeros(max_sweeps, dtype=np.float32)
    pf = None if pi_fixed is None else np.ascontiguousarray(pi_fixed, dtype=np.uint8)
    k = lib().mgo_vi_sweeps(_p(succ), C.c_long(S), C.c_int(A), C.c_float(gamma), C.c_float(tol),
                            C.c_int(max_sweeps), _p(V), _p(pi), _p(pf), _p(hist))
    return V, pi, int(k), hist[:k].copy()


def value_iteration_numpy(succ, gamma=0.99, tol=1e-6, max_sweeps=100000):
    """The same Jacobi iteration written with numpy gathers (the 'numpy DP loop' CPU baseline)."""
    succ = np.asarray(succ)
    S, A = succ.shape
    g = np.float32(gamma)
    win, lose = succ == WIN, succ == LOSE
    idx = np.where(succ < 0, 0, succ)
    V = np.zeros(S, dtype=np.float32)
    for k in range(1, max_sweeps + 1):
        q = g * V[idx]
        q[win] = np.float32(1.0)
        q[lose] = np.float32(0.0)
        Vn = q.max(axis=1)
        resid = float(np.max(np.abs(Vn - V))) if S else 0.0
        pi = q.argmax(axis=1).astype(np.uint8)
        V = Vn
        if resid < tol:
            break
    return V, pi, k


# ---------------------------------------------------------------- RGB rendering (mg_oracle_render.c)

def render_tile(type_, color, state, agent_dir=None, highlight=False, tile_size=8):
    """Grid.render_tile (core/grid.py:145-198) as uint8[ts, ts, 3]; type_ == 1 is None, agent_dir None = no agent."""
    out = np.zeros((tile_size, tile_size, 3), dtype=np.uint8)
    lib().mgo_render_tile(int(type_), int(color), int(state), -1 if agent_dir is None else int(agent_dir),
                          int(bool(highlight)), int(tile_size), _p(out))
    return out


def render_pov(cells, cfg, st, tile_size=8):
    """get_pov_render (minigrid_env.py:647-661): uint8[7*ts, 7*ts, 3]."""
    cfg = np.ascontiguousarray(cfg, dtype=np.int32)
    V = int(cfg[5])
    out = np.zeros((V * tile_size, V * tile_size, 3), dtype=np.uint8)
    lib().mgo_render_pov(_p(np.ascontiguousarray(cells, dtype=np.uint8)), _p(cfg),
                         _p(np.ascontiguousarray(st, dtype=np.int32)), int(tile_size), _p(out))
    return out


def render_full(cells, cfg, st, highlight=True, tile_size=8):
    """get_full_render (minigrid_env.py:663-709): uint8[H*ts, W*ts, 3]."""
    cfg = np.ascontiguousarray(cfg, dtype=np.int32)
    W, H = int(cfg[0]), int(cfg[1])
    out = np.zeros((H * tile_size, W * tile_size, 3), dtype=np.uint8)
    lib().mgo_render_full(_p(np.ascontiguousarray(cells, dtype=np.uint8)), _p(cfg),
                          _p(np.ascontiguousarray(st, dtype=np.int32)), int(bool(highlight)), int(tile_size), _p(out))
    return out


def finite_horizon_numpy(succ, max_steps, gamma=1.0, return_policy=False):
    """Backward induction over step_count with the true reward (minigrid_env.py:235-240, 549-551, 582-583), fp64.
    succ int32[S, A] (-1 WIN, -2 LOSE).  V_t(s) = max_a {WIN: 1 - 0.9*((t+1)/T); LOSE or t+1 >= T: 0; else gamma*V_{t+1}(s')}."""
    succ = np.asarray(succ)
    S, A = succ.shape
    V = np.zeros(S, np.float64)
    pis = np.zeros((max_steps, S), np.uint8) if return_policy else None
    win, lose = succ == -1, succ == -2
    safe = np.where(succ >= 0, succ, 0)
    for t in range(max_steps - 1, -1, -1):
        r = 1 - 0.9 * ((t + 1) / max_steps)
        if t + 1 >= max_steps:
            q = np.where(win, r, 0.0)
        else:
            q = np.where(win, r, np.where(lose, 0.0, gamma * V[safe]))
        V = q.max(axis=1)
        if return_policy:
            pis[t] = q.argmax(axis=1)
    return (V, pis) if return_policy else V


def step_no_death(env: OracleEnv, action, no_death_types, death_cost):
    """NoDeath.step (minigrid/wrappers.py:842-865) around OracleEnv.step; no_death_types = iterable of type indices."""
    W, H = int(env.cfg[0]), int(env.cfg[1])
    dx, dy = ((1, 0), (0, 1), (-1, 0), (0, -1))[int(env.st[2])]
    fx, fy = int(env.st[0]) + dx, int(env.st[1]) + dy
    front_t = int(env.cells.reshape(4, -1)[0, fy * W + fx]) if (0 <= fx < W and 0 <= fy < H) else 2
    going = action == 2 and front_t != 1 and front_t in no_death_types
    img, d, r, te, tr = env.step(action)
    cur_t = int(env.cells.reshape(4, -1)[0, int(env.st[1]) * W + int(env.st[0])])
    in_death = cur_t != 1 and cur_t in no_death_types
    if te and (going or in_death):
        te = False
        r += death_cost
    return img, d, r, te, tr
