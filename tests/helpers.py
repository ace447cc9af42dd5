"""Shared helpers for the parity tests (oracle side)."""
import os

import numpy as np

from oracle import oracle as O
from minigrid_dynamicprogramming_b200 import levelgen as L

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CONFIG_ENVS = ["MiniGrid-Empty-8x8-v0", "MiniGrid-DoorKey-8x8-v0", "MiniGrid-DoorKey-16x16-v0",
               "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0", "MiniGrid-KeyCorridorS6R3-v0"]


def short(eid):
    return eid.replace("MiniGrid-", "").replace("-v0", "")


def golden(name):
    return np.load(os.path.join(GOLDEN, name))


def oracle_env(layout):
    cfg = O.make_cfg(layout.width, layout.height, layout.max_steps, layout.see_through_walls, layout.rule,
                     layout.agent_view_size)
    st = np.array([*layout.agent_pos, layout.agent_dir, 0, 0, 0, 0, 0], dtype=np.int32)
    return O.OracleEnv(layout.cells, st, cfg, dyn=layout.dyn)


def oracle_env_from_seed(eid, seed):
    return oracle_env(L.generate(eid, seed))


def oracle_batch(layouts, plane_stride=None):
    """Flat oracle arrays for a list of layouts: cells uint8[n, 4*ps], st int32[n, 8], class ids, cfg table."""
    ps = plane_stride or max(((l.width * l.height + 15) // 16) * 16 for l in layouts)
    n = len(layouts)
    cells = np.zeros((n, 4, ps), np.uint8)
    cells[:, 0, :] = 1
    st = np.zeros((n, 8), np.int32)
    keys, cls = [], np.zeros(n, np.uint8)
    for i, l in enumerate(layouts):
        k = (l.width, l.height, l.max_steps, int(l.see_through_walls), l.rule, l.agent_view_size)
        if k not in keys:
            keys.append(k)
        cls[i] = keys.index(k)
        cells[i, :, : l.width * l.height] = l.cells
        st[i, :3] = (*l.agent_pos, l.agent_dir)
    # the oracle indexes planes with W*H, the engine with plane_stride: repack per env for the oracle
    return cells, st, cls, np.array(keys, np.int32), ps


def oracle_rollout(layouts, actions, autoreset=False, pool=None, nthreads=8, final_obs=False):
    """Runs the C oracle on `layouts` (one env each).  The oracle wants planes packed at W*H, so envs are
    packed to the oracle's own stride (max W*H) and classes must share W*H when autoreset mixes them."""
    n = len(layouts)
    cellsz = max(l.width * l.height for l in (pool or layouts))
    stride = 4 * cellsz

    def pack(ls):
        c = np.zeros((len(ls), stride), np.uint8)
        s = np.zeros((len(ls), 8), np.int32)
        for i, l in enumerate(ls):
            m = l.width * l.height
            c[i, : 4 * m] = l.cells.reshape(-1)
            s[i, :3] = (*l.agent_pos, l.agent_dir)
        return c, s

    keys = []

    def cls_of(ls):
        out = np.zeros(len(ls), np.uint8)
        for i, l in enumerate(ls):
            k = (l.width, l.height, l.max_steps, int(l.see_through_walls), l.rule, l.agent_view_size)
            if k not in keys:
                keys.append(k)
            out[i] = keys.index(k)
        return out

    def dyn_of(ls):          # DynamicObstacles: generator + obstacle records (levelgen.dyn_record)
        if not any(l.dyn is not None for l in ls):
            return None
        d = np.zeros((len(ls), 64), np.uint8)
        for i, l in enumerate(ls):
            if l.dyn is not None:
                d[i] = l.dyn
        return d

    cells, st = pack(layouts)
    cls = cls_of(layouts)
    kw = {}
    dyn = dyn_of(layouts)
    if autoreset:
        pc, pst = pack(pool)
        kw = dict(autoreset=int(autoreset), pool_cells=pc, pool_st=pst, pool_class=cls_of(pool),     # 1 same step, 2 next step
                  layout_id=np.arange(n, dtype=np.int32), pool_dyn=dyn_of(pool))
    fin = None
    if final_obs:
        V = layouts[0].agent_view_size
        fin = np.zeros((len(actions), n, V, V, 3), np.uint8)
        kw["final_image"] = fin
    cfgs = np.array(keys, np.int32)
    img, d, r, te, tr = O.rollout(cells, st, cls, cfgs, np.ascontiguousarray(actions, np.uint8), nthreads=nthreads, dyn=dyn, **kw)
    return dict(image=img, direction=d, reward=r, terminated=te, truncated=tr, st=st, cells=cells, dyn=dyn, final_image=fin)


# env ids of the N3 row: part a (tests/golden/n3.npz) and part b (n3b.npz), oracle/gen_golden_n3.py
N3_IDS = ["MiniGrid-Unlock-v0", "MiniGrid-UnlockPickup-v0", "MiniGrid-BlockedUnlockPickup-v0", "MiniGrid-DistShift1-v0",
          "MiniGrid-DistShift2-v0", "MiniGrid-LavaGapS5-v0", "MiniGrid-LavaGapS6-v0", "MiniGrid-LavaGapS7-v0",
          "MiniGrid-Fetch-5x5-N2-v0", "MiniGrid-Fetch-6x6-N2-v0", "MiniGrid-Fetch-8x8-N3-v0", "MiniGrid-GoToDoor-5x5-v0",
          "MiniGrid-GoToDoor-6x6-v0", "MiniGrid-GoToDoor-8x8-v0", "MiniGrid-GoToObject-6x6-N2-v0", "MiniGrid-GoToObject-8x8-N2-v0",
          "MiniGrid-RedBlueDoors-6x6-v0", "MiniGrid-RedBlueDoors-8x8-v0"]
N3B_IDS = ["MiniGrid-ObstructedMaze-1Dl-v0", "MiniGrid-ObstructedMaze-1Dlh-v0", "MiniGrid-ObstructedMaze-1Dlhb-v0",
           "MiniGrid-ObstructedMaze-2Dl-v0", "MiniGrid-ObstructedMaze-2Dlh-v0", "MiniGrid-ObstructedMaze-2Dlhb-v0",
           "MiniGrid-ObstructedMaze-1Q-v0", "MiniGrid-ObstructedMaze-2Q-v0", "MiniGrid-ObstructedMaze-Full-v0",
           "MiniGrid-ObstructedMaze-2Dlhb-v1", "MiniGrid-ObstructedMaze-1Q-v1", "MiniGrid-ObstructedMaze-2Q-v1",
           "MiniGrid-ObstructedMaze-Full-v1", "MiniGrid-LockedRoom-v0", "MiniGrid-MultiRoom-N2-S4-v0", "MiniGrid-MultiRoom-N4-S5-v0",
           "MiniGrid-MultiRoom-N6-v0", "MiniGrid-Playground-v0", "MiniGrid-MemoryS17Random-v0", "MiniGrid-MemoryS13Random-v0",
           "MiniGrid-MemoryS13-v0", "MiniGrid-MemoryS11-v0", "MiniGrid-MemoryS9-v0", "MiniGrid-MemoryS7-v0",
           "MiniGrid-PutNear-6x6-N2-v0", "MiniGrid-PutNear-8x8-N3-v0"]


DYN_IDS = ["MiniGrid-Dynamic-Obstacles-5x5-v0", "MiniGrid-Dynamic-Obstacles-Random-5x5-v0", "MiniGrid-Dynamic-Obstacles-6x6-v0",
           "MiniGrid-Dynamic-Obstacles-Random-6x6-v0", "MiniGrid-Dynamic-Obstacles-8x8-v0", "MiniGrid-Dynamic-Obstacles-16x16-v0"]


def n3_golden(eid):
    return golden("n3b.npz" if eid in N3B_IDS else "n3.npz")
