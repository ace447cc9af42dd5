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
    return O.OracleEnv(layout.cells, st, cfg)


def oracle_env_from_seed(eid, seed):
    return oracle_env(L.generate(eid, seed))
