"""CPU: the debugging surface (minigrid_dynamicprogramming_b200/debug.py: hash, pprint_grid, agent_sees) against what
the reference returned for real states (tests/golden/debug.json, recorded by oracle/gen_golden_debug.py).  agent_sees
needs the env's observation: the oracle's gen_obs supplies it here, the engine's own image on the GPU (test_gpu_parity)."""
import json
import os

import numpy as np

from helpers import GOLDEN, O
from minigrid_dynamicprogramming_b200 import debug as D


def _style(pos_repr):
    return "np_tuple" if pos_repr.startswith("(np.") else "array" if pos_repr.startswith("[") else "tuple"


def test_hash_pprint_agent_sees_match_reference():
    with open(os.path.join(GOLDEN, "debug.json")) as f:
        gold = json.load(f)
    assert len(gold) == 7
    nsees = 0
    for eid, snaps in gold.items():
        for k, s in enumerate(snaps):
            cells, st, cfg = np.array(s["cells"], np.uint8), np.array(s["st"], np.int32), np.array(s["cfg"], np.int32)
            W, H = int(cfg[0]), int(cfg[1])
            pos, d = (st[0], st[1]), int(st[2])
            assert D.state_hash(cells, W, H, pos, d, pos_repr=_style(s["pos_repr"])) == s["hash"], (eid, k)
            assert D.state_hash(cells, W, H, pos, d, size=8, pos_repr=_style(s["pos_repr"])) == s["hash8"], (eid, k)
            assert D.pprint_grid(cells, W, H, pos, d) == s["pprint"], (eid, k)
            image, _ = O.OracleEnv(cells, st, cfg).gen_obs()
            for x, y, want in s["sees"]:
                assert D.agent_sees(cells, W, H, pos, d, image, x, y) == want, (eid, k, x, y)
                nsees += want
    assert nsees > 100


def test_grid_encode_decode_roundtrip():
    rng = np.random.default_rng(0)
    W, H = 9, 6
    cells = np.zeros((4, W * H), np.uint8)
    cells[0] = rng.integers(1, 10, W * H)
    cells[1] = rng.integers(0, 6, W * H)
    enc = D.grid_encode(cells, W, H)
    assert enc.shape == (W, H, 3) and enc[3, 2, 0] == cells[0, 2 * W + 3] and enc[3, 2, 1] == cells[1, 2 * W + 3]
    types, vis = D.decode_image(enc)
    assert vis.all() and ((types == 0) == (enc[..., 0] == 1)).all()
