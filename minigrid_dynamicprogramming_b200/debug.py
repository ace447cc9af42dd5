"""State digests and the debugging surface of MiniGridEnv on downloaded engine state — host side.

  grid_encode    Grid.encode()           core/grid.py:244-268
  decode_image   Grid.decode()           core/grid.py:270-289
  state_hash     MiniGridEnv.hash()      minigrid_env.py:159-169
  pprint_grid    MiniGridEnv.pprint_grid minigrid_env.py:175-229
  view_coords / agent_sees               minigrid_env.py:421-446, 481-518

cells: uint8[4, >= W*H] planes (type, color, state, aux) as mg_get_state returns them for one env.
"""
from __future__ import annotations

import hashlib

import numpy as np

DIR_TO_VEC = ((1, 0), (0, 1), (-1, 0), (0, -1))           # core/constants.py:49-58
_OBJ_STR = {2: "W", 3: "F", 4: "D", 5: "K", 6: "A", 7: "B", 8: "G", 9: "V"}
_COLOR_INITIAL = "RGBPYG"                                   # red green blue purple yellow grey -> color[0].upper()
_DIR_STR = ">V<^"


def grid_encode(cells, W, H):
    """uint8[W, H, 3], indexed [i][j]."""
    c = np.asarray(cells)[:3, : W * H].reshape(3, H, W)
    return np.ascontiguousarray(c.transpose(2, 1, 0))


def decode_image(image):
    """Grid.decode: (types int[W, H] with 0 where the decoded object is None, vis_mask bool[W, H])."""
    image = np.asarray(image)
    t = image[..., 0].astype(np.int64)
    vis = t != 0
    types = np.where((t == 0) | (t == 1), 0, t)             # WorldObj.decode returns None for unseen / empty
    return types, vis


def _pos_str(pos, pos_repr):
    x, y = int(pos[0]), int(pos[1])
    if pos_repr == "tuple":
        return f"({x}, {y})"
    if pos_repr == "np_tuple":
        return f"(np.int64({x}), np.int64({y}))"
    if pos_repr == "array":
        return f"[{x} {y}]"
    raise ValueError("pos_repr must be 'tuple', 'np_tuple' or 'array'")


def state_hash(cells, W, H, agent_pos, agent_dir, size=16, pos_repr="np_tuple"):
    """MiniGridEnv.hash.  The reference hashes str(self.agent_pos), whose text depends on the Python type the env
    happens to hold — a tuple of ints (Empty before the agent first moves), a tuple of numpy ints (after place_agent
    or a forward move; printed 'np.int64(3)' by numpy >= 2) or an ndarray (Crossing's start) — so the digest is not
    a function of the state alone; `pos_repr` selects the spelling."""
    h = hashlib.sha256()
    for item in (str(grid_encode(cells, W, H).tolist()), _pos_str(agent_pos, pos_repr), str(int(agent_dir))):
        h.update(item.encode("utf8"))
    return h.hexdigest()[:size]


def pprint_grid(cells, W, H, agent_pos, agent_dir):
    c = np.asarray(cells)
    lines = []
    for j in range(H):
        row = ""
        for i in range(W):
            if i == agent_pos[0] and j == agent_pos[1]:
                row += 2 * _DIR_STR[agent_dir]
                continue
            t, col, s = (int(v) for v in c[:3, j * W + i])
            if t == 1:
                row += "  "
            elif t == 4:
                row += "__" if s == 0 else ("L" if s == 2 else "D") + _COLOR_INITIAL[col]
            else:
                row += _OBJ_STR[t] + _COLOR_INITIAL[col]
        lines.append(row)
    return "\n".join(lines)


def view_coords(agent_pos, agent_dir, i, j, view=7):
    """get_view_coords: may lie outside [0, view)."""
    ax, ay = int(agent_pos[0]), int(agent_pos[1])
    dx, dy = DIR_TO_VEC[agent_dir]
    rx, ry = -dy, dx
    tx = ax + dx * (view - 1) - rx * (view // 2)
    ty = ay + dy * (view - 1) - ry * (view // 2)
    lx, ly = i - tx, j - ty
    return rx * lx + ry * ly, -(dx * lx + dy * ly)


def agent_sees(cells, W, H, agent_pos, agent_dir, image, x, y, view=7):
    """MiniGridEnv.agent_sees for a non-empty world cell (x, y), given the env's current observation image."""
    vx, vy = view_coords(agent_pos, agent_dir, x, y, view)
    if vx < 0 or vy < 0 or vx >= view or vy >= view:
        return False
    world_t = int(np.asarray(cells)[0, y * W + x])
    assert world_t != 1, "agent_sees is defined for non-empty cells (minigrid_env.py:516)"
    types, _ = decode_image(image)
    return bool(types[vx, vy] != 0 and types[vx, vy] == world_t)
