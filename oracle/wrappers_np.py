"""numpy restatement of the reference's observation / reward wrappers (minigrid/wrappers.py) on the oracle's
flat state (cells uint8[4, H*W], st int32[8], cfg int32[6]).  TEST INFRASTRUCTURE ONLY (same rules as oracle.py).

Parity pin: tests/golden/wrappers.npz, recorded from the unmodified reference by oracle/gen_golden_wrappers.py;
tests/test_oracle_golden.py compares every function below with it.
"""
from __future__ import annotations

import math

import numpy as np

N_TYPES, N_COLORS, N_STATES = 11, 6, 3          # len(OBJECT_TO_IDX), len(COLOR_TO_IDX), len(STATE_TO_IDX)


def fully_obs(cells, cfg, st):
    """FullyObsWrapper.observation (wrappers.py:415-423): Grid.encode() is indexed [i][j] (core/grid.py:252-266)."""
    W, H = int(cfg[0]), int(cfg[1])
    out = np.zeros((W, H, 3), np.uint8)
    for i in range(W):
        for j in range(H):
            out[i, j] = cells[:3, j * W + i]
    out[st[0], st[1]] = (10, 0, st[2])            # OBJECT_TO_IDX["agent"], COLOR_TO_IDX["red"], agent_dir
    return out


def onehot_obs(image):
    """OneHotPartialObsWrapper.observation (wrappers.py:269-287) on an image uint8[..., 3]."""
    out = np.zeros(image.shape[:-1] + (N_TYPES + N_COLORS + N_STATES,), np.uint8)
    for idx in np.ndindex(*image.shape[:-1]):
        t, c, s = image[idx]
        out[idx + (t,)] = 1
        out[idx + (N_TYPES + c,)] = 1
        out[idx + (N_TYPES + N_COLORS + s,)] = 1
    return out


def symbolic_obs(cells, cfg, st):
    """SymbolicObsWrapper.observation (wrappers.py:750-766): int64[W, H, 3] = (x, y, type index or -1 for None)."""
    W, H = int(cfg[0]), int(cfg[1])
    out = np.zeros((W, H, 3), np.int64)
    for i in range(W):
        for j in range(H):
            t = int(cells[0, j * W + i])
            out[i, j] = (i, j, -1 if t == 1 else t)      # empty list slot -> -1 (there is no "empty" object)
    out[st[0], st[1], 2] = 10
    return out


def mission_onehot(mission, max_len=96):
    """The cached strArray of FlatObsWrapper (wrappers.py:585-613): float32[max_len, 28]."""
    assert len(mission) <= max_len
    arr = np.zeros((max_len, 28), np.float32)
    for k, ch in enumerate(mission.lower()):
        if "a" <= ch <= "z":
            n = ord(ch) - ord("a")
        elif ch == " ":
            n = 26
        elif ch == ",":
            n = 27
        else:
            raise ValueError(f"Character {ch} is not available in mission string.")
        arr[k, n] = 1
    return arr


def flat_obs(image, mission, max_len=96):
    """FlatObsWrapper.observation (wrappers.py:615-617): float32[147 + 28*max_len]."""
    return np.concatenate((image.flatten(), mission_onehot(mission, max_len).flatten()))


MINIGRID_WORDS = (["red", "green", "blue", "yellow", "purple", "grey"]
                  + ["unseen", "empty", "wall", "floor", "box", "key", "ball", "door", "goal", "agent", "lava"]
                  + ["pick", "avoid", "get", "find", "put", "use", "open", "go", "fetch", "reach", "unlock", "traverse"]
                  + ["up", "the", "a", "at", ",", "square", "and", "then", "to", "of", "rooms", "near", "opening",
                     "must", "you", "matching", "end", "hallway", "object", "from", "room"])


def mission_words(mission, max_words=50):
    """DictObservationSpaceWrapper.observation (wrappers.py:527-550): word indices + 1, zero padded."""
    d = {w: i for i, w in enumerate(MINIGRID_WORDS)}
    idx = [d[w] + 1 for w in mission.replace(",", " , ").split()]
    assert len(idx) < max_words
    return np.array(idx + [0] * (max_words - len(idx)), np.int64)


class BonusCounter:
    """ActionBonus (kind='action', wrappers.py:98-121) / PositionBonus (kind='position', :159-187) for ONE env."""

    def __init__(self, kind):
        self.kind, self.counts = kind, {}

    def bonus(self, st_after, action):
        key = (int(st_after[0]), int(st_after[1]), int(st_after[2]), int(action)) if self.kind == "action" \
            else (int(st_after[0]), int(st_after[1]))
        n = self.counts.get(key, 0) + 1
        self.counts[key] = n
        return 1 / math.sqrt(n)


def goal_position(cells, cfg):
    """The tuple DirectionObsWrapper.reset caches (wrappers.py:699-709): first Goal in grid order, (int(idx / height), idx % width)."""
    W, H = int(cfg[0]), int(cfg[1])
    idx = [k for k in range(W * H) if cells[0, k] == 8]
    return (int(idx[0] / H), idx[0] % W) if idx else None


def goal_direction(goal_pos, st, type="slope"):
    """DirectionObsWrapper.observation (wrappers.py:714-725)."""
    with np.errstate(divide="ignore", invalid="ignore"):
        slope = np.divide(goal_pos[1] - int(st[1]), goal_pos[0] - int(st[0]))
    return float(np.arctan(slope)) if type == "angle" else float(slope)


def stochastic_action(action, uniform, replacement, prob, random_action):
    """StochasticActionWrapper.action (wrappers.py:786-796) with the two draws made explicit."""
    if uniform < prob:
        return action
    return replacement if random_action is None else random_action
