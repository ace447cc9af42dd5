"""The reference's wrappers (minigrid/wrappers.py) over a BatchedMiniGrid: same names, same constructor arguments,
same observation contents — for all N envs at once, computed by kernels of libmgb200.so (csrc/mg_wrappers.cu,
csrc/mg_render.cu) on the device.  Wrappers compose like the reference's (`FlatObsWrapper(env)`,
`ImgObsWrapper(RGBImgPartialObsWrapper(env))`, `ActionBonus(PositionBonus(env))` ...).

Every `step(actions)` takes int[N] actions and returns batched (obs, reward[N], terminated[N], truncated[N], info);
observations are device tensors with a leading N axis.  With autoreset="same_step" a finished env's observation is the
first one of its next episode (env.py); the wrappers transform whatever state the engine holds after the step.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _capi


class Wrapper:
    """gymnasium.Wrapper for the batched engine: forwards everything it does not override."""

    def __init__(self, env):
        self.env = env

    def __getattr__(self, name):
        if name.startswith("_"):
            raise AttributeError(name)
        return getattr(self.env, name)

    @property
    def unwrapped(self):
        return getattr(self.env, "unwrapped", self.env)

    def reset(self, *, seed=None, options=None):
        return self.env.reset(seed=seed, options=options)

    def step(self, actions):
        return self.env.step(actions)

    def _h(self):
        return self.unwrapped._handle

    def _s(self):
        return self.unwrapped._stream()


class ObservationWrapper(Wrapper):
    def reset(self, *, seed=None, options=None):
        obs, info = self.env.reset(seed=seed, options=options)
        return self.observation(obs), info

    def step(self, actions):
        obs, r, te, tr, info = self.env.step(actions)
        return self.observation(obs), r, te, tr, info

    def observation(self, obs):
        raise NotImplementedError


class ReseedWrapper(Wrapper):
    """wrappers.py:17-67.  Every reset() regenerates the batch from the next entry of `seeds`: entry s seeds env e with
    s + e (the SyncVectorEnv convention BatchedMiniGrid.reset(seed=int) follows); an entry may also be a list of N seeds."""

    def __init__(self, env, seeds=(0,), seed_idx=0):
        super().__init__(env)
        self.seeds, self.seed_idx = list(seeds), seed_idx

    def reset(self, *, seed=None, options=None):
        seed = self.seeds[self.seed_idx]
        self.seed_idx = (self.seed_idx + 1) % len(self.seeds)
        return self.env.reset(seed=seed, options=options)


class NoDeath(Wrapper):
    """wrappers.py:799-872: no death in cells of the given types (e.g. ("lava",)); the agent pays `death_cost` instead.
    Implemented as a mode of the step kernel (mg_set_no_death), so a cancelled termination does not autoreset either."""

    def __init__(self, env, no_death_types, death_cost=-1.0):
        from .constants import OBJECT_TO_IDX
        assert "goal" not in no_death_types, "goal cannot be a death cell"
        super().__init__(env)
        self.death_cost, self.no_death_types = death_cost, tuple(no_death_types)
        mask = 0
        for t in self.no_death_types:
            mask |= 1 << OBJECT_TO_IDX[t]
        _capi.check(_capi.lib.mg_set_no_death(self._h(), mask, float(death_cost)), "mg_set_no_death")


class _Bonus(Wrapper):
    KIND, SLOTS = 0, 1

    def __init__(self, env):
        super().__init__(env)
        u = self.unwrapped
        self.counts = torch.zeros((u.num_envs, u.plane_stride * self.SLOTS), dtype=torch.int32, device=u.device)

    def step(self, actions):
        u = self.unwrapped
        a = torch.as_tensor(actions).to(device=u.device, dtype=torch.uint8).contiguous()
        obs, r, te, tr, info = self.env.step(a)
        r = r.clone() if r.data_ptr() == u._reward.data_ptr() else r        # do not accumulate into the engine's buffer
        _capi.check(_capi.lib.mg_bonus(self._h(), self.KIND, a.data_ptr(), self.counts.data_ptr(), r.data_ptr(), self._s()),
                    "mg_bonus")
        return obs, r, te, tr, info


class ActionBonus(_Bonus):
    """wrappers.py:70-121: reward += 1/sqrt(visits of (agent_pos, agent_dir, action))."""
    KIND, SLOTS = 0, 28


class PositionBonus(_Bonus):
    """wrappers.py:124-187 (formerly StateBonus): reward += 1/sqrt(visits of agent_pos)."""
    KIND, SLOTS = 1, 1


class ImgObsWrapper(ObservationWrapper):
    """wrappers.py:190-217: the image is the whole observation."""

    def observation(self, obs):
        return obs["image"]


class OneHotPartialObsWrapper(ObservationWrapper):
    """wrappers.py:220-287: uint8[N, 7, 7, 20] one-hot of (type, color, state)."""

    def __init__(self, env, tile_size=8):
        super().__init__(env)
        self.tile_size = tile_size

    def observation(self, obs):
        img = obs["image"].contiguous()
        out = torch.empty(img.shape[:-1] + (20,), dtype=torch.uint8, device=img.device)
        _capi.check(_capi.lib.mg_obs_onehot(self._h(), img.data_ptr(), img.numel() // 147, out.data_ptr(), self._s()),
                    "mg_obs_onehot")
        return {**obs, "image": out}


class RGBImgObsWrapper(ObservationWrapper):
    """wrappers.py:290-335: the fully observable RGB frame, uint8[N, H*ts, W*ts, 3]."""

    def __init__(self, env, tile_size=8):
        super().__init__(env)
        self.tile_size = tile_size

    def observation(self, obs):
        u = self.unwrapped
        return {**obs, "image": u.get_frame(highlight=getattr(u, "highlight", True), tile_size=self.tile_size)}


class RGBImgPartialObsWrapper(ObservationWrapper):
    """wrappers.py:337-381: the agent's view as an RGB frame, uint8[N, 7*ts, 7*ts, 3]."""

    def __init__(self, env, tile_size=8):
        super().__init__(env)
        self.tile_size = tile_size

    def observation(self, obs):
        return {**obs, "image": self.unwrapped.get_frame(tile_size=self.tile_size, agent_pov=True)}


class ViewSizeWrapper(ObservationWrapper):
    """wrappers.py:626-670: the partial observation at another (odd) agent_view_size, uint8[N, V, V, 3]."""

    def __init__(self, env, agent_view_size=7):
        super().__init__(env)
        assert agent_view_size % 2 == 1 and agent_view_size >= 3
        self.agent_view_size = agent_view_size

    def observation(self, obs):
        u, v = self.unwrapped, self.agent_view_size
        out = torch.empty((u.num_envs, v, v, 3), dtype=torch.uint8, device=u.device)
        _capi.check(_capi.lib.mg_obs_view(self._h(), v, out.data_ptr(), self._s()), "mg_obs_view")
        return {**obs, "image": out}


def _per_env_views(flat, classes, cls, shape_of):
    """Split a [N, stride] buffer into per-env views when the batch mixes grid sizes."""
    return [flat[e, :int(np.prod(shape_of(classes[c])))].view(*shape_of(classes[c])) for e, c in enumerate(cls)]


class FullyObsWrapper(ObservationWrapper):
    """wrappers.py:384-423: uint8[N, W, H, 3] grid encoding with the agent as (10, 0, dir)."""

    def observation(self, obs):
        u = self.unwrapped
        buf = torch.empty((u.num_envs, u.plane_stride * 3), dtype=torch.uint8, device=u.device)
        _capi.check(_capi.lib.mg_obs_fully(self._h(), buf.data_ptr(), self._s()), "mg_obs_fully")
        if len(u._classes) == 1:
            w, h = u._classes[0][:2]
            return {**obs, "image": buf[:, :w * h * 3].view(u.num_envs, w, h, 3)}
        cls = u.get_state(cells=False)["class_id"]
        return {**obs, "image": _per_env_views(buf, u._classes, cls, lambda c: (c[0], c[1], 3))}


class SymbolicObsWrapper(ObservationWrapper):
    """wrappers.py:728-766: int64[N, W, H, 3] = (x, y, object index or -1), 10 at the agent."""

    def observation(self, obs):
        u = self.unwrapped
        buf = torch.empty((u.num_envs, u.plane_stride * 3), dtype=torch.int64, device=u.device)
        _capi.check(_capi.lib.mg_obs_symbolic(self._h(), buf.data_ptr(), self._s()), "mg_obs_symbolic")
        if len(u._classes) == 1:
            w, h = u._classes[0][:2]
            return {**obs, "image": buf[:, :w * h * 3].view(u.num_envs, w, h, 3)}
        cls = u.get_state(cells=False)["class_id"]
        return {**obs, "image": _per_env_views(buf, u._classes, cls, lambda c: (c[0], c[1], 3))}


class DictObservationSpaceWrapper(ObservationWrapper):
    """wrappers.py:426-550: the mission becomes int64[N, max_words_in_mission] word indices (+1, zero padded).
    Missions only change when layouts do, so the table is built on the host per distinct mission and gathered
    on the device by layout id."""

    def __init__(self, env, max_words_in_mission=50, word_dict=None):
        super().__init__(env)
        self.max_words_in_mission = max_words_in_mission
        self.word_dict = word_dict if word_dict is not None else self.get_minigrid_words()
        self._pool_key, self._table, self._mid = None, None, None

    @staticmethod
    def get_minigrid_words():
        colors = ["red", "green", "blue", "yellow", "purple", "grey"]
        objects = ["unseen", "empty", "wall", "floor", "box", "key", "ball", "door", "goal", "agent", "lava"]
        verbs = ["pick", "avoid", "get", "find", "put", "use", "open", "go", "fetch", "reach", "unlock", "traverse"]
        extra = ["up", "the", "a", "at", ",", "square", "and", "then", "to", "of", "rooms", "near", "opening", "must",
                 "you", "matching", "end", "hallway", "object", "from", "room"]
        words = colors + objects + verbs + extra
        return {w: i for i, w in enumerate(words)}

    def string_to_indices(self, string, offset=1):
        out = []
        for word in string.replace(",", " , ").split():
            if word not in self.word_dict:
                raise ValueError(f"Unknown word: {word}")
            out.append(self.word_dict[word] + offset)
        return out

    def _tables(self):
        u = self.unwrapped
        if self._pool_key is not u._pool_missions:
            uniq, mid = _unique_missions(u._pool_missions)
            rows = []
            for m in uniq:
                idx = self.string_to_indices(m)
                assert len(idx) < self.max_words_in_mission
                rows.append(idx + [0] * (self.max_words_in_mission - len(idx)))
            self._table = torch.tensor(rows, dtype=torch.int64, device=u.device)
            self._mid = torch.tensor(mid, dtype=torch.int64, device=u.device)
            self._pool_key = u._pool_missions
        return self._table, self._mid

    def observation(self, obs):
        table, mid = self._tables()
        lid = self.unwrapped.layout_ids_device()
        return {**obs, "mission": table[mid[lid]]}


def _unique_missions(missions):
    index, uniq, mid = {}, [], []
    for m in missions:
        if m not in index:
            index[m] = len(uniq)
            uniq.append(m)
        mid.append(index[m])
    return uniq, mid


class FlatObsWrapper(ObservationWrapper):
    """wrappers.py:553-623: float32[N, 147 + 28*maxStrLen] = image.flatten() ++ one-hot characters of the mission."""

    def __init__(self, env, maxStrLen=96):
        super().__init__(env)
        self.maxStrLen, self.numCharCodes = maxStrLen, 28
        self._pool_key = None

    def _encode(self, mission):
        assert len(mission) <= self.maxStrLen, f"mission string too long ({len(mission)} chars)"
        arr = np.zeros((self.maxStrLen, self.numCharCodes), dtype=np.float32)
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
        return arr.reshape(-1)

    def observation(self, obs):
        u = self.unwrapped
        if self._pool_key is not u._pool_missions:
            uniq, mid = _unique_missions(u._pool_missions)
            self._table = torch.from_numpy(np.stack([self._encode(m) for m in uniq])).to(u.device)
            self._mid = torch.tensor(mid, dtype=torch.int32, device=u.device)
            self._pool_key = u._pool_missions
        img = obs["image"].contiguous()
        mlen = self.maxStrLen * self.numCharCodes
        out = torch.empty((u.num_envs, 147 + mlen), dtype=torch.float32, device=u.device)
        _capi.check(_capi.lib.mg_obs_flat(self._h(), img.data_ptr(), self._table.data_ptr(), self._mid.data_ptr(), mlen,
                                          out.data_ptr(), self._s()), "mg_obs_flat")
        return out


class DirectionObsWrapper(ObservationWrapper):
    """wrappers.py:673-725: obs["goal_direction"] = slope (type="slope") or arctan of it (type="angle") from the agent to
    the goal.  Like the reference, the goal is located ONCE, at the first reset, and kept for later episodes, and its
    cached tuple is (int(idx / height), idx % width) of the first Goal in grid order — (row, column), which the slope
    formula then reads as (x, y).  Envs without a goal get nan."""

    def __init__(self, env, type="slope"):
        super().__init__(env)
        self.type = type
        self.goal_position = None

    def _locate(self):
        from .constants import T_GOAL
        u = self.unwrapped
        agents, cells = u.get_state()
        gp = np.zeros((u.num_envs, 2), dtype=np.int32)
        self._has_goal = np.zeros(u.num_envs, dtype=bool)
        for e in range(u.num_envs):
            w, h = u._classes[agents["class_id"][e]][:2]
            idx = np.flatnonzero(cells[e, 0, :w * h] == T_GOAL)
            if len(idx):
                gp[e] = (int(idx[0] / h), idx[0] % w)
                self._has_goal[e] = True
        self.goal_position = gp
        self._gp_dev = torch.from_numpy(gp).to(u.device)

    def reset(self, *, seed=None, options=None):
        obs, info = self.env.reset(seed=seed, options=options)   # (the reference drops the seed here, wrappers.py:697)
        if self.goal_position is None:
            self._locate()
        return self.observation(obs), info

    def observation(self, obs):
        u = self.unwrapped
        out = torch.empty(u.num_envs, dtype=torch.float64, device=u.device)
        _capi.check(_capi.lib.mg_obs_goal_direction(self._h(), self._gp_dev.data_ptr(), int(self.type == "angle"),
                                                    out.data_ptr(), self._s()), "mg_obs_goal_direction")
        if not self._has_goal.all():
            out[torch.from_numpy(~self._has_goal).to(u.device)] = float("nan")
        return {**obs, "goal_direction": out}


class StochasticActionWrapper(Wrapper):
    """wrappers.py:769-796: every env's action is kept with probability `prob`, otherwise replaced by `random_action`
    (or, if None, a uniform draw from 0..5 — `np_random.integers(0, high=6)`, as the reference writes it).
    The reference draws the coin from the process-global numpy state; a batch has no such global order, so the draws come
    from a device generator (seed it with `seed=`), or are handed in by the caller (`uniform=`, `replacement=` in
    `action()`), which is how the parity test replays the reference's draws."""

    def __init__(self, env=None, prob=0.9, random_action=None, seed=0):
        super().__init__(env)
        self.prob, self.random_action = prob, random_action
        u = self.unwrapped
        self._gen = torch.Generator(device=u.device)
        self._gen.manual_seed(seed)

    def action(self, actions, uniform=None, replacement=None):
        u = self.unwrapped
        a = torch.as_tensor(actions).to(device=u.device, dtype=torch.uint8).contiguous()
        n = u.num_envs
        if uniform is None:
            uniform = torch.rand(n, dtype=torch.float64, device=u.device, generator=self._gen)
        uniform = torch.as_tensor(uniform, dtype=torch.float64).to(u.device).contiguous()
        if self.random_action is None:
            if replacement is None:
                replacement = torch.randint(0, 6, (n,), device=u.device, generator=self._gen)
            replacement = torch.as_tensor(replacement).to(device=u.device, dtype=torch.uint8).contiguous()
        out = torch.empty(n, dtype=torch.uint8, device=u.device)
        _capi.check(_capi.lib.mg_stochastic_actions(self._h(), a.data_ptr(), uniform.data_ptr(),
                                                    None if self.random_action is not None else replacement.data_ptr(),
                                                    float(self.prob), -1 if self.random_action is None else int(self.random_action),
                                                    out.data_ptr(), self._s()), "mg_stochastic_actions")
        return out

    def step(self, actions):
        return self.env.step(self.action(actions))
