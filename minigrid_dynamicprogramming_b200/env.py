"""BatchedMiniGrid — the reference's Gymnasium-style reset/step surface for N envs on one B200.

Mirrors `MiniGridEnv.reset` / `MiniGridEnv.step` (minigrid/minigrid_env.py:119-157, 520-590) with the
batched shapes `gymnasium.vector.SyncVectorEnv` stacks (tests/test_envs.py:318-330):

    obs = {"image": uint8[N,7,7,3], "direction": uint8[N], "mission": sequence of N str}
    reward float64[N], terminated bool[N], truncated bool[N], info {}

All arrays are CUDA tensors written by the sm_100a kernels behind the C ABI (include/mgb200.h); there is
no CPU path.  Level layouts come from the host generators (levelgen.py, or layouts extracted from a
reference env) and are uploaded once as a pool.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Sequence

import numpy as np
import torch

from . import _capi
from . import levelgen
from .constants import Actions, IDX_TO_COLOR, IDX_TO_OBJECT

AUTORESET = {"off": 0, "same_step": 1, "next_step": 2}

AGENT_DTYPE = np.dtype([("x", "u1"), ("y", "u1"), ("dir", "u1"), ("class_id", "u1"), ("carry_type", "u1"),
                        ("carry_color", "u1"), ("carry_state", "u1"), ("carry_aux", "u1"), ("step_count", "<u2"),
                        ("pad", "<u2"), ("layout_id", "<u4")])
LAYOUT_AGENT_DTYPE = np.dtype([("x", "u1"), ("y", "u1"), ("dir", "u1"), ("class_id", "u1")])


class _Discrete:
    def __init__(self, n):
        self.n = n


class _Box:
    """gymnasium.spaces.Box stand-in (shape / dtype / bounds are what callers read)."""

    def __init__(self, low, high, shape, dtype):
        self.low, self.high, self.shape, self.dtype = low, high, tuple(shape), np.dtype(dtype)


class _Missions(Sequence):
    """obs["mission"]: resolves env -> pool slot -> string only when somebody looks (needs a device sync)."""

    def __init__(self, env):
        self._env = env

    def _ids(self):
        return self._env.get_state(cells=False)["layout_id"]

    def __len__(self):
        return self._env.num_envs

    def __getitem__(self, i):
        ids = self._ids()
        if isinstance(i, slice):
            return tuple(self._env._pool_missions[j] for j in ids[i])
        return self._env._pool_missions[int(ids[i])]

    def __iter__(self):
        return iter(self._env._pool_missions[j] for j in self._ids())


class BatchedMiniGrid:
    """N reference envs stepped by one kernel launch.

    env_id      a registered id ("MiniGrid-DoorKey-8x8-v0") or a list of ids (env e uses ids[e % len]) —
                the mixed-layout batch of BASELINE config 4.
    num_envs    N on this device.
    seeds       reset seeds of the initial layout pool (default range(pool_size)); layout of pool slot i is
                `gym.make(id).reset(seed=seeds[i])` (levelgen.py).
    pool_size   P layouts kept in HBM for autoreset (default N).
    layouts     alternatively, ready-made `levelgen.Layout`s (e.g. extracted from your own reference envs).
    autoreset   "same_step" (default): a terminated/truncated env restarts inside the same step from pool slot
                (layout_id + reset_stride) % P (reset_stride: N mod P moved up to the next value coprime to P, so every env
                walks through the whole pool) and the returned obs is the first of the new episode; reward/terminated/
                truncated belong to the finished one.  "next_step" (gymnasium >= 1.0's default): the step that ends an
                episode returns its last obs; the env's next step ignores its action and only resets (reward 0, flags
                clear, first obs of the new episode).  "off": like the bare reference env.
    final_obs   with autoreset="same_step": keep the finished episodes' last observations; step() then returns
                info = {"final_observation": {"image", "direction"}, "_final_observation": terminated | truncated}
                (gymnasium 0.29's vector-env convention; entries outside the mask are stale).
    """

    def __init__(self, env_id, num_envs, *, seeds=None, pool_size=None, layouts=None, device="cuda:0",
                 autoreset="same_step", final_obs=False, agent_view_size=7, **env_kwargs):
        if not torch.cuda.is_available():
            raise RuntimeError("BatchedMiniGrid needs a CUDA device (there is no CPU fallback)")
        self.device = torch.device(device)
        self.num_envs = int(num_envs)
        self.env_ids = [env_id] if isinstance(env_id, str) else list(env_id)
        self._env_kwargs = env_kwargs
        self.autoreset = autoreset
        self.final_obs = bool(final_obs) and autoreset == "same_step"
        self.actions = Actions
        self.action_space = _Discrete(len(Actions))            # minigrid_env.py:60-63
        self.reward_range = (0, 1)                             # minigrid_env.py:87
        # MiniGridEnv(agent_view_size=...) (minigrid_env.py:42,66-68: odd, >= 3).  Every registered id uses 7, which is what the
        # fused step kernel emits; another size steps with the same kernel (observation output off) and takes its view from
        # the ViewSizeWrapper kernel (mg_obs_view) after every step — one launch pair per env step instead of one per T steps.
        v = int(agent_view_size)
        if v % 2 != 1 or v < 3 or v > 15:
            raise ValueError("agent_view_size must be odd, 3..15")
        if v != 7 and final_obs:
            raise ValueError("final_obs needs the 7x7 view of the fused kernel")
        self.agent_view_size = v
        # minigrid_env.py:72-84 (per env; the batch adds a leading N axis): image Box(0, 255, (V, V, 3), uint8), direction
        # Discrete(4), mission = the mission space's strings
        self.observation_space = {"image": _Box(0, 255, (v, v, 3), np.uint8), "direction": _Discrete(4), "mission": str}
        self.single_observation_space, self.single_action_space = self.observation_space, self.action_space
        self._handle = None
        if layouts is None:
            P = int(pool_size) if pool_size is not None else self.num_envs
            seeds = list(range(P)) if seeds is None else list(seeds)
            layouts = self._generate(seeds)
        self._set_pool(layouts)
        self._alloc_outputs(1)
        self._T_alloc = 1

    # ------------------------------------------------------------------ layouts / handle
    def _generate(self, seeds):
        gens = [levelgen.make_generator(e, **self._env_kwargs) for e in self.env_ids]
        k = len(gens)
        return [gens[i % k].generate(self.env_ids[i % k], int(s)) for i, s in enumerate(seeds)]

    def _set_pool(self, layouts):
        classes, class_of = [], {}
        for lay in layouts:
            key = (lay.width, lay.height, lay.max_steps, int(lay.see_through_walls), lay.rule)
            if key not in class_of:
                class_of[key] = len(classes)
                classes.append(key)
        if self._handle is None or classes != getattr(self, "_classes", None):
            self._destroy()
            arr = (_capi.ClassDesc * len(classes))()
            for i, (w, h, ms, st, rule) in enumerate(classes):
                arr[i].width, arr[i].height, arr[i].max_steps, arr[i].see_through_walls, arr[i].rule = w, h, ms, st, rule
            cfg = _capi.Config(self.device.index or 0, self.num_envs, len(classes), AUTORESET[self.autoreset], arr)
            h = C.c_void_p()
            _capi.check(_capi.lib.mg_create(C.byref(h), C.byref(cfg)), "mg_create")
            self._handle = h
            self._classes = classes
            self.plane_stride = _capi.lib.mg_plane_stride(h)
        P = len(layouts)
        cells = np.zeros((P, 4, self.plane_stride), dtype=np.uint8)
        cells[:, 0, :] = 1
        agents = np.zeros(P, dtype=LAYOUT_AGENT_DTYPE)
        for i, lay in enumerate(layouts):
            n = lay.width * lay.height
            cells[i, :, :n] = lay.cells
            key = (lay.width, lay.height, lay.max_steps, int(lay.see_through_walls), lay.rule)
            agents[i] = (lay.agent_pos[0], lay.agent_pos[1], lay.agent_dir, class_of[key])
        _capi.check(_capi.lib.mg_upload_layouts(self._handle, cells.ctypes.data, agents.ctypes.data, P), "mg_upload_layouts")
        if any(lay.dyn is not None for lay in layouts):      # DynamicObstacles: generator state + obstacle list per layout
            dyn = np.zeros((P, 64), dtype=np.uint8)
            dyn[:, 16] = 1                                   # odd increment for layouts of other classes in a mixed pool
            for i, lay in enumerate(layouts):
                if lay.dyn is not None:
                    dyn[i] = lay.dyn
            _capi.check(_capi.lib.mg_upload_dyn(self._handle, dyn.ctypes.data, P), "mg_upload_dyn")
        self._has_dyn = any(lay.dyn is not None for lay in layouts)
        self.pool_size = P
        self._pool_missions = [lay.mission for lay in layouts]
        self._pool_layouts = layouts
        self.width = max(c[0] for c in classes)
        self.height = max(c[1] for c in classes)
        self.max_steps = classes[0][2] if len(classes) == 1 else [c[2] for c in classes]

    def _alloc_outputs(self, T):
        N, dev, v = self.num_envs, self.device, self.agent_view_size
        self._image = torch.empty((T, N, v, v, 3), dtype=torch.uint8, device=dev)
        self._direction = torch.empty((T, N), dtype=torch.uint8, device=dev)
        self._reward = torch.empty((T, N), dtype=torch.float64, device=dev)
        self._terminated = torch.empty((T, N), dtype=torch.uint8, device=dev)
        self._truncated = torch.empty((T, N), dtype=torch.uint8, device=dev)
        if self.final_obs:
            self._final_image = torch.zeros((T, N, 7, 7, 3), dtype=torch.uint8, device=dev)
            self._final_direction = torch.zeros((T, N), dtype=torch.uint8, device=dev)
        self._T_alloc = T

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _destroy(self):
        if getattr(self, "_handle", None) is not None:
            _capi.lib.mg_destroy(self._handle)
            self._handle = None

    def close(self):
        self._destroy()

    def __del__(self):
        try:
            self._destroy()
        except Exception:
            pass

    # ------------------------------------------------------------------ reset / step
    def reset(self, *, seed=None, options=None):
        """MiniGridEnv.reset for every env.  seed: None -> pool slot e % P; int s -> seeds s+e (SyncVectorEnv
        convention); sequence of N ints -> those seeds (the pool is regenerated from them)."""
        N = self.num_envs
        if seed is not None:
            seeds = [int(seed) + e for e in range(N)] if np.isscalar(seed) else [int(s) for s in seed]
            if len(seeds) != N:
                raise ValueError(f"need {N} seeds, got {len(seeds)}")
            self._set_pool(self._generate(seeds))
        ids = torch.arange(N, dtype=torch.int32, device=self.device) % self.pool_size
        with torch.cuda.device(self.device):
            _capi.check(_capi.lib.mg_reset(self._handle, ids.data_ptr(), None, self._image7(0),
                                           self._direction[0].data_ptr(), self._stream()), "mg_reset")
            self._view(0)
        return self._obs(self._image[0], self._direction[0]), {}

    def reset_envs(self, layout_ids, mask=None):
        """Partial reset: env e <- pool[layout_ids[e]] where mask[e] (all if None).  Returns the obs of every env."""
        ids = torch.as_tensor(layout_ids, dtype=torch.int32, device=self.device).contiguous()
        m = None if mask is None else torch.as_tensor(mask, device=self.device).to(torch.uint8).contiguous()
        with torch.cuda.device(self.device):
            _capi.check(_capi.lib.mg_reset(self._handle, ids.data_ptr(), None if m is None else m.data_ptr(),
                                           self._image7(0), self._direction[0].data_ptr(), self._stream()),
                        "mg_reset")
            self._view(0)
        return self._obs(self._image[0], self._direction[0]), {}

    def _image7(self, t):
        """Where the fused kernel writes its 7x7 observation of step t (None when the env has another agent_view_size)."""
        return self._image[t].data_ptr() if self.agent_view_size == 7 else None

    def _view(self, t):
        """agent_view_size != 7: the observation of the CURRENT state into image[t] (mg_obs_view, wrappers.py:626-670's kernel)."""
        if self.agent_view_size != 7:
            _capi.check(_capi.lib.mg_obs_view(self._handle, self.agent_view_size, self._image[t].data_ptr(), self._stream()), "mg_obs_view")

    def _obs(self, image, direction):
        return {"image": image, "direction": direction, "mission": _Missions(self)}

    def _actions_tensor(self, actions, shape):
        host = not (isinstance(actions, torch.Tensor) and actions.is_cuda)
        a = torch.as_tensor(actions)
        if tuple(a.shape) != shape:
            raise ValueError(f"actions must have shape {shape}, got {tuple(a.shape)}")
        if a.dtype != torch.uint8:
            a = a.clamp(-1, 255).to(torch.uint8)          # -1 -> 255: still invalid, still reported
        return a.to(self.device, non_blocking=True).contiguous(), host

    def step(self, actions):
        """MiniGridEnv.step for every env: actions int[N] -> (obs, reward, terminated, truncated, info).
        The returned tensors are views of buffers the next step overwrites."""
        obs, r, te, tr, info = self.step_many(torch.as_tensor(actions)[None] if not isinstance(actions, torch.Tensor)
                                              else actions[None])
        if info:
            info = {"final_observation": {k: v[0] for k, v in info["final_observation"].items()},
                    "_final_observation": info["_final_observation"][0]}
        return ({"image": obs["image"][0], "direction": obs["direction"][0], "mission": obs["mission"]},
                r[0], te[0], tr[0], info)

    def step_many(self, actions):
        """T consecutive steps in ONE kernel launch: actions int[T, N]; every output gains a leading T axis."""
        T = int(actions.shape[0])
        a, host = self._actions_tensor(actions, (T, self.num_envs))
        if T != self._T_alloc:
            self._alloc_outputs(T)
        info = {}
        with torch.cuda.device(self.device):
            if self.final_obs:
                _capi.check(_capi.lib.mg_step_many_final(self._handle, a.data_ptr(), T, self._image.data_ptr(),
                                                         self._direction.data_ptr(), self._reward.data_ptr(),
                                                         self._terminated.data_ptr(), self._truncated.data_ptr(),
                                                         self._final_image.data_ptr(), self._final_direction.data_ptr(),
                                                         self._stream()), "mg_step_many_final")
                info = {"final_observation": {"image": self._final_image, "direction": self._final_direction},
                        "_final_observation": (self._terminated | self._truncated).view(torch.bool)}
            elif self.agent_view_size != 7:
                for t in range(T):                   # one step + one view launch per env step
                    _capi.check(_capi.lib.mg_step_many(self._handle, a[t].data_ptr(), 1, None, self._direction[t].data_ptr(),
                                                       self._reward[t].data_ptr(), self._terminated[t].data_ptr(),
                                                       self._truncated[t].data_ptr(), self._stream()), "mg_step_many")
                    self._view(t)
            else:
                _capi.check(_capi.lib.mg_step_many(self._handle, a.data_ptr(), T, self._image.data_ptr(),
                                                   self._direction.data_ptr(), self._reward.data_ptr(),
                                                   self._terminated.data_ptr(), self._truncated.data_ptr(), self._stream()),
                            "mg_step_many")
        if host:
            self.check_errors()       # host actions: raise ValueError now, like minigrid_env.py:579-580
        return (self._obs(self._image, self._direction), self._reward, self._terminated.view(torch.bool),
                self._truncated.view(torch.bool), info)

    def check_errors(self):
        """Raises ValueError if a kernel saw an action outside 0..6 since the last check (synchronises)."""
        env, act = C.c_int64(0), C.c_int32(0)
        rc = _capi.lib.mg_poll_error(self._handle, C.byref(env), C.byref(act))
        if rc == _capi.MG_EACTION:
            raise ValueError(f"Unknown action: {act.value} (env {env.value})")
        _capi.check(rc, "mg_poll_error")

    def launch_shape(self):
        """(resident CTAs per SM, threads per CTA, dynamic shared memory bytes) of the step kernel for this batch."""
        b, t, sm = C.c_int32(0), C.c_int32(0), C.c_int64(0)
        _capi.check(_capi.lib.mg_step_occupancy(self._handle, C.byref(b), C.byref(t), C.byref(sm)), "mg_step_occupancy")
        return b.value, t.value, sm.value

    @property
    def reset_stride(self):
        """Autoreset: a finished env restarts from pool slot (layout_id + reset_stride) % pool_size (mg_reset_stride)."""
        return int(_capi.lib.mg_reset_stride(self._handle))

    @property
    def launch_count(self):
        return int(_capi.lib.mg_launch_count(self._handle))

    # ------------------------------------------------------------------ RGB rendering
    def get_frame(self, highlight=True, tile_size=8, agent_pov=False, out=None):
        """MiniGridEnv.get_frame (minigrid_env.py:711-734) for every env: one kernel launch, frames stay on the device.

        agent_pov=True  -> uint8[N, 7*ts, 7*ts, 3]   (get_pov_render; what RGBImgPartialObsWrapper returns)
        agent_pov=False -> uint8[N, H*ts, W*ts, 3]   (get_full_render; RGBImgObsWrapper / env.render()); for a batch that
                           mixes grid sizes a list of N tensors (views into one buffer) is returned instead.
        Frames are bit-identical to the reference's numpy rasteriser.  `out`: a tensor from an earlier call to reuse."""
        ts = int(tile_size)
        N = self.num_envs
        if agent_pov:
            img = out if out is not None else torch.empty((N, 7 * ts, 7 * ts, 3), dtype=torch.uint8, device=self.device)
            assert img.is_contiguous() and img.dtype == torch.uint8 and img.numel() == N * 147 * ts * ts
            _capi.check(_capi.lib.mg_render_pov(self._handle, ts, img.data_ptr(), self._stream()), "mg_render_pov")
            return img
        stride = int(_capi.lib.mg_render_full_stride(self._handle, ts))
        buf = out if out is not None else torch.empty((N, stride), dtype=torch.uint8, device=self.device)
        assert buf.is_contiguous() and buf.dtype == torch.uint8 and buf.numel() == N * stride
        _capi.check(_capi.lib.mg_render_full(self._handle, ts, int(bool(highlight)), buf.data_ptr(), self._stream()),
                    "mg_render_full")
        flat = buf.view(N, stride)
        if len(self._classes) == 1:
            w, h = self._classes[0][0], self._classes[0][1]
            return flat[:, :h * w * ts * ts * 3].view(N, h * ts, w * ts, 3)
        cls = self.get_state(cells=False)["class_id"]
        return [flat[e, :self._classes[c][1] * self._classes[c][0] * ts * ts * 3]
                .view(self._classes[c][1] * ts, self._classes[c][0] * ts, 3) for e, c in enumerate(cls)]

    def render(self, highlight=True, tile_size=32):
        """MiniGridEnv.render() in `rgb_array` mode for every env (minigrid_env.py:736-737: get_frame(self.highlight,
        self.tile_size); the reference's default tile size is 32).  There is no `human` mode: it needs a window."""
        return self.get_frame(highlight=highlight, tile_size=tile_size, agent_pov=False)

    def grid_encode(self, e=0):
        """`env.grid.encode()` of env e (core/grid.py:244-268): uint8[W, H, 3], indexed [x][y], host array."""
        from . import debug
        c, w, h, _, _ = self._env_state(e)
        return debug.grid_encode(c, w, h)

    def render_atlas(self, tile_size=8):
        """The tile atlas the frame kernels gather from (Grid.render_tile for every key; core/grid.py:145-198) as
        uint8[5, 2, 128, ts, ts, 3]: [agent slot (none, dir 0..3)][highlight][packed code]."""
        ts = int(tile_size)
        a = np.zeros((5, 2, 128, ts, ts, 3), dtype=np.uint8)
        _capi.check(_capi.lib.mg_render_atlas(self._handle, ts, a.ctypes.data), "mg_render_atlas")
        return a

    # ------------------------------------------------------------------ debugging surface (host side; debug.py)
    def _env_state(self, e):
        agents, cells = self.get_state()
        w, h = self._classes[agents["class_id"][e]][:2]
        return cells[e], w, h, (int(agents["x"][e]), int(agents["y"][e])), int(agents["dir"][e])

    def hash(self, e=0, size=16, pos_repr="np_tuple"):
        """MiniGridEnv.hash of env e (minigrid_env.py:159-169); see debug.state_hash for pos_repr."""
        from . import debug
        c, w, h, pos, d = self._env_state(e)
        return debug.state_hash(c, w, h, pos, d, size, pos_repr)

    def pprint_grid(self, e=0):
        """MiniGridEnv.pprint_grid of env e (minigrid_env.py:175-229)."""
        from . import debug
        c, w, h, pos, d = self._env_state(e)
        return debug.pprint_grid(c, w, h, pos, d)

    def agent_sees(self, e, x, y):
        """MiniGridEnv.agent_sees(x, y) of env e (minigrid_env.py:500-518), from the observation of the last step."""
        from . import debug
        c, w, h, pos, d = self._env_state(e)
        return debug.agent_sees(c, w, h, pos, d, self._image[-1, e].cpu().numpy(), x, y)

    def layout_ids_device(self):
        """int64[N] device tensor: the pool slot every env's current episode started from (indexes the missions)."""
        ids = torch.empty(self.num_envs, dtype=torch.int32, device=self.device)
        _capi.check(_capi.lib.mg_layout_ids(self._handle, ids.data_ptr(), self._stream()), "mg_layout_ids")
        return ids.long()

    # ------------------------------------------------------------------ state access (host side, synchronises)
    def get_state(self, cells=True):
        """Structured array of per-env state (+ 'cells' uint8[N,4,plane_stride] when cells=True)."""
        agents = np.zeros(self.num_envs, dtype=AGENT_DTYPE)
        cbuf = np.zeros((self.num_envs, 4, self.plane_stride), dtype=np.uint8) if cells else None
        _capi.check(_capi.lib.mg_get_state(self._handle, None if cbuf is None else cbuf.ctypes.data, agents.ctypes.data),
                    "mg_get_state")
        return (agents, cbuf) if cells else agents

    def get_dyn(self):
        """uint8[N, 64]: the DynamicObstacles records (levelgen.dyn_record layout: PCG64 state + obstacle list) of every env."""
        d = np.zeros((self.num_envs, 64), dtype=np.uint8)
        _capi.check(_capi.lib.mg_get_dyn(self._handle, d.ctypes.data), "mg_get_dyn")
        return d

    def set_dyn(self, dyn):
        d = np.ascontiguousarray(dyn, dtype=np.uint8).reshape(self.num_envs, 64)
        _capi.check(_capi.lib.mg_set_dyn(self._handle, d.ctypes.data), "mg_set_dyn")

    def set_state(self, agents=None, cells=None):
        a = None if agents is None else np.ascontiguousarray(agents, dtype=AGENT_DTYPE)
        c = None if cells is None else np.ascontiguousarray(cells, dtype=np.uint8)
        _capi.check(_capi.lib.mg_set_state(self._handle, None if c is None else c.ctypes.data,
                                           None if a is None else a.ctypes.data), "mg_set_state")

    @property
    def agent_pos(self):
        s = self.get_state(cells=False)
        return np.stack([s["x"], s["y"]], axis=1).astype(np.int64)

    @property
    def agent_dir(self):
        return self.get_state(cells=False)["dir"].astype(np.int64)

    @property
    def step_count(self):
        return self.get_state(cells=False)["step_count"].astype(np.int64)

    @property
    def carrying(self):
        """List of None | (type name, color name) per env (MiniGridEnv.carrying, minigrid_env.py:111)."""
        s = self.get_state(cells=False)
        return [None if t == 0 else (IDX_TO_OBJECT[int(t)], IDX_TO_COLOR[int(c)])
                for t, c in zip(s["carry_type"], s["carry_color"])]

    @property
    def mission(self):
        return tuple(_Missions(self))
