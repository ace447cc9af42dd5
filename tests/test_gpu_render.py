"""GPU parity of the RGB renderers through the C ABI: the device-built tile atlas against the tiles
the reference's Grid.render_tile produced, whole frames against the reference's get_frame goldens and against the
oracle (mg_oracle_render.c) on random states — bit-exact."""
import numpy as np
import pytest

from helpers import L, O, golden, oracle_rollout

pytestmark = pytest.mark.gpu

torch = pytest.importorskip("torch")


def _engine():
    from minigrid_dynamicprogramming_b200.env import BatchedMiniGrid
    return BatchedMiniGrid


def _code(t, c, s):
    return (13 if (t == 4 and s == 1) else 14 if (t == 4 and s == 2) else t) | (c << 4)


@pytest.mark.parametrize("ts", [8, 5, 12, 32])
def test_atlas_matches_reference_tiles(ts):
    env = _engine()("MiniGrid-Empty-8x8-v0", 4, autoreset="off")
    atlas = env.render_atlas(ts)                       # [slot][highlight][code][ts][ts][3]
    d = golden("atlas.npz")
    keys, tiles = d[f"ts{ts}/keys"], d[f"ts{ts}/tiles"]
    assert len(keys) == 510
    for (t, c, s, ad, hl), want in zip(keys, tiles):
        got = atlas[ad + 1, hl, _code(int(t), int(c), int(s))]
        assert np.array_equal(got, want), (t, c, s, ad, hl)


def _layout_from_state(cells, st, cfg, name="golden"):
    W, H = int(cfg[0]), int(cfg[1])
    return L.Layout(name, 0, W, H, int(cfg[2]), bool(cfg[3]), int(cfg[4]), np.ascontiguousarray(cells[:, : W * H]),
                    (int(st[0]), int(st[1])), int(st[2]), name)


def _load_states(env, cells, sts):
    """Put the given (cells, st) into the envs of a freshly reset engine (carrying included)."""
    env.reset()
    agents = env.get_state(cells=False)
    agents["x"], agents["y"], agents["dir"] = sts[:, 0], sts[:, 1], sts[:, 2]
    agents["step_count"] = sts[:, 3]
    agents["carry_type"], agents["carry_color"], agents["carry_state"], agents["carry_aux"] = sts[:, 4], sts[:, 5], sts[:, 6], sts[:, 7]
    n = cells.shape[-1]
    buf = np.zeros((env.num_envs, 4, env.plane_stride), np.uint8)
    buf[:, 0, :] = 1
    buf[:, :, :n] = cells
    env.set_state(agents=agents, cells=buf)


def test_frames_match_reference_goldens():
    """get_frame of reference episodes (tests/golden/frames.npz): POV, full with highlight, full without."""
    f = golden("frames.npz")
    tags = sorted({k.rsplit("/", 1)[0] for k in f.files if k.endswith("/pov") and not k.startswith("ts32")})
    assert len(tags) == 11
    for tag in tags:
        cfg, cells, sts = f[tag + "/cfg"], f[tag + "/cells"], f[tag + "/st"]
        n = len(sts)
        lays = [_layout_from_state(cells[i], sts[i], cfg) for i in range(n)]
        env = _engine()("golden", n, layouts=lays, autoreset="off")
        _load_states(env, cells, sts)
        assert np.array_equal(env.get_frame(agent_pov=True, tile_size=8).cpu().numpy(), f[tag + "/pov"]), tag
        assert np.array_equal(env.get_frame(highlight=True, tile_size=8).cpu().numpy(), f[tag + "/full"]), tag
        assert np.array_equal(env.get_frame(highlight=False, tile_size=8).cpu().numpy(), f[tag + "/plain"]), tag
    lay = _layout_from_state(f["ts32/cells"], f["ts32/st"], f["ts32/cfg"])
    env = _engine()("golden", 1, layouts=[lay], autoreset="off")
    _load_states(env, f["ts32/cells"][None], f["ts32/st"][None])
    assert np.array_equal(env.get_frame(agent_pov=True, tile_size=32).cpu().numpy()[0], f["ts32/pov"])
    assert np.array_equal(env.get_frame(highlight=True, tile_size=32).cpu().numpy()[0], f["ts32/full"])


@pytest.mark.parametrize("W,H,ts", [(9, 9, 8), (13, 7, 8), (5, 5, 5), (22, 22, 4), (9, 9, 3), (7, 11, 16), (9, 9, 1)])
def test_frames_vs_oracle_on_playground_states(W, H, ts):
    """Random grids with every object type, stepped for a while (doors toggled, objects carried), then rendered at
    several tile sizes — exercises every piece width of the copy loop (16 / 8 / 4 / 1 bytes) and ragged batch sizes."""
    from test_gpu_parity import _random_playground
    rng = np.random.default_rng(W * 1000 + H * 10 + ts)
    N, T = 77, 40
    lays = [_random_playground(rng, W, H, s) for s in range(N)]
    acts = rng.integers(0, 7, (T, N)).astype(np.uint8)
    acts[rng.random((T, N)) < 0.3] = 2
    ref = oracle_rollout(lays, acts, autoreset=False, pool=lays)
    env = _engine()("playground", N, layouts=lays, autoreset="off")
    env.reset()
    env.step_many(torch.from_numpy(acts).cuda())
    pov = env.get_frame(agent_pov=True, tile_size=ts).cpu().numpy()
    full = env.get_frame(highlight=True, tile_size=ts).cpu().numpy()
    plain = env.get_frame(highlight=False, tile_size=ts).cpu().numpy()
    cfg = O.make_cfg(W, H, 200, False, 1)
    n = W * H
    for e in range(N):
        c = ref["cells"][e, : 4 * n].reshape(4, n)
        st = ref["st"][e]
        assert np.array_equal(pov[e], O.render_pov(c, cfg, st, ts)), e
        assert np.array_equal(full[e], O.render_full(c, cfg, st, True, ts)), e
        assert np.array_equal(plain[e], O.render_full(c, cfg, st, False, ts)), e


def test_frames_mixed_classes_and_see_through():
    """A batch mixing grid sizes (Empty-8x8 is see_through_walls=True: everything in view is highlighted)."""
    ids = ["MiniGrid-Empty-8x8-v0", "MiniGrid-LavaCrossingS11N5-v0", "MiniGrid-FourRooms-v0"]
    N, T = 96, 30
    env = _engine()(ids, N, autoreset="off")
    env.reset()
    acts = np.random.default_rng(5).integers(0, 3, (T, N)).astype(np.uint8)
    env.step_many(torch.from_numpy(acts).cuda())
    agents, cells = env.get_state()
    pov = env.get_frame(agent_pov=True).cpu().numpy()
    full = env.get_frame(highlight=True)
    assert isinstance(full, list) and len(full) == N
    for e in range(N):
        w, h, ms, st_, rule = env._classes[agents["class_id"][e]]
        cfg = O.make_cfg(w, h, ms, st_, rule)
        st = np.array([agents["x"][e], agents["y"][e], agents["dir"][e], agents["step_count"][e], agents["carry_type"][e],
                       agents["carry_color"][e], agents["carry_state"][e], agents["carry_aux"][e]], np.int32)
        c = cells[e, :, : w * h]
        assert np.array_equal(pov[e], O.render_pov(c, cfg, st, 8)), e
        assert np.array_equal(full[e].cpu().numpy(), O.render_full(c, cfg, st, True, 8)), e


def test_render_before_reset_is_an_error():
    from minigrid_dynamicprogramming_b200._capi import MgError
    env = _engine()("MiniGrid-Empty-8x8-v0", 4, autoreset="off")
    with pytest.raises((MgError, AssertionError)):
        env.get_frame(agent_pov=True, tile_size=0)
