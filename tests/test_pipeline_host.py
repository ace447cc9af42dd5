"""CPU: the kernel's register pipeline (csrc/mg_packed.cuh: packed tile, column fetch, carry-chain
process_vis, SIMD decode, interleave) compiled for the HOST and compared with the oracle."""
import ctypes as C
import os
import shutil
import subprocess

import numpy as np
import pytest

from helpers import CONFIG_ENVS, L, oracle_env

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "host", "pipeline_host.cu")
SO = os.path.join(HERE, "host", "_build", "libmgpipe_host.so")


@pytest.fixture(scope="module")
def pipe():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        pytest.skip("nvcc not available")
    os.makedirs(os.path.dirname(SO), exist_ok=True)
    subprocess.check_call([nvcc, "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared",
                           "-o", SO, SRC])
    lib = C.CDLL(SO)
    lib.mgp_obs.argtypes = [C.c_void_p] + [C.c_int] * 10 + [C.c_void_p]
    lib.mgp_process_vis.argtypes = [C.c_void_p, C.c_void_p]
    lib.mgp_process_vis_packed.argtypes = [C.c_uint32, C.c_uint32, C.c_void_p]
    lib.mgp_tile.argtypes = [C.c_int] * 6 + [C.c_void_p]
    lib.mgp_rng_draws.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    return lib


def _vis_reference(t):
    """Grid.process_vis (core/grid.py:291-328) written out literally on a 7x7 boolean transparency grid."""
    mask = np.zeros((7, 7), bool)
    mask[3, 6] = True
    for j in reversed(range(7)):
        for i in range(6):
            if mask[i, j] and t[i, j]:
                mask[i + 1, j] = True
                if j > 0:
                    mask[i + 1, j - 1] = True
                    mask[i, j - 1] = True
        for i in reversed(range(1, 7)):
            if mask[i, j] and t[i, j]:
                mask[i - 1, j] = True
                if j > 0:
                    mask[i - 1, j - 1] = True
                    mask[i, j - 1] = True
    return mask


def test_process_vis_closed_form(pipe):
    rng = np.random.default_rng(0)
    for it in range(5000):
        t = rng.random((7, 7)) < rng.choice([0.3, 0.6, 0.85])
        rows = np.array([sum(int(t[a, b]) << a for a in range(7)) for b in range(7)], np.uint32)
        vis = np.zeros(7, np.uint32)
        pipe.mgp_process_vis(rows.ctypes.data, vis.ctypes.data)
        got = np.array([[(vis[b] >> a) & 1 for b in range(7)] for a in range(7)], bool)
        assert (got == _vis_reference(t)).all(), it


def test_process_vis_packed_alternating_orders(pipe):
    """process_vis_packed (the kernel's form: packed opacity words in, one bit reversal per row, rows 6/4/2/0 carried
    bit-reversed and put right at the end) against the literal double loop — also with garbage in the unused top byte."""
    rng = np.random.default_rng(5)
    for it in range(5000):
        t = rng.random((7, 7)) < rng.choice([0.2, 0.5, 0.7, 0.9])
        if it == 0:
            t[:] = True
        if it == 1:
            t[:] = False
        op = [sum(int(not t[a, b]) << a for a in range(7)) for b in range(7)]            # opacity bits of row b
        mlo = op[0] | op[1] << 8 | op[2] << 16 | op[3] << 24
        mhi = op[4] | op[5] << 8 | op[6] << 16 | int(rng.integers(0, 256)) << 24
        mlo |= int(rng.integers(0, 2)) * 0x80808080                                       # bit 7 of a row byte is not a cell
        v = np.zeros(2, np.uint32)
        pipe.mgp_process_vis_packed(mlo & 0xFFFFFFFF, mhi & 0xFFFFFFFF, v.ctypes.data)
        vis = [(int(v[0]) >> (8 * b)) & 0xFF for b in range(4)] + [(int(v[1]) >> (8 * b)) & 0xFF for b in range(4)]
        assert vis[7] == 0 and all(x < 128 for x in vis)
        got = np.array([[(vis[b] >> a) & 1 for b in range(7)] for a in range(7)], bool)
        assert (got == _vis_reference(t)).all(), it


@pytest.mark.parametrize("eid", CONFIG_ENVS + ["MiniGrid-KeyCorridorS3R3-v0", "MiniGrid-Empty-5x5-v0"])
def test_packed_obs_pipeline_vs_oracle(pipe, eid):
    rng = np.random.default_rng(1)
    n = 0
    for seed in range(4):
        lay = L.generate(eid, seed)
        env = oracle_env(lay)
        W, H = lay.width, lay.height
        for t in range(300):
            a = int(rng.integers(0, 7)) if rng.random() > 0.3 else 2
            img, d, r, te, tr = env.step(a)
            if te:
                break
            out = np.zeros(147, np.uint8)
            rc = pipe.mgp_obs(env.cells.ctypes.data, W * H, W, H, int(env.st[0]), int(env.st[1]), int(env.st[2]),
                              int(env.st[4]), int(env.st[5]), int(env.st[6]), int(lay.see_through_walls), out.ctypes.data)
            assert (out.reshape(7, 7, 3) == img).all(), (eid, seed, t)
            assert rc == 0, ("rows-only layout differs from the dual layout", eid, seed, t)
            n += 1
    assert n > 30


@pytest.mark.parametrize("ts", [8, 5, 12, 32])
def test_rasteriser_matches_reference_tiles(pipe, ts):
    """csrc/mg_raster.cuh (the atlas kernel's per-pixel rasteriser, host build) against every tile the reference's
    Grid.render_tile produced (tests/golden/atlas.npz, recorded by oracle/gen_golden_render.py)."""
    d = np.load(os.path.join(HERE, "golden", "atlas.npz"))
    keys, tiles = d[f"ts{ts}/keys"], d[f"ts{ts}/tiles"]
    assert len(keys) == 510
    for k, want in zip(keys, tiles):
        got = np.zeros((ts, ts, 3), np.uint8)
        pipe.mgp_tile(int(k[0]), int(k[1]), int(k[2]), int(k[3]), int(k[4]), ts, got.ctypes.data)
        assert np.array_equal(got, want), k


def _rng_rec(gen):
    """numpy Generator(PCG64) state -> the six slots of mgp_rng_draws (= the generator part of a mg_dyn_record)."""
    st = gen.bit_generator.state
    s, inc, m = st["state"]["state"], st["state"]["inc"], (1 << 64) - 1
    return np.array([s & m, s >> 64, inc & m, inc >> 64, st["has_uint32"], st["uinteger"]], dtype=np.uint64)


def _rng_run(pipe, rec, mode, low, high, n):
    rec = rec.copy()
    low, high = np.ascontiguousarray(low, np.int32), np.ascontiguousarray(high, np.int32)
    out = np.zeros(n if mode == 0 else 2 * n, np.int32)
    pipe.mgp_rng_draws(rec.ctypes.data, mode, low.ctypes.data, high.ctypes.data, n, out.ctypes.data)
    return out, rec


@pytest.mark.parametrize("seed", [0, 1, 12345])
def test_device_generator_equals_numpy(pipe, seed):
    """csrc/mg_dynrng.cuh (host build) against numpy itself: Generator(PCG64).integers(low, high), the call behind
    MiniGridEnv._rand_int (minigrid_env.py:244-249) — values and the generator state afterwards, on ranges from the obstacle
    loop's 2..3 up to ones whose Lemire rejection zone is hit, with single-value ranges (no draw) mixed in, from both phases of
    the 32-bit buffer."""
    gen = np.random.Generator(np.random.PCG64(seed))
    r = np.random.default_rng(seed + 99)
    if seed & 1:
        gen.integers(0, 5)                                  # leaves a buffered half word: has_uint32 = 1
    n = 4000
    span = np.where(r.random(n) < 0.1, 1, np.where(r.random(n) < 0.5, r.integers(2, 4, n), r.integers(2, 3_000_000_000, n)))
    low = r.integers(-5, 20, n)
    high = np.minimum(low + span, 2**31 - 1)
    rec0 = _rng_rec(gen)
    want = np.array([gen.integers(int(a), int(b)) for a, b in zip(low, high)], dtype=np.int64)
    got, rec1 = _rng_run(pipe, rec0, 0, low, high, n)
    assert np.array_equal(got, want)
    assert np.array_equal(rec1, _rng_rec(gen))


def test_draw_pair_equals_two_draws(pipe):
    """The kernel's one-generator-step placement try (draw_pair) against two rand_int calls: same cells, same generator state,
    from both buffer phases; 3x3 boxes (the obstacle loop), clipped 1-wide boxes (general path) and wide ranges whose products
    land in the rejection zone (general path from the untouched state)."""
    r = np.random.default_rng(7)
    for phase in (0, 1):
        gen = np.random.Generator(np.random.PCG64(42 + phase))
        if phase:
            gen.integers(0, 3)
        assert gen.bit_generator.state["has_uint32"] == phase
        n = 20000
        low = r.integers(0, 16, 2 * n)
        u = r.random(2 * n)
        span = np.where(u < 0.05, 1, np.where(u < 0.9, r.integers(2, 4, 2 * n), r.integers(1 << 28, 1 << 31, 2 * n)))
        high = np.minimum(low + span, 2**31 - 1)
        rec0 = _rng_rec(gen)
        fast, rf = _rng_run(pipe, rec0, 1, low, high, n)
        slow, rs = _rng_run(pipe, rec0, 2, low, high, n)
        assert np.array_equal(fast, slow) and np.array_equal(rf, rs)
        assert not np.array_equal(rf, rec0)
