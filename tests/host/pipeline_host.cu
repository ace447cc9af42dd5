// Host build of the register pipeline in csrc/mg_packed.cuh so that the CPU test-suite can compare the
// exact kernel arithmetic (fetch / vis / decode / interleave) with the oracle, without a GPU.
#include "../../minigrid_dynamicprogramming_b200/csrc/mg_packed.cuh"
#include <vector>
#include <cstring>

extern "C" int mgp_obs(const uint8_t* cells, int ps, int W, int H, int ax, int ay, int dir,
                       int ct, int cc, int cs, int see_through, uint8_t* image147)
{
    mg::TileGeom g = mg::tile_geom(W, H);
    std::vector<uint8_t> tile((size_t)g.bytes + 16);
    mg::pack_tile_serial(cells, ps, g, tile.data());
    uint32_t out[37];
    uint32_t lut[mg::LUT_ENTRIES];
    for (int i = 0; i < mg::LUT_ENTRIES; ++i) lut[i] = mg::lut_entry((uint32_t)i);
    const uint32_t carry = ct ? mg::encode_code(ct, cc, cs) : mg::CODE_EMPTY;
    mg::obs_words(tile.data(), g, ax, ay, dir, carry, see_through != 0, lut, out);
    std::memcpy(image147, out, 147);
    // the rows-only layout (large grids): only the R part of the record is visible; the rest is poisoned
    std::vector<uint8_t> rows(tile.begin(), tile.begin() + mg::tile_rows_bytes(g));
    for (int i = g.offC + mg::PAD; i < (int)rows.size(); ++i) rows[i] = 0x5A;
    uint32_t out2[37];
    mg::obs_words<true>(rows.data(), g, ax, ay, dir, carry, see_through != 0, lut, out2);
    return std::memcmp(out, out2, 147) == 0 ? 0 : 1;
}

extern "C" void mgp_process_vis(const uint32_t* t, uint32_t* vis) { mg::process_vis_rows(t, vis); }
extern "C" void mgp_process_vis_packed(uint32_t mlo, uint32_t mhi, uint32_t* v) { mg::process_vis_packed(mlo, mhi, v[0], v[1]); }

// ---- csrc/mg_raster.cuh on the host: one tile of the atlas ----
#include "../../minigrid_dynamicprogramming_b200/csrc/mg_raster.cuh"
#include <cmath>

extern "C" void mgp_tile(int type, int color, int state, int agent_dir, int highlight, int ts, uint8_t* out)
{
    mg::raster::Rotations rot;
    for (int k = 0; k < 4; ++k) {
        const double theta = 0.5 * M_PI * k;
        rot.cs[k] = std::cos(-theta); rot.sn[k] = std::sin(-theta);
    }
    for (int py = 0; py < ts; ++py)
        for (int px = 0; px < ts; ++px) {
            const uint32_t v = mg::raster::tile_pixel(type, color, state, agent_dir, highlight != 0, rot, ts, px, py);
            uint8_t* o = out + (py * ts + px) * 3;
            o[0] = (uint8_t)v; o[1] = (uint8_t)(v >> 8); o[2] = (uint8_t)(v >> 16);
        }
}

// ---- csrc/mg_dynrng.cuh on the host: the step kernel's PCG64 / bounded draws (DynamicObstacles) ----
#include "../../minigrid_dynamicprogramming_b200/csrc/mg_dynrng.cuh"

// rec: s_lo, s_hi, inc_lo, inc_hi, has32, u32 (six 64-bit slots, updated in place).  mode 0: out[i] = rand_int(low[i], high[i]);
// mode 1: (out[2i], out[2i+1]) = draw_pair over [low[2i], high[2i]) x [low[2i+1], high[2i+1]) — the kernel's try;
// mode 2: the same pair as two rand_int calls (what draw_pair must equal, values and generator state).
extern "C" void mgp_rng_draws(uint64_t* rec, int mode, const int32_t* low, const int32_t* high, int n, int32_t* out)
{
    mg::DynRng d;
    d.s_lo = rec[0]; d.s_hi = rec[1]; d.inc_lo = rec[2]; d.inc_hi = rec[3]; d.has32 = (uint32_t)rec[4]; d.u32 = (uint32_t)rec[5];
    for (int i = 0; i < n; ++i) {
        if (mode == 0) out[i] = mg::rand_int(d, low[i], high[i]);
        else if (mode == 1) mg::draw_pair(d, low[2 * i], high[2 * i], low[2 * i + 1], high[2 * i + 1], out[2 * i], out[2 * i + 1]);
        else { out[2 * i] = mg::rand_int(d, low[2 * i], high[2 * i]); out[2 * i + 1] = mg::rand_int(d, low[2 * i + 1], high[2 * i + 1]); }
    }
    rec[0] = d.s_lo; rec[1] = d.s_hi; rec[4] = d.has32; rec[5] = d.u32;
}
