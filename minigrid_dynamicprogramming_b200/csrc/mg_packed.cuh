// mg_packed.cuh — the packed per-env tile the step kernel works on, and the register-resident
// observation pipeline (gen_obs in closed form on packed bytes).
//
// Besides the authoritative struct-of-arrays planes (type, color, state, aux), every env keeps a
// derived "packed tile" in HBM that the kernel stages in shared memory:
//
//   code byte = type' | color << 4 | opaque << 7      type' = type, except closed door 13, locked door 14
//   R  : H rows, pitch PR = W + PAD, cell (x,y) at PAD + y*PR + x          (contiguous along x)
//   Cm : W rows, pitch PC = H + PAD, cell (x,y) at offC + PAD + x*PC + y   (contiguous along y)
//   the PAD = 6 bytes in front of every line (and after the last) hold the wall code — what Grid.slice
//   substitutes out of bounds (core/grid.py:136-139); one gap pads the end of a line and the start of the next
//
// A view column (fixed a, b = 0..6) is 7 cells along the agent's forward axis, i.e. 7 CONTIGUOUS bytes of
// R (facing left/right) or Cm (facing up/down): three aligned 32-bit loads + a funnel shift fetch it, a
// byte permute reverses it when the agent faces +x / +y.  The 49 cells never leave registers:
// opacity bits are transposed into 7-bit row masks for process_vis, the visibility masks are expanded
// back to byte masks, the codes are decoded four cells per instruction, and the (type,color,state)
// triples are interleaved into the 147-byte image record [a][b][c] (core/grid.py:244-268).
//
// Everything here is __host__ __device__ so that tests/ can run the exact same code on the CPU.
#pragma once
#include "mg_common.cuh"

namespace mg {

constexpr int PAD = 6;                       // the view reaches 6 cells ahead of the agent
constexpr uint32_t CODE_WALL = 0xD2;         // wall | grey<<4 | opaque
constexpr uint32_t CODE_EMPTY = 0x01;
constexpr uint32_t WALL4 = 0xD2D2D2D2u;

struct TileGeom {                            // per class
    int W, H, PR, PC, offC, bytes;           // bytes = size of the packed record (multiple of 16, incl. 16 slack)
};
__host__ __device__ inline TileGeom tile_geom(int W, int H)
{
    TileGeom g;
    // one PAD-wide run of wall codes between consecutive lines serves as the high-side padding of one line and the
    // low-side padding of the next: line pitch W + PAD (not W + 2*PAD), one extra PAD in front of line 0
    g.W = W; g.H = H; g.PR = W + PAD; g.PC = H + PAD;
    g.offC = (PAD + H * g.PR + 15) & ~15;
    g.bytes = ((g.offC + PAD + W * g.PC + 15) & ~15) + 16;     // +16: fetch7 may read up to 5 bytes past a line end
    return g;
}

// bytes of a record the ROWS layout of the step kernel stages: R, the wall gap behind it, and 16 bytes of over-read slack
__host__ __device__ inline int tile_rows_bytes(const TileGeom& g) { return g.offC + 16; }

__host__ __device__ __forceinline__ uint32_t encode_code(int t, int c, int s)
{
    int tt = t;
    if (t == T_DOOR && s == S_CLOSED) tt = 13;
    else if (t == T_DOOR && s == S_LOCKED) tt = 14;
    return (uint32_t)tt | ((uint32_t)c << 4) | (is_opaque(t, s) ? 0x80u : 0u);
}
__host__ __device__ __forceinline__ uint32_t cell_code(cell_t c) { return encode_code(c_type(c), c_color(c), c_state(c)); }
__host__ __device__ __forceinline__ cell_t decode_code(uint32_t code)      // aux is not in the code
{
    int t = code & 15, c = (code >> 4) & 7, s = 0;
    if (t >= 13) { s = t - 12; t = T_DOOR; }
    return make_cell(t, c, s, 0);
}
// Can this (type,color,state) be packed?  (checked on the host for every uploaded cell)
__host__ __device__ inline bool code_representable(int t, int c, int s)
{
    if (t < 1 || t > 10 || c > 7) return false;
    if (t == T_DOOR) return s <= 2;
    return s == 0;
}

// ---- intrinsics with host twins --------------------------------------------------------------
#ifndef MG_SHR_FMA
#define MG_SHR_FMA 1         // 1: off-chain constant right shifts as mul.hi (see mg_shr_fma)
#endif
__host__ __device__ __forceinline__ uint32_t mg_prmt(uint32_t a, uint32_t b, uint32_t sel)
{
#ifdef __CUDA_ARCH__
    return __byte_perm(a, b, sel);
#else
    const uint64_t v = ((uint64_t)b << 32) | a;
    uint32_t r = 0;
    for (int i = 0; i < 4; ++i) r |= (uint32_t)((v >> (8 * ((sel >> (4 * i)) & 7))) & 0xff) << (8 * i);
    return r;
#endif
}
// every byte -> 0xFF if its top bit is set, else 0x00
__host__ __device__ __forceinline__ uint32_t mg_sign_bytes(uint32_t w)
{
#ifdef __CUDA_ARCH__
    // PTX prmt (default mode): selector nibble 8 + i = replicate the sign of byte i.  Not __byte_perm(): the intrinsic is
    // specified to look at 3 bits per nibble only and the compiler folds it that way (round-2 GPU run: every observation wrong).
    uint32_t r;
    asm("prmt.b32 %0, %1, %2, 0xba98;" : "=r"(r) : "r"(w), "r"(0u));
    return r;
#else
    return ((w >> 7) & 0x01010101u) * 0xFFu;
#endif
}
__host__ __device__ __forceinline__ uint32_t mg_shf_r(uint32_t lo, uint32_t hi, uint32_t sh)   // (hi:lo) >> sh, sh < 32
{
#ifdef __CUDA_ARCH__
    return __funnelshift_r(lo, hi, sh);
#else
    return (uint32_t)(((((uint64_t)hi) << 32) | lo) >> (sh & 31));
#endif
}
// x >> K (0 < K < 32) as the high half of a multiply: IMAD.HI runs on the FMA pipe, which this kernel leaves almost idle, while
// SHF competes with LOP3 / PRMT / ISETP for the ALU pipe — the step kernel's busiest unit.  For shifts that are not on a
// dependency chain (the multiply's latency is longer).
template <int K>
__host__ __device__ __forceinline__ uint32_t mg_shr_fma(uint32_t x)
{
#if defined(__CUDA_ARCH__) && MG_SHR_FMA
    uint32_t r;
    asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(x), "n"(1u << (32 - K)));
    return r;
#else
    return x >> K;
#endif
}
__host__ __device__ __forceinline__ uint32_t mg_brev32(uint32_t m)
{
#ifdef __CUDA_ARCH__
    return __brev(m);
#else
    uint32_t r = 0;
    for (int i = 0; i < 32; ++i) r |= ((m >> i) & 1u) << (31 - i);
    return r;
#endif
}
__host__ __device__ __forceinline__ uint32_t mg_brev7(uint32_t m)     // reverse the low 7 bits
{
#ifdef __CUDA_ARCH__
    return __brev(m) >> 25;
#else
    uint32_t r = 0;
    for (int i = 0; i < 7; ++i) r |= ((m >> i) & 1u) << (6 - i);
    return r;
#endif
}

// Word stride WS of a tile: 1 = the packed record as it lies in HBM; the step kernel keeps the tiles of its 32 envs
// interleaved word by word in shared memory (WS = 32: word w of lane l's tile is word 32 w + l of the block), so every
// access of lane l falls into bank l whatever the agent's position — with contiguous tiles the 21 fetches of an observation
// took 3.5 wavefronts each (ncu: 48 % of the kernel's shared-memory wavefronts were bank conflicts).
// tile_at<WS>(i): byte offset of logical byte i from the tile's base.
template <int WS>
__host__ __device__ __forceinline__ int tile_at(int i) { return WS == 1 ? i : (i & ~3) * WS + (i & 3); }

// 7 bytes at byte offset p of `tile` (4-byte aligned base) -> lo (bytes 0-3), hi (bytes 4-6, top byte garbage)
template <int WS = 1>
__host__ __device__ __forceinline__ void fetch7(const uint8_t* tile, int p, uint32_t& lo, uint32_t& hi)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(tile + (p & ~3) * WS);
    const uint32_t w0 = w[0], w1 = w[WS], w2 = w[2 * WS];
    const uint32_t sh = (uint32_t)(p & 3) * 8;
    lo = mg_shf_r(w0, w1, sh);
    hi = mg_shf_r(w1, w2, sh);
}

// Grid.process_vis (core/grid.py:291-328) on 7-bit row masks; t[j] bit a = view cell (a, j) is transparent.
// In-row floods are carry chains: (t + (m & t)) ^ t marks each lit transparent run and the cell after it.
__host__ __device__ __forceinline__ void process_vis_rows(const uint32_t t[VIEW], uint32_t vis[VIEW])
{
    uint32_t m = 1u << (VIEW / 2);                       // agent cell (3, 6), grid.py:294
#pragma unroll
    for (int j = VIEW - 1; j >= 0; --j) {
        const uint32_t tj = t[j];
        m = (m | ((tj + (m & tj)) ^ tj)) & 0x7fu;        // left-to-right sweep, grid.py:297-308
        uint32_t mr = mg_brev7(m);
        const uint32_t tr = mg_brev7(tj);
        mr |= ((tr + (mr & tr)) ^ tr);                   // right-to-left sweep, grid.py:310-321
        m = mg_brev7(mr & 0x7fu);
        vis[j] = m;
        const uint32_t e = m & tj;                       // lit transparent cells light the 3 cells above
        m = (e | (e << 1) | (e >> 1)) & 0x7fu;
    }
}

#ifndef MG_VIS_ALTERNATE
#define MG_VIS_ALTERNATE 1   // 1: process_vis_packed (one bit reversal per view row instead of two + one per row for t); 0: process_vis_rows
#endif
// process_vis on the packed opacity words (mlo / mhi: byte b = opacity bits of view row b, bit a; the top byte of mhi is
// garbage) -> vlo / vhi (byte b = visible bits of row b).  Same result as process_vis_rows, fewer instructions on the serial
// chain: the two in-row floods of a row commute (each only lights cells on its own side of the cells that were lit when the row
// started), and the step to the row above is symmetric under bit reversal — so a row floods low-to-high in the bit order it
// arrives in, is reversed ONCE, floods low-to-high again and hands the result to the next row in that reversed order.  Rows
// 6, 4, 2, 0 therefore end up bit-reversed; they are put right together by one 32-bit reversal (byte = (row << 1)), and the
// reversed transparency masks of all rows come from two more.
__host__ __device__ __forceinline__ void process_vis_packed(uint32_t mlo, uint32_t mhi, uint32_t& vlo, uint32_t& vhi)
{
    const uint32_t Tlo = ~mlo & 0x7f7f7f7fu, Thi = ~mhi & 0x007f7f7fu;          // byte b = transparency of row b / row 4 + b
    const uint32_t Rlo = mg_brev32(Tlo << 1), Rhi = mg_brev32(Thi << 1);        // byte 3 - b = the same, bit-reversed (7 bits)
    uint32_t t[VIEW], r[VIEW];
    t[0] = mg_prmt(Tlo, 0u, 0x4440u); t[1] = mg_prmt(Tlo, 0u, 0x4441u); t[2] = mg_prmt(Tlo, 0u, 0x4442u); t[3] = mg_prmt(Tlo, 0u, 0x4443u);
    t[4] = mg_prmt(Thi, 0u, 0x4440u); t[5] = mg_prmt(Thi, 0u, 0x4441u); t[6] = mg_prmt(Thi, 0u, 0x4442u);
    r[0] = mg_prmt(Rlo, 0u, 0x4443u); r[1] = mg_prmt(Rlo, 0u, 0x4442u); r[2] = mg_prmt(Rlo, 0u, 0x4441u); r[3] = mg_prmt(Rlo, 0u, 0x4440u);
    r[4] = mg_prmt(Rhi, 0u, 0x4443u); r[5] = mg_prmt(Rhi, 0u, 0x4442u); r[6] = mg_prmt(Rhi, 0u, 0x4441u);
    uint32_t vis[VIEW];
    uint32_t m = 1u << (VIEW / 2);                       // agent cell (3, 6), grid.py:294 — the same bit in both orders
#pragma unroll
    for (int j = VIEW - 1; j >= 0; --j) {
        const bool fwd = ((VIEW - 1 - j) & 1) == 0;      // row arrives in normal bit order (rows 6, 4, 2, 0) or reversed
        const uint32_t a = fwd ? t[j] : r[j], b = fwd ? r[j] : t[j];
        m = (m | ((a + (m & a)) ^ a)) & 0x7fu;           // flood towards the high bits
        m = mg_brev7(m);
        m = (m | ((b + (m & b)) ^ b)) & 0x7fu;           // the other direction
        vis[j] = m;                                      // rows 6, 4, 2, 0: bit-reversed
        const uint32_t e = m & b;                        // lit transparent cells light the 3 cells above
        m = (e | (e << 1) | (e >> 1)) & 0x7fu;
    }
    // rows 0, 2, 4, 6 (reversed): one word, byte k = row 2k << 1; reversing it puts the right-way-round row 2k in byte 3 - k
    const uint32_t R = mg_brev32((vis[0] << 1) + (vis[2] << 9) + (vis[4] << 17) + (vis[6] << 25));
    const uint32_t N = vis[1] + (vis[3] << 8) + (vis[5] << 16);                  // rows 1, 3, 5 are the right way round
    vlo = mg_prmt(R, N, 0x5243u);                        // rows 0 (R byte 3), 1 (N byte 0), 2 (R byte 2), 3 (N byte 1)
    vhi = mg_prmt(R, N, 0x7061u);                        // rows 4 (R byte 1), 5 (N byte 2), 6 (R byte 0), 0 (N byte 3)
}

#ifndef MG_OBS_PRMT
#define MG_OBS_PRMT 1        // 1: byte-permute forms of the visibility masks and of the record merges (A/B: tools/ab_bench.sh)
#endif
#ifndef MG_DECODE_LUT
#define MG_DECODE_LUT 1      // 1: shared-memory decode table (LSU pipe, +8..12 % on B200); 0: SIMD-in-register decode (ALU pipe)
#endif

// Four packed codes -> byte planes type / color / state (SIMD within a register).
__host__ __device__ __forceinline__ void decode4(uint32_t x, uint32_t& ty, uint32_t& co, uint32_t& st)
{
    const uint32_t t = x & 0x0F0F0F0Fu;
    co = (x >> 4) & 0x07070707u;
    const uint32_t dm = (((t + 0x03030303u) >> 4) & 0x01010101u) * 0xFFu;   // 0xFF where type' >= 13 (closed/locked door)
    ty = (t & ~dm) | (dm & 0x04040404u);
    st = t & dm & 0x03030303u;
}

// Decode table: code byte -> type | color << 8 | state << 16 (entry 0 = invisible -> (0,0,0)).
// The kernel keeps it in shared memory: the table lookups run on the load/store pipe, which is idle, instead of
// ~10 ALU-pipe instructions per 4 cells (LOP3 / SHF / PRMT issue at half rate on sm_100).
#ifndef MG_LUT_BYTE_INDEX
#define MG_LUT_BYTE_INDEX 1  // 1: the index is the code byte itself, extracted with ONE byte permute (the opaque bit is cleared for
#endif                       //    free by the visibility mask's AND); 0: shift + and per lookup
constexpr int LUT_ENTRIES = 128;         // indexed by code & 0x7f (the opaque bit does not change the decoded triple)
__host__ __device__ __forceinline__ uint32_t lut_entry(uint32_t code)
{
    if ((code & 0x7f) == 0) return 0;
    return decode_code(code) & 0x00FFFFFFu;
}
// table index of byte b of a word of packed codes
template <int B>
__host__ __device__ __forceinline__ uint32_t lut_index(uint32_t x)
{
#if MG_LUT_BYTE_INDEX
    return mg_prmt(x, 0u, 0x4440u | (uint32_t)B);            // byte B (bit 7 already clear), zero-extended: one ALU instruction
#else
    return (x >> (8 * B)) & 0x7Fu;
#endif
}

// Stages A + B of gen_obs_grid (minigrid_env.py:592-627) for one env: lo[a] / hi[a] = packed codes of view column a
// (byte b = view cell (a, b); the top byte of hi is garbage), carried object already placed at (3, 6);
// vlo / vhi = vis_mask, byte b = row b, bit a.  ax, ay, dir: agent pose; carry_code: packed code of the carried
// object or CODE_EMPTY.  Shared by the observation encoder below and the RGB renderers (mg_render.cu).
// 4x4 byte transpose: o[i] byte j = w[j] byte i  (8 byte permutes)
__host__ __device__ __forceinline__ void transpose4x4(uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3,
                                                      uint32_t& o0, uint32_t& o1, uint32_t& o2, uint32_t& o3)
{
    const uint32_t t0 = mg_prmt(w0, w1, 0x5140u), t1 = mg_prmt(w2, w3, 0x5140u);     // bytes 0,1 of the four words, pairwise
    const uint32_t t2 = mg_prmt(w0, w1, 0x7362u), t3 = mg_prmt(w2, w3, 0x7362u);     // bytes 2,3
    o0 = mg_prmt(t0, t1, 0x5410u); o1 = mg_prmt(t0, t1, 0x7632u);
    o2 = mg_prmt(t2, t3, 0x5410u); o3 = mg_prmt(t2, t3, 0x7632u);
}

// ROWS = true: the tile holds only the row-major copy R (half the shared memory: grids above 8x8 keep twice as many CTAs
// resident).  An agent facing +-y then fetches the 7 view ROWS (7 contiguous bytes of R each, one per b) and transposes
// the 7x7 byte matrix in registers (30 byte permutes); facing +-x is the same fetch as before.
template <bool ROWS = false, int WS = 1>
__host__ __device__ __forceinline__ void view_codes(const uint8_t* tile, const TileGeom& g, int ax, int ay, int dir,
                                                    uint32_t carry_code, bool see_through,
                                                    uint32_t lo[VIEW], uint32_t hi[VIEW], uint32_t& vlo, uint32_t& vhi)
{
    // ---- stage A: fetch the 7 view columns -------------------------------------------------
    const bool horiz = (dir & 1) == 0;                   // facing +-x: columns run along x -> R
    const bool rev = (dir == 0) || (dir == 1);           // b grows against the memory order
    if (ROWS) {
        // line k of the fetch is y = ay + c0 + c1*k; facing +-x: k = a (view column), 7 bytes from x = ax (reversed) or ax-6;
        // facing +-y: k = b (view row), 7 bytes from x = ax-3, reversed (a = 6 - i) when facing +y
        const int sgn = (dir == 0 || dir == 3) ? 1 : -1;
        const int c1 = horiz ? sgn : (dir == 1 ? -1 : 1);
        const int c0 = horiz ? -3 * sgn : (dir == 1 ? 6 : -6);
        const int start = PAD + (horiz ? (rev ? ax : ax - (VIEW - 1)) : ax - VIEW / 2);
        const uint32_t sel_lo = rev ? 0x3456u : 0x3210u, sel_hi = rev ? 0x7012u : 0x7654u;
        uint32_t ml[8], mh[8];
#pragma unroll
        for (int k = 0; k < VIEW; ++k) {
            const int line = ay + c0 + c1 * k;
            const bool ok = (unsigned)line < (unsigned)g.H;
            uint32_t l, h;
            fetch7<WS>(tile, start + (ok ? line : ay) * g.PR, l, h);
            l = ok ? l : WALL4; h = ok ? h : WALL4;
            ml[k] = mg_prmt(l, h, sel_lo);
            mh[k] = mg_prmt(l, h, sel_hi);
        }
        ml[7] = 0; mh[7] = 0;
        uint32_t tl[8], th[8];                             // transposed: tl[a] bytes b = 0..3, th[a] bytes b = 4..6
        transpose4x4(ml[0], ml[1], ml[2], ml[3], tl[0], tl[1], tl[2], tl[3]);
        transpose4x4(ml[4], ml[5], ml[6], ml[7], th[0], th[1], th[2], th[3]);
        transpose4x4(mh[0], mh[1], mh[2], mh[3], tl[4], tl[5], tl[6], tl[7]);
        transpose4x4(mh[4], mh[5], mh[6], mh[7], th[4], th[5], th[6], th[7]);
#pragma unroll
        for (int a = 0; a < VIEW; ++a) {
            lo[a] = horiz ? ml[a] : tl[a];
            hi[a] = horiz ? mh[a] : th[a];
        }
    } else {
    const int sgn = (dir == 0 || dir == 3) ? 1 : -1;     // line index grows with a?
    const int pitch = horiz ? g.PR : g.PC;
    const int base = horiz ? 0 : g.offC;
    const int u = horiz ? ax : ay, v = horiz ? ay : ax;
    const int nlines = horiz ? g.H : g.W;
    const int start = base + (rev ? u : u - (VIEW - 1)) + PAD;
    const uint32_t sel_lo = rev ? 0x3456u : 0x3210u, sel_hi = rev ? 0x7012u : 0x7654u;
#pragma unroll
    for (int a = 0; a < VIEW; ++a) {
        const int line = v + sgn * (a - VIEW / 2);
        const bool ok = (unsigned)line < (unsigned)nlines;           // out-of-grid line -> a column of walls (branch-free:
        uint32_t l, h;                                               // fetch a clamped line, then select)
        fetch7<WS>(tile, start + (ok ? line : v) * pitch, l, h);
        l = ok ? l : WALL4; h = ok ? h : WALL4;
        lo[a] = mg_prmt(l, h, sel_lo);
        hi[a] = mg_prmt(l, h, sel_hi);                               // top byte: a neighbour's cell, never used
    }
    }
    // ---- stage B: visibility ---------------------------------------------------------------
    vlo = 0x7f7f7f7fu; vhi = 0x007f7f7fu;                // byte b = visible-bits of view row b
    if (!see_through) {
        uint32_t mlo = 0, mhi = 0;                       // byte b = opacity bits of view row b (bit a)
        // bit 7 of byte b of column a -> bit a of byte b
        mlo |= mg_shr_fma<7>(lo[0]) & 0x01010101u; mhi |= mg_shr_fma<7>(hi[0]) & 0x01010101u;
        mlo |= mg_shr_fma<6>(lo[1]) & 0x02020202u; mhi |= mg_shr_fma<6>(hi[1]) & 0x02020202u;
        mlo |= mg_shr_fma<5>(lo[2]) & 0x04040404u; mhi |= mg_shr_fma<5>(hi[2]) & 0x04040404u;
        mlo |= mg_shr_fma<4>(lo[3]) & 0x08080808u; mhi |= mg_shr_fma<4>(hi[3]) & 0x08080808u;
        mlo |= mg_shr_fma<3>(lo[4]) & 0x10101010u; mhi |= mg_shr_fma<3>(hi[4]) & 0x10101010u;
        mlo |= mg_shr_fma<2>(lo[5]) & 0x20202020u; mhi |= mg_shr_fma<2>(hi[5]) & 0x20202020u;
        mlo |= mg_shr_fma<1>(lo[6]) & 0x40404040u; mhi |= mg_shr_fma<1>(hi[6]) & 0x40404040u;
        static_assert(VIEW == 7, "the opacity gather is written out for a 7x7 view");
#if MG_VIS_ALTERNATE
        process_vis_packed(mlo, mhi, vlo, vhi);
#else
        uint32_t t[VIEW], vis[VIEW];
#pragma unroll
        for (int b = 0; b < 4; ++b) t[b] = (~(mlo >> (8 * b))) & 0x7fu;
#pragma unroll
        for (int b = 4; b < VIEW; ++b) t[b] = (~(mhi >> (8 * (b - 4)))) & 0x7fu;
        process_vis_rows(t, vis);
        vlo = vis[0] | (vis[1] << 8) | (vis[2] << 16) | (vis[3] << 24);
        vhi = vis[4] | (vis[5] << 8) | (vis[6] << 16);
#endif
    }
    // carried object (or empty) is shown at the agent's own cell (3, 6): minigrid_env.py:618-625
    hi[VIEW / 2] = (hi[VIEW / 2] & 0xFF00FFFFu) | (carry_code << 16);
}

// The whole observation of one env as 37 little-endian words (147 bytes + 1 zero byte): image[a][b][c]
// (gen_obs, minigrid_env.py:629-645 = gen_obs_grid + Grid.encode(vis_mask), core/grid.py:244-268).
template <bool ROWS = false, int WS = 1>
__host__ __device__ __forceinline__ void obs_words(const uint8_t* tile, const TileGeom& g, int ax, int ay, int dir,
                                                   uint32_t carry_code, bool see_through, const uint32_t* lut,
                                                   uint32_t out[37])
{
    uint32_t lo[VIEW], hi[VIEW], vlo, vhi;
    view_codes<ROWS, WS>(tile, g, ax, ay, dir, carry_code, see_through, lo, hi, vlo, vhi);
#if MG_DECODE_LUT
    // ---- stage C/D: mask, table-decode, concatenate ------------------------------------------
    // cell i = a*7 + b owns bytes 3i..3i+2 of the record; four cells fill three words.
    uint32_t tr[VIEW * VIEW];
#pragma unroll
    for (int a = 0; a < VIEW; ++a) {
        // byte mask of column a: bit a of every row byte moved to the byte's sign bit, then replicated over the byte (PRMT with
        // the sign-replication flag: 3 ALU instructions per word instead of shift / and / multiply / and)
#if MG_OBS_PRMT
        // (& 0x7f7f7f7f: drops the opaque bits so that a code byte is its own table index; same LOP3 as the mask's AND)
        const uint32_t xl = lo[a] & mg_sign_bytes(vlo << (7 - a)) & 0x7f7f7f7fu;   // invisible -> code 0 -> (0,0,0)
        const uint32_t xh = hi[a] & mg_sign_bytes(vhi << (7 - a)) & 0x7f7f7f7fu;
#else
        const uint32_t xl = lo[a] & (((vlo >> a) & 0x01010101u) * 0x7Fu);
        const uint32_t xh = hi[a] & (((vhi >> a) & 0x01010101u) * 0x7Fu);
#endif
        tr[a * VIEW + 0] = lut[lut_index<0>(xl)]; tr[a * VIEW + 1] = lut[lut_index<1>(xl)];
        tr[a * VIEW + 2] = lut[lut_index<2>(xl)]; tr[a * VIEW + 3] = lut[lut_index<3>(xl)];
        tr[a * VIEW + 4] = lut[lut_index<0>(xh)]; tr[a * VIEW + 5] = lut[lut_index<1>(xh)];
        tr[a * VIEW + 6] = lut[lut_index<2>(xh)];
    }
#pragma unroll
    for (int gq = 0; gq < 12; ++gq) {
        const uint32_t t0 = tr[4 * gq], t1 = tr[4 * gq + 1], t2 = tr[4 * gq + 2], t3 = tr[4 * gq + 3];
        out[3 * gq] = t0 + (t1 << 24);                       // disjoint bit ranges: + is |, and maps to ONE IMAD
#if MG_OBS_PRMT
        out[3 * gq + 1] = mg_prmt(t1, t2, 0x5421u);          // t1 bytes 1,2 | t2 bytes 0,1   (one PRMT instead of SHF + IMAD)
        out[3 * gq + 2] = mg_prmt(t2, t3, 0x6542u);          // t2 byte 2   | t3 bytes 0,1,2
#else
        out[3 * gq + 1] = (t1 >> 8) + (t2 << 16);
        out[3 * gq + 2] = (t2 >> 16) + (t3 << 8);
#endif
    }
    out[36] = tr[48];
#else
    (void)lut;
    // ---- stage C/D: mask, decode, interleave, concatenate ------------------------------------
#pragma unroll
    for (int i = 0; i < 37; ++i) out[i] = 0;
#pragma unroll
    for (int a = 0; a < VIEW; ++a) {
        const uint32_t xl = lo[a] & (((vlo >> a) & 0x01010101u) * 0xFFu);      // invisible -> code 0 -> (0,0,0)
        const uint32_t xh = hi[a] & (((vhi >> a) & 0x01010101u) * 0xFFu);
        uint32_t ty, co, st, c[6];
        decode4(xl, ty, co, st);
        c[0] = mg_prmt(mg_prmt(ty, co, 0x1040u), st, 0x3410u);      // ty0 c0 s0 ty1
        c[1] = mg_prmt(mg_prmt(co, st, 0x2051u), ty, 0x3610u);      // c1 s1 ty2 c2
        c[2] = mg_prmt(mg_prmt(st, ty, 0x3072u), co, 0x3710u);      // s2 ty3 c3 s3
        decode4(xh, ty, co, st);
        c[3] = mg_prmt(mg_prmt(ty, co, 0x1040u), st, 0x3410u);      // ty4 c4 s4 ty5
        c[4] = mg_prmt(mg_prmt(co, st, 0x2051u), ty, 0x3610u);      // c5 s5 ty6 c6
        c[5] = (st >> 16) & 0xFFu;                                   // s6
        const int o = 21 * a, wi = o >> 2, sh = (o & 3) * 8;        // static after unrolling
        if (sh == 0) {
#pragma unroll
            for (int j = 0; j < 6; ++j) out[wi + j] |= c[j];
        } else {
            out[wi] |= c[0] << sh;
#pragma unroll
            for (int j = 1; j < 6; ++j) out[wi + j] |= mg_shf_r(c[j - 1], c[j], 32 - sh);
            if (wi + 6 < 37) out[wi + 6] |= c[5] >> (32 - sh);
        }
    }
#endif
}

}  // namespace mg

namespace mg {

// Build the packed tile of one env from its SoA planes (serial; used by the host tests and as the
// definition the pack kernel must agree with).
__host__ __device__ inline void pack_tile_serial(const uint8_t* cells, int ps, const TileGeom& g, uint8_t* tile)
{
    for (int i = 0; i < g.bytes; ++i) tile[i] = (uint8_t)CODE_WALL;
    for (int y = 0; y < g.H; ++y)
        for (int x = 0; x < g.W; ++x) {
            const int k = y * g.W + x;
            const uint8_t code = (uint8_t)encode_code(cells[k], cells[ps + k], cells[2 * ps + k]);
            tile[y * g.PR + x + PAD] = code;
            tile[g.offC + x * g.PC + y + PAD] = code;
        }
}

}  // namespace mg
