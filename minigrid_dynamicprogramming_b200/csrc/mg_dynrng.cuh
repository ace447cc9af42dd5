// The env's own numpy Generator(PCG64) on the device, for DynamicObstaclesEnv's in-step draws (envs/dynamicobstacles.py:136-167
// -> MiniGridEnv.place_obj -> _rand_int, minigrid_env.py:244-249,308-373).  numpy is third-party and absent from
// /root/reference: the algorithms are restated from its published sources — pcg64.h pcg_setseq_128_xsl_rr_64 (128-bit LCG,
// XSL-RR output), next_uint32's half-word buffer (has_uint32 / uinteger), distributions.c buffered_bounded_lemire_uint32.
// Everything is __host__ __device__: tests/host builds the same arithmetic for the CPU suite, which pins it against numpy itself.
#pragma once
#include <cstdint>

namespace mg {

#ifndef MG_DYN_FAST_DRAW
#define MG_DYN_FAST_DRAW 1      // one generator step per placement try in the common case (draw_pair)
#endif

// the generator part of a mg_dyn_record (include/mgb200.h), held in registers
struct DynRng {
    uint64_t s_lo, s_hi, inc_lo, inc_hi;
    uint32_t has32, u32;
};

__host__ __device__ __forceinline__ uint64_t mul64hi(uint64_t a, uint64_t b)
{
#ifdef __CUDA_ARCH__
    return __umul64hi(a, b);
#else
    return (uint64_t)(((unsigned __int128)a * b) >> 64);
#endif
}

// state' = state * mult + inc (mod 2^128); returns the XSL-RR output of the NEW state without committing it
__host__ __device__ __forceinline__ uint64_t pcg64_peek(const DynRng& d, uint64_t& lo, uint64_t& hi)
{
    constexpr uint64_t M_HI = 0x2360ED051FC65DA4ull, M_LO = 0x4385DF649FCCF645ull;
    lo = d.s_lo * M_LO;
    hi = mul64hi(d.s_lo, M_LO) + d.s_lo * M_HI + d.s_hi * M_LO;
    lo += d.inc_lo;
    hi += d.inc_hi + (lo < d.inc_lo ? 1ull : 0ull);
    const uint64_t x = hi ^ lo;
    const unsigned rot = (unsigned)(hi >> 58);
    return (x >> rot) | (x << ((64u - rot) & 63u));
}
__host__ __device__ __forceinline__ uint64_t pcg64_next64(DynRng& d)
{
    uint64_t lo, hi;
    const uint64_t v = pcg64_peek(d, lo, hi);
    d.s_lo = lo; d.s_hi = hi;
    return v;
}
__host__ __device__ __forceinline__ uint32_t pcg64_next32(DynRng& d)
{
    if (d.has32) { d.has32 = 0; return d.u32; }
    const uint64_t v = pcg64_next64(d);
    d.has32 = 1; d.u32 = (uint32_t)(v >> 32);
    return (uint32_t)v;
}
// MiniGridEnv._rand_int (minigrid_env.py:244-249) = Generator.integers(low, high): Lemire rejection on 32-bit draws
__host__ __device__ __forceinline__ int rand_int(DynRng& d, int low, int high)
{
    const uint32_t rng = (uint32_t)(high - 1 - low);
    if (rng == 0) return low;                                                  // no draw
    const uint32_t rng_excl = rng + 1;
    uint64_t m = (uint64_t)pcg64_next32(d) * rng_excl;
    uint32_t leftover = (uint32_t)m;
    if (leftover < rng_excl) {
        const uint32_t threshold = (0xFFFFFFFFu - rng) % rng_excl;
        while (leftover < threshold) { m = (uint64_t)pcg64_next32(d) * rng_excl; leftover = (uint32_t)m; }
    }
    return low + (int)(m >> 32);
}

// (x, y) = (_rand_int(tx, hx), _rand_int(ty, hy)): one placement try of place_obj (minigrid_env.py:340-345).
// The two bounded draws are the two halves of ONE 64-bit output, whichever half the generator's buffer is at — as long as both
// coordinates are drawn (ranges wider than one value) and neither Lemire product lands in the rejection zone (low word < range:
// 3 in 2^32 for the 3x3 box).  That case is computed straight-line with ONE generator step.  With rand_int alone the lanes of a
// warp sit at both buffer phases, so the warp executed the 128-bit step at both draw sites of every try.  Anything else takes
// the general path, from the untouched state.  Same values and same generator state either way (tests/test_pipeline_host.py).
// draw_pair_from: the same with the generator step already taken by the caller — (v, lo, hi) = pcg64_peek(d) — so that a loop
// can compute the NEXT try's output while this try's cell lookup is in flight (move_obstacles).
__host__ __device__ __forceinline__ void draw_pair_from(DynRng& d, uint64_t v, uint64_t lo, uint64_t hi,
                                                        int tx, int hx, int ty, int hy, int& px, int& py)
{
    const uint32_t v_lo = (uint32_t)v, v_hi = (uint32_t)(v >> 32);
    const uint32_t ex = (uint32_t)(hx - tx), ey = (uint32_t)(hy - ty);
    const uint64_t ma = (uint64_t)(d.has32 ? d.u32 : v_lo) * ex;
    const uint64_t mb = (uint64_t)(d.has32 ? v_lo : v_hi) * ey;
    if ((ex > 1u) & (ey > 1u) & ((uint32_t)ma >= ex) & ((uint32_t)mb >= ey)) {
        d.s_lo = lo; d.s_hi = hi; d.u32 = v_hi;                                 // has32 keeps its phase
        px = tx + (int)(ma >> 32); py = ty + (int)(mb >> 32);
        return;
    }
    px = rand_int(d, tx, hx);
    py = rand_int(d, ty, hy);
}
__host__ __device__ __forceinline__ void draw_pair(DynRng& d, int tx, int hx, int ty, int hy, int& px, int& py)
{
#if MG_DYN_FAST_DRAW
    uint64_t lo, hi;
    const uint64_t v = pcg64_peek(d, lo, hi);
    draw_pair_from(d, v, lo, hi, tx, hx, ty, hy, px, py);
#else
    px = rand_int(d, tx, hx);
    py = rand_int(d, ty, hy);
#endif
}

}  // namespace mg
