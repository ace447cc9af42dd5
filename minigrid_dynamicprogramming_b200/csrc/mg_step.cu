// mg_step.cu — fused step + partial-observation kernel (K1+K2+K3 of SURVEY §8.7), reset and pack kernels.
//
// Replaces MiniGridEnv.step (minigrid/minigrid_env.py:520-590), gen_obs_grid / gen_obs (:592-645),
// Grid.slice / rotate_left / process_vis / encode (minigrid/core/grid.py:110-143, 244-328),
// _reward (:235-240), KeyCorridorEnv.step (minigrid/envs/keycorridor.py:129-137) and the
// state-copy part of reset (:119-157).
//
// Design (one thread = one env, one CTA = one warp = 32 envs, T steps per launch):
//   * the CTA stages its packed tiles (mg_packed.cuh) in shared memory with coalesced 16-byte
//     loads and keeps them there for all T steps; mutations (pickup / drop / toggle — rare) are
//     written through to the HBM copies (SoA planes + packed tile) so HBM stays authoritative;
//   * per step a thread reads ONE byte (front cell) for the action switch, then builds the whole
//     7x7x3 observation in registers from 7 contiguous 7-byte column fetches (obs_words);
//   * the 147-byte records are laid out back to back in a staging buffer (147-byte stride, so the
//     CTA's block is exactly the bytes of image[t][env0 .. env0+31]) and leave as ONE bulk copy
//     per step (cp.async.bulk, issued by an elected lane; ragged CTAs: 16-byte streaming stores):
//     HBM sees only full-sector writes;
//   * same-step autoreset reloads a done env from the layout pool warp-cooperatively.
#include "mg_step.cuh"
#include "mg_dynrng.cuh"

namespace mg {

__device__ __forceinline__ uint32_t funnel_l(uint32_t lo, uint32_t hi, uint32_t sh) { return __funnelshift_l(lo, hi, sh); }
// the same with the shift given as a multiplier m = 1 << sh (sh < 32): hi * m + (lo * m >> 32) — two FMA-pipe instructions
// (IMAD.HI + IMAD) instead of one SHF on the ALU pipe, the step kernel's busiest unit (A/B: -DMG_STAGE_FMA=0)
#ifndef MG_STAGE_FMA
#define MG_STAGE_FMA 0
#endif
__device__ __forceinline__ uint32_t funnel_l_mul(uint32_t lo, uint32_t hi, uint32_t m) { return hi * m + __umulhi(lo, m); }

// Store this thread's 37-word record at byte offset 147*rec of the staging buffer (not word aligned: 147*rec mod 4 =
// (3*rec) mod 4).  The record's first and last word are shared with the neighbouring records.  MG_STAGE_SHFL (default): a
// lane hands its zero-padded tail fragment to the next lane (one shuffle), which ORs it into its own zero-padded head
// fragment and stores the whole word — instead of up to three byte stores from either side (~25 instructions and six STS.U8
// per record).  Every lane named in `mask` must call it (records of inactive envs: S = 0), consecutive lanes = consecutive
// records; record 0 and every 16th record start word-aligned (16 * 147 = 588 words), so no fragment crosses a half-warp or a
// warp, and the last record of a half-warp ends on a word boundary.
#ifndef MG_STAGE_SHFL
#define MG_STAGE_SHFL 1
#endif
__device__ __forceinline__ void stage_record(uint8_t* staging, int rec, const uint32_t S[37], unsigned mask)
{
    const int A = rec * OBS_BYTES, s = A & 3;
    uint8_t* base = staging + (A - s);
    uint32_t* wbase = reinterpret_cast<uint32_t*>(base);
    const uint32_t sh = 8u * s;
    const uint32_t f0 = S[0] << sh;
#if MG_STAGE_FMA
    uint32_t mul;                                       // 1 << sh, opaque to the compiler (it would turn hi * mul back into SHF)
    asm("shl.b32 %0, 1, %1;" : "=r"(mul) : "r"(sh));
#pragma unroll
    for (int k = 1; k <= 35; ++k) wbase[k] = funnel_l_mul(S[k - 1], S[k], mul);
    const uint32_t f36 = funnel_l_mul(S[35], S[36], mul);
    const uint32_t f37 = __umulhi(S[36], mul);
#else
#pragma unroll
    for (int k = 1; k <= 35; ++k) wbase[k] = funnel_l(S[k - 1], S[k], sh);
    const uint32_t f36 = funnel_l(S[35], S[36], sh);
    const uint32_t f37 = funnel_l(S[36], 0u, sh);
#endif
#if MG_STAGE_SHFL
    // tail fragment: s == 0: word 36 holds the record's last 3 bytes (S[36]'s top byte is zero); s == 1: the record ends with
    // word 36 (f37 = 0); s >= 2: word 37 holds its last 1 or 2 bytes.  All zero-padded towards the next record's bytes.
    const uint32_t tail = s == 0 ? f36 : f37;
    const uint32_t prev = __shfl_up_sync(mask, tail, 1);
    wbase[0] = s == 0 ? f0 : (prev | f0);               // f0 is zero-padded towards the previous record's bytes
    if (s != 0) wbase[36] = f36;
#else
    (void)mask;
    if (s == 0) {
        wbase[0] = f0;
        base[144] = (uint8_t)f36; base[145] = (uint8_t)(f36 >> 8); base[146] = (uint8_t)(f36 >> 16);
    } else {
        wbase[36] = f36;
#pragma unroll
        for (int j = 1; j < 4; ++j) if (j >= s) base[j] = (uint8_t)(f0 >> (8 * j));
        if (s >= 2) base[148] = (uint8_t)f37;
        if (s == 3) base[149] = (uint8_t)(f37 >> 8);
    }
#endif
}

// Copy `nbytes` from shared memory to global memory with the whole CTA, 16 bytes per thread when aligned.
__device__ __forceinline__ void cta_copy_out(uint8_t* __restrict__ dst, const uint8_t* src, int nbytes)
{
    if ((((uintptr_t)dst) & 15) == 0) {
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        if (nbytes == STEP_THREADS * OBS_BYTES) {            // full CTA: 588 chunks, unrolled (9 full rounds + 12 threads)
            constexpr int N16 = STEP_THREADS * OBS_BYTES / 16, ROUNDS = N16 / STEP_THREADS;
            static_assert(STEP_THREADS * OBS_BYTES % 16 == 0, "CTA block must be a whole number of 16-byte chunks");
            uint4 v[ROUNDS];
#pragma unroll
            for (int j = 0; j < ROUNDS; ++j) v[j] = s4[threadIdx.x + j * STEP_THREADS];
#pragma unroll
            for (int j = 0; j < ROUNDS; ++j) __stcs(d4 + threadIdx.x + j * STEP_THREADS, v[j]);
            if (threadIdx.x < N16 - ROUNDS * STEP_THREADS)
                __stcs(d4 + threadIdx.x + ROUNDS * STEP_THREADS, s4[threadIdx.x + ROUNDS * STEP_THREADS]);
            return;
        }
        if (nbytes == 16 * OBS_BYTES && STEP_THREADS == 32) {   // half a warp's records (LAYOUT_ROWS_HALF): 147 chunks
            uint4 v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = s4[threadIdx.x + j * 32];
#pragma unroll
            for (int j = 0; j < 4; ++j) __stcs(d4 + threadIdx.x + j * 32, v[j]);
            if (threadIdx.x < 147 - 128) __stcs(d4 + threadIdx.x + 128, s4[threadIdx.x + 128]);
            return;
        }
        const int n16 = nbytes >> 4;
        for (int i = threadIdx.x; i < n16; i += blockDim.x) __stcs(d4 + i, s4[i]);
        for (int i = (n16 << 4) + threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    } else {
        for (int i = threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    }
}

// ---- observation block -> HBM as ONE bulk copy (TMA engine, cp.async.bulk shared::cta -> global) ----
// The CTA's 32 records are contiguous both in the staging buffer and in image[t][env0 ..], so instead of every lane moving
// 18 x 16 bytes through registers (LDS.128 + STG.128), one elected lane hands the 4,704-byte block to the copy engine and the
// warp goes on with the next step; the staging buffer is only waited for (wait_group.read) right before the next step's
// records are staged, a whole step of compute later, so the wait is free.  Needs 16-byte aligned source, destination and
// size: a full CTA (32 x 147 = 16 x 294) and N * 147 divisible by 16; otherwise cta_copy_out below.
#ifndef MG_STEP_BULK_STORE
#define MG_STEP_BULK_STORE 1
#endif
__device__ __forceinline__ void bulk_store_fence()          // generic-proxy writes of this lane -> visible to the async proxy
{
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_issue(uint8_t* dst, const uint8_t* src, uint32_t bytes, uint64_t policy)
{
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;"
                 ::"l"(__cvta_generic_to_global(dst)), "r"((uint32_t)__cvta_generic_to_shared(src)), "r"(bytes), "l"(policy) : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
}
__device__ __forceinline__ void bulk_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }

// CTA-wide barrier; a single-warp CTA only needs warp convergence + shared-memory ordering
__device__ __forceinline__ void cta_sync()
{
    if (STEP_THREADS == 32) __syncwarp(); else __syncthreads();
}

// The CTA's tiles are INTERLEAVED word by word in shared memory (mg_packed.cuh, tile_at / fetch7<WS>): word w of lane l's
// tile is word STEP_THREADS * w + l of the block, lane l's tile pointer is block + 4 l, and logical byte i of the tile is
// TILE(i).  Every tile access of lane l then falls into bank l, wherever the env's agent stands (contiguous tiles: equal
// offsets were conflict-free thanks to an odd pitch, but agents at different cells made every fetch ~3.5 wavefronts).
// -DMG_TILE_INTERLEAVE=0 is the contiguous layout (A/B: tools/ab_bench.sh).
#ifndef MG_STEP_PDL
#define MG_STEP_PDL 1           // programmatic dependent launch between consecutive step kernels (see the kernel's prologue)
#endif
#ifndef MG_TILE_INTERLEAVE
#define MG_TILE_INTERLEAVE 0
#endif
#ifndef MG_DYN_PIPE
#define MG_DYN_PIPE 1           // DynamicObstacles: next try's generator step overlaps this try's cell lookup (move_obstacles)
#endif
#ifndef MG_DYN_DEFER
#define MG_DYN_DEFER 1          // DynamicObstacles: generator records live in shared memory during a launch, HBM is patched once at its end
#endif
#ifndef MG_CHAIN_ONE_WORD
#define MG_CHAIN_ONE_WORD 1     // chaining counters of a CTA index in one 64-bit word: one atomic takes the ticket and reads `finished`
#endif
#ifndef MG_STEP_CHAIN
#define MG_STEP_CHAIN 1         // consecutive step launches of an engine ordered CTA by CTA instead of grid by grid (kernel prologue)
#endif
#ifndef MG_STAGE_PB
#define MG_STAGE_PB 8           // tile staging at the start of a launch: 16-byte loads in flight per lane before the first store (1 = the old loop)
#endif
#ifndef MG_RESET_MID
#define MG_RESET_MID 1          // same-step autoreset of records up to 96 chunks per part: all loads of an env before the first store
#endif
#ifndef MG_RESET_PAIRS
#define MG_RESET_PAIRS 1        // same-step autoreset of small records: the loads of TWO done envs in flight before anything is stored
#endif
constexpr int TILE_WS = MG_TILE_INTERLEAVE ? STEP_THREADS : 1;
#define TILE(i) tile_at<TILE_WS>(i)

// Shared-memory pitch of one tile: the packed record without its 16 slack bytes, padded to an ODD number of words (the
// contiguous layout needs that: equal offsets in the tiles of 32 different envs fall into 32 different banks; the interleaved
// one just keeps the size).  (Not 16-byte aligned: tiles are staged with 4-byte words; that happens once per launch.)
__host__ __device__ inline int tile_smem_stride(int staged_bytes)
{
    const int words = staged_bytes >> 2;
    return ((words & 1) ? words : words + 1) << 2;
}
// bytes of a packed record the kernel stages: everything but the slack (dual layout R + Cm), or only R (+ gap + over-read
// slack; rows_bytes = max over classes of tile_rows_bytes) in the ROWS layout
__host__ __device__ inline int staged_bytes(int packed_stride, int rows_bytes, bool rows) { return rows ? rows_bytes : packed_stride - 16; }

// shared-memory layouts of the step kernel (template parameter LAYOUT)
enum : int { LAYOUT_DUAL = 0,        // R + Cm tiles: every view column is one contiguous fetch
             LAYOUT_ROWS = 1,        // R only; views of agents facing +-y are fetched as rows and transposed in registers
             LAYOUT_ROWS_HALF = 2 }; // R only + an observation staging buffer for 16 envs, filled and drained twice per step

// tiles (fetch7's few over-read bytes past the last tile land in the staging buffer and are discarded) | staging | class table (only the classes in use) | decode table (LUT build only).
// Every byte counts: the CTA must stay under 228 KB / 14 (incl. the 1 KB the driver reserves per CTA) for 65,536 envs to be
// ONE wave of 2,048 one-warp CTAs.
// (total per CTA: the decode table is static shared memory, the rest dynamic)
constexpr size_t STEP_STATIC_SMEM = MG_DECODE_LUT ? LUT_ENTRIES * sizeof(uint32_t) : 0;
size_t step_smem_bytes(int packed_stride, int num_classes, int rows_bytes, int layout)
{
    const int staged_envs = layout == LAYOUT_ROWS_HALF ? STEP_THREADS / 2 : STEP_THREADS;
    return (size_t)STEP_THREADS * tile_smem_stride(staged_bytes(packed_stride, rows_bytes, layout != LAYOUT_DUAL)) +
           (size_t)staged_envs * OBS_BYTES + (size_t)num_classes * sizeof(ClassDev) + (MG_DECODE_LUT ? LUT_ENTRIES * sizeof(uint32_t) : 0);
}

// The dual layout is kept while it leaves room for one wave of 65,536 envs (>= 14 one-warp CTAs per SM: grids up to 8x8).
// Larger grids stage only R — at 16x16 that is 12 CTAs per SM instead of 7, at 19x19 10 instead of 6 — and pay 30 byte
// permutes per observation for the transposed fetch.  If halving the observation staging buffer then brings the CTA back
// under the one-wave size (16x16: 17.7 KB -> 15.3 KB, 14 CTAs per SM), that is taken too: the record staging and the
// copy-out run twice per step on half a warp each (+7 % instructions) but the grid no longer has a 15 % second wave.
int step_layout(int packed_stride, int num_classes, int rows_bytes)
{
#ifdef MG_STEP_FORCE_LAYOUT
    (void)packed_stride; (void)num_classes; (void)rows_bytes;
    return MG_STEP_FORCE_LAYOUT;
#else
    auto fits14 = [&](int layout) { return (step_smem_bytes(packed_stride, num_classes, rows_bytes, layout) + 1024) * 14 <= 228 * 1024; };
    if (STEP_THREADS != 32) return fits14(LAYOUT_DUAL) ? LAYOUT_DUAL : LAYOUT_ROWS;
    if (fits14(LAYOUT_DUAL)) return LAYOUT_DUAL;
    if (fits14(LAYOUT_ROWS)) return LAYOUT_ROWS;
    return fits14(LAYOUT_ROWS_HALF) ? LAYOUT_ROWS_HALF : LAYOUT_ROWS;
#endif
}

// ---- DynamicObstaclesEnv (envs/dynamicobstacles.py:136-167): the blue-ball obstacles move INSIDE step with draws from the
// env's own numpy Generator(PCG64) (gymnasium seeding.np_random; numpy is third-party — algorithms restated from its
// published sources: pcg64.h pcg_setseq_128_xsl_rr_64 + next_uint32's half-word buffer, distributions.c
// buffered_bounded_lemire_uint32).  One 64-byte record per env (mg_dyn_record, include/mgb200.h) holds the generator state
// and `self.obstacles` in creation order.
// Record layout (mg_dyn_record): bytes 0-31 state / increment, 32-39 has_uint32 / uinteger, 40 n_obstacles, 48-55 ox[8],
// 56-63 oy[8].  The kernel keeps it in REGISTERS for the duration of the obstacle loop (four 16-byte loads, four stores; the
// eight positions of an axis are one 64-bit word) — a struct on the local-memory stack made every draw a chain of local loads.
struct DynRegs : DynRng {     // generator (mg_dynrng.cuh) + the obstacle positions, eight per 64-bit word
    uint64_t ox, oy;
};

// The obstacle loop (dynamicobstacles.py:146-157) with place_obj (minigrid_env.py:308-373) inlined: every obstacle is
// re-placed uniformly in the 3x3 box around its old cell (clipped to the grid), rejecting non-empty cells and the agent's
// cell, at most 101 samples (`max_tries=100`; the RecursionError is swallowed by `except Exception: pass`).
__device__ __noinline__ void move_obstacles(uint8_t* rec, uint8_t* tile, uint8_t* gpacked, uint8_t* gcells, int ps,
                                            int W, int H, int PR, int PC, int offC, int ax, int ay, bool rows_only)
{
    uint4* g4 = reinterpret_cast<uint4*>(rec);
    const uint4 r0 = g4[0], r1 = g4[1], r2 = g4[2], r3 = g4[3];
    DynRegs d;
    d.s_lo = (uint64_t)r0.x | ((uint64_t)r0.y << 32); d.s_hi = (uint64_t)r0.z | ((uint64_t)r0.w << 32);
    d.inc_lo = (uint64_t)r1.x | ((uint64_t)r1.y << 32); d.inc_hi = (uint64_t)r1.z | ((uint64_t)r1.w << 32);
    d.has32 = r2.x; d.u32 = r2.y;
    d.ox = (uint64_t)r3.x | ((uint64_t)r3.y << 32); d.oy = (uint64_t)r3.z | ((uint64_t)r3.w << 32);
    const int n = min((int)(r2.z & 0xffu), 8);
    // ONE loop over tries for all obstacles: a lane moves on to its next obstacle as soon as it has placed the current one, so
    // the warp pays the maximum over its lanes of the TOTAL number of tries of a step, not the sum over the obstacles of the
    // per-obstacle maxima (round 1: ~36 warp iterations per step at 8x8 with 4 obstacles; now ~14).  Inside the loop only the
    // shared-memory tile is updated (later tries must see the move); the write-through to HBM is replayed afterwards,
    // obstacle by obstacle in the same order, with the warp converged.  (Positions stay in two packed 64-bit registers:
    // reading / writing them in the record with the dynamic index was measured at 1.72 instead of 1.94 G env-steps/s.)
    // MG_DYN_DEFER (default): `rec` is the launch's SHARED-MEMORY copy of the record and nothing here touches HBM — planes, packed
    // record and generator record are brought up to date once, when the launch ends (sync_obstacles below).
#if !MG_DYN_DEFER
    const uint64_t ox0 = d.ox, oy0 = d.oy;
    uint32_t moved = 0;
#endif
    int i = 0, tries = 0;
    int ox = (int)(d.ox & 0xff), oy = (int)(d.oy & 0xff);
#if MG_DYN_PIPE
    // the generator step of the NEXT try is taken while this try's cell lookup is in flight: the output does not depend on the
    // lookup (the state advances whether the cell is free or not), only the box it is scaled to does
    uint64_t nlo, nhi;
    uint64_t nv = pcg64_peek(d, nlo, nhi);
#endif
    while (i < n) {
        const int tx = max(ox - 1, 0), ty = max(oy - 1, 0);                    // :331
        const int hx = min(tx + 3, W), hy = min(ty + 3, H);
        int px, py;
#if MG_DYN_PIPE
        draw_pair_from(d, nv, nlo, nhi, tx, hx, ty, hy, px, py);            // mg_dynrng.cuh
        MG_ASSERT((unsigned)px < (unsigned)W && (unsigned)py < (unsigned)H && i < 8);
        const uint8_t at = tile[TILE(py * PR + px + PAD)];
        nv = pcg64_peek(d, nlo, nhi);
        ++tries;
        const bool found = ((at & 0x0Fu) == T_EMPTY) &                         // :355-356 (type' of an empty slot is 1)
                           !(px == ax && py == ay);                            // :359-360
#else
        draw_pair(d, tx, hx, ty, hy, px, py);                               // mg_dynrng.cuh
        ++tries;
        MG_ASSERT((unsigned)px < (unsigned)W && (unsigned)py < (unsigned)H && i < 8);
        const bool found = ((tile[TILE(py * PR + px + PAD)] & 0x0Fu) == T_EMPTY) &  // :355-356 (type' of an empty slot is 1)
                           !(px == ax && py == ay);                            // :359-360
#endif
        if (found) {
            const int sh = 8 * i;
            const int ro = oy * PR + ox + PAD, rn = py * PR + px + PAD;
            const uint8_t code = tile[TILE(ro)];
            tile[TILE(rn)] = code;
            tile[TILE(ro)] = (uint8_t)CODE_EMPTY;
            if (!rows_only) { tile[TILE(offC + px * PC + py + PAD)] = code; tile[TILE(offC + ox * PC + oy + PAD)] = (uint8_t)CODE_EMPTY; }   // no Cm in the ROWS layout
            d.ox = (d.ox & ~(0xffull << sh)) | ((uint64_t)px << sh);
            d.oy = (d.oy & ~(0xffull << sh)) | ((uint64_t)py << sh);
#if !MG_DYN_DEFER
            moved |= 1u << i;
#endif
        }
        if (found | (tries > 100)) {                                            // 101 samples at most (`max_tries=100`), then give up
            ++i; tries = 0;
            const int sn = 8 * (i & 7);                                         // (i == 8 ends the loop: keep the shift in range)
            ox = (int)((d.ox >> sn) & 0xff); oy = (int)((d.oy >> sn) & 0xff);
        }
    }
#if !MG_DYN_DEFER
    for (i = 0; i < n; ++i) {
        if (!((moved >> i) & 1u)) continue;
        const int sh = 8 * i;
        const int ox = (int)((ox0 >> sh) & 0xff), oy = (int)((oy0 >> sh) & 0xff);
        const int px = (int)((d.ox >> sh) & 0xff), py = (int)((d.oy >> sh) & 0xff);
        const int ko = oy * W + ox, kn = py * W + px;
        const int ro = oy * PR + ox + PAD, co = offC + ox * PC + oy + PAD;
        const int rn = py * PR + px + PAD, cn = offC + px * PC + py + PAD;
        const uint8_t code = tile[TILE(rn)];        // the obstacle sits at its final cell: nothing moves onto an occupied cell
        gpacked[rn] = code; gpacked[cn] = code;
        gpacked[ro] = (uint8_t)CODE_EMPTY; gpacked[co] = (uint8_t)CODE_EMPTY;
        // the planes' bytes come from the code.  Only type and colour change hands: the obstacle is a Ball (state 0, no aux
        // bits: dynamicobstacles.py:129-132) and the cell it moved onto was empty (type 1, colour / state / aux 0), so the state
        // and aux planes hold 0 at both cells before and after — four byte stores per move instead of eight.
        const cell_t ob = decode_code(code);
        MG_ASSERT(c_state(ob) == 0);
        gcells[kn] = (uint8_t)c_type(ob); gcells[ps + kn] = (uint8_t)c_color(ob);
        gcells[ko] = (uint8_t)T_EMPTY; gcells[ps + ko] = 0;
    }
#endif
    g4[0] = make_uint4((uint32_t)d.s_lo, (uint32_t)(d.s_lo >> 32), (uint32_t)d.s_hi, (uint32_t)(d.s_hi >> 32));
    g4[2] = make_uint4(d.has32, d.u32, r2.z, r2.w);
    g4[3] = make_uint4((uint32_t)d.ox, (uint32_t)(d.ox >> 32), (uint32_t)d.oy, (uint32_t)(d.oy >> 32));       // the increment never changes
}

#if MG_DYN_DEFER
// End of a launch, DynamicObstacles env: HBM still shows the obstacles where they stood when the launch began — or where the pool
// record put them, if the env was reset on the way (the autoreset copies planes, packed record AND generator record, so the HBM
// record always names the cells the HBM planes show).  `rec_s` is the live record in shared memory.  Every obstacle whose cell
// differs is cleared at the old cell and set at the new one, in two passes (an obstacle may now stand where another one stood),
// then the live record goes to HBM.  Per launch, not per step: round 2's per-step write-through was 8 scattered byte stores per
// move plus the record's 4 loads and 3 stores, and the step waited for them (2.6 -> see DESIGN 3.1 DynamicObstacles).
// Only type and colour change hands: the obstacle is a Ball (state 0, no aux bits: dynamicobstacles.py:129-132) and the cell it
// moved onto was empty (type 1, colour / state / aux 0), so the state and aux planes hold 0 at both cells before and after.
__device__ __noinline__ void sync_obstacles(const uint8_t* rec_s, uint8_t* rec_g, const uint8_t* tile, uint8_t* gpacked, uint8_t* gcells,
                                            int ps, int W, int PR, int PC, int offC)
{
    const uint4* s4 = reinterpret_cast<const uint4*>(rec_s);
    uint4* g4 = reinterpret_cast<uint4*>(rec_g);
    const uint4 n0 = s4[0], n2 = s4[2], n3 = s4[3];
    const uint4 o3 = __ldcg(g4 + 3);                              // (L2: other lanes of the warp may have reset it in the last step)
    const int n = min((int)(n2.z & 0xffu), 8);
    const uint64_t nx = (uint64_t)n3.x | ((uint64_t)n3.y << 32), ny = (uint64_t)n3.z | ((uint64_t)n3.w << 32);
    const uint64_t ox = (uint64_t)o3.x | ((uint64_t)o3.y << 32), oy = (uint64_t)o3.z | ((uint64_t)o3.w << 32);
    const uint64_t dxy = (nx ^ ox) | (ny ^ oy);
    for (int i = 0; i < n; ++i) {
        const int sh = 8 * i;
        if (!((dxy >> sh) & 0xff)) continue;
        const int x = (int)((ox >> sh) & 0xff), y = (int)((oy >> sh) & 0xff), k = y * W + x;
        MG_ASSERT(x < W && k < ps);
        gpacked[y * PR + x + PAD] = (uint8_t)CODE_EMPTY; gpacked[offC + x * PC + y + PAD] = (uint8_t)CODE_EMPTY;
        gcells[k] = (uint8_t)T_EMPTY; gcells[ps + k] = 0;
    }
    for (int i = 0; i < n; ++i) {
        const int sh = 8 * i;
        if (!((dxy >> sh) & 0xff)) continue;
        const int x = (int)((nx >> sh) & 0xff), y = (int)((ny >> sh) & 0xff), k = y * W + x;
        MG_ASSERT(x < W && k < ps);
        const uint8_t code = tile[TILE(y * PR + x + PAD)];
        const cell_t ob = decode_code(code);
        MG_ASSERT(c_type(ob) == T_BALL && c_state(ob) == 0);
        gpacked[y * PR + x + PAD] = code; gpacked[offC + x * PC + y + PAD] = code;
        gcells[k] = (uint8_t)c_type(ob); gcells[ps + k] = (uint8_t)c_color(ob);
    }
    g4[0] = n0; g4[2] = n2; g4[3] = n3;                       // the increment never changes
}
#endif

// LAYOUT: see step_layout.
template <bool EXT, int LAYOUT>
__global__ void __launch_bounds__(STEP_THREADS, MG_STEP_MINBLOCKS)
step_many_kernel(const StepParams p)
{
    constexpr bool ROWS = LAYOUT != LAYOUT_DUAL, HALF = LAYOUT == LAYOUT_ROWS_HALF;
#if MG_STEP_CHAIN
    // Per-CTA launch chaining.  CTA i of a step launch touches only the state of envs [32 i, 32 i + 32) — tiles, planes, agent and
    // generator records, its slots of the output arrays — and the read-only pool, so all it has to wait for is CTA i of the
    // PREVIOUS step launch of this engine, not that whole grid: the slots that the previous launch's early finishers (or its
    // part-filled last wave: 4,096 CTAs on 1,480 slots at 19x19) leave idle are taken by this launch's CTAs right away.
    //   ticket = taken[i]++ (atomic)          — BEFORE dependents are released: the next launch's CTA i cannot exist until every
    //                                           CTA of this grid has passed launch_dependents, so tickets follow launch order
    //   wait for finished[i] == ticket        — acquire at gpu scope by one thread + CTA barrier (bounded: ~2 s, then error bit 61)
    //   ... the launch ...                    — ends with every thread's fence, a CTA barrier, finished[i]++ with release semantics
    // No griddepcontrol.wait: ordering against OTHER kernels is untouched, because only step kernels release their dependents
    // early — anything else in the stream (a policy writing actions, reset, copies) completes before the next step launch begins,
    // and was itself ordered after the whole previous step launch.  Counters live in the engine (zeroed at creation) and count on
    // their own, so a captured CUDA graph of step launches can be replayed.  One-warp CTAs only (the default).
    uint32_t chain_ticket = 0;
    const bool chained = STEP_THREADS == 32 && p.chain != nullptr;
    if (chained) {
#if MG_CHAIN_ONE_WORD
        // one 64-bit word per CTA index: tickets taken in the high half, launches finished in the low half — ONE atomic takes the
        // ticket and says whether the predecessor is through already (the usual case), so a short launch (T = 1) pays one L2
        // round trip for the chain, not two
        unsigned long long* word = reinterpret_cast<unsigned long long*>(p.chain) + blockIdx.x;
        uint32_t fin = 0;
        if (threadIdx.x == 0) {
            unsigned long long old;
            asm volatile("atom.acquire.gpu.global.add.u64 %0, [%1], %2;" : "=l"(old) : "l"(word), "l"(1ull << 32) : "memory");
            chain_ticket = (uint32_t)(old >> 32);
            fin = (uint32_t)old;
        }
        chain_ticket = __shfl_sync(0xffffffffu, chain_ticket, 0);
        asm volatile("griddepcontrol.launch_dependents;" ::"r"(chain_ticket) : "memory");     // (after the add has RETURNED)
        if (threadIdx.x == 0 && fin != chain_ticket) {
            const long long t0 = clock64();
            for (;;) {
                unsigned long long v;
                asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(word) : "memory");
                if ((uint32_t)v == chain_ticket) break;
                if (clock64() - t0 > (1ll << 32)) {
                    atomicCAS(p.error_word, 0ull, 0x2000000000000000ull | (unsigned long long)blockIdx.x);
                    break;
                }
                __nanosleep(40);
            }
        }
        __syncwarp();
#else
        if (threadIdx.x == 0) chain_ticket = atomicAdd(p.chain + blockIdx.x, 1u);
        chain_ticket = __shfl_sync(0xffffffffu, chain_ticket, 0);           // (the add has returned before anyone goes on)
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
        if (threadIdx.x == 0) {
            const uint32_t* fin = p.chain + gridDim.x + blockIdx.x;
            uint32_t v;
            const long long t0 = clock64();
            for (;;) {
                asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(fin) : "memory");
                if (v == chain_ticket) break;
                if (clock64() - t0 > (1ll << 32)) {
                    atomicCAS(p.error_word, 0ull, 0x2000000000000000ull | (unsigned long long)blockIdx.x);
                    break;
                }
                __nanosleep(100);
            }
        }
        __syncwarp();
#endif
    } else {
        asm volatile("griddepcontrol.wait;" ::: "memory");
        asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    }
#elif MG_STEP_PDL
    // Programmatic dependent launch (launch_step_many_t): this grid may have been made resident while the previous kernel of
    // the stream was still running.  Everything below waits for that kernel to have completed and flushed; the next kernel of
    // the stream is released for the same early residency right away (it waits here too).  What is saved is the launch
    // latency and the CTA ramp between back-to-back step launches.  (First statements of the kernel on purpose: further down
    // the two barriers-to-the-compiler cost the ROWS_HALF instantiation a 16-byte spill inside the step loop.)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
    extern __shared__ __align__(16) uint8_t smem[];
    const int tid = threadIdx.x, lane = tid & 31;
    const int SB = staged_bytes(p.packed_stride, p.rows_bytes, ROWS);
    const int TS = tile_smem_stride(SB);
    // dynamic shared memory: tiles | staging | class table   (the decode table is static, below)
    // The decode table is STATIC shared memory: its address is a link-time constant, so a lookup is SHF + LOP3 + LDS [R + imm]
    // (with the table in the dynamic block the compiler had to add the block's base first: 49 more IADDs per observation).
#if MG_DECODE_LUT
    __shared__ uint32_t lut_s[LUT_ENTRIES];
    uint32_t* lut = lut_s;
#else
    uint32_t* lut = nullptr;
#endif
    uint8_t* tiles = smem;
    uint8_t* staging = tiles + (size_t)STEP_THREADS * TS;           // STEP_THREADS * TS is a multiple of 128
    ClassDev* cls_tab = reinterpret_cast<ClassDev*>(staging + (HALF ? STEP_THREADS / 2 : STEP_THREADS) * OBS_BYTES);
#if MG_DYN_DEFER
    // DynamicObstacles: the CTA's 64-byte generator records (launch_step_many_t adds the room when p.dyn is set)
    uint4* dynrec = reinterpret_cast<uint4*>(cls_tab + p.num_classes);
    const bool dyn_live = EXT && p.dyn && p.actions;
#endif

    const int env0 = blockIdx.x * STEP_THREADS;
    const int env = env0 + tid;
    const int nvalid = min(STEP_THREADS, p.N - env0);
    const bool active = tid < nvalid;
    const int ps = p.plane_stride;

    // ---- stage the CTA's packed tiles and the class table ----
    // (plain loads, not the read-only path: the previous step kernel wrote these records, and with dependent launch its
    // lifetime overlaps this grid's)
    {
        const int cpt = SB >> 4, gc = p.packed_stride >> 4;                          // 16-byte chunks per tile: smem / HBM
        const uint4* src = reinterpret_cast<const uint4*>(p.packed + (size_t)env0 * p.packed_stride);
#if MG_TILE_INTERLEAVE
        // every lane copies ITS env's tile, chunk by chunk: the four word stores of a chunk go to the lane's own bank (lanes
        // sharing a tile — the coalesced order — would all hit one bank: 16-way conflicts, 7 us per launch with 14 CTAs per SM
        // doing it at once); PB loads are issued before the first store (one HBM round trip per PB chunks)
        if (tid < nvalid) {
            constexpr int PB = 8;
            const uint4* mine = src + (size_t)tid * gc;
            uint32_t* d = reinterpret_cast<uint32_t*>(tiles) + tid;
            for (int w0 = 0; w0 < cpt; w0 += PB) {
                uint4 v[PB];
#pragma unroll
                for (int j = 0; j < PB; ++j) if (w0 + j < cpt) v[j] = mine[w0 + j];
#pragma unroll
                for (int j = 0; j < PB; ++j)
                    if (w0 + j < cpt) {
                        uint32_t* dj = d + 4 * (w0 + j) * TILE_WS;
                        dj[0] = v[j].x; dj[TILE_WS] = v[j].y; dj[2 * TILE_WS] = v[j].z; dj[3 * TILE_WS] = v[j].w;
                    }
            }
        }
#else
        // coalesced 16-byte loads, word stores (tiles are only 4-byte aligned), MG_STAGE_PB loads in flight per lane before the
        // first store: written as load -> store per chunk the loop was one HBM round trip per chunk — 9 at 8x8, 31 at 19x19,
        // where the launch's first 8 % went (profiles/r2/ncu_mixed_lines.txt)
        const int nchunk = nvalid * cpt;
        for (int c0 = tid; c0 < nchunk; c0 += STEP_THREADS * MG_STAGE_PB) {
            uint4 v[MG_STAGE_PB];
            int dw[MG_STAGE_PB];
#pragma unroll
            for (int j = 0; j < MG_STAGE_PB; ++j) {
                const int c = c0 + j * STEP_THREADS;
                dw[j] = -1;
                if (c < nchunk) {
                    const int ee = c / cpt, w = c - ee * cpt;
                    v[j] = src[(size_t)ee * gc + w];
                    dw[j] = ee * (TS >> 2) + 4 * w;
                }
            }
#pragma unroll
            for (int j = 0; j < MG_STAGE_PB; ++j)
                if (dw[j] >= 0) {
                    uint32_t* d = reinterpret_cast<uint32_t*>(tiles) + dw[j];
                    d[0] = v[j].x; d[1] = v[j].y; d[2] = v[j].z; d[3] = v[j].w;
                }
        }
#endif
        if (MG_DECODE_LUT) for (int i = tid; i < LUT_ENTRIES; i += STEP_THREADS) lut[i] = lut_entry((uint32_t)i);
#if MG_DYN_DEFER
        if (dyn_live && tid < nvalid) {
            const uint4* g = reinterpret_cast<const uint4*>(p.dyn + (size_t)(env0 + tid) * 64);
            const uint4 a = g[0], b = g[1], c = g[2], d = g[3];
            dynrec[4 * tid] = a; dynrec[4 * tid + 1] = b; dynrec[4 * tid + 2] = c; dynrec[4 * tid + 3] = d;
        }
#endif
        const int nw = p.num_classes * (int)(sizeof(ClassDev) / 4);
        const uint32_t* ct = reinterpret_cast<const uint32_t*>(&p.classes);
        for (int i = tid; i < nw; i += STEP_THREADS) reinterpret_cast<uint32_t*>(cls_tab)[i] = ct[i];
    }
    int x = 0, y = 0, dir = 0, cls = 0, steps = 0;
    cell_t carry = 0;
    uint32_t layout_id = 0;
    if (active) {
        const AgentState a = p.agents[env];
        x = a.x; y = a.y; dir = a.dir; cls = a.cls; steps = a.steps; layout_id = a.layout_id;
        carry = a.ct ? make_cell(a.ct, a.cc, a.cs, a.ca) : 0;
    }
    __syncthreads();

    uint8_t* tile = MG_TILE_INTERLEAVE ? tiles + 4 * tid : tiles + (size_t)tid * TS;
    uint8_t* gcells = p.cells + (size_t)(active ? env : env0) * 4 * ps;
    uint8_t* gpacked = p.packed + (size_t)(active ? env : env0) * p.packed_stride;
    ClassDev cd = cls_tab[cls];
    bool has_aux = gpacked[p.packed_stride - 1] != 0;        // see pack_kernel
    // whole-CTA observation blocks leave through the copy engine (see bulk_store_issue)
    const bool bulk = MG_STEP_BULK_STORE && STEP_THREADS == 32 && p.image_bulk && nvalid == STEP_THREADS;
    uint64_t bulk_policy = 0;
    bool bulk_pending = false;                                   // lane 0: a bulk copy may still be reading the staging buffer
    if (bulk && tid == 0) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(bulk_policy));
    int action = (active && p.actions) ? p.actions[env] : A_DONE;
    bool pending = EXT && p.autoreset == 2 && active && (p.agents[env].pad & 1);     // next-step autoreset owed from the last launch

    for (int t = 0; t < p.T; ++t) {
        const size_t k = (size_t)t * p.N + env;
        const int next_action = (active && p.actions && t + 1 < p.T) ? p.actions[k + p.N] : A_DONE;   // prefetch
        bool done = false;
        if (active && p.actions) {
            double reward = 0.0; bool terminated = false, truncated = false;
            if (EXT && pending) {
                // next-step autoreset: the step after an episode ended ignores the action and only resets the env
                // (reward 0, flags clear, the returned observation is the new episode's first)
                pending = false;
                done = true;
            } else {
            bool bumped = false;
            int pre_front_t = -1;                                               // DYNOBS: type in front BEFORE the obstacles move
            if (EXT && cd.rule == RULE_DYNOBS) {                                // DynamicObstaclesEnv.step, dynamicobstacles.py:137-157
                if (action >= 3) action = A_LEFT;                               // `if action >= self.action_space.n: action = 0`
                const int t0 = tile[TILE((y + dir_dy(dir)) * cd.PR + x + dir_dx(dir) + PAD)] & 0x0F;   // front cell BEFORE the obstacles move
                bumped = (action == A_FORWARD) & (t0 != T_EMPTY) & (t0 != T_GOAL);
                pre_front_t = t0 >= 13 ? T_DOOR : t0;
#if MG_DYN_DEFER
                if (p.dyn) move_obstacles(reinterpret_cast<uint8_t*>(dynrec + 4 * tid), tile, gpacked, gcells, ps, cd.W, cd.H, cd.PR, cd.PC, cd.offC, x, y, ROWS);
#else
                if (p.dyn) move_obstacles(p.dyn + (size_t)env * 64, tile, gpacked, gcells, ps, cd.W, cd.H, cd.PR, cd.PC, cd.offC, x, y, ROWS);
#endif
            }
            steps = min(steps + 1, 65535);                                      // minigrid_env.py:523 (the 16-bit counter of the
                                                                                // agent record saturates: max_steps <= 65535, so
                                                                                // `truncated` stays true once it was reached)
            if (EXT && cd.rule == RULE_MEMORY && action == A_PICKUP) action = A_TOGGLE;         // memory.py:154-155
            const cell_t pre_carry = carry;                                     // putnear.py:177
            const int fx = x + dir_dx(dir), fy = y + dir_dy(dir);               // front_pos :413-419
            const int fk = fy * cd.W + fx;
            MG_ASSERT((unsigned)x < (unsigned)cd.W && (unsigned)y < (unsigned)cd.H && (unsigned)dir < 4u && (unsigned)cls < (unsigned)p.num_classes);
            // GoToDoor puts doors IN the border wall, so once one is open the agent can stand on the border and face
            // off-grid (the reference's Grid.get asserts there); engine and oracle both read a wall.  Other envs never do.
            // (The -DMG_BOUNDS_CHECK build caught round 1's form reading tile[-PR + ...] of the NEIGHBOURING env's tile there
            // — inside the CTA's shared memory and discarded, but out of the env's own tile: the index is clamped now.)
            const bool off_grid = EXT && ((unsigned)fx >= (unsigned)cd.W || (unsigned)fy >= (unsigned)cd.H);
            const int fidx = off_grid ? PAD : fy * cd.PR + fx + PAD;
            MG_ASSERT(fidx >= 0 && fidx < SB);
            cell_t front = decode_code(tile[TILE(fidx)]);                             // :533
            if (off_grid) front = make_cell(T_WALL, 5, 0, 0);
            if (has_aux && ((action == A_PICKUP && can_pickup(front)) || (action == A_TOGGLE && c_type(front) == T_BOX)))
                front |= (cell_t)gcells[3 * ps + fk] << 24;                     // hidden contents / target bit live in the aux plane
            const Transition tr = apply_action(action, x, y, dir, carry, fx, fy, front, cd.rule);
            if (tr.bad_action) {
                atomicCAS(p.error_word, 0ull, 0x8000000000000000ull | ((unsigned long long)action << 40) | (unsigned long long)env);
            } else {
                x = tr.x; y = tr.y; dir = tr.dir; carry = tr.carry;
                if (tr.wrote_front) {
                    MG_ASSERT((unsigned)fx < (unsigned)cd.W && (unsigned)fy < (unsigned)cd.H && fk < ps);
                    const uint8_t code = (uint8_t)cell_code(tr.front);
                    const int pr = fy * cd.PR + fx + PAD, pc = cd.offC + fx * cd.PC + fy + PAD;
                    tile[TILE(pr)] = code;
                    if (!ROWS) tile[TILE(pc)] = code;
                    gpacked[pr] = code; gpacked[pc] = code;                     // write-through: HBM stays authoritative
                    gcells[fk] = (uint8_t)c_type(tr.front); gcells[ps + fk] = (uint8_t)c_color(tr.front);
                    gcells[2 * ps + fk] = (uint8_t)c_state(tr.front); gcells[3 * ps + fk] = (uint8_t)c_aux(tr.front);
                }
                terminated = tr.terminated;
                bool success = tr.success;
                if (EXT) {
                    // slack bytes of the packed record: R-offset of the mission's target cell / of the blue target door (pack_kernel)
                    // (read only by the rules that use it: hoisted above the rule switch it was a dependent HBM/L2 round trip in
                    // every step of every EXT env — DynamicObstacles, Fetch, PutNear never look at it)
                    const bool need_offA = cd.rule == RULE_UNLOCK || cd.rule == RULE_GOTO || cd.rule == RULE_MEMORY || cd.rule == RULE_REDBLUE;
                    const unsigned offA = need_offA ? ((unsigned)gpacked[p.packed_stride - 4] | ((unsigned)gpacked[p.packed_stride - 3] << 8)) : 0xFFFFu;
                    if (cd.rule == RULE_UNLOCK && action == A_TOGGLE) {         // UnlockEnv.step, envs/unlock.py:92-95
                        if (offA != 0xFFFFu && (tile[TILE(offA)] & 0x0Fu) == T_DOOR) { terminated = true; success = true; }   // type' 4 = OPEN door
                    } else if (cd.rule == RULE_FETCH) {                         // FetchEnv.step, envs/fetch.py:165-174
                        if (carry != 0) { terminated = true; success = (c_aux(carry) & 0x80) != 0; }
                    } else if (cd.rule == RULE_GOTO) {                          // gotodoor.py:136-148, gotoobject.py:148-160
                        if (action == A_TOGGLE) terminated = true;
                        if (action == A_DONE) {
                            terminated = true;
                            const int o = (int)offA - PAD, ty = o / cd.PR, tx = o - ty * cd.PR;
                            success = offA != 0xFFFFu && ((x == tx && abs(y - ty) == 1) || (y == ty && abs(x - tx) == 1));
                        }
                    } else if (cd.rule == RULE_MEMORY) {                        // memory.py:158-163
                        const unsigned offB = (unsigned)gpacked[p.packed_stride - 6] | ((unsigned)gpacked[p.packed_stride - 5] << 8);
                        const unsigned here = (unsigned)(y * cd.PR + x + PAD);
                        if (here == offA) { terminated = true; success = true; }
                        if (here == offB) { terminated = true; success = false; }
                    } else if (cd.rule == RULE_PUTNEAR) {                       // putnear.py:181-199
                        if (action == A_PICKUP && carry != 0 && !((c_aux(carry) & 0x80) && (c_aux(carry) & 0x0f) != 1))
                            terminated = true;                                  // not the object to move
                        if (action == A_DROP && pre_carry != 0) {
                            terminated = true;
                            const unsigned offB = (unsigned)gpacked[p.packed_stride - 6] | ((unsigned)gpacked[p.packed_stride - 5] << 8);
                            const int o = (int)offB - PAD, ty = o / cd.PR, tx = o - ty * cd.PR;
                            success = carry == 0 && offB != 0xFFFFu && abs(fx - tx) <= 1 && abs(fy - ty) <= 1;   // the drop happened, near target_pos
                        }
                    } else if (cd.rule == RULE_REDBLUE) {                       // redbluedoors.py:114-131
                        const unsigned offB = (unsigned)gpacked[p.packed_stride - 6] | ((unsigned)gpacked[p.packed_stride - 5] << 8);
                        const bool red_after = offA != 0xFFFFu && (tile[TILE(offA)] & 0x0Fu) == T_DOOR;
                        const bool blue_after = offB != 0xFFFFu && (tile[TILE(offB)] & 0x0Fu) == T_DOOR;
                        // "before" = "after" unless this very toggle changed that door (the front cell is the only cell a step writes)
                        const unsigned offF = (unsigned)(fy * cd.PR + fx + PAD);
                        const bool front_was_open = c_type(front) == T_DOOR && c_state(front) == S_OPEN;
                        const bool red_before = (tr.wrote_front && offF == offA) ? front_was_open : red_after;
                        const bool blue_before = (tr.wrote_front && offF == offB) ? front_was_open : blue_after;
                        if (blue_after) { terminated = true; success = red_before; }
                        else if (red_after && blue_before) { terminated = true; success = false; }
                    }
                }
                if (success) reward = reward_fn(steps, cd.max_steps);
                if (EXT && bumped) { reward = -1.0; terminated = true; }       // dynamicobstacles.py:163-166
                if (EXT && p.no_death_mask && terminated) {                     // NoDeath.step, wrappers.py:842-865
                    // the wrapper samples the front cell BEFORE env.step (wrappers.py:852-858), i.e. before DynamicObstacles
                    // moves its balls: a ball that was in front and then moved away still cancels the death
                    const int ft_nd = pre_front_t >= 0 ? pre_front_t : c_type(front);
                    const bool going_to_death = (action == A_FORWARD) & ((p.no_death_mask >> ft_nd) & 1u);
                    const int here = c_type(decode_code(tile[TILE(y * cd.PR + x + PAD)]));      // cell under the agent AFTER the step
                    if (going_to_death || ((p.no_death_mask >> here) & 1u)) { terminated = false; reward = __dadd_rn(reward, p.death_cost); }
                }
                truncated = steps >= cd.max_steps;                              // :582-583
                done = p.autoreset == 1 && (terminated || truncated);
                if (EXT && p.autoreset == 2) pending = terminated || truncated;
                if (EXT && done && p.final_image) {
                    // same-step autoreset: keep the finished episode's last observation (gymnasium's info["final_observation"])
                    TileGeom g;
                    g.W = cd.W; g.H = cd.H; g.PR = cd.PR; g.PC = cd.PC; g.offC = cd.offC; g.bytes = 0;
                    uint32_t S[37];
                    obs_words<ROWS, TILE_WS>(tile, g, x, y, dir, carry ? cell_code(carry) : CODE_EMPTY, cd.see_through != 0, lut, S);
                    uint8_t* fo = p.final_image + k * OBS_BYTES;
                    for (int i = 0; i < OBS_BYTES; ++i) fo[i] = (uint8_t)(S[i >> 2] >> (8 * (i & 3)));
                    if (p.final_direction) p.final_direction[k] = (uint8_t)dir;
                }
            }
            }
            if (p.reward) p.reward[k] = reward;
            if (p.terminated) p.terminated[k] = terminated;
            if (p.truncated) p.truncated[k] = truncated;
        }
        // ---- same-step autoreset: the warp reloads each done env from pool slot (layout_id + reset_stride) % P ----
        // reset_stride (mg_reset_stride): N mod P moved up to the next value coprime to P, so that an env walks through ALL P
        // layouts before it sees one again — also when P divides N (the default pool of N layouts, 65,536 envs on 4,096)
        if (p.autoreset) {
            __syncwarp();                                    // order the done lane's write-through before the reload
            unsigned m = __ballot_sync(0xffffffffu, done);
            uint32_t my_new = 0;
            LayoutAgent la = {0, 0, 0, 0};
            bool new_aux = false;
            if (done) {                                      // only done lanes pay the modulo
                const uint32_t P = (uint32_t)max(p.P, 1);
                my_new = (layout_id % P + (uint32_t)p.reset_stride) % P;
                la = p.pool_agents[my_new];                  // (issued here: the round trip overlaps the copy below)
                new_aux = p.pool_packed[(size_t)my_new * p.packed_stride + p.packed_stride - 1] != 0;
            }
            // Records of up to 32 16-byte chunks per part (grids up to 11x11): lane c holds chunk c of the planes, of the packed
            // record (whose first SB bytes are also the shared-memory tile) and of the generator record — of TWO done envs — before
            // anything is stored: one L2 round trip per pair of resets instead of four per reset.  DynamicObstacles-8x8 ends an
            // episode every ~9 steps under random actions (3.5 resets per warp and step).  Larger records keep the plain loops: cut
            // into chunks with per-chunk source selects they were 15 % SLOWER on the 11x11 + 19x19 pool (profiles/r2/ab_step_variants.txt).
            const int nc = (4 * ps) >> 4, np = p.packed_stride >> 4;
            const bool small_rec = MG_RESET_PAIRS && nc <= 32 && np <= 32 && (SB & 15) == 0;
#if MG_RESET_MID
            const bool mid_rec = nc <= 96 && np <= 96 && (SB & 15) == 0;
#endif
            while (m) {
                const int l = __ffs(m) - 1;
                m &= m - 1;
                const uint32_t lid = __shfl_sync(0xffffffffu, my_new, l);
                if (small_rec) {
                    const int l1 = __ffs(m) - 1;                                    // -1: no partner
                    m &= m - 1;                                                     // (0 stays 0)
                    const uint32_t lid1 = __shfl_sync(0xffffffffu, my_new, l1 & 31);
                    const bool two = l1 >= 0, hd = EXT && p.dyn && lane < 4;
                    MG_ASSERT(lid < (uint32_t)p.P && lid1 < (uint32_t)p.P);
                    uint4 c0 = make_uint4(0, 0, 0, 0), p0 = c0, d0 = c0, c1 = c0, p1 = c0, d1 = c0;
                    if (lane < nc) c0 = __ldg(reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid * 4 * ps) + lane);
                    if (lane < np) p0 = __ldg(reinterpret_cast<const uint4*>(p.pool_packed + (size_t)lid * p.packed_stride) + lane);
                    if (hd) d0 = __ldg(reinterpret_cast<const uint4*>(p.pool_dyn + (size_t)lid * 64) + lane);
                    if (two) {
                        if (lane < nc) c1 = __ldg(reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid1 * 4 * ps) + lane);
                        if (lane < np) p1 = __ldg(reinterpret_cast<const uint4*>(p.pool_packed + (size_t)lid1 * p.packed_stride) + lane);
                        if (hd) d1 = __ldg(reinterpret_cast<const uint4*>(p.pool_dyn + (size_t)lid1 * 64) + lane);
                    }
                    auto put = [&](int le, const uint4& vc, const uint4& vp, const uint4& vd) {
                        const int we = tid - lane + le, e = env0 + we;
                        MG_ASSERT(e < p.N);
                        if (lane < nc) reinterpret_cast<uint4*>(p.cells + (size_t)e * 4 * ps)[lane] = vc;
                        if (lane < np) reinterpret_cast<uint4*>(p.packed + (size_t)e * p.packed_stride)[lane] = vp;
                        if (lane < (SB >> 4)) {
                            uint32_t* dt = (MG_TILE_INTERLEAVE ? reinterpret_cast<uint32_t*>(tiles) + we
                                                               : reinterpret_cast<uint32_t*>(tiles + (size_t)we * TS)) + 4 * lane * TILE_WS;
                            dt[0] = vp.x; dt[TILE_WS] = vp.y; dt[2 * TILE_WS] = vp.z; dt[3 * TILE_WS] = vp.w;
                        }
                        if (hd) {
                            reinterpret_cast<uint4*>(p.dyn + (size_t)e * 64)[lane] = vd;
#if MG_DYN_DEFER
                            dynrec[4 * we + lane] = vd;                             // (the launch's live copy; HBM keeps the cells its planes show)
#endif
                        }
                    };
                    put(l, c0, p0, d0);
                    if (two) put(l1, c1, p1, d1);
                    continue;
                }
#if MG_RESET_MID
                if (mid_rec) {
                    // records of up to 96 chunks per part (grids up to 19x19): ONE env, three chunks of each part per lane, again all
                    // loads before the first store (fixed unrolled rounds with a bound check each — no per-chunk source selects)
                    const int we = tid - lane + l, e = env0 + we;
                    MG_ASSERT(lid < (uint32_t)p.P && e < p.N);
                    const uint4* sc = reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid * 4 * ps);
                    const uint4* sp = reinterpret_cast<const uint4*>(p.pool_packed + (size_t)lid * p.packed_stride);
                    uint4 vc[3], vp[3], vd = make_uint4(0, 0, 0, 0);
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
                        vc[r] = vp[r] = make_uint4(0, 0, 0, 0);
                        if (lane + 32 * r < nc) vc[r] = __ldg(sc + lane + 32 * r);
                        if (lane + 32 * r < np) vp[r] = __ldg(sp + lane + 32 * r);
                    }
                    const bool hd = EXT && p.dyn && lane < 4;
                    if (hd) vd = __ldg(reinterpret_cast<const uint4*>(p.pool_dyn + (size_t)lid * 64) + lane);
                    uint4* dc = reinterpret_cast<uint4*>(p.cells + (size_t)e * 4 * ps);
                    uint4* dp = reinterpret_cast<uint4*>(p.packed + (size_t)e * p.packed_stride);
                    uint32_t* dt = MG_TILE_INTERLEAVE ? reinterpret_cast<uint32_t*>(tiles) + we : reinterpret_cast<uint32_t*>(tiles + (size_t)we * TS);
#pragma unroll
                    for (int r = 0; r < 3; ++r) {
                        const int i = lane + 32 * r;
                        if (i < nc) dc[i] = vc[r];
                        if (i < np) dp[i] = vp[r];
                        if (i < (SB >> 4)) {
                            uint32_t* d4 = dt + 4 * i * TILE_WS;
                            d4[0] = vp[r].x; d4[TILE_WS] = vp[r].y; d4[2 * TILE_WS] = vp[r].z; d4[3 * TILE_WS] = vp[r].w;
                        }
                    }
                    if (hd) {
                        reinterpret_cast<uint4*>(p.dyn + (size_t)e * 64)[lane] = vd;
#if MG_DYN_DEFER
                        dynrec[4 * we + lane] = vd;
#endif
                    }
                    continue;
                }
#endif
                const int e = env0 + (tid - lane) + l;
                MG_ASSERT(lid < (uint32_t)p.P && e < p.N);
                const uint4* sc = reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid * 4 * ps);
                uint4* dc = reinterpret_cast<uint4*>(p.cells + (size_t)e * 4 * ps);
                for (int i = lane; i < (4 * ps) >> 4; i += 32) dc[i] = __ldg(sc + i);
                const uint4* sp = reinterpret_cast<const uint4*>(p.pool_packed + (size_t)lid * p.packed_stride);
                uint4* dp = reinterpret_cast<uint4*>(p.packed + (size_t)e * p.packed_stride);
                for (int i = lane; i < p.packed_stride >> 4; i += 32) dp[i] = __ldg(sp + i);
                uint32_t* dt = MG_TILE_INTERLEAVE ? reinterpret_cast<uint32_t*>(tiles) + (tid - lane + l)
                                                  : reinterpret_cast<uint32_t*>(tiles + (size_t)(tid - lane + l) * TS);
                const uint32_t* sw = reinterpret_cast<const uint32_t*>(sp);
                for (int i = lane; i < SB >> 2; i += 32) dt[i * TILE_WS] = __ldg(sw + i);     // (one bank: 32-way conflict, once per episode)
                if (EXT && p.dyn && lane < 4) {
                    const uint4 v = __ldg(reinterpret_cast<const uint4*>(p.pool_dyn + (size_t)lid * 64) + lane);
                    reinterpret_cast<uint4*>(p.dyn + (size_t)e * 64)[lane] = v;
#if MG_DYN_DEFER
                    dynrec[4 * (tid - lane + l) + lane] = v;          // (the launch's live copy; HBM keeps the cells its planes show)
#endif
                }
            }
            __syncwarp();
            if (done) {
                x = la.x; y = la.y; dir = la.dir; cls = la.cls; steps = 0; carry = 0; layout_id = my_new;
                cd = cls_tab[cls];
                has_aux = new_aux;
            }
        }
        if (active && p.direction) p.direction[k] = (uint8_t)dir;
        if (p.image) {
            uint32_t S[37];
            if (active) {
                MG_ASSERT((unsigned)x < (unsigned)cd.W && (unsigned)y < (unsigned)cd.H && (unsigned)dir < 4u);
                MG_ASSERT(PAD + (cd.H - 1) * cd.PR + cd.W - 1 + PAD < SB + 16 && (ROWS || cd.offC + PAD + (cd.W - 1) * cd.PC + cd.H - 1 + PAD < SB + 16));
                TileGeom g;
                g.W = cd.W; g.H = cd.H; g.PR = cd.PR; g.PC = cd.PC; g.offC = cd.offC; g.bytes = 0;
                obs_words<ROWS, TILE_WS>(tile, g, x, y, dir, carry ? cell_code(carry) : CODE_EMPTY, cd.see_through != 0, lut, S);
            } else if (MG_STAGE_SHFL) {
#pragma unroll
                for (int i = 0; i < 37; ++i) S[i] = 0;          // ragged last CTA: the lane still takes part in stage_record
            }
            uint8_t* dst = p.image + ((size_t)t * p.N + env0) * OBS_BYTES;
            if (bulk) {
                if (tid == 0 && bulk_pending) bulk_store_wait_read();           // the previous block has left the staging buffer
                cta_sync();
            }
            if (!HALF) {
                if (MG_STAGE_SHFL || active) stage_record(staging, tid, S, 0xffffffffu);
                if (bulk) {
                    bulk_store_fence();
                    cta_sync();
                    if (tid == 0) { bulk_store_issue(dst, staging, STEP_THREADS * OBS_BYTES, bulk_policy); bulk_pending = true; }
                } else {
                    cta_sync();
                    cta_copy_out(dst, staging, nvalid * OBS_BYTES);
                    cta_sync();
                }
            } else {
                // 16 records at a time through the half-size buffer; with the copy engine the SECOND half leaves asynchronously
                // (its wait falls behind the next step's compute), the first half is drained by the lanes as before
                if ((MG_STAGE_SHFL || active) && (tid >> 4) == 0) stage_record(staging, tid & 15, S, 0x0000ffffu);
                cta_sync();
                const int nv0 = min(nvalid, 16);
                if (nv0 > 0) cta_copy_out(dst, staging, nv0 * OBS_BYTES);
                cta_sync();
                if ((MG_STAGE_SHFL || active) && (tid >> 4) == 1) stage_record(staging, tid & 15, S, 0xffff0000u);
                if (bulk) {
                    bulk_store_fence();
                    cta_sync();
                    if (tid == 0) { bulk_store_issue(dst + 16 * OBS_BYTES, staging, 16 * OBS_BYTES, bulk_policy); bulk_pending = true; }
                } else {
                    cta_sync();
                    const int nv1 = min(max(nvalid - 16, 0), 16);
                    if (nv1 > 0) cta_copy_out(dst + 16 * OBS_BYTES, staging, nv1 * OBS_BYTES);
                    cta_sync();
                }
            }
        }
        action = next_action;
    }
#if MG_STEP_CHAIN
    // (chained: the block must have REACHED global memory, not just left shared memory, before the next launch's CTA is let go)
    if (bulk && tid == 0 && bulk_pending) { if (chained) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); else bulk_store_wait_read(); }
#else
    if (bulk && tid == 0 && bulk_pending) bulk_store_wait_read();       // shared memory must outlive the copy engine's reads
#endif
#if MG_DYN_DEFER
    if (dyn_live) {
        cta_sync();                                                      // a reset of the last step wrote the lane's record from another lane
        if (active && cd.rule == RULE_DYNOBS)
            sync_obstacles(reinterpret_cast<const uint8_t*>(dynrec + 4 * tid), p.dyn + (size_t)env * 64, tile, gpacked, gcells, ps, cd.W, cd.PR, cd.PC, cd.offC);
    }
#endif
    if (active && p.actions) {
        AgentState a;
        a.x = (uint8_t)x; a.y = (uint8_t)y; a.dir = (uint8_t)dir; a.cls = (uint8_t)cls;
        a.ct = (uint8_t)c_type(carry); a.cc = (uint8_t)c_color(carry); a.cs = (uint8_t)c_state(carry); a.ca = (uint8_t)c_aux(carry);
        a.steps = (uint16_t)steps; a.pad = pending ? 1 : 0; a.layout_id = layout_id;
        p.agents[env] = a;
    }
#if MG_STEP_CHAIN
    if (chained) {
        __threadfence();
        __syncwarp();
        if (threadIdx.x == 0)
#if MG_CHAIN_ONE_WORD
        {
            // finished++ in the low half.  Once in 2^32 launches the low half wraps and its carry would land in the ticket half:
            // that one add takes the carry back out (1 - 2^32), so both halves count modulo 2^32 for ever
            const unsigned long long inc = chain_ticket == 0xffffffffu ? 1ull - (1ull << 32) : 1ull;
            asm volatile("red.release.gpu.global.add.u64 [%0], %1;" ::"l"(reinterpret_cast<unsigned long long*>(p.chain) + blockIdx.x), "l"(inc) : "memory");
        }
#else
            asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p.chain + gridDim.x + blockIdx.x), "r"(chain_ticket + 1u) : "memory");
#endif
    }
#endif
}

// reset: env e <- pool[layout_ids[e]] (where mask says so); one warp per env
__global__ void __launch_bounds__(256)
reset_kernel(const ResetParams p)
{
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= p.N) return;
    if (p.mask && !p.mask[warp]) return;
    uint32_t lid = (uint32_t)p.layout_ids[warp];
    if (lid >= (uint32_t)p.P) {                               // an id outside the pool: report it (mg_poll_error) and leave the env alone
        if (lane == 0) atomicCAS(p.error_word, 0ull, 0x4000000000000000ull | (unsigned long long)warp);
        return;
    }
    const uint4* s = reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid * 4 * p.plane_stride);
    uint4* d = reinterpret_cast<uint4*>(p.cells + (size_t)warp * 4 * p.plane_stride);
    for (int i = lane; i < (4 * p.plane_stride) >> 4; i += 32) d[i] = __ldg(s + i);
    const uint4* sp = reinterpret_cast<const uint4*>(p.pool_packed + (size_t)lid * p.packed_stride);
    uint4* dp = reinterpret_cast<uint4*>(p.packed + (size_t)warp * p.packed_stride);
    for (int i = lane; i < p.packed_stride >> 4; i += 32) dp[i] = __ldg(sp + i);
    if (p.dyn && lane < 4)
        reinterpret_cast<uint4*>(p.dyn + (size_t)warp * 64)[lane] = __ldg(reinterpret_cast<const uint4*>(p.pool_dyn + (size_t)lid * 64) + lane);
    if (lane == 0) {
        const LayoutAgent la = p.pool_agents[lid];
        AgentState a;
        a.x = la.x; a.y = la.y; a.dir = la.dir; a.cls = la.cls; a.ct = a.cc = a.cs = a.ca = 0;
        a.steps = 0; a.pad = 0; a.layout_id = lid;
        p.agents[warp] = a;
    }
}

// pack: SoA planes -> packed tile (R + Cm, wall padding); one warp per record
__global__ void __launch_bounds__(256)
pack_kernel(const PackParams p)
{
    const int rec = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (rec >= p.n) return;
    const ClassDev cd = p.classes.c[p.cls[(size_t)rec * p.cls_stride]];
    uint8_t* tile = p.packed + (size_t)rec * p.packed_stride;
    const uint8_t* cells = p.cells + (size_t)rec * 4 * p.plane_stride;
    uint4* t4 = reinterpret_cast<uint4*>(tile);
    const uint4 w4 = make_uint4(WALL4, WALL4, WALL4, WALL4);
    for (int i = lane; i < p.packed_stride >> 4; i += 32) t4[i] = w4;
    __syncwarp();
    unsigned any_aux = 0, target = 0xFFFFu, target_b = 0xFFFFu;
    for (int k = lane; k < cd.W * cd.H; k += 32) {
        const int yy = k / cd.W, xx = k - yy * cd.W;
        const uint8_t code = (uint8_t)encode_code(cells[k], cells[p.plane_stride + k], cells[2 * p.plane_stride + k]);
        tile[yy * cd.PR + xx + PAD] = code;
        tile[cd.offC + xx * cd.PC + yy + PAD] = code;
        any_aux |= cells[3 * p.plane_stride + k];
        if (cells[3 * p.plane_stride + k] & 0x80) {               // mission target: object / door / position (levelgen.py)
            const unsigned off = (unsigned)(yy * cd.PR + xx + PAD);
            const bool second = (cd.rule == RULE_REDBLUE && cells[p.plane_stride + k] == 2) ||  // the blue door of RedBlueDoors
                                ((cd.rule == RULE_MEMORY || cd.rule == RULE_PUTNEAR) && (cells[3 * p.plane_stride + k] & 0x0f) == 1);   // aux 0x81
            if (second) target_b = min(target_b, off);
            else target = min(target, off);
        }
    }
    // slack bytes [-4, -3]: byte offset (in R) of the first cell carrying the target bit when the record was packed —
    // `self.door` (envs/unlock.py:86), `self.target_pos` (gotodoor.py:125, gotoobject.py:139), the red door; [-6, -5]: the
    // blue door of RedBlueDoors, failure_pos of Memory (memory.py:149), target_pos of PutNear (putnear.py:167); 0xFFFF = none.  Pool records are packed from the initial layouts, which is when
    // target_pos is defined.
    {
        target = __reduce_min_sync(0xffffffffu, target);
        target_b = __reduce_min_sync(0xffffffffu, target_b);
        if (lane == 0) {
            tile[p.packed_stride - 4] = (uint8_t)target; tile[p.packed_stride - 3] = (uint8_t)(target >> 8);
            tile[p.packed_stride - 6] = (uint8_t)target_b; tile[p.packed_stride - 5] = (uint8_t)(target_b >> 8);
        }
    }
    if (p.cls_stride == 16) any_aux |= p.cls[(size_t)rec * 16 + 4];      // env records: the carried object's aux byte
    any_aux = __any_sync(0xffffffffu, any_aux != 0);
    // last slack byte of the record: "this layout has hidden state in the aux plane" (box contents / target bit);
    // layouts without any (DoorKey, Empty, Crossing, FourRooms) never touch the aux plane in the step kernel
    if (lane == 0) tile[p.packed_stride - 1] = any_aux ? 1 : 0;
}

static bool needs_ext(const StepParams& p)
{
    if (p.no_death_mask || p.autoreset == 2 || p.final_image) return true;
    for (int i = 0; i < p.num_classes; ++i) if (p.classes.c[i].rule >= RULE_UNLOCK) return true;
    return false;
}

template <bool EXT, int LAYOUT>
static cudaError_t launch_step_many_t(const StepParams& p, cudaStream_t stream)
{
    const int grid = (p.N + STEP_THREADS - 1) / STEP_THREADS;
    const size_t smem = step_smem_bytes(p.packed_stride, p.num_classes, p.rows_bytes, LAYOUT) - STEP_STATIC_SMEM     // dynamic part
                        + ((MG_DYN_DEFER && EXT && p.dyn) ? (size_t)STEP_THREADS * 64 : 0);                           // + the generator records
    static size_t configured[64] = {0};                      // per device: the attribute belongs to the device's copy of the kernel
    int dev = 0;
    cudaGetDevice(&dev);
    if (smem > 48 * 1024 && (dev < 0 || dev >= 64 || smem > configured[dev])) {
        cudaError_t e = cudaFuncSetAttribute(step_many_kernel<EXT, LAYOUT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
        if (dev >= 0 && dev < 64) configured[dev] = smem;
    }
#if MG_STEP_PDL
    // back-to-back step launches (a rollout is hundreds of them): the next grid becomes resident while this one drains and
    // starts at its griddepcontrol.wait the moment this one has completed, instead of paying launch latency in between
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)grid); cfg.blockDim = dim3(STEP_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, step_many_kernel<EXT, LAYOUT>, p);
#else
    step_many_kernel<EXT, LAYOUT><<<grid, STEP_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
#endif
}

template <bool EXT>
static cudaError_t launch_step_many_l(const StepParams& p, cudaStream_t stream)
{
    switch (step_layout(p.packed_stride, p.num_classes, p.rows_bytes)) {
    case LAYOUT_ROWS: return launch_step_many_t<EXT, LAYOUT_ROWS>(p, stream);
    case LAYOUT_ROWS_HALF: return launch_step_many_t<EXT, LAYOUT_ROWS_HALF>(p, stream);
    default: return launch_step_many_t<EXT, LAYOUT_DUAL>(p, stream);
    }
}

cudaError_t launch_step_many(const StepParams& p_in, cudaStream_t stream)
{
    StepParams p = p_in;
    p.image_bulk = p.image && (reinterpret_cast<uintptr_t>(p.image) & 15) == 0 && ((size_t)p.N * OBS_BYTES) % 16 == 0;
    return needs_ext(p) ? launch_step_many_l<true>(p, stream) : launch_step_many_l<false>(p, stream);
}

// resident CTAs per SM of the step kernel at this tile size (for DESIGN.md / bench.py: is the grid one wave?)
cudaError_t step_occupancy(int packed_stride, int num_classes, int rows_bytes, int* blocks_per_sm, int* threads, size_t* smem)
{
    const int layout = step_layout(packed_stride, num_classes, rows_bytes);
    *threads = STEP_THREADS;
    *smem = step_smem_bytes(packed_stride, num_classes, rows_bytes, layout) - STEP_STATIC_SMEM;              // dynamic part
    const void* k = layout == LAYOUT_ROWS ? (const void*)step_many_kernel<false, LAYOUT_ROWS>
                  : layout == LAYOUT_ROWS_HALF ? (const void*)step_many_kernel<false, LAYOUT_ROWS_HALF>
                                               : (const void*)step_many_kernel<false, LAYOUT_DUAL>;
    if (*smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)*smem);
        if (e != cudaSuccess) return e;
    }
    return cudaOccupancyMaxActiveBlocksPerMultiprocessor(blocks_per_sm, k, STEP_THREADS, *smem);
}

cudaError_t launch_reset(const ResetParams& p, cudaStream_t stream)
{
    const int threads = 256;
    const int grid = (int)(((size_t)p.N * 32 + threads - 1) / threads);
    reset_kernel<<<grid, threads, 0, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_pack(const PackParams& p, cudaStream_t stream)
{
    const int threads = 256;
    const int grid = (int)(((size_t)p.n * 32 + threads - 1) / threads);
    pack_kernel<<<grid, threads, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mg
