// mg_dp.cu — Bellman sweep (K4); the state-space enumeration (K3b) is mg_enum.cu.  C ABI in include/mgb200_dp.h.
//
// The reference has no DP code.  The MDP is the one MiniGridEnv.step defines
// (minigrid/minigrid_env.py:520-590 + envs/keycorridor.py:129-137): K3b expands frontiers with the
// SAME apply_action() the step kernel uses (mg_common.cuh), K4 sweeps the successor table.
//
// K4 is an HBM stream: per state it reads A successor ids (action-major table: a warp reads one 128-byte line per
// action), gathers A values of V (L2 / L1 resident: BFS numbering keeps successors close), and writes 4 bytes.
// 36 B/state at A = 7.  No tensor cores: 0.5 flop/byte.
#include "../../include/mgb200.h"
#include "../../include/mgb200_dp.h"
#include "mg_common.cuh"

#include <cooperative_groups.h>
#include <cstdio>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <new>
#include <vector>

namespace cg = cooperative_groups;

namespace mgdp {

int64_t g_launches = 0;                   // also counted by mg_enum.cu
int fail(int code, const char* msg);      // mg_enum.cu (writes mg_last_error)

#ifndef MGDP_SWEEP_THREADS
#define MGDP_SWEEP_THREADS 256
#endif
// Launch shape, measured on B200 with tools/ab_sweep.sh (KeyCorridorS6R3, 16.8 M states x 7 actions; us per sweep):
//   4 consecutive states per thread (16-byte table loads), 6 CTAs/SM, grid 32 CTAs/SM      144.7   (round-1 v10)
//   same, persistent grid (one wave: resident CTAs only, grid-stride)                      130.4
//   NS strided states per thread (below), persistent: NS=4 @6 CTAs 128.3, NS=3 @8 117.6, NS=1 @8 123.5,
//   NS=2 @8 CTAs/SM (64 warps, 32 registers)                                               112.3   <- default
//   NS=2 with 128 / 512 / 1024-thread CTAs at full occupancy                               111.3 / 112.8 / 118.4
//   software-pipelined table loads (next tile's rows requested before the gathers)         166.7   (registers -> 32 warps)
//   TMA-staged table (cp.async.bulk + mbarrier ring, 24-28 warps/SM; MGDP_TMA=1)           195.9   (see sweep_tma_kernel)
// What bounds the sweep is the number of warps with gathers in flight, not the table stream: every variant that trades
// resident warps for deeper per-thread or asynchronous prefetch loses.
#ifndef MGDP_NS
#define MGDP_NS 2               // states per thread
#endif
#ifndef MGDP_SWEEP_MINB
#define MGDP_SWEEP_MINB 8       // <= 32 registers: 64 resident warps per SM
#endif
#ifndef MGDP_SOLVE_PUSH_MINB
#define MGDP_SOLVE_PUSH_MINB 6
#endif
#ifndef MGDP_SOLVE_MINB
#define MGDP_SOLVE_MINB 8
#endif
#ifndef MGDP_EVEN_TAIL
#define MGDP_EVEN_TAIL 0        // sweep_tiles: 1 = split the last, incomplete round of tiles evenly over the CTAs.  Measured and
#endif                          // rejected (profiles/r2/dp_tail_ab.txt): a tile's time is latency, not work — a partial tile with a
                                // third of its threads costs a whole tile-time, so every CTA pays what only some paid before:
                                // 1 GPU 103.1 vs 100.6 us per masked sweep, fused 2-GPU solve 90.0 vs 74.7 us per sweep
constexpr int SWEEP_THREADS = MGDP_SWEEP_THREADS;
constexpr int MAXA = MGDP_MAX_ACTIONS;

// V loads.  Single sweeps read V through the non-coherent path (ld.global.nc: V_in is constant for the launch).  The
// cooperative solves rewrite V between grid barriers, so they use ordinary weak loads: L1-cached like .nc (the gathers
// hit L1 half of the time), and made coherent by the barrier itself — grid.sync() is a gpu-scope release/acquire pattern,
// after which a weak load must observe the other CTAs' stores (the acquire invalidates L1).  ld.global.cg (L2 only) cost
// 150 us per sweep against 134 us with plain loads in the quad layout.
template <bool COHERENT> __device__ __forceinline__ float ld_v(const float* p) { return COHERENT ? *p : __ldg(p); }
template <bool COHERENT> __device__ __forceinline__ float4 ld_v4(const float* p)
{
    return COHERENT ? *reinterpret_cast<const float4*>(p) : __ldg(reinterpret_cast<const float4*>(p));
}

// q(s,a) = 1 | 0 | gamma * V[s']  — a single rounded multiply, identical to the oracle's C expression.
// Self-loops (pickup / drop / toggle / done are usually no-ops) reuse the V[s] the thread already holds.
// The compiler wraps every gather below in its own divergence region (BSSY / BRA / BSYNC: 131 regions in sweep_kernel<7>).
// Two branch-free forms were measured against it on B200 (tools/ab_sweep.sh, us per sweep): always-load
// `V[term ? s : sp]` 109.8, predicated `@p ld.global.nc` in inline PTX 117.9, this one 109.0 — with 64 warps per SM the
// serialisation inside a thread does not matter, and the branchy form skips the most instructions for self-loops.
template <bool COHERENT>
__device__ __forceinline__ float q_value(int32_t sp, int32_t s, float v_self, const float* V, float gamma)
{
    if (sp == MGDP_WIN) return 1.0f;
    if (sp == MGDP_LOSE) return 0.0f;
    MG_ASSERT(sp >= 0);
    return __fmul_rn(gamma, sp == s ? v_self : ld_v<COHERENT>(V + sp));
}

// Peer copies of V' for the fused multi-GPU sweep (mgdp_sweep_push / mgdp_solve_push): a value a rank computes is stored
// into the V' buffer of every other rank as well (P2P stores over NVLink, fire-and-forget) — but only if it differs
// from what that buffer already holds — so the exchange rides inside the sweep instead of following it as an all-gather.
// mc != NULL: an NVLS multicast mapping of the same buffer (all ranks' copies at once): ONE multimem.st through the NVSwitch
// replaces the n unicast stores — the rank's outbound NVLink traffic drops n-fold, the switch does the replication.
struct PeerOut { float* p[MGDP_MAX_PEERS]; int n; int changed_only; float* mc; };

__device__ __forceinline__ void multimem_st(float* mc_addr, float v)
{
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(mc_addr), "f"(v) : "memory");
}

// Policy evaluation (pi_fixed): one successor per state.  One thread backs up FOUR consecutive states s4 .. s4+3
// (s4 % 4 == 0); states outside [s_lo, s_hi) are computed on padding and discarded.  Returns the residual of the valid ones.
template <bool COHERENT>
__device__ __forceinline__ float backup4_fixed(const int32_t* __restrict__ succ, int64_t S_pad, int64_t s4,
                                               int64_t s_lo, int64_t s_hi, const float* V, float* V_out,
                                               uint8_t* pi_out, const uint8_t* __restrict__ pi_fixed, float gamma)
{
    const float4 vs4 = ld_v4<COHERENT>(V + s4);
    const float vs[4] = {vs4.x, vs4.y, vs4.z, vs4.w};
    float best[4]; int ba[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t s = s4 + j;
        ba[j] = (s >= s_lo && s < s_hi) ? pi_fixed[s] : 0;
        best[j] = q_value<COHERENT>(__ldcs(succ + (int64_t)ba[j] * S_pad + s), (int32_t)s, vs[j], V, gamma);
    }
    float r = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int64_t s = s4 + j;
        if (s >= s_lo && s < s_hi) {
            V_out[s] = best[j];
            if (pi_out) pi_out[s] = (uint8_t)ba[j];
            r = fmaxf(r, fabsf(best[j] - vs[j]));
        }
    }
    return r;
}

// One thread backs up NS states s0, s0 + STRIDE, ... (STRIDE = CTA size): every warp-wide instruction covers 32
// CONSECUTIVE states — table loads and V' stores are single 128-byte lines, and a gather whose successors are near the
// states themselves (turns, blocked moves: BFS numbering keeps them close) touches 4-5 sectors.  NS * A gathers in flight
// per thread.  Self-loops reuse V[s].  PUSH: see PeerOut; in changed-only mode every rank holds identical full copies of
// both buffers, so the stale content of the local V_out says what the peers' V_out holds — only values that differ from it
// cross NVLink (a value that changed in sweep k is pushed in sweeps k and k+1, once per buffer, and never again while it
// stays put).  The stale value is loaded early: its latency hides behind the gathers.
// MODE: PI_NONE values only (the max is one FMNMX per action, no argmax bookkeeping), PI_BOTH values + policy,
// PI_ONLY policy only (no V', no residual: the extraction pass the solves run once after convergence).
enum : int { PI_NONE = 0, PI_BOTH = 1, PI_ONLY = 2 };
// MASKED: rowmask[s >> 5] bit a = "action row a holds at least one non-self successor among the 32 states s & ~31 .. + 31"
// (mgdp_rowmask).  A warp covers exactly those 32 states, so the test is warp-uniform: rows whose bit is clear are not read at
// all — every lane's successor is the state itself (pickup / drop / toggle / done are no-ops in most states; `done` always) —
// and the sweep moves 20-27 bytes per backup instead of 36.  Same arithmetic, bit-identical V and pi.
template <int A_T, bool COHERENT, bool FULL, bool PUSH, int MODE, bool MC = false, bool MASKED = false>
__device__ __forceinline__ float backup_strided(const int32_t* __restrict__ succ, int64_t S_pad, int A, int32_t s0,
                                                int32_t s_hi, const float* V, float* V_out, uint8_t* pi_out, float gamma,
                                                const PeerOut* peers, const uint8_t* __restrict__ rowmask = nullptr)
{
    constexpr int NS = MGDP_NS, STRIDE = SWEEP_THREADS;       // compile-time stride: the NS addresses are immediate offsets
    const int AA = A_T > 0 ? A_T : A;
    // a partial (last) tile recomputes state s0 in its out-of-range slots and drops the result
    int32_t off[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) off[j] = (FULL || s0 + j * STRIDE < s_hi) ? j * STRIDE : 0;
    int32_t sp[MAXA][NS];
    const int32_t* row = succ + s0;
    uint32_t rm[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) rm[j] = MASKED ? (uint32_t)__ldg(rowmask + ((s0 + off[j]) >> 5)) : 0xffu;
#pragma unroll
    for (int a = 0; a < MAXA; ++a)
        if (a < AA) {
#pragma unroll
            for (int j = 0; j < NS; ++j)
                sp[a][j] = (!MASKED || ((rm[j] >> a) & 1u)) ? __ldcs(row + off[j]) : s0 + off[j];       // streamed: evict-first
            row += S_pad;
        }
    float vs[NS], old[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) vs[j] = ld_v<COHERENT>(V + s0 + off[j]);
    const bool filter = PUSH && MODE != PI_ONLY && peers->changed_only;
#pragma unroll
    for (int j = 0; j < NS; ++j) old[j] = filter ? __ldcg(V_out + s0 + off[j]) : 0.0f;
    float best[NS]; int ba[NS];
#pragma unroll
    for (int j = 0; j < NS; ++j) { best[j] = -1.0f; ba[j] = 0; }
    // values only + row mask: a skipped row contributes gamma * V[s] for every lane of the warp — fold ALL skipped rows of a
    // state into ONE max after the loop and branch (warp-uniformly) around the rest: the masked sweep is issue-bound, not
    // HBM-bound (ncu: 76 % issue slots busy at 40 % DRAM), so the instructions of the skipped rows are what is worth saving
    constexpr bool SKIP = MASKED && MODE == PI_NONE;
#pragma unroll
    for (int a = 0; a < MAXA; ++a)
        if (a < AA) {
#pragma unroll
            for (int j = 0; j < NS; ++j) {
                if (SKIP && !((rm[j] >> a) & 1u)) continue;
                const float q = q_value<COHERENT>(sp[a][j], s0 + off[j], vs[j], V, gamma);
                if (MODE == PI_NONE) best[j] = fmaxf(best[j], q);
                else if (q > best[j]) { best[j] = q; ba[j] = a; }         // lowest index wins ties
            }
        }
    if (SKIP) {
#pragma unroll
        for (int j = 0; j < NS; ++j)
            if ((~rm[j]) & ((1u << AA) - 1u)) best[j] = fmaxf(best[j], __fmul_rn(gamma, vs[j]));
    }
    float r = 0.0f;
#pragma unroll
    for (int j = 0; j < NS; ++j)
        if (FULL || j == 0 || off[j] != 0) {
            const int32_t s = s0 + off[j];
            if (MODE != PI_ONLY) {
                if (PUSH && (!filter || old[j] != best[j])) {
                    if (MC) multimem_st(peers->mc + s, best[j]);
                    else for (int k = 0; k < peers->n; ++k) peers->p[k][s] = best[j];
                }
                V_out[s] = best[j];
                r = fmaxf(r, fabsf(best[j] - vs[j]));
            }
            if (MODE != PI_NONE) pi_out[s] = (uint8_t)ba[j];
        }
    return r;
}

// the CTA's loop over tiles of NS * SWEEP_THREADS states of [s_begin, s_end): full tiles take the branch-free path
template <int A_T, bool COHERENT, bool PUSH = false, int MODE = PI_NONE, bool MC = false, bool MASKED = false>
__device__ __forceinline__ float sweep_tiles(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                                             const float* V, float* V_out, uint8_t* pi_out, float gamma,
                                             const PeerOut* peers = nullptr, const uint8_t* __restrict__ rowmask = nullptr)
{
    constexpr int64_t T = (int64_t)MGDP_NS * SWEEP_THREADS;
    float r = 0.0f;
#if MGDP_EVEN_TAIL
    // Full rounds (every CTA one whole tile), then the remainder — fewer tiles than CTAs — is split EVENLY over all CTAs in
    // 32-state-aligned chunks (<= one tile each, partial-tile path) instead of giving some CTAs one more whole tile: with
    // 4.6 tiles per CTA (a rank of the 8-GPU solve) the slowest CTA does 4.6 tile-times instead of 5; a barrier follows
    // every sweep, so the slowest CTA is what counts.  (32-bit arithmetic: ranges are below 2^31 states.)
    const uint32_t n = (uint32_t)(s_end - s_begin), G = gridDim.x, GT = G * (uint32_t)T;
    const uint32_t full = n / GT * GT;                                   // states covered by the full rounds
    const int32_t b0 = (int32_t)s_begin;
    for (uint32_t o = blockIdx.x * (uint32_t)T; o < full; o += GT) {
        const int32_t s0 = b0 + (int32_t)o + (int32_t)threadIdx.x;
        r = fmaxf(r, backup_strided<A_T, COHERENT, true, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, (int32_t)s_end, V, V_out, pi_out, gamma, peers, rowmask));
    }
    if (full < n) {
        const uint32_t chunk = ((n - full + G - 1) / G + 31u) & ~31u;
        const uint32_t lo = full + blockIdx.x * chunk, hi = min(lo + chunk, n);
        const int32_t s0 = b0 + (int32_t)lo + (int32_t)threadIdx.x;
        if (lo + threadIdx.x < hi)
            r = fmaxf(r, backup_strided<A_T, COHERENT, false, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, b0 + (int32_t)hi, V, V_out, pi_out, gamma, peers, rowmask));
    }
#else
    for (int64_t base = s_begin + (int64_t)blockIdx.x * T; base < s_end; base += (int64_t)gridDim.x * T) {
        const int32_t s0 = (int32_t)base + (int32_t)threadIdx.x;
        if (base + T <= s_end) r = fmaxf(r, backup_strided<A_T, COHERENT, true, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, (int32_t)s_end, V, V_out, pi_out, gamma, peers, rowmask));
        else if (s0 < s_end) r = fmaxf(r, backup_strided<A_T, COHERENT, false, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, (int32_t)s_end, V, V_out, pi_out, gamma, peers, rowmask));
    }
#endif
    return r;
}

// The same tiles handed out DYNAMICALLY inside a persistent (cooperative) kernel: a CTA takes tile blockIdx.x first and then
// whatever index the next atomicAdd on `counter` returns (+ gridDim.x).  A barrier follows every sweep of the fused multi-GPU
// solve, so what counts is when the LAST CTA finishes; with the static round-robin split a rank of the 8-GPU solve has 5.5 tiles
// per CTA — some do 6, and a tile's duration is HBM latency, which varies — and the first CTA waited 8 us of a 21 us sweep for
// the last one (profiles/r2/dp_fused_8gpu_timing.txt).  The counter is reset by the CTA that arrives last at the sweep's
// barrier (nobody is taking tiles then).  Which CTA computes a tile does not change what it computes: same V, pi, residual.
// MEASURED AND NOT TAKEN (8 GPUs, profiles/r2/dp_tail_ab.txt): first-to-last arrival 8.4 -> 6.1 us, but the work phase grows from
// 21 to 25 us (an atomic + a CTA barrier per tile, and consecutive tiles of a CTA are no longer a fixed stride apart): 36.0 against
// 33.2 us per sweep.  Kept behind -DMGDP_PUSH_DYN_TILES=1.
#ifndef MGDP_PUSH_DYN_TILES
#define MGDP_PUSH_DYN_TILES 0
#endif
template <int A_T, bool PUSH, int MODE, bool MC, bool MASKED>
__device__ __forceinline__ float sweep_tiles_dyn(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                                                 const float* V, float* V_out, float gamma, const PeerOut* peers,
                                                 const uint8_t* __restrict__ rowmask, uint32_t* counter, uint32_t* next_sh /*[2] shared*/)
{
    constexpr int64_t T = (int64_t)MGDP_NS * SWEEP_THREADS;
    const uint32_t ntiles = (uint32_t)((s_end - s_begin + T - 1) / T);
    float r = 0.0f;
    uint32_t tile = blockIdx.x, par = 0;
    while (tile < ntiles) {
        uint32_t nxt = 0;
        if (threadIdx.x == 0) nxt = gridDim.x + atomicAdd(counter, 1u);        // in flight while the tile is worked on
        const int64_t base = s_begin + (int64_t)tile * T;
        const int32_t s0 = (int32_t)base + (int32_t)threadIdx.x;
        if (base + T <= s_end) r = fmaxf(r, backup_strided<A_T, true, true, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, (int32_t)s_end, V, V_out, nullptr, gamma, peers, rowmask));
        else if (s0 < s_end) r = fmaxf(r, backup_strided<A_T, true, false, PUSH, MODE, MC, MASKED>(succ, S_pad, A, s0, (int32_t)s_end, V, V_out, nullptr, gamma, peers, rowmask));
        if (threadIdx.x == 0) next_sh[par] = nxt;
        __syncthreads();                                                        // (the two slots alternate: one barrier per tile)
        tile = next_sh[par];
        par ^= 1u;
    }
    return r;
}

// the same loop for policy evaluation (quads)
template <bool COHERENT>
__device__ __forceinline__ float sweep_quads_fixed(const int32_t* __restrict__ succ, int64_t S_pad, int64_t s_begin, int64_t s_end,
                                                   const float* V, float* V_out, uint8_t* pi_out, const uint8_t* pi_fixed, float gamma)
{
    float r = 0.0f;
    const int64_t q0 = s_begin >> 2, q1 = (s_end + 3) >> 2;
    for (int64_t q = q0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < q1; q += (int64_t)gridDim.x * blockDim.x)
        r = fmaxf(r, backup4_fixed<COHERENT>(succ, S_pad, q << 2, s_begin, s_end, V, V_out, pi_out, pi_fixed, gamma));
    return r;
}

__device__ __forceinline__ float block_max(float v, float* sh)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) sh[w] = v;
    __syncthreads();
    if (w == 0) {
        v = (l < (int)(blockDim.x >> 5)) ? sh[l] : 0.0f;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    }
    __syncthreads();
    return v;      // valid in warp 0
}

// residuals are >= 0, so their float bit patterns order like unsigned ints
__device__ __forceinline__ void atomic_max_nonneg(float* addr, float v)
{
    atomicMax(reinterpret_cast<unsigned int*>(addr), __float_as_uint(v));
}

// WITH_PI: the caller wants the policy too (or policy evaluation) — its own instantiation, so that the values-only sweep (the
// one the solves and the bench time) does not carry the argmax path's registers
template <int A_T, bool MASKED = false, bool WITH_PI = false>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SWEEP_MINB)
sweep_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
             const float* __restrict__ V_in, float* __restrict__ V_out, uint8_t* __restrict__ pi_out,
             const uint8_t* __restrict__ pi_fixed, float gamma, float* resid, const uint8_t* __restrict__ rowmask)
{
    __shared__ float sh[32];
    float r = 0.0f;
    if (WITH_PI && !MASKED && pi_fixed) r = sweep_quads_fixed<false>(succ, S_pad, s_begin, s_end, V_in, V_out, pi_out, pi_fixed, gamma);
    else if (WITH_PI) r = sweep_tiles<A_T, false, false, PI_BOTH, false, MASKED>(succ, S_pad, A, s_begin, s_end, V_in, V_out, pi_out, gamma, nullptr, rowmask);
    else r = sweep_tiles<A_T, false, false, PI_NONE, false, MASKED>(succ, S_pad, A, s_begin, s_end, V_in, V_out, nullptr, gamma, nullptr, rowmask);
    r = block_max(r, sh);
    if (threadIdx.x == 0 && r > 0.0f) atomic_max_nonneg(resid, r);
}

// ---- TMA-staged sweep: the table rows reach shared memory as 1-D bulk copies (cp.async.bulk, completion on an mbarrier) ----
// A persistent CTA walks tiles of TMA_THREADS quads; one thread keeps MGDP_TMA_STAGES tiles of A rows in flight, so the
// HBM stream of the successor table never waits for the V gathers of the warps (which hold no table data in registers
// while they wait).  Rows are fetched with an L2 evict-first policy: the table is read once per sweep, V stays resident.
#ifndef MGDP_TMA
#define MGDP_TMA 0
#endif
#ifndef MGDP_TMA_THREADS
#define MGDP_TMA_THREADS 256
#endif
#ifndef MGDP_TMA_STAGES
#define MGDP_TMA_STAGES 2
#endif
#ifndef MGDP_TMA_MINB
#define MGDP_TMA_MINB 3
#endif
[[maybe_unused]] constexpr int TMA_THREADS = MGDP_TMA_THREADS, TMA_STAGES = MGDP_TMA_STAGES;

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* b, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* b, uint64_t policy)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(b)), "l"(policy) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t parity)
{
    uint32_t ok = 0;
    const uint32_t a = smem_u32(b);
    do {
        asm volatile("{\n .reg .pred p;\n mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n selp.u32 %0, 1, 0, p;\n}"
                     : "=r"(ok) : "r"(a), "r"(parity) : "memory");
    } while (!ok);
}

template <int A_T>
__global__ void __launch_bounds__(TMA_THREADS, MGDP_TMA_MINB)
sweep_tma_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                 const float* __restrict__ V_in, float* __restrict__ V_out, uint8_t* __restrict__ pi_out,
                 float gamma, float* resid)
{
    extern __shared__ __align__(128) unsigned char tma_smem[];
    __shared__ float sh[32];
    __shared__ __align__(8) uint64_t full[TMA_STAGES];
    const int AA = A_T > 0 ? A_T : A;
    int4* buf = reinterpret_cast<int4*>(tma_smem);                    // [stage][action][TMA_THREADS] quads
    const int tid = threadIdx.x;
    const int64_t q0 = s_begin >> 2, q1 = (s_end + 3) >> 2;
    const int64_t ntiles = (q1 - q0 + TMA_THREADS - 1) / TMA_THREADS;
    uint64_t policy = 0;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < TMA_STAGES; ++i) mbar_init(full + i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
    }
    __syncthreads();
    auto issue = [&](int64_t tile, int stage) {                       // thread 0 only
        const int64_t qs = q0 + tile * TMA_THREADS;
        const uint32_t bytes = (uint32_t)min((int64_t)TMA_THREADS, q1 - qs) * 16u;
        mbar_expect_tx(full + stage, bytes * (uint32_t)AA);
        for (int a = 0; a < AA; ++a)
            bulk_g2s(buf + ((size_t)stage * AA + a) * TMA_THREADS, succ + (int64_t)a * S_pad + (qs << 2), bytes, full + stage, policy);
    };
    int64_t tile = blockIdx.x;
    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < TMA_STAGES; ++i) {
            const int64_t t = tile + (int64_t)i * gridDim.x;
            if (t < ntiles) issue(t, i);
        }
    }
    float r = 0.0f;
    int stage = 0; uint32_t parity = 0;
    for (; tile < ntiles; tile += gridDim.x) {
        mbar_wait(full + stage, parity);
        const int64_t q = q0 + tile * TMA_THREADS + tid;
        if (q < q1) {
            // four consecutive states per thread (the 16-byte unit of the bulk copies); rows come from shared memory
            const int64_t s4 = q << 2;
            const float4 vs4 = __ldg(reinterpret_cast<const float4*>(V_in + s4));
            const float vs[4] = {vs4.x, vs4.y, vs4.z, vs4.w};
            float best[4] = {-1.0f, -1.0f, -1.0f, -1.0f}; int ba[4] = {0, 0, 0, 0};
#pragma unroll
            for (int a = 0; a < MAXA; ++a)
                if (a < AA) {
                    const int4 row = buf[((size_t)stage * AA + a) * TMA_THREADS + tid];
                    const int32_t spj[4] = {row.x, row.y, row.z, row.w};
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float qv = q_value<false>(spj[j], (int32_t)(s4 + j), vs[j], V_in, gamma);
                        if (qv > best[j]) { best[j] = qv; ba[j] = a; }
                    }
                }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int64_t sx = s4 + j;
                if (sx >= s_begin && sx < s_end) {
                    V_out[sx] = best[j];
                    if (pi_out) pi_out[sx] = (uint8_t)ba[j];
                    r = fmaxf(r, fabsf(best[j] - vs[j]));
                }
            }
        }
        __syncthreads();                                              // the stage's rows are consumed: refill it
        if (tid == 0) {
            const int64_t tn = tile + (int64_t)TMA_STAGES * gridDim.x;
            if (tn < ntiles) issue(tn, stage);
        }
        if (++stage == TMA_STAGES) { stage = 0; parity ^= 1u; }
    }
    r = block_max(r, sh);
    if (tid == 0 && r > 0.0f) atomic_max_nonneg(resid, r);
}

template <int A_T>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SWEEP_MINB)
sweep_push_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                  const float* __restrict__ V_in, float* __restrict__ V_out, uint8_t* __restrict__ pi_out,
                  float gamma, float* resid, const __grid_constant__ PeerOut peers)
{
    __shared__ float sh[32];
    float r = 0.0f;
    if (pi_out) r = sweep_tiles<A_T, false, true, PI_BOTH>(succ, S_pad, A, s_begin, s_end, V_in, V_out, pi_out, gamma, &peers);
    else r = sweep_tiles<A_T, false, true, PI_NONE>(succ, S_pad, A, s_begin, s_end, V_in, V_out, nullptr, gamma, &peers);
    r = block_max(r, sh);
    if (threadIdx.x == 0 && r > 0.0f) atomic_max_nonneg(resid, r);
    __threadfence_system();            // peer stores are ordered before whatever the stream runs next (the residual all-reduce)
}

#ifdef MGDP_TIMING
// instrumentation build only (tools/solve_timing.py): per sweep, first / last arrival at the grid barrier and first release
__device__ unsigned long long g_timing[3][128];
__device__ __forceinline__ unsigned long long timing_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
#endif

// All sweeps in one cooperative launch: grid-wide barrier between sweeps, convergence test on the device.
// resid_slots: float[3] rotating residual accumulators, all zero on entry (sweep k uses slot k%3 and re-zeroes slot (k+1)%3).
// V is read with ordinary (coherent at the barrier) loads: it is rewritten every sweep, so the .nc path would go stale.
// FIXED: policy evaluation (pi_fixed) — its own instantiation, so that the quad path does not cost the value-iteration loop
// registers
template <int A_T, bool FIXED, bool MASKED = false>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SOLVE_MINB)
solve_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t S, float* Va, float* Vb,
             uint8_t* __restrict__ pi, const uint8_t* __restrict__ pi_fixed, float gamma, float tol,
             int max_sweeps, float* resid_slots, int32_t* result, const uint8_t* __restrict__ rowmask)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ float sh[32];
    float* Vin = Va; float* Vout = Vb;
    int k = 0;
    float last = 0.0f;
    for (; k < max_sweeps;) {
        float r = FIXED ? sweep_quads_fixed<true>(succ, S_pad, 0, S, Vin, Vout, pi, pi_fixed, gamma)
                        : sweep_tiles<A_T, true, false, PI_NONE, false, MASKED>(succ, S_pad, A, 0, S, Vin, Vout, nullptr, gamma, nullptr, rowmask);
        r = block_max(r, sh);
        float* slot = resid_slots + (k % 3);
        if (threadIdx.x == 0) {
            if (r > 0.0f) atomic_max_nonneg(slot, r);
            if (blockIdx.x == 0) resid_slots[(k + 1) % 3] = 0.0f;       // last read before barrier k-1; next used in sweep k+1
        }
#ifdef MGDP_TIMING
        if (threadIdx.x == 0 && k < 128) {                  // when does each CTA reach the barrier of sweep k?
            const unsigned long long t = timing_ns();
            atomicMin(&g_timing[0][k], t); atomicMax(&g_timing[1][k], t);
        }
#endif
        __threadfence();
        grid.sync();
#ifdef MGDP_TIMING
        if (threadIdx.x == 0 && k < 128) atomicMin(&g_timing[2][k], timing_ns());     // first CTA released
#endif
        last = __ldcg(slot);
        ++k;
        float* tmp = Vin; Vin = Vout; Vout = tmp;
        if (last < tol) break;
    }
    // After k sweeps the newest values are in Va (k even) or Vb (k odd) and the other buffer still holds V_{k-1}, the input of
    // the last sweep: extract_kernel reads the policy out of it, final_copy_kernel then moves the result into Va (mgdp_solve).
    if (blockIdx.x == 0 && threadIdx.x == 0) { result[0] = k; result[1] = __float_as_int(last); }
}

// The policy is extracted ONCE per solve, from the input of the last sweep: exactly the argmax that sweep computed, without
// argmax bookkeeping and 1 B/state of writes in every sweep.  k = result[0] (on the device: no host round trip).
template <int A_T, bool MASKED = false>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SWEEP_MINB)
extract_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
               const float* Va, const float* Vb, const int32_t* __restrict__ result, uint8_t* __restrict__ pi, float gamma,
               const uint8_t* __restrict__ rowmask)
{
    const int k = result[0];
    if (k <= 0) return;
    const float* prev = (k & 1) ? Va : Vb;
    sweep_tiles<A_T, false, false, PI_ONLY, false, MASKED>(succ, S_pad, A, s_begin, s_end, prev, nullptr, pi, gamma, nullptr, rowmask);
}

// ---- streamed solve (mgdp_solve_stream): one launch per sweep, convergence decided on the device ----
// In the HBM regime the cooperative solve is ~9 % slower per sweep than back-to-back sweep launches (109.6 us against 100.5 us
// on the 16.8 M-state table): its V gathers must be ordinary loads (V is rewritten between the grid barriers), its tiles are
// split statically, and it pays the barrier.  Here every sweep is its own launch of the sweep_kernel body — V through the
// read-only path, the hardware CTA scheduler doing the balancing — behind a GATE: sweep k first reads the residual sweep k-1
// left in its slot of a small ring; below tol (or with the stop flag already up) the launch returns at once, and block 0 of
// the FIRST launch that sees it records the result and raises the flag.  A launch that returns leaves its own slot at zero,
// so every launch behind it returns too.  There is nothing after the tiles but the residual's atomic max, exactly like
// sweep_kernel (an in-kernel epilogue that found the last CTA and tested the residual there cost the 32-register tile loop a
// 40-byte spill).  The host enqueues sweeps in chunks, clears the ring slots of the chunk after the next, and looks at the
// flag one chunk late, so the device never waits for the host.
// ctrl[0]: stop flag (zero on entry).  ring: STREAM_RING floats, slot k % STREAM_RING = residual of sweep k.
constexpr int STREAM_RING = 32, STREAM_CHUNK = 8;
template <int A_T, bool MASKED>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SWEEP_MINB)
sweep_gated_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                   const float* __restrict__ Vin, float* __restrict__ Vout, float gamma, float tol, int k, unsigned int* ctrl,
                   float* ring, int32_t* result, const uint8_t* __restrict__ rowmask)
{
    __shared__ float sh[32];
    if (k > 0) {
        const float prev = __ldcg(ring + ((k - 1) & (STREAM_RING - 1)));
        const unsigned int stopped = __ldcg(ctrl);
        if (stopped != 0u || prev < tol) {                                  // converged k sweeps ago or earlier
            if (stopped == 0u && blockIdx.x == 0 && threadIdx.x == 0) {
                result[0] = k; result[1] = __float_as_int(prev); *ctrl = 1u;
            }
            return;
        }
    }
    // (s_begin is a parameter although it is always 0: with the literal the compiler's tile loop spilled 40 bytes)
    float r = sweep_tiles<A_T, false, false, PI_NONE, false, MASKED>(succ, S_pad, A, s_begin, s_end, Vin, Vout, nullptr, gamma, nullptr, rowmask);
    r = block_max(r, sh);
    if (threadIdx.x == 0 && r > 0.0f) atomic_max_nonneg(ring + (k & (STREAM_RING - 1)), r);
}
// after the last launch the host enqueues: nothing stopped the stream, so all `k` sweeps ran
__global__ void stream_finish_kernel(int k, const unsigned int* ctrl, const float* ring, int32_t* result)
{
    if (*ctrl == 0u) { result[0] = k; result[1] = __float_as_int(ring[(k - 1) & (STREAM_RING - 1)]); }
}

// rowmask[w] bit a = some state of [32 w, 32 w + 32) below S has succ[a][s] != s   (one warp per 32 states)
__global__ void __launch_bounds__(256)
rowmask_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t S, uint8_t* __restrict__ rowmask)
{
    const int64_t w = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (w >= (S_pad >> 5)) return;
    const int64_t s = (w << 5) + (threadIdx.x & 31);
    uint32_t m = 0;
    for (int a = 0; a < A; ++a) {
        const bool moves = s < S && __ldg(succ + (int64_t)a * S_pad + s) != (int32_t)s;
        if (__ballot_sync(0xffffffffu, moves)) m |= 1u << a;
    }
    if ((threadIdx.x & 31) == 0) rowmask[w] = (uint8_t)m;
}

__global__ void __launch_bounds__(256) final_copy_kernel(float* Va, const float* __restrict__ Vb, const int32_t* __restrict__ result, int64_t S)
{
    if (!(result[0] & 1)) return;
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < S; s += (int64_t)gridDim.x * blockDim.x) Va[s] = Vb[s];
}

template <int A_T>
static cudaError_t launch_extract(const int32_t* succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end, const float* Va,
                                  const float* Vb, const int32_t* result, uint8_t* pi, float gamma, int sms, cudaStream_t st,
                                  const uint8_t* rowmask = nullptr)
{
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = (s_end - s_begin + TILE - 1) / TILE;
    if (blocks > (int64_t)sms * MGDP_SWEEP_MINB) blocks = (int64_t)sms * MGDP_SWEEP_MINB;
    if (blocks < 1) return cudaSuccess;
    if (rowmask) extract_kernel<A_T, true><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, rowmask);
    else extract_kernel<A_T, false><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, nullptr);
    return cudaGetLastError();
}
static cudaError_t launch_extract_any(const int32_t* succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end, const float* Va,
                                      const float* Vb, const int32_t* result, uint8_t* pi, float gamma, int sms, cudaStream_t st,
                                      const uint8_t* rowmask = nullptr)
{
    switch (A) {
    case 3: return launch_extract<3>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, sms, st, rowmask);
    case 5: return launch_extract<5>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, sms, st, rowmask);
    case 7: return launch_extract<7>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, sms, st, rowmask);
    default: return launch_extract<0>(succ, S_pad, A, s_begin, s_end, Va, Vb, result, pi, gamma, sms, st, rowmask);
    }
}


// ---- all sweeps of the state-sharded solve in ONE cooperative launch per GPU (mgdp_solve_push) ----
// Per sweep: back up the own state range, storing V' locally and into every peer's buffer (P2P stores over NVLink, or one
// multimem.st through the NVSwitch); then ONE synchronisation round:
//   * every CTA: block_max (bar.sync: the CTA's stores are ordered before thread 0), thread 0 adds its residual with an
//     atomicMax and ARRIVES with one atom.add.acq_rel.gpu on a monotonically growing counter;
//   * the CTA whose add completes the sweep's count is the publisher: its acquire has synchronised with every other CTA's
//     release, so ONE fence.acq_rel.sys + st.release.sys of (residual, epoch flag) into every peer's control block makes
//     all of this GPU's peer stores visible to whoever acquires the flag (release cumulativity; the pattern of symmetric-
//     memory barriers) — no per-thread __threadfence_system().  Lanes 0..world-1 of its first warp then poll the peers'
//     flags (ld.acquire.sys, bounded wait on %globaltimer: a missing peer is an error, never a hang), reduce the residuals
//     with shuffles, and release the local generation word (st.release.gpu);
//   * everybody else spins on the generation word with one thread per CTA (ld.acquire.gpu) and rejoins with bar.sync.
// That replaces round 1's protocol (a sys-scope fence in EVERY thread, two grid.sync() and a serial flag poll by one
// thread; kept under MGDP_PUSH_SYNC=0 for A/B runs).  No host round trip, no NCCL call per sweep.
// Control block (symmetric memory, zeroed and barrier-ed by the host before the launch), 32-bit words:
//   [0..15] flags[r] = last epoch rank r finished      [16..31], [32..47] resid[epoch & 1][r] (float bits)
// scratch (local, zeroed by the launcher): float [0..2] rotating local residuals, [3] global residual, [4] abort; [7] tile counter (sweep_tiles_dyn);
//   uint32 [5] arrival counter, [6] generation
struct PeerCtrl { uint32_t* p[MGDP_MAX_PEERS]; int n; };
constexpr int CTRL_FLAGS = 0, CTRL_RESID = 16;
#ifndef MGDP_PUSH_SYNC
#define MGDP_PUSH_SYNC 1
#endif
#ifndef MGDP_PUSH_REL_FENCE
#define MGDP_PUSH_REL_FENCE 1
#endif

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) { asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(uint32_t* p, uint32_t v) { asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p)
{
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint32_t atom_add_acq_rel_gpu(uint32_t* p, uint32_t v)
{
    uint32_t old;
    asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
    return old;
}
__device__ __forceinline__ void fence_acq_rel_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ unsigned long long global_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}

#ifdef MGDP_TIMING
// instrumentation build (tools/solve_timing.py --push): per sweep k < 128 of solve_push_kernel, %globaltimer at
//   [0] first CTA arrival   [1] last CTA arrival (= the publisher takes over)   [2] own flags + residual published (after the
//   sys fence)   [3] all peers' flags seen   [4] generation released
__device__ unsigned long long g_push_timing[5][128];
#define PUSH_T(i, k, op) do { if ((k) < 128) op(&g_push_timing[i][k], global_ns()); } while (0)
__device__ __forceinline__ void t_set(unsigned long long* p, unsigned long long v) { *p = v; }
__device__ __forceinline__ void t_min(unsigned long long* p, unsigned long long v) { atomicMin(p, v); }
#else
#define PUSH_T(i, k, op) do { } while (0)
#endif

template <int A_T, bool MC, bool MASKED = false>
__global__ void __launch_bounds__(SWEEP_THREADS, MGDP_SOLVE_PUSH_MINB)
solve_push_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t s_begin, int64_t s_end,
                  float* Va, float* Vb, float gamma, float tol, int max_sweeps,
                  const __grid_constant__ PeerOut peersA, const __grid_constant__ PeerOut peersB,
                  uint32_t* ctrl, const __grid_constant__ PeerCtrl peer_ctrl, int rank, int world,
                  float* scratch, int32_t* result, const uint8_t* __restrict__ rowmask)
{
#if !MGDP_PUSH_SYNC
    cg::grid_group grid = cg::this_grid();
#endif
    __shared__ float sh[32];
    float* Vin = Va; float* Vout = Vb;
    const PeerOut* pout = &peersB;                       // peers' copies of Vout
    const PeerOut* pin = &peersA;
#if MGDP_PUSH_SYNC
    uint32_t* arrive = reinterpret_cast<uint32_t*>(scratch) + 5;
    uint32_t* gen = reinterpret_cast<uint32_t*>(scratch) + 6;
#if MGDP_PUSH_DYN_TILES
    uint32_t* tile_ctr = reinterpret_cast<uint32_t*>(scratch) + 7;             // tiles handed out so far in this sweep (beyond the first gridDim.x)
    __shared__ uint32_t next_sh[2];
#endif
#endif
    int k = 0;
    float g = 0.0f;
    bool aborted = false;
    for (; k < max_sweeps;) {
#if MGDP_PUSH_SYNC && MGDP_PUSH_DYN_TILES
        float r = sweep_tiles_dyn<A_T, true, PI_NONE, MC, MASKED>(succ, S_pad, A, s_begin, s_end, Vin, Vout, gamma, pout, rowmask, tile_ctr, next_sh);
#else
        float r = sweep_tiles<A_T, true, true, PI_NONE, MC, MASKED>(succ, S_pad, A, s_begin, s_end, Vin, Vout, nullptr, gamma, pout, rowmask);
#endif
        r = block_max(r, sh);
        float* slot = scratch + (k % 3);
        const uint32_t epoch = (uint32_t)k + 1u;
#if MGDP_PUSH_SYNC
        if (threadIdx.x < 32) {                          // warp 0 (block_max left the CTA's maximum in it)
            bool publisher = false;
            if (threadIdx.x == 0) {
                if (r > 0.0f) atomic_max_nonneg(slot, r);
                PUSH_T(0, k, t_min);
                publisher = atom_add_acq_rel_gpu(arrive, 1u) == epoch * gridDim.x - 1u;
            }
            publisher = __shfl_sync(0xffffffffu, publisher, 0);
            if (publisher) {
                const int lane = threadIdx.x;
                const int rs = CTRL_RESID + (int)(epoch & 1u) * 16;
                PUSH_T(1, k, t_set);
                if (lane == 0) {
                    scratch[(k + 1) % 3] = 0.0f;         // last read two sweeps ago; next used after the generation release below
#if MGDP_PUSH_DYN_TILES
                    *tile_ctr = 0u;                      // every CTA has arrived: nobody takes tiles until the release below
#endif
                }
                __syncwarp();                            // lane 0's acquire of every CTA's arrival is ordered before the other lanes' releases
                const uint32_t mine = __float_as_uint(__ldcg(slot));
                if (lane < world) {                      // lane r talks to rank r (its own control block included)
                    uint32_t* c = lane == rank ? ctrl : peer_ctrl.p[lane < rank ? lane : lane - 1];
                    st_relaxed_sys(c + rs + rank, mine);
                    st_release_sys(c + CTRL_FLAGS + rank, epoch);     // sys-scope release: cumulative over every CTA's peer stores
                }
                PUSH_T(2, k, t_set);
                float gm = 0.0f;
                bool late = false;
                if (lane < world) {
                    const unsigned long long t0 = global_ns();
                    while (ld_acquire_sys(ctrl + CTRL_FLAGS + lane) < epoch)
                        if (global_ns() - t0 > 4000000000ull) { late = true; break; }     // a peer is gone: give up, never hang
                    gm = __uint_as_float(ld_relaxed_sys(ctrl + rs + lane));
                }
                PUSH_T(3, k, t_set);
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) gm = fmaxf(gm, __shfl_xor_sync(0xffffffffu, gm, o));
                late = __any_sync(0xffffffffu, late);
                __syncwarp();                            // the other lanes' acquires of the peers' flags are ordered before lane 0's release
                if (lane == 0) {
                    scratch[3] = gm; scratch[4] = late ? 1.0f : 0.0f;
#if MGDP_PUSH_REL_FENCE
                    fence_acq_rel_sys();                 // belt and braces: the peers' V' stores before the release below (the
#endif                                                   // acquire -> __syncwarp -> release chain already orders them)
                    st_release_gpu(gen, epoch);
                    PUSH_T(4, k, t_set);
                }
            }
            if (threadIdx.x == 0) while (ld_acquire_gpu(gen) < epoch) { }
        }
        __syncthreads();
#else
        if (threadIdx.x == 0) {
            if (r > 0.0f) atomic_max_nonneg(slot, r);
            if (blockIdx.x == 0) scratch[(k + 1) % 3] = 0.0f;
        }
        __threadfence_system();                          // this thread's peer stores are performed before the flag below
        PUSH_T(0, k, t_min);
        grid.sync();
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            PUSH_T(1, k, t_set);
            const uint32_t mine = __float_as_uint(__ldcg(slot));
            const int rs = CTRL_RESID + (int)(epoch & 1u) * 16;
            for (int p = 0; p < peer_ctrl.n; ++p) *reinterpret_cast<volatile uint32_t*>(peer_ctrl.p[p] + rs + rank) = mine;
            *reinterpret_cast<volatile uint32_t*>(ctrl + rs + rank) = mine;
            __threadfence_system();
            for (int p = 0; p < peer_ctrl.n; ++p) st_release_sys(peer_ctrl.p[p] + CTRL_FLAGS + rank, epoch);
            st_release_sys(ctrl + CTRL_FLAGS + rank, epoch);
            PUSH_T(2, k, t_set);
            const unsigned long long t0 = global_ns();
            float abort_flag = 0.0f;
            for (int rr = 0; rr < world && abort_flag == 0.0f; ++rr)
                while (ld_acquire_sys(ctrl + CTRL_FLAGS + rr) < epoch)
                    if (global_ns() - t0 > 4000000000ull) { abort_flag = 1.0f; break; }     // a peer is gone: give up, never hang
            PUSH_T(3, k, t_set);
            float gm = 0.0f;
            for (int rr = 0; rr < world; ++rr) gm = fmaxf(gm, __uint_as_float(*reinterpret_cast<volatile uint32_t*>(ctrl + rs + rr)));
            scratch[3] = gm; scratch[4] = abort_flag;
            __threadfence();
            PUSH_T(4, k, t_set);
        }
        grid.sync();
#endif
        g = __ldcg(scratch + 3);
        aborted = __ldcg(scratch + 4) != 0.0f;
        ++k;
        float* tmp = Vin; Vin = Vout; Vout = tmp;
        const PeerOut* tp = pin; pin = pout; pout = tp;
        if (aborted || g < tol) break;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        result[0] = k; result[1] = __float_as_int(g); result[2] = (Vin == Va) ? 0 : 1; result[3] = aborted ? 1 : 0;
    }
}

// ---- exact finite-horizon backward induction ----
// State (s, t): t = step_count.  Acting at t moves the clock to t+1; a success pays the reference's time-dependent
// _reward() = 1 - 0.9*((t+1)/max_steps) (minigrid_env.py:235-240) as three separately rounded fp64 operations, lava
// pays 0, and the episode is truncated once t+1 reaches max_steps (:582-583).  V_T = 0.
//   V_t(s) = max_a  { WIN: R(t+1);  LOSE: 0;  t+1 >= T: 0;  else gamma * V_{t+1}(s') },   lowest action index on ties
__global__ void __launch_bounds__(256) fh_kernel(const int32_t* __restrict__ succ, int64_t S_pad, int A, int64_t S,
                                                 const double* __restrict__ V_next, double* __restrict__ V_out,
                                                 uint8_t* __restrict__ pi_out, double r_win, double gamma, int last)
{
    for (int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; s < S; s += (int64_t)gridDim.x * blockDim.x) {
        double best = 0.0;
        int arg = 0;
        for (int a = 0; a < A; ++a) {
            const int32_t sp = __ldg(succ + (int64_t)a * S_pad + s);
            double q;
            if (sp == MGDP_WIN) q = r_win;
            else if (sp == MGDP_LOSE || last) q = 0.0;
            else q = __dmul_rn(gamma, V_next[sp]);
            if (a == 0 || q > best) { best = q; arg = a; }
        }
        V_out[s] = best;
        if (pi_out) pi_out[s] = (uint8_t)arg;
    }
}

}  // namespace mgdp

// ------------------------------------------------------------------------------------------------ C ABI: K4
extern "C" {

int64_t mgdp_launch_count(void) { return mgdp::g_launches; }

static int sweep_impl(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                      const float* V_in_dev, float* V_out_dev, uint8_t* pi_out_dev, const uint8_t* pi_fixed_dev,
                      float gamma, float* resid_dev, void* stream, const uint8_t* rowmask)
{
    using namespace mgdp;
    if (rowmask && (pi_fixed_dev || (s_begin & 31)))
        return fail(MG_EINVAL, "mgdp_sweep_masked: s_begin must be a multiple of 32 and pi_fixed is not supported");
    if (!succ_dev || !V_in_dev || !V_out_dev || !resid_dev || A < 1 || A > MAXA || s_begin < 0 || s_end < s_begin || s_end > S_pad)
        return fail(MG_EINVAL, "mgdp_sweep: bad argument");
    const int64_t n = s_end - s_begin;
    if (n == 0) return MG_OK;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // persistent grid: one wave of resident CTAs, each walking tiles grid-stride (130 us vs 145 us with a 32-CTA/SM grid)
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = pi_fixed_dev ? ((n + 3) / 4 + SWEEP_THREADS - 1) / SWEEP_THREADS : (n + TILE - 1) / TILE;
    const int64_t cap = (int64_t)sms * MGDP_SWEEP_MINB;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = (cudaStream_t)stream;
#if MGDP_TMA
    if (!rowmask && !pi_fixed_dev && (A == 7 || A == 5 || A == 3)) {
        const size_t smem = (size_t)TMA_STAGES * A * TMA_THREADS * 16;
        static int occ[8] = {0};
        auto setup = [&](auto kern) -> cudaError_t {
            if (!occ[A]) {
                cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return e;
                e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ[A], kern, TMA_THREADS, smem);
                if (e != cudaSuccess) return e;
                if (occ[A] < 1) occ[A] = 1;
            }
            int64_t tiles = ((n + 3) / 4 + TMA_THREADS - 1) / TMA_THREADS, g = (int64_t)sms * occ[A];
            if (g > tiles) g = tiles;
            kern<<<(unsigned)g, TMA_THREADS, smem, st>>>(succ_dev, S_pad, A, s_begin, s_end, V_in_dev, V_out_dev, pi_out_dev, gamma, resid_dev);
            return cudaGetLastError();
        };
        cudaError_t e = A == 7 ? setup(sweep_tma_kernel<7>) : (A == 5 ? setup(sweep_tma_kernel<5>) : setup(sweep_tma_kernel<3>));
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        g_launches++;
        return MG_OK;
    }
#endif
#define MGDP_SWEEP_ARGS succ_dev, S_pad, A, s_begin, s_end, V_in_dev, V_out_dev, pi_out_dev, pi_fixed_dev, gamma, resid_dev, rowmask
#define MGDP_LAUNCH_SWEEP(AT) do { const bool wp = pi_out_dev || pi_fixed_dev; \
        if (rowmask && wp) sweep_kernel<AT, true, true><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(MGDP_SWEEP_ARGS); \
        else if (rowmask) sweep_kernel<AT, true, false><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(MGDP_SWEEP_ARGS); \
        else if (wp) sweep_kernel<AT, false, true><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(MGDP_SWEEP_ARGS); \
        else sweep_kernel<AT, false, false><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(MGDP_SWEEP_ARGS); } while (0)
    switch (A) {
    case 3: MGDP_LAUNCH_SWEEP(3); break;
    case 5: MGDP_LAUNCH_SWEEP(5); break;
    case 7: MGDP_LAUNCH_SWEEP(7); break;
    default: MGDP_LAUNCH_SWEEP(0); break;
    }
#undef MGDP_LAUNCH_SWEEP
#undef MGDP_SWEEP_ARGS
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    return MG_OK;
}

int mgdp_sweep(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
               const float* V_in_dev, float* V_out_dev, uint8_t* pi_out_dev, const uint8_t* pi_fixed_dev,
               float gamma, float* resid_dev, void* stream)
{
    return sweep_impl(succ_dev, S_pad, A, s_begin, s_end, V_in_dev, V_out_dev, pi_out_dev, pi_fixed_dev, gamma, resid_dev, stream, nullptr);
}

int mgdp_sweep_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                      const float* V_in_dev, float* V_out_dev, uint8_t* pi_out_dev, float gamma, float* resid_dev, void* stream)
{
    if (!rowmask_dev) return mgdp::fail(MG_EINVAL, "mgdp_sweep_masked: null row mask (mgdp_rowmask builds it)");
    return sweep_impl(succ_dev, S_pad, A, s_begin, s_end, V_in_dev, V_out_dev, pi_out_dev, nullptr, gamma, resid_dev, stream, rowmask_dev);
}

int mgdp_rowmask(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S, uint8_t* rowmask_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !rowmask_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || (S_pad & 31))
        return fail(MG_EINVAL, "mgdp_rowmask: bad argument (S_pad must be a multiple of 32)");
    const int64_t threads = S_pad;                                   // one warp per 32 states
    rowmask_kernel<<<(unsigned)((threads + 255) / 256), 256, 0, (cudaStream_t)stream>>>(succ_dev, S_pad, A, S, rowmask_dev);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    return MG_OK;
}

// Push only the values that changed (default).  It costs one extra 4-byte read per state; without it every rank receives
// all of V' every sweep, which at 8 GPUs is 59 MB per rank per sweep and bounds the solve (126 us/sweep vs 56 us with the
// filter; neutral at 2 GPUs: profiles/r1_dp_multi_gpu.txt).  MGDP_PUSH_FILTER=0 turns it off for A/B runs.
static int push_filter(int num_peers)
{
    (void)num_peers;
    const char* e = getenv("MGDP_PUSH_FILTER");
    return !(e && e[0] == '0');
}

int mgdp_sweep_push(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                    const float* V_in_dev, float* V_out_dev, float* const* peer_V_out, int32_t num_peers,
                    uint8_t* pi_out_dev, float gamma, float* resid_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !V_in_dev || !V_out_dev || !resid_dev || A < 1 || A > MAXA || s_begin < 0 || s_end < s_begin || s_end > S_pad ||
        num_peers < 0 || num_peers > MGDP_MAX_PEERS || (num_peers > 0 && !peer_V_out))
        return fail(MG_EINVAL, "mgdp_sweep_push: bad argument");
    const int64_t n = s_end - s_begin;
    if (n == 0) return MG_OK;
    PeerOut peers{};
    peers.n = num_peers;
    peers.changed_only = push_filter(num_peers);
    for (int k = 0; k < num_peers; ++k) peers.p[k] = peer_V_out[k];
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = (n + TILE - 1) / TILE;
    const int64_t cap = (int64_t)sms * MGDP_SWEEP_MINB;             // persistent: one wave
    if (blocks > cap) blocks = cap;
    cudaStream_t st = (cudaStream_t)stream;
#define MGDP_LAUNCH_PUSH(AT) sweep_push_kernel<AT><<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(succ_dev, S_pad, A, s_begin, s_end, V_in_dev, V_out_dev, pi_out_dev, gamma, resid_dev, peers)
    switch (A) {
    case 3: MGDP_LAUNCH_PUSH(3); break;
    case 5: MGDP_LAUNCH_PUSH(5); break;
    case 7: MGDP_LAUNCH_PUSH(7); break;
    default: MGDP_LAUNCH_PUSH(0); break;
    }
#undef MGDP_LAUNCH_PUSH
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    return MG_OK;
}

int mgdp_solve_push(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                    float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                    uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                    uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream)
{
    return mgdp_solve_push_mc(succ_dev, S_pad, A, s_begin, s_end, Va_dev, Vb_dev, peer_Va, peer_Vb, nullptr, nullptr, ctrl_dev, peer_ctrl,
                              num_peers, rank, pi_dev, gamma, tol, max_sweeps, scratch_dev, result_dev, stream);
}

static int solve_push_impl(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                           float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                           float* mc_Va, float* mc_Vb,
                           uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                           uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream,
                           const uint8_t* rowmask)
{
    using namespace mgdp;
    if (rowmask && (s_begin & 31)) return fail(MG_EINVAL, "mgdp_solve_push_masked: s_begin must be a multiple of 32");
    if (!succ_dev || !Va_dev || !Vb_dev || !ctrl_dev || !scratch_dev || !result_dev || A < 1 || A > MAXA || s_begin < 0 ||
        s_end < s_begin || s_end > S_pad || (s_begin & 3) || num_peers < 0 || num_peers > MGDP_MAX_PEERS || rank < 0 || rank > num_peers ||
        (num_peers > 0 && (!peer_Va || !peer_Vb || !peer_ctrl)) || max_sweeps < 1)
        return fail(MG_EINVAL, "mgdp_solve_push: bad argument");
    PeerOut pa{}, pb{};
    PeerCtrl pc{};
    pa.n = pb.n = pc.n = num_peers;
    pa.changed_only = pb.changed_only = push_filter(num_peers);
    if (num_peers > 0 && mc_Va && mc_Vb) { pa.mc = mc_Va; pb.mc = mc_Vb; }
    for (int k = 0; k < num_peers; ++k) { pa.p[k] = peer_Va[k]; pb.p[k] = peer_Vb[k]; pc.p[k] = peer_ctrl[k]; }
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const bool mc = pa.mc != nullptr;
    const void* kern;
#define MGDP_PUSH_KERN(AT) (rowmask ? (mc ? (const void*)solve_push_kernel<AT, true, true> : (const void*)solve_push_kernel<AT, false, true>) \
                                    : (mc ? (const void*)solve_push_kernel<AT, true, false> : (const void*)solve_push_kernel<AT, false, false>))
    switch (A) {
    case 3: kern = MGDP_PUSH_KERN(3); break;
    case 5: kern = MGDP_PUSH_KERN(5); break;
    case 7: kern = MGDP_PUSH_KERN(7); break;
    default: kern = MGDP_PUSH_KERN(0); break;
    }
#undef MGDP_PUSH_KERN
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SWEEP_THREADS, 0);
    if (e != cudaSuccess || per_sm < 1) return fail(MG_ECUDA, "mgdp_solve_push: occupancy query failed");
    const int64_t n = s_end - s_begin;
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = (n + TILE - 1) / TILE;
    const int64_t cap = (int64_t)sms * per_sm;               // cooperative launch: every CTA must be resident
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;                              // an empty range still takes part in the barriers
    cudaStream_t st = (cudaStream_t)stream;
    e = cudaMemsetAsync(scratch_dev, 0, 8 * sizeof(float), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(result_dev, 0, 4 * sizeof(int32_t), st);
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    int A_ = A, ms = max_sweeps, world = num_peers + 1, rk = rank;
    void* args[] = {(void*)&succ_dev, (void*)&S_pad, (void*)&A_, (void*)&s_begin, (void*)&s_end, (void*)&Va_dev, (void*)&Vb_dev,
                    (void*)&gamma, (void*)&tol, (void*)&ms, (void*)&pa, (void*)&pb, (void*)&ctrl_dev, (void*)&pc,
                    (void*)&rk, (void*)&world, (void*)&scratch_dev, (void*)&result_dev, (void*)&rowmask};
    e = cudaLaunchCooperativeKernel(kern, dim3((unsigned)blocks), dim3(SWEEP_THREADS), args, 0, st);
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    if (pi_dev && n > 0) {            // policy of the rank's own range, from the input of the last sweep (see extract_kernel)
        e = launch_extract_any(succ_dev, S_pad, A, s_begin, s_end, Va_dev, Vb_dev, result_dev, pi_dev, gamma, sms, st, rowmask);
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        g_launches++;
    }
    return MG_OK;
}

int mgdp_solve_push_mc(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                       float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                       float* mc_Va, float* mc_Vb,
                       uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                       uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream)
{
    return solve_push_impl(succ_dev, S_pad, A, s_begin, s_end, Va_dev, Vb_dev, peer_Va, peer_Vb, mc_Va, mc_Vb, ctrl_dev, peer_ctrl,
                           num_peers, rank, pi_dev, gamma, tol, max_sweeps, scratch_dev, result_dev, stream, nullptr);
}

int mgdp_solve_push_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                           float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                           float* mc_Va, float* mc_Vb,
                           uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                           uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream)
{
    if (!rowmask_dev) return mgdp::fail(MG_EINVAL, "mgdp_solve_push_masked: null row mask (mgdp_rowmask builds it)");
    return solve_push_impl(succ_dev, S_pad, A, s_begin, s_end, Va_dev, Vb_dev, peer_Va, peer_Vb, mc_Va, mc_Vb, ctrl_dev, peer_ctrl,
                           num_peers, rank, pi_dev, gamma, tol, max_sweeps, scratch_dev, result_dev, stream, rowmask_dev);
}

int mgdp_finite_horizon(const int32_t* succ_dev, int64_t S, int32_t A, int64_t S_pad, int32_t max_steps, double gamma,
                        double* V0_dev, double* V_tmp_dev, uint8_t* pi_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !V0_dev || !V_tmp_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || max_steps < 1)
        return fail(MG_EINVAL, "mgdp_finite_horizon: bad argument");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    int64_t blocks = (S + 255) / 256;
    if (blocks > (int64_t)sms * 32) blocks = (int64_t)sms * 32;
    cudaStream_t st = (cudaStream_t)stream;
    // ping-pong so that the last sweep (t = 0) lands in V0_dev
    double* bufs[2] = { V0_dev, V_tmp_dev };
    for (int t = max_steps - 1; t >= 0; --t) {
        double* out = bufs[t & 1];
        const double* next = bufs[(t + 1) & 1];
        volatile double q = (double)(t + 1) / (double)max_steps;          // _reward(): three rounded operations
        volatile double m = 0.9 * q;
        volatile double r = 1.0 - m;
        fh_kernel<<<(unsigned)blocks, 256, 0, st>>>(succ_dev, S_pad, A, S, next, out,
                                                   pi_dev ? pi_dev + (size_t)t * S : nullptr, r, gamma, t + 1 >= max_steps);
        g_launches++;
    }
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    return MG_OK;
}

#ifdef MGDP_TIMING
int mgdp_debug_push_timing(unsigned long long* host /*[5][128]*/, int32_t reset)
{
    if (reset) {
        static unsigned long long init[5][128];
        for (int i = 0; i < 128; ++i) { init[0][i] = ~0ull; init[1][i] = init[2][i] = init[3][i] = init[4][i] = 0; }
        return cudaMemcpyToSymbol(mgdp::g_push_timing, init, sizeof(init)) == cudaSuccess ? 0 : -2;
    }
    return cudaMemcpyFromSymbol(host, mgdp::g_push_timing, sizeof(unsigned long long) * 5 * 128) == cudaSuccess ? 0 : -2;
}
int mgdp_debug_timing(unsigned long long* host /*[3][128]*/, int32_t reset)
{
    if (reset) {
        static unsigned long long init[3][128];
        for (int i = 0; i < 128; ++i) { init[0][i] = ~0ull; init[1][i] = 0; init[2][i] = ~0ull; }
        return cudaMemcpyToSymbol(mgdp::g_timing, init, sizeof(init)) == cudaSuccess ? 0 : -2;
    }
    return cudaMemcpyFromSymbol(host, mgdp::g_timing, sizeof(unsigned long long) * 3 * 128) == cudaSuccess ? 0 : -2;
}
#endif

// the cooperative solve behind mgdp_solve / mgdp_solve_masked / mgdp_solve_static: V_tmp and the three residual accumulators
// (`slots`) are the caller's (mgdp_solve: result_dev + 2) or the library's static scratch
static int solve_impl(const char* who, const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, float* V_tmp_dev, uint8_t* pi_dev, const uint8_t* pi_fixed_dev,
                      float gamma, float tol, int32_t max_sweeps, float* slots, int32_t* result_dev, void* stream, const uint8_t* rowmask)
{
    using namespace mgdp;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    void (*kern)(const int32_t*, int64_t, int, int64_t, float*, float*, uint8_t*, const uint8_t*, float, float, int, float*, int32_t*, const uint8_t*);
    if (pi_fixed_dev) kern = solve_kernel<0, true, false>;
    else if (rowmask) switch (A) {
    case 3: kern = solve_kernel<3, false, true>; break;
    case 5: kern = solve_kernel<5, false, true>; break;
    case 7: kern = solve_kernel<7, false, true>; break;
    default: kern = solve_kernel<0, false, true>; break;
    }
    else switch (A) {
    case 3: kern = solve_kernel<3, false, false>; break;
    case 5: kern = solve_kernel<5, false, false>; break;
    case 7: kern = solve_kernel<7, false, false>; break;
    default: kern = solve_kernel<0, false, false>; break;
    }
    int per_sm = 0;
    cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kern, SWEEP_THREADS, 0);
    if (e != cudaSuccess || per_sm < 1) return fail(MG_ECUDA, "mgdp_solve: occupancy query failed");
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = pi_fixed_dev ? ((S + 3) / 4 + SWEEP_THREADS - 1) / SWEEP_THREADS : (S + TILE - 1) / TILE;
    const int64_t cap = (int64_t)sms * per_sm;               // cooperative launch: every CTA must be resident
    if (blocks > cap) blocks = cap;
    cudaStream_t st = (cudaStream_t)stream;
    e = cudaMemsetAsync(result_dev, 0, 2 * sizeof(int32_t), st);
    if (e == cudaSuccess) e = cudaMemsetAsync(slots, 0, 3 * sizeof(float), st);          // what must start at zero
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    (void)who;
    int A_ = A; int ms = max_sweeps;
    void* args[] = {(void*)&succ_dev, (void*)&S_pad, (void*)&A_, (void*)&S, (void*)&V_dev, (void*)&V_tmp_dev, (void*)&pi_dev,
                    (void*)&pi_fixed_dev, (void*)&gamma, (void*)&tol, (void*)&ms, (void*)&slots, (void*)&result_dev, (void*)&rowmask};
    e = cudaLaunchCooperativeKernel((const void*)kern, dim3((unsigned)blocks), dim3(SWEEP_THREADS), args, 0, st);
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    if (pi_dev && !pi_fixed_dev) {
        e = launch_extract_any(succ_dev, S_pad, A, 0, S, V_dev, V_tmp_dev, result_dev, pi_dev, gamma, sms, st, rowmask);
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        g_launches++;
    }
    final_copy_kernel<<<sms * 4, 256, 0, st>>>(V_dev, V_tmp_dev, result_dev, S);       // newest values -> V_dev when k is odd
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    return MG_OK;
}

int mgdp_solve(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S,
               float* V_dev, float* V_tmp_dev, uint8_t* pi_dev, const uint8_t* pi_fixed_dev,
               float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !V_dev || !V_tmp_dev || !result_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || max_sweeps < 1)
        return fail(MG_EINVAL, "mgdp_solve: bad argument");
    // the three residual slots live right behind the two result words
    return solve_impl("mgdp_solve", succ_dev, S_pad, A, S, V_dev, V_tmp_dev, pi_dev, pi_fixed_dev, gamma, tol, max_sweeps,
                      reinterpret_cast<float*>(result_dev + 2), result_dev, stream, nullptr);
}

int mgdp_solve_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, float* V_tmp_dev, uint8_t* pi_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !rowmask_dev || !V_dev || !V_tmp_dev || !result_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || max_sweeps < 1)
        return fail(MG_EINVAL, "mgdp_solve_masked: bad argument");
    return solve_impl("mgdp_solve_masked", succ_dev, S_pad, A, S, V_dev, V_tmp_dev, pi_dev, nullptr, gamma, tol, max_sweeps,
                      reinterpret_cast<float*>(result_dev + 2), result_dev, stream, rowmask_dev);
}

// mgdp_solve_stream: value iteration as a stream of gated sweep launches (see sweep_gated_kernel).  HOST-SYNCHRONOUS up to
// convergence (the host follows the stop flag one chunk behind the device), then the policy extraction and the final copy
// are enqueued like in mgdp_solve.  Same sweeps, same arithmetic: V, pi, sweep count and residual are bit-identical.
}  // extern "C"
namespace mgdp {
static __device__ float g_stream_ring[STREAM_RING];          // residual ring of the streamed solve (one solve at a time per device)
}
extern "C" {
int mgdp_solve_stream(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, float* V_tmp_dev, uint8_t* pi_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !V_dev || !V_tmp_dev || !result_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || max_sweeps < 1 ||
        (rowmask_dev && (S_pad & 31)))
        return fail(MG_EINVAL, "mgdp_solve_stream: bad argument");
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    void (*kern)(const int32_t*, int64_t, int, int64_t, int64_t, const float*, float*, float, float, int, unsigned int*, float*, int32_t*, const uint8_t*);
#define MGDP_PICK_GATED(M) switch (A) { \
    case 3: kern = sweep_gated_kernel<3, M>; break; \
    case 5: kern = sweep_gated_kernel<5, M>; break; \
    case 7: kern = sweep_gated_kernel<7, M>; break; \
    default: kern = sweep_gated_kernel<0, M>; break; }
    if (rowmask_dev) MGDP_PICK_GATED(true) else MGDP_PICK_GATED(false)
#undef MGDP_PICK_GATED
    constexpr int64_t TILE = (int64_t)MGDP_NS * SWEEP_THREADS;
    int64_t blocks = (S + TILE - 1) / TILE;
    if (blocks > (int64_t)sms * MGDP_SWEEP_MINB) blocks = (int64_t)sms * MGDP_SWEEP_MINB;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned int* ctrl = reinterpret_cast<unsigned int*>(result_dev + 2);
    // per device, made on first use: the ring's address, pinned landing words for the stop flag and the events that say they
    // landed (not thread-safe: one streamed solve at a time per device)
    static float* rings[64] = {nullptr};
    static unsigned int* h_flags[64] = {nullptr};
    static cudaEvent_t evs[64][2];
    if (dev < 0 || dev >= 64) return fail(MG_EINVAL, "mgdp_solve_stream: device ordinal out of range");
    cudaError_t e = cudaSuccess;
    if (!h_flags[dev]) {
        unsigned int* h = nullptr;
        void* ring_p = nullptr;
        e = cudaGetSymbolAddress(&ring_p, g_stream_ring);
        if (e == cudaSuccess) e = cudaHostAlloc(reinterpret_cast<void**>(&h), 2 * sizeof(unsigned int), cudaHostAllocDefault);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&evs[dev][0], cudaEventDisableTiming);
        if (e == cudaSuccess) e = cudaEventCreateWithFlags(&evs[dev][1], cudaEventDisableTiming);
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        rings[dev] = static_cast<float*>(ring_p);
        h_flags[dev] = h;
    }
    float* ring = rings[dev];
    unsigned int* h_flag = h_flags[dev];
    cudaEvent_t* ev = evs[dev];
    e = cudaMemsetAsync(result_dev, 0, 5 * sizeof(int32_t), st);                 // what must start at zero: result, stop flag,
    if (e == cudaSuccess) e = cudaMemsetAsync(ring, 0, STREAM_RING * sizeof(float), st);   // and the residual ring
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    static_assert(STREAM_RING % STREAM_CHUNK == 0 && STREAM_RING >= 4 * STREAM_CHUNK, "ring = a whole number (>= 4) of chunks");
    int A_ = A;
    int k = 0;
    for (int c = 0;; ++c) {
        // the slots of this chunk were last used STREAM_RING sweeps ago and last read by the sweep after that one
        if (k >= STREAM_RING) {
            e = cudaMemsetAsync(ring + (k & (STREAM_RING - 1)), 0, STREAM_CHUNK * sizeof(float), st);
            if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        }
        const int k_end = (int)std::min<int64_t>((int64_t)k + STREAM_CHUNK, max_sweeps);
        for (; k < k_end; ++k) {
            const float* vin = (k & 1) ? V_tmp_dev : V_dev;                      // sweep k reads V_k: in V_dev for even k
            float* vout = (k & 1) ? V_dev : V_tmp_dev;
            kern<<<(unsigned)blocks, SWEEP_THREADS, 0, st>>>(succ_dev, S_pad, A_, (int64_t)0, S, vin, vout, gamma, tol, k, ctrl, ring, result_dev, rowmask_dev);
            e = cudaGetLastError();
            if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
            g_launches++;
        }
        e = cudaMemcpyAsync(h_flag + (c & 1), ctrl, sizeof(unsigned int), cudaMemcpyDeviceToHost, st);
        if (e == cudaSuccess) e = cudaEventRecord(ev[c & 1], st);
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        if (c > 0) {                                     // the PREVIOUS chunk's flag, while this chunk runs
            e = cudaEventSynchronize(ev[(c - 1) & 1]);
            if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
            if (h_flag[(c - 1) & 1]) break;              // converged: what is enqueued behind it returns at once
        }
        if (k >= max_sweeps) break;                      // nothing left to enqueue
    }
    stream_finish_kernel<<<1, 1, 0, st>>>(k, ctrl, ring, result_dev);           // all k sweeps ran and none was followed by a gate
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    if (pi_dev) {
        e = launch_extract_any(succ_dev, S_pad, A, 0, S, V_dev, V_tmp_dev, result_dev, pi_dev, gamma, sms, st, rowmask_dev);
        if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
        g_launches++;
    }
    final_copy_kernel<<<sms * 4, 256, 0, st>>>(V_dev, V_tmp_dev, result_dev, S);       // newest values -> V_dev when k is odd
    e = cudaGetLastError();
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    g_launches++;
    return MG_OK;
}

}  // extern "C"

// ------------------------------------------------------------------------------------------------ no-allocation solve
// mgdp_solve_static: mgdp_solve for callers that cannot allocate (code under CUDA graph capture, tight serving loops,
// harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The scratch mgdp_solve takes from the caller — the second V buffer of the Jacobi ping-pong and the three rotating residual
// accumulators — lives in ONE file-scope device array of a fixed size, so the call allocates nothing (no cudaMalloc / cudaFree,
// no allocating library routine).  Every GPU that loads the library holds the array (4 MiB + 256 B) until the process exits.
// Every call on a device shares it and it keeps the last call's contents: the accumulators are cleared at the start of each
// call (V_tmp needs no clearing: sweep 1 writes every state before anything reads it), and two calls that OVERLAP on the same
// GPU (different streams, no dependency between them) race on it — serialise them.
namespace mgdp {

constexpr int64_t STATIC_MAX_STATES = MGDP_STATIC_MAX_STATES;
constexpr int64_t STATIC_SLOTS = STATIC_MAX_STATES;                      // float index of the three residual accumulators
static __device__ __align__(256) float g_static_scratch[STATIC_MAX_STATES + 64];

// In host code `g_static_scratch` is only a stand-in for the symbol, not a device address: the address is looked up once per
// device, with that device current, and cached.
static float* static_scratch_ptr(int dev, cudaError_t* err)
{
    static float* ptrs[64] = {nullptr};
    *err = cudaSuccess;
    if (dev >= 0 && dev < 64 && ptrs[dev]) return ptrs[dev];
    float* p = nullptr;
    *err = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch);
    if (*err != cudaSuccess) return nullptr;
    if (dev >= 0 && dev < 64) ptrs[dev] = p;
    return p;
}

}  // namespace mgdp

extern "C" {

int64_t mgdp_static_max_states(void) { return mgdp::STATIC_MAX_STATES; }

int mgdp_solve_static(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, uint8_t* pi_dev, const uint8_t* pi_fixed_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream)
{
    using namespace mgdp;
    if (!succ_dev || !V_dev || !result_dev || A < 1 || A > MAXA || S <= 0 || S > S_pad || max_sweeps < 1)
        return fail(MG_EINVAL, "mgdp_solve_static: bad argument");
    if (S > STATIC_MAX_STATES) {
        static thread_local char msg[160];
        snprintf(msg, sizeof(msg), "mgdp_solve_static: %lld states exceed the static scratch (MGDP_STATIC_MAX_STATES = %lld); use mgdp_solve "
                 "with caller-owned scratch", (long long)S, (long long)STATIC_MAX_STATES);
        return fail(MG_EINVAL, msg);
    }
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(MG_ECUDA, cudaGetErrorString(e));
    float* scratch = static_scratch_ptr(dev, &e);
    if (e != cudaSuccess || !scratch) return fail(MG_ECUDA, "mgdp_solve_static: cudaGetSymbolAddress failed for the static scratch");
    return solve_impl("mgdp_solve_static", succ_dev, S_pad, A, S, V_dev, scratch, pi_dev, pi_fixed_dev, gamma, tol, max_sweeps,
                      scratch + STATIC_SLOTS, result_dev, stream, nullptr);
}

}  // extern "C"
