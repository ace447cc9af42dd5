/*
 * mgb200_dp.h — C ABI of the dynamic-programming path (part of libmgb200.so).
 *
 * The reference ships no DP code: the MDP is the one MiniGridEnv.step defines
 * (minigrid/minigrid_env.py:520-590, envs/keycorridor.py:129-137) with step_count dropped from the state.
 * These entry points enumerate that MDP on the GPU with the step kernel's own transition function and
 * run Bellman sweeps over the resulting successor table.  Conventions:
 *
 *   state ids   serial-FIFO-BFS discovery order from the start state (id 0), actions in the given order
 *   succ        int32, ACTION-MAJOR: succ[a * S_pad + s]  (coalesced for one thread per state);
 *               terminal transitions hold sentinels MGDP_WIN (-1, reward 1) / MGDP_LOSE (-2, reward 0)
 *   backup      fp32 Jacobi:  V'[s] = max_a q(s,a),  q = 1 | 0 | gamma * V[succ]   (one mul, no FMA)
 *   stop        after the first sweep with max|V'-V| < tol;  pi[s] = lowest action index attaining the max
 *
 * Pointers marked dev are device memory owned by the caller; stream is a cudaStream_t as void*.
 * Errors and mg_last_error() as in mgb200.h.
 */
#ifndef MGB200_DP_H
#define MGB200_DP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MGDP_WIN   (-1)
#define MGDP_LOSE  (-2)
#define MGDP_MAX_ACTIONS 8

/* ---- K4: one Bellman sweep over states [s_begin, s_end) ----
 * V_in is indexed by GLOBAL state id (length >= max successor id + 1); V_out / pi_out are indexed by global id too
 * (so a rank that owns [s_begin, s_end) writes its slice of the full-length buffer that is then all-gathered).
 * pi_fixed != NULL: policy evaluation, q(s, pi_fixed[s]).  pi_out may be NULL.
 * resid_dev: one float, max|V'-V| over the range is atomically max-ed into it (caller zeroes it).  Asynchronous. */
int mgdp_sweep(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
               const float* V_in_dev, float* V_out_dev, uint8_t* pi_out_dev, const uint8_t* pi_fixed_dev,
               float gamma, float* resid_dev, void* stream);

/* ---- K4 driver: sweeps until max|V'-V| < tol, entirely on the device (one cooperative launch, grid-wide
 * barrier between sweeps, no host round trip per sweep).
 * V_dev: in = start values (zeros), out = result; V_tmp_dev: scratch of the same length S.
 * result_dev: int32[5]; on completion [0] = sweeps performed, [1] = float bits of the last residual
 * ([2..4] are scratch residual accumulators).  Asynchronous. */
int mgdp_solve(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S,
               float* V_dev, float* V_tmp_dev, uint8_t* pi_dev, const uint8_t* pi_fixed_dev,
               float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream);

/* ---- K4 with a row mask: skip the action rows that are all self-loops (the sweep is an HBM stream of successor ids) ----
 * pickup / drop / toggle / done are no-ops in most states (`done` in all of them: minigrid_env.py:576-577), i.e. succ[a][s] == s.
 * mgdp_rowmask writes one byte per 32 consecutive states: bit a set = action row a holds at least one non-self successor among
 * them.  The masked sweep / solve read a row's 128-byte line only where the bit is set (a warp covers exactly those 32 states, so
 * the test is warp-uniform) and use the state's own value otherwise: 20-27 bytes per backup instead of 36 on the BASELINE tables,
 * identical arithmetic, bit-identical V, pi and sweep counts.  rowmask_dev: dev uint8[S_pad / 32]; S_pad % 32 == 0;
 * s_begin % 32 == 0.  Policy evaluation (pi_fixed) reads one row per state anyway and has no masked form. */
int mgdp_rowmask(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S, uint8_t* rowmask_dev, void* stream);
int mgdp_sweep_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                      const float* V_in_dev, float* V_out_dev, uint8_t* pi_out_dev, float gamma, float* resid_dev, void* stream);
int mgdp_solve_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, float* V_tmp_dev, uint8_t* pi_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream);

/* ---- K4 driver as a stream of launches: one gated launch per sweep instead of one cooperative launch for all of them ----
 * For tables that do not fit L2 (the HBM regime: KeyCorridorS6R3's 16.8 M states), where the cooperative solve pays ~9 us per
 * sweep for its static tile split and the grid barrier.  The convergence test stays on the device (the last CTA of a sweep to
 * finish reads the residual and raises a stop flag; the launches queued behind it return at once); consecutive launches are
 * chained with programmatic dependent launch; the host enqueues chunks of 8 sweeps and reads the flag one chunk late.
 * HOST-SYNCHRONOUS until convergence; the policy extraction and the final copy are then enqueued as in mgdp_solve.
 * rowmask_dev: NULL or mgdp_rowmask's output.  Value iteration only (no pi_fixed).  result_dev: int32[5], [0] = sweeps
 * performed, [1] = float bits of the last residual, [2..4] scratch.  V, pi, sweeps and residual are bit-identical to mgdp_solve.
 * One call at a time per device (a pinned flag word per device is shared by the calls). */
int mgdp_solve_stream(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, float* V_tmp_dev, uint8_t* pi_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream);

/* ---- K4 driver without allocations (NO counterpart in the reference) ----
 * mgdp_solve for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid it): the scratch mgdp_solve takes from the caller (V_tmp and the residual accumulators) is a file-scope device array
 * inside the library, so this call allocates nothing — no cudaMalloc / cudaFree, no allocating library routine.
 * Limit: S <= MGDP_STATIC_MAX_STATES (1,048,576 states; the array is 4 MiB + 256 B, held by every GPU that loads the library
 * until the process exits); larger S -> MG_EINVAL with a message naming the limit.
 * The array is shared by every call on a device and keeps the last call's contents; what must start at zero is cleared by each
 * call.  Calls that overlap on the same GPU (different streams without a dependency) RACE on it: serialise them.
 * Arguments as mgdp_solve, minus V_tmp_dev; result_dev: dev int32[2] = sweeps performed, float bits of the last residual.
 * Results are bit-identical to mgdp_solve.  Asynchronous. */
#define MGDP_STATIC_MAX_STATES (1 << 20)
int mgdp_solve_static(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t S,
                      float* V_dev, uint8_t* pi_dev, const uint8_t* pi_fixed_dev,
                      float gamma, float tol, int32_t max_sweeps, int32_t* result_dev, void* stream);
int64_t mgdp_static_max_states(void);

/* ---- fused multi-GPU sweep: backup + exchange in ONE kernel ----
 * Like mgdp_sweep on the state range [s_begin, s_end) this rank owns, but every V'[s] is ALSO stored into the V' buffer
 * of each peer rank (peer_V_out: host array of num_peers device pointers to the peers' full-length V' buffers, mapped
 * into this process — CUDA IPC / symmetric memory; the stores travel over NVLink while the sweep is still computing).
 * After the kernel, one small all-reduce(MAX) of the residual on the same stream both decides convergence and orders
 * the peers' stores before the next sweep reads them; no all-gather of V' is needed.  s_begin % 4 == 0. */
#define MGDP_MAX_PEERS 15
int  mgdp_sweep_push(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                     const float* V_in_dev, float* V_out_dev, float* const* peer_V_out, int32_t num_peers,
                     uint8_t* pi_out_dev, float gamma, float* resid_dev, void* stream);

/* ---- the whole state-sharded solve in ONE cooperative launch per GPU ----
 * Every sweep: back up [s_begin, s_end) storing V' locally and into every peer's buffer, then a cross-GPU barrier built
 * from epoch flags in symmetric memory (each rank publishes its residual + flag into every peer's control block and
 * waits, at most 4 s, for the peers' flags) — no host round trip and no NCCL call per sweep.
 *   Va/Vb        this rank's two full-length V buffers (symmetric memory), peer_Va/peer_Vb the peers' mappings of theirs;
 *                all zero and barrier-ed before the launch
 *   ctrl         dev uint32[48] control block in symmetric memory (zeroed + barrier-ed before EVERY launch), peer_ctrl
 *                the peers' control blocks; rank = this rank's index in 0..num_peers
 *   scratch      dev float[8] (local), result dev int32[4]: sweeps, residual bits, 0/1 = final V is in Va/Vb, 1 = aborted
 * Every rank of the group must launch it (a missing peer makes the others give up after the timeout; nothing hangs). */
int  mgdp_solve_push(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                     float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                     uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                     uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream);
/* The same with NVLS multicast: mc_Va / mc_Vb are multicast mappings of the two V buffers (one address = every rank's copy,
 * e.g. torch symmetric memory's multicast_ptr); each changed value then leaves the GPU once (multimem.st) and the NVSwitch
 * replicates it, instead of num_peers unicast stores.  NULL mc pointers = mgdp_solve_push. */
int  mgdp_solve_push_mc(const int32_t* succ_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                        float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                        float* mc_Va, float* mc_Vb,
                        uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                        uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream);

/* mgdp_solve_push_mc with a row mask (mgdp_rowmask): the rank's sweep skips the all-self-loop rows; same results. */
int  mgdp_solve_push_masked(const int32_t* succ_dev, const uint8_t* rowmask_dev, int64_t S_pad, int32_t A, int64_t s_begin, int64_t s_end,
                            float* Va_dev, float* Vb_dev, float* const* peer_Va, float* const* peer_Vb,
                            float* mc_Va, float* mc_Vb,
                            uint32_t* ctrl_dev, uint32_t* const* peer_ctrl, int32_t num_peers, int32_t rank,
                            uint8_t* pi_dev, float gamma, float tol, int32_t max_sweeps, float* scratch_dev, int32_t* result_dev, void* stream);

/* ---- exact finite-horizon DP ----
 * Backward induction over the step counter with the reference's true, time-dependent reward: acting at step_count t
 * pays _reward() = 1 - 0.9*((t+1)/max_steps) on success (minigrid_env.py:235-240, 549-551), 0 on lava, and the episode is
 * truncated when t+1 reaches max_steps (:582-583).  V_t(s) = max_a q, lowest action index on ties, fp64; gamma = 1 gives
 * the returns the reference's episodes actually pay.  max_steps kernel launches.
 *   V0_dev    dev double[S]: V_0 on return        V_tmp_dev  dev double[S] scratch
 *   pi_dev    NULL or dev uint8[max_steps][S]: the time-dependent greedy policy pi_t(s) (action INDEX) */
int  mgdp_finite_horizon(const int32_t* succ_dev, int64_t S, int32_t A, int64_t S_pad, int32_t max_steps, double gamma,
                         double* V0_dev, double* V_tmp_dev, uint8_t* pi_dev, void* stream);

/* ---- K3b: enumerate the reachable state space of ONE layout on the GPU ----
 * cells: host uint8[4][W*H] planes of the start layout; (ax, ay, adir) start pose; rule = MG_RULE_*;
 * actions: host int32[A].  Up to 20 movable objects (keys, balls, boxes; a box may hide a key, a ball or nothing), 32 doors,
 * 4094 cells, as long as the state fits a 256-bit key (agent cell + dir + carried object + a position per movable + 2 bits per
 * door + 1 bit per movable; ObstructedMaze-Full and Playground do); rule = any MG_RULE_* but MG_RULE_DYNOBS (else MG_EINVAL).
 * More reachable states than max_states -> MG_ENOMEM.  On success *out is a handle owning the device tables. */
typedef struct mgdp_mdp mgdp_mdp;
int  mgdp_enumerate(mgdp_mdp** out, int32_t device, const uint8_t* cells, int32_t W, int32_t H,
                    int32_t ax, int32_t ay, int32_t adir, int32_t rule,
                    const int32_t* actions, int32_t A, int64_t max_states);
/* ---- K3b, batched: B layouts of the SAME env class (same W, H, rule, action set) in one level-synchronous BFS — five kernel
 * launches per BFS level whatever B is (thousands of DoorKey-16x16 layouts).
 * cells: host uint8[B][4][W*H]; agents: host int32[B][3] = x, y, dir; max_states bounds the TOTAL.
 * The result is ONE table: layout l owns the contiguous ids [offsets[l], offsets[l+1]) (mgdp_layout_offsets) in that
 * layout's own serial-BFS order — the rows of layout l equal mgdp_enumerate(layout l) with offsets[l] added to the
 * successor ids — so one mgdp_solve / mgdp_sweep covers the whole batch. */
int  mgdp_enumerate_batch(mgdp_mdp** out, int32_t device, int32_t B, const uint8_t* cells, int32_t W, int32_t H,
                          const int32_t* agents, int32_t rule, const int32_t* actions, int32_t A, int64_t max_states);
int32_t mgdp_num_layouts(const mgdp_mdp* m);
int32_t mgdp_key_words(const mgdp_mdp* m);             /* 64-bit words of the state key this table was enumerated with */
int  mgdp_layout_offsets(const mgdp_mdp* m, int64_t* offsets_host /* [B + 1] */);
void mgdp_free(mgdp_mdp* m);
int64_t mgdp_num_states(const mgdp_mdp* m);
int64_t mgdp_padded_states(const mgdp_mdp* m);         /* S_pad of the succ table */
int32_t mgdp_num_levels(const mgdp_mdp* m);            /* BFS depth reached */
const int32_t* mgdp_succ_dev(const mgdp_mdp* m);       /* dev int32[A][S_pad] */
/* per-state description, dev uint8[S][8] = x, y, dir, carry type, carry color, carry state, carry aux, #modified cells */
const uint8_t* mgdp_info_dev(const mgdp_mdp* m);
/* copies the tables into caller-owned device buffers (int32[A][S_pad], uint8[S][8]; either may be NULL); synchronises */
int mgdp_copy_out(const mgdp_mdp* m, int32_t* succ_dst_dev, uint8_t* info_dst_dev, void* stream);
int64_t mgdp_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* MGB200_DP_H */
