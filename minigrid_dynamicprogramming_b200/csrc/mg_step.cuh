// mg_step.cuh — launch parameters of the step / reset kernels (host <-> device contract).
#pragma once
#include "mg_packed.cuh"

namespace mg {

#ifndef MG_STEP_THREADS
#define MG_STEP_THREADS 32      // one warp per CTA: 65,536 envs -> 2,048 CTAs, 14 resident per SM at 8x8 (15,616 B of
#endif                          // shared memory each) = ONE wave on 148 SMs, and no CTA-wide barriers
#ifndef MG_STEP_MINBLOCKS
#define MG_STEP_MINBLOCKS 14
#endif
constexpr int STEP_THREADS = MG_STEP_THREADS;

struct LayoutAgent { uint8_t x, y, dir, cls; };

// Class table as the kernels see it: constructor constants + packed-tile geometry.
struct ClassDev {
    int32_t W, H, max_steps, see_through, rule;
    int32_t PR, PC, offC;            // packed tile: row pitch, column pitch, byte offset of the column-major copy
};
struct ClassTableDev { ClassDev c[MAX_CLASSES]; };

struct StepParams {
    // env state (engine-owned, HBM)
    uint8_t* cells;                 // [N][4][plane_stride]   authoritative SoA planes
    uint8_t* packed;                // [N][packed_stride]     derived packed tiles (mg_packed.cuh)
    AgentState* agents;             // [N]
    uint8_t* dyn;                   // [N][64]  DynamicObstacles records (mg_dyn_record; NULL unless mg_upload_dyn was called)
    const uint8_t* pool_dyn;        // [P][64]
    // layout pool (engine-owned)
    const uint8_t* pool_cells;      // [P][4][plane_stride]
    const uint8_t* pool_packed;     // [P][packed_stride]
    const LayoutAgent* pool_agents; // [P]
    // inputs / outputs (caller-owned); any output may be NULL
    const uint8_t* actions;         // [T][N]; NULL = write observations only
    uint8_t* image;                 // [T][N][147]
    uint8_t* direction;             // [T][N]
    double* reward;                 // [T][N]
    uint8_t* terminated;            // [T][N]
    uint8_t* truncated;             // [T][N]
    uint8_t* final_image;           // [T][N][147] or NULL: last observation of an episode that ended (and was reset) in that step
    uint8_t* final_direction;       // [T][N] or NULL
    unsigned long long* error_word; // first invalid action: bit63 | action<<40 | env  (bit62 | env: mg_reset got a layout id outside the pool;
                                    // bit61 | cta: a step CTA waited ~2 s for its predecessor, see `chain`)
    uint32_t* chain;                // [2][ceil(N / STEP_THREADS)] tickets taken / launches finished per CTA index (engine-owned, zeroed at
                                    // creation): consecutive step launches are ordered CTA by CTA, not grid by grid (mg_step.cu); NULL = grid-wide
    int N, T, P, plane_stride, packed_stride, autoreset, num_classes;
    int image_bulk;                 // image + N * 147 are 16-byte aligned: whole-CTA observation blocks may leave as bulk copies
    int reset_stride;               // autoreset: next pool slot = (layout_id + reset_stride) % P  (reset_stride_for)
    int rows_bytes;                 // max over classes of tile_rows_bytes(): what the ROWS layout stages per env
    // NoDeath (minigrid/wrappers.py:799-872): bit t set = a termination in / into a cell of type t is cancelled and costs
    // death_cost instead; 0 = off
    uint32_t no_death_mask;
    double death_cost;
    ClassTableDev classes;
};

struct ResetParams {
    uint8_t* cells;
    uint8_t* packed;
    AgentState* agents;
    const uint8_t* pool_cells;
    const uint8_t* pool_packed;
    const LayoutAgent* pool_agents;
    const int32_t* layout_ids;      // [N]
    const uint8_t* mask;            // [N] or NULL
    int N, P, plane_stride, packed_stride;
    uint8_t* dyn;                   // [N][64] or NULL
    const uint8_t* pool_dyn;        // [P][64] or NULL
    unsigned long long* error_word;
};

// Autoreset walks the pool with a stride coprime to P: N mod P (rank r of a sharded batch owns slots (rN + e) mod P, so N
// is the stride that keeps the shards' sequences disjoint) moved up to the next value with gcd(stride, P) == 1 — every env
// sees all P layouts before one repeats, also when P divides N.
inline int reset_stride_for(int N, int P)
{
    if (P <= 1) return 0;
    auto gcd = [](int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; };
    int s = N % P;
    if (s == 0) s = 1;
    while (gcd(s, P) != 1) ++s;
    return s % P;
}

// (re)build packed tiles from SoA planes: n records; cls[i*cls_stride] is the class id of record i
struct PackParams {
    const uint8_t* cells;           // [n][4][plane_stride]
    uint8_t* packed;                // [n][packed_stride]
    const uint8_t* cls;             // class id bytes
    int cls_stride;                 // 4 for LayoutAgent arrays (offset 3), 16 for AgentState arrays (offset 3)
    int n, plane_stride, packed_stride;
    ClassTableDev classes;
};

cudaError_t launch_step_many(const StepParams& p, cudaStream_t stream);
cudaError_t launch_reset(const ResetParams& p, cudaStream_t stream);
cudaError_t launch_pack(const PackParams& p, cudaStream_t stream);
size_t step_smem_bytes(int packed_stride, int num_classes, int rows_bytes, int layout);
int step_layout(int packed_stride, int num_classes, int rows_bytes);
cudaError_t step_occupancy(int packed_stride, int num_classes, int rows_bytes, int* blocks_per_sm, int* threads, size_t* smem);

}  // namespace mg
