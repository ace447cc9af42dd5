// mg_step.cuh — launch parameters of the step / reset kernels (host <-> device contract).
#pragma once
#include "mg_common.cuh"

namespace mg {

constexpr int STEP_THREADS = 128;

struct LayoutAgent { uint8_t x, y, dir, cls; };

struct StepParams {
    // env state (engine-owned, HBM)
    uint8_t* cells;                 // [N][4][plane_stride]
    AgentState* agents;             // [N]
    // layout pool (engine-owned)
    const uint8_t* pool_cells;      // [P][4][plane_stride]
    const LayoutAgent* pool_agents; // [P]
    // inputs / outputs (caller-owned); any output may be NULL
    const uint8_t* actions;         // [T][N]; NULL = write observations only
    uint8_t* image;                 // [T][N][147]
    uint8_t* direction;             // [T][N]
    double* reward;                 // [T][N]
    uint8_t* terminated;            // [T][N]
    uint8_t* truncated;             // [T][N]
    unsigned long long* error_word; // first invalid action: bit63 | action<<40 | env
    int N, T, P, plane_stride, autoreset;
    ClassTable classes;
};

struct ResetParams {
    uint8_t* cells;
    AgentState* agents;
    const uint8_t* pool_cells;
    const LayoutAgent* pool_agents;
    const int32_t* layout_ids;      // [N]
    const uint8_t* mask;            // [N] or NULL
    int N, P, plane_stride;
};

cudaError_t launch_step_many(const StepParams& p, cudaStream_t stream);
cudaError_t launch_reset(const ResetParams& p, cudaStream_t stream);

}  // namespace mg
