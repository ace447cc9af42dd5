// mg_engine.cuh — the engine handle behind the C ABI (include/mgb200.h), shared by mg_capi.cu and mg_render.cu.
#pragma once
#include "mg_step.cuh"

struct mg_atlas { int tile_size; int tile_pitch; uint8_t* dev; };   // one rendered tile atlas (mg_render.cu)

struct mg_engine {
    int device = 0, N = 0, C = 0, P = 0, plane_stride = 0, packed_stride = 0, rows_bytes = 0, autoreset = 0;
    mg::ClassTableDev classes{};
    uint8_t* cells = nullptr;
    uint8_t* packed = nullptr;
    mg::AgentState* agents = nullptr;
    uint8_t* pool_cells = nullptr;
    uint8_t* pool_packed = nullptr;
    mg::LayoutAgent* pool_agents = nullptr;
    uint8_t* dyn = nullptr;                                   // [N][64] DynamicObstacles records (mg_upload_dyn)
    uint8_t* pool_dyn = nullptr;                              // [P][64]
    unsigned long long* error_word = nullptr;
    uint32_t* chain = nullptr;                                // [2][number of step CTAs]: per-CTA launch chaining (StepParams::chain)
    // staging for the host-buffer entry point
    cudaStream_t hstream = nullptr, hstream2 = nullptr;       // two streams: chunk k's copies overlap chunk k+1's kernel
    cudaEvent_t hev[3] = {nullptr, nullptr, nullptr};         // [0] actions uploaded, [1], [2] last kernel of stream 0 / 1
    uint8_t* stage = nullptr; size_t stage_bytes = 0;
    uint32_t no_death_mask = 0; double death_cost = 0.0;      // mg_set_no_death
    int64_t launches = 0;
    // tile atlases built so far, one per tile size asked for (mg_render.cu)
    mg_atlas atlases[8] = {};
    int num_atlases = 0;
};

namespace mg { int set_error(int code, const char* msg); }           // fills mg_last_error(); defined in mg_capi.cu
