// mg_capi.cu — the C ABI declared in include/mgb200.h: handle lifetime, layout pool, reset, step.
#include "../../include/mgb200.h"
#include "mg_engine.cuh"

#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, const char* a = "", const char* b = "")
{
    snprintf(g_err, sizeof(g_err), fmt, a, b);
    return code;
}

#define MG_CUDA(call)                                                                     \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess) return fail(MG_ECUDA, "%s: %s", #call, cudaGetErrorString(_e)); \
    } while (0)

// every cell must be representable in the packed code (mg_packed.cuh)
int validate_cells(const uint8_t* cells, int ps, int ncell)
{
    for (int k = 0; k < ncell; ++k)
        if (!mg::code_representable(cells[k], cells[ps + k], cells[2 * ps + k]) || ((cells[3 * ps + k] & 0x0f) > 10))
            return fail(MG_EINVAL, "cell not representable: type 1..10, color 0..7, state 0 (doors 0..2)");
    return MG_OK;
}

}  // namespace

namespace mg { int set_error(int code, const char* msg) { return fail(code, "%s", msg); } }   // used by mg_dp.cu


static_assert(sizeof(mg::AgentState) == sizeof(mg_agent_state), "agent state ABI");
static_assert(sizeof(mg::LayoutAgent) == sizeof(mg_layout_agent), "layout agent ABI");
static_assert(sizeof(mg_dyn_record) == 64, "dyn record ABI");

extern "C" {

const char* mg_last_error(void) { return g_err; }
int mg_version(void) { return 100; }

int mg_create(mg_engine** out, const mg_config* cfg)
{
    if (!out || !cfg || !cfg->classes) return fail(MG_EINVAL, "mg_create: null argument");
    if (cfg->num_envs <= 0) return fail(MG_EINVAL, "mg_create: num_envs must be positive");
    if (cfg->num_classes <= 0 || cfg->num_classes > MG_MAX_CLASSES) return fail(MG_EINVAL, "mg_create: 1..16 classes");
    if (cfg->autoreset < MG_AUTORESET_OFF || cfg->autoreset > MG_AUTORESET_NEXT_STEP) return fail(MG_EINVAL, "mg_create: autoreset must be MG_AUTORESET_*");
    mg_engine* h = new (std::nothrow) mg_engine();
    if (!h) return fail(MG_ENOMEM, "mg_create: out of host memory");
    h->device = cfg->device; h->N = cfg->num_envs; h->C = cfg->num_classes; h->autoreset = cfg->autoreset;
    int maxcells = 0;
    for (int i = 0; i < h->C; ++i) {
        const mg_class_desc& c = cfg->classes[i];
        if (c.width < 3 || c.height < 3 || c.width > 255 || c.height > 255 || c.max_steps <= 0 || c.max_steps > 65535) {
            delete h; return fail(MG_EINVAL, "mg_create: class out of range (3..255 cells, max_steps 1..65535)");
        }
        const mg::TileGeom g = mg::tile_geom(c.width, c.height);
        h->classes.c[i] = mg::ClassDev{c.width, c.height, c.max_steps, c.see_through_walls, c.rule, g.PR, g.PC, g.offC};
        if (c.width * c.height > maxcells) maxcells = c.width * c.height;
        if (g.bytes > h->packed_stride) h->packed_stride = g.bytes;
        if (mg::tile_rows_bytes(g) > h->rows_bytes) h->rows_bytes = mg::tile_rows_bytes(g);
    }
    h->plane_stride = (maxcells + 15) & ~15;
    if (mg::step_smem_bytes(h->packed_stride, h->C, h->rows_bytes, mg::step_layout(h->packed_stride, h->C, h->rows_bytes)) > 227 * 1024) {
        delete h; return fail(MG_EINVAL, "mg_create: grid too large for the shared-memory tile (max about 80x80 cells)");
    }
    cudaError_t e = cudaSetDevice(h->device);
    if (e == cudaSuccess) e = cudaMalloc(&h->cells, (size_t)h->N * 4 * h->plane_stride);
    if (e == cudaSuccess) e = cudaMalloc(&h->packed, (size_t)h->N * h->packed_stride);
    if (e == cudaSuccess) e = cudaMalloc(&h->agents, (size_t)h->N * sizeof(mg::AgentState));
    if (e == cudaSuccess) e = cudaMalloc(&h->error_word, sizeof(unsigned long long));
    if (e == cudaSuccess) e = cudaMemset(h->error_word, 0, sizeof(unsigned long long));
    const size_t nchain = 2 * (((size_t)h->N + mg::STEP_THREADS - 1) / mg::STEP_THREADS) * sizeof(uint32_t);
    if (e == cudaSuccess) e = cudaMalloc(&h->chain, nchain);
    if (e == cudaSuccess) e = cudaMemset(h->chain, 0, nchain);
    if (const char* cs = getenv("MG_CHAIN_START")) {          // test hook: start the chaining counters near their 2^32 wrap
        const std::vector<uint32_t> v(nchain / sizeof(uint32_t), (uint32_t)strtoul(cs, nullptr, 0));
        if (e == cudaSuccess) e = cudaMemcpy(h->chain, v.data(), nchain, cudaMemcpyHostToDevice);
    }
    if (e == cudaSuccess) e = cudaMemset(h->agents, 0, (size_t)h->N * sizeof(mg::AgentState));
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->hstream, cudaStreamNonBlocking);
    if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&h->hstream2, cudaStreamNonBlocking);
    for (int i = 0; i < 3 && e == cudaSuccess; ++i) e = cudaEventCreateWithFlags(&h->hev[i], cudaEventDisableTiming);
    if (e != cudaSuccess) {
        fail(MG_ECUDA, "mg_create: %s", cudaGetErrorString(e));
        mg_destroy(h);
        return MG_ECUDA;
    }
    *out = h;
    return MG_OK;
}

void mg_destroy(mg_engine* h)
{
    if (!h) return;
    cudaSetDevice(h->device);
    cudaFree(h->cells); cudaFree(h->packed); cudaFree(h->agents);
    cudaFree(h->pool_cells); cudaFree(h->pool_packed); cudaFree(h->pool_agents);
    cudaFree(h->dyn); cudaFree(h->pool_dyn);
    cudaFree(h->error_word); cudaFree(h->chain);
    for (int i = 0; i < h->num_atlases; ++i) cudaFree(h->atlases[i].dev);
    if (h->stage) cudaFree(h->stage);
    if (h->hstream) cudaStreamDestroy(h->hstream);
    if (h->hstream2) cudaStreamDestroy(h->hstream2);
    for (int i = 0; i < 3; ++i) if (h->hev[i]) cudaEventDestroy(h->hev[i]);
    delete h;
}

int mg_num_envs(const mg_engine* h) { return h ? h->N : 0; }
int mg_plane_stride(const mg_engine* h) { return h ? h->plane_stride : 0; }
int64_t mg_launch_count(const mg_engine* h) { return h ? h->launches : 0; }
int mg_reset_stride(const mg_engine* h) { return h ? mg::reset_stride_for(h->N, h->P) : 0; }

int mg_step_occupancy(mg_engine* h, int32_t* blocks_per_sm, int32_t* threads_per_block, int64_t* smem_bytes)
{
    if (!h) return fail(MG_EINVAL, "mg_step_occupancy: null handle");
    MG_CUDA(cudaSetDevice(h->device));
    int b = 0, t = 0; size_t s = 0;
    MG_CUDA(mg::step_occupancy(h->packed_stride, h->C, h->rows_bytes, &b, &t, &s));
    if (blocks_per_sm) *blocks_per_sm = b;
    if (threads_per_block) *threads_per_block = t;
    if (smem_bytes) *smem_bytes = (int64_t)s;
    return MG_OK;
}

int mg_upload_layouts(mg_engine* h, const uint8_t* cells, const mg_layout_agent* agents, int32_t P)
{
    if (!h || !cells || !agents || P <= 0) return fail(MG_EINVAL, "mg_upload_layouts: bad argument");
    for (int i = 0; i < P; ++i)
        if (agents[i].class_id >= h->C || agents[i].dir > 3) return fail(MG_EINVAL, "mg_upload_layouts: bad class id / dir");
    for (int i = 0; i < P; ++i) {
        const mg::ClassDev& cd = h->classes.c[agents[i].class_id];
        if (agents[i].x >= cd.W || agents[i].y >= cd.H) return fail(MG_EINVAL, "mg_upload_layouts: agent outside the grid");
        int rc = validate_cells(cells + (size_t)i * 4 * h->plane_stride, h->plane_stride, cd.W * cd.H);
        if (rc) return rc;
    }
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    cudaFree(h->pool_cells); cudaFree(h->pool_packed); cudaFree(h->pool_agents); cudaFree(h->pool_dyn);
    h->pool_cells = nullptr; h->pool_packed = nullptr; h->pool_agents = nullptr; h->pool_dyn = nullptr; h->P = 0;
    const size_t cb = (size_t)P * 4 * h->plane_stride;
    MG_CUDA(cudaMalloc(&h->pool_cells, cb));
    MG_CUDA(cudaMalloc(&h->pool_agents, (size_t)P * sizeof(mg::LayoutAgent)));
    MG_CUDA(cudaMemcpy(h->pool_cells, cells, cb, cudaMemcpyHostToDevice));
    MG_CUDA(cudaMemcpy(h->pool_agents, agents, (size_t)P * sizeof(mg::LayoutAgent), cudaMemcpyHostToDevice));
    MG_CUDA(cudaMalloc(&h->pool_packed, (size_t)P * h->packed_stride));
    mg::PackParams pk{h->pool_cells, h->pool_packed, reinterpret_cast<const uint8_t*>(h->pool_agents) + 3, 4, P,
                      h->plane_stride, h->packed_stride, h->classes};
    MG_CUDA(mg::launch_pack(pk, nullptr));
    MG_CUDA(cudaDeviceSynchronize());
    h->launches++;
    h->P = P;
    return MG_OK;
}

int mg_upload_dyn(mg_engine* h, const mg_dyn_record* dyn, int32_t P)
{
    if (!h || !dyn) return fail(MG_EINVAL, "mg_upload_dyn: null argument");
    if (h->P == 0 || P != h->P) return fail(MG_ESTATE, "mg_upload_dyn: call after mg_upload_layouts with the same P");
    for (int i = 0; i < P; ++i) {
        if (dyn[i].n_obstacles > 8) return fail(MG_EINVAL, "mg_upload_dyn: at most 8 obstacles per env");
        if (!(dyn[i].inc_lo & 1)) return fail(MG_EINVAL, "mg_upload_dyn: PCG64 increment must be odd");
    }
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    if (!h->dyn) {
        MG_CUDA(cudaMalloc(&h->dyn, (size_t)h->N * sizeof(mg_dyn_record)));
        MG_CUDA(cudaMemset(h->dyn, 0, (size_t)h->N * sizeof(mg_dyn_record)));
    }
    cudaFree(h->pool_dyn); h->pool_dyn = nullptr;
    MG_CUDA(cudaMalloc(&h->pool_dyn, (size_t)P * sizeof(mg_dyn_record)));
    MG_CUDA(cudaMemcpy(h->pool_dyn, dyn, (size_t)P * sizeof(mg_dyn_record), cudaMemcpyHostToDevice));
    return MG_OK;
}

int mg_get_dyn(mg_engine* h, mg_dyn_record* dyn_host)
{
    if (!h || !dyn_host) return fail(MG_EINVAL, "mg_get_dyn: null argument");
    if (!h->dyn) return fail(MG_ESTATE, "mg_get_dyn: no records (mg_upload_dyn was never called)");
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    MG_CUDA(cudaMemcpy(dyn_host, h->dyn, (size_t)h->N * sizeof(mg_dyn_record), cudaMemcpyDeviceToHost));
    return MG_OK;
}

int mg_set_dyn(mg_engine* h, const mg_dyn_record* dyn_host)
{
    if (!h || !dyn_host) return fail(MG_EINVAL, "mg_set_dyn: null argument");
    if (!h->dyn) return fail(MG_ESTATE, "mg_set_dyn: no records (mg_upload_dyn was never called)");
    for (int i = 0; i < h->N; ++i)
        if (dyn_host[i].n_obstacles > 8) return fail(MG_EINVAL, "mg_set_dyn: at most 8 obstacles per env");
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    MG_CUDA(cudaMemcpy(h->dyn, dyn_host, (size_t)h->N * sizeof(mg_dyn_record), cudaMemcpyHostToDevice));
    return MG_OK;
}

static mg::StepParams step_params(mg_engine* h)
{
    mg::StepParams p{};
    p.cells = h->cells; p.packed = h->packed; p.agents = h->agents;
    p.dyn = h->pool_dyn ? h->dyn : nullptr; p.pool_dyn = h->pool_dyn;
    p.pool_cells = h->pool_cells; p.pool_packed = h->pool_packed; p.pool_agents = h->pool_agents;
    p.error_word = h->error_word; p.chain = h->chain; p.N = h->N; p.P = h->P; p.plane_stride = h->plane_stride;
    p.packed_stride = h->packed_stride; p.num_classes = h->C; p.rows_bytes = h->rows_bytes;
    p.autoreset = h->autoreset; p.classes = h->classes; p.T = 1;
    p.reset_stride = mg::reset_stride_for(h->N, h->P);
    p.no_death_mask = h->no_death_mask; p.death_cost = h->death_cost;
    return p;
}

int mg_set_no_death(mg_engine* h, uint32_t type_mask, double death_cost)
{
    if (!h) return fail(MG_EINVAL, "mg_set_no_death: null handle");
    if (type_mask & ((1u << mg::T_GOAL) | (1u << mg::T_EMPTY) | (1u << mg::T_UNSEEN) | ~0x7FFu))
        return fail(MG_EINVAL, "mg_set_no_death: goal cannot be a death cell (wrappers.py:831); types are 2..10 except 8");
    h->no_death_mask = type_mask; h->death_cost = death_cost;
    return MG_OK;
}

int mg_reset(mg_engine* h, const int32_t* layout_ids_dev, const uint8_t* mask_dev,
             uint8_t* image_dev, uint8_t* direction_dev, void* stream)
{
    if (!h || !layout_ids_dev) return fail(MG_EINVAL, "mg_reset: null argument");
    if (h->P == 0) return fail(MG_ESTATE, "mg_reset: no layouts uploaded");
    MG_CUDA(cudaSetDevice(h->device));
    mg::ResetParams r{h->cells, h->packed, h->agents, h->pool_cells, h->pool_packed, h->pool_agents,
                      layout_ids_dev, mask_dev, h->N, h->P, h->plane_stride, h->packed_stride,
                      h->pool_dyn ? h->dyn : nullptr, h->pool_dyn, h->error_word};
    MG_CUDA(mg::launch_reset(r, (cudaStream_t)stream));
    h->launches++;
    if (image_dev || direction_dev) {
        mg::StepParams p = step_params(h);
        p.actions = nullptr; p.image = image_dev; p.direction = direction_dev;
        MG_CUDA(mg::launch_step_many(p, (cudaStream_t)stream));
        h->launches++;
    }
    return MG_OK;
}

int mg_step_many(mg_engine* h, const uint8_t* actions_dev, int32_t T,
                 uint8_t* image_dev, uint8_t* direction_dev, double* reward_dev,
                 uint8_t* terminated_dev, uint8_t* truncated_dev, void* stream)
{
    if (!h || !actions_dev || T <= 0) return fail(MG_EINVAL, "mg_step_many: bad argument");
    if (h->P == 0) return fail(MG_ESTATE, "mg_step_many: no layouts uploaded / envs never reset");
    if (!h->pool_dyn)
        for (int i = 0; i < h->C; ++i)
            if (h->classes.c[i].rule == mg::RULE_DYNOBS)
                return fail(MG_ESTATE, "mg_step_many: a MG_RULE_DYNOBS class needs mg_upload_dyn after mg_upload_layouts");
    MG_CUDA(cudaSetDevice(h->device));
    mg::StepParams p = step_params(h);
    p.actions = actions_dev; p.T = T; p.image = image_dev; p.direction = direction_dev; p.reward = reward_dev;
    p.terminated = terminated_dev; p.truncated = truncated_dev;
    MG_CUDA(mg::launch_step_many(p, (cudaStream_t)stream));
    h->launches++;
    return MG_OK;
}

int mg_step_many_final(mg_engine* h, const uint8_t* actions_dev, int32_t T,
                       uint8_t* image_dev, uint8_t* direction_dev, double* reward_dev,
                       uint8_t* terminated_dev, uint8_t* truncated_dev,
                       uint8_t* final_image_dev, uint8_t* final_direction_dev, void* stream)
{
    if (!h || !actions_dev || T <= 0) return fail(MG_EINVAL, "mg_step_many_final: bad argument");
    if (h->P == 0) return fail(MG_ESTATE, "mg_step_many_final: no layouts uploaded / envs never reset");
    if (!h->pool_dyn)
        for (int i = 0; i < h->C; ++i)
            if (h->classes.c[i].rule == mg::RULE_DYNOBS)
                return fail(MG_ESTATE, "mg_step_many_final: a MG_RULE_DYNOBS class needs mg_upload_dyn after mg_upload_layouts");
    MG_CUDA(cudaSetDevice(h->device));
    mg::StepParams p = step_params(h);
    p.actions = actions_dev; p.T = T; p.image = image_dev; p.direction = direction_dev; p.reward = reward_dev;
    p.terminated = terminated_dev; p.truncated = truncated_dev;
    p.final_image = final_image_dev; p.final_direction = final_direction_dev;
    MG_CUDA(mg::launch_step_many(p, (cudaStream_t)stream));
    h->launches++;
    return MG_OK;
}

int mg_step(mg_engine* h, const uint8_t* actions_dev, uint8_t* image_dev, uint8_t* direction_dev,
            double* reward_dev, uint8_t* terminated_dev, uint8_t* truncated_dev, void* stream)
{
    return mg_step_many(h, actions_dev, 1, image_dev, direction_dev, reward_dev, terminated_dev, truncated_dev, stream);
}

// Host-buffer entry point.  The T steps are cut into chunks that alternate between two streams: chunk k's five
// device->host copies (the 147-byte observations first: 93 % of the bytes) run while chunk k+1's kernel computes, so apart
// from the first chunk's kernel the call takes exactly as long as the PCIe transfers.  Chunk k+1's kernel waits for chunk
// k's (env state carries over); the staging buffer holds all T steps, so no copy ever waits for a buffer.
// MG_HOST_CHUNKS (default 4) sets the number of chunks; 1 = one kernel followed by the copies (round 1's form).
int mg_step_many_host(mg_engine* h, const uint8_t* actions_host, int32_t T,
                      uint8_t* image_host, uint8_t* direction_host, double* reward_host,
                      uint8_t* terminated_host, uint8_t* truncated_host)
{
    if (!h || !actions_host || T <= 0) return fail(MG_EINVAL, "mg_step_many_host: bad argument");
    MG_CUDA(cudaSetDevice(h->device));
    const size_t N = (size_t)h->N, n = (size_t)T * N;
    // device staging: reward (8n, first for alignment) | image (147n) | actions, direction, terminated, truncated (n each)
    const size_t need = n * (8 + MG_OBS_BYTES + 4) + 64;
    if (need > h->stage_bytes) {
        if (h->stage) cudaFree(h->stage);
        h->stage = nullptr; h->stage_bytes = 0;
        MG_CUDA(cudaMalloc(&h->stage, need));
        h->stage_bytes = need;
    }
    double* d_rew = reinterpret_cast<double*>(h->stage);
    uint8_t* d_img = h->stage + 8 * n;
    uint8_t* d_act = d_img + (size_t)MG_OBS_BYTES * n;
    uint8_t* d_dir = d_act + n; uint8_t* d_term = d_dir + n; uint8_t* d_trunc = d_term + n;
    static const int want_chunks = [] { const char* e = getenv("MG_HOST_CHUNKS"); const int v = e ? atoi(e) : 4; return v < 1 ? 1 : v; }();
    const int Tc = (T + want_chunks - 1) / want_chunks;
    cudaStream_t st[2] = {h->hstream, h->hstream2};
    int rc = MG_OK;
    cudaError_t e = cudaMemcpyAsync(d_act, actions_host, n, cudaMemcpyHostToDevice, st[0]);
    if (e == cudaSuccess) e = cudaEventRecord(h->hev[0], st[0]);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(st[1], h->hev[0], 0);
    int c = 0;
    for (int t0 = 0; t0 < T && e == cudaSuccess && rc == MG_OK; t0 += Tc, ++c) {
        const int tc = T - t0 < Tc ? T - t0 : Tc, k = c & 1;
        const size_t o = (size_t)t0 * N, m = (size_t)tc * N;
        if (c > 0) e = cudaStreamWaitEvent(st[k], h->hev[1 + (k ^ 1)], 0);            // the previous chunk's kernel
        if (e != cudaSuccess) break;
        rc = mg_step_many(h, d_act + o, tc, image_host ? d_img + o * MG_OBS_BYTES : nullptr, direction_host ? d_dir + o : nullptr,
                          reward_host ? d_rew + o : nullptr, terminated_host ? d_term + o : nullptr,
                          truncated_host ? d_trunc + o : nullptr, st[k]);
        if (rc) break;
        e = cudaEventRecord(h->hev[1 + k], st[k]);
        if (e == cudaSuccess && image_host) e = cudaMemcpyAsync(image_host + o * MG_OBS_BYTES, d_img + o * MG_OBS_BYTES, m * MG_OBS_BYTES, cudaMemcpyDeviceToHost, st[k]);
        if (e == cudaSuccess && reward_host) e = cudaMemcpyAsync(reward_host + o, d_rew + o, 8 * m, cudaMemcpyDeviceToHost, st[k]);
        if (e == cudaSuccess && direction_host) e = cudaMemcpyAsync(direction_host + o, d_dir + o, m, cudaMemcpyDeviceToHost, st[k]);
        if (e == cudaSuccess && terminated_host) e = cudaMemcpyAsync(terminated_host + o, d_term + o, m, cudaMemcpyDeviceToHost, st[k]);
        if (e == cudaSuccess && truncated_host) e = cudaMemcpyAsync(truncated_host + o, d_trunc + o, m, cudaMemcpyDeviceToHost, st[k]);
    }
    const cudaError_t s0 = cudaStreamSynchronize(st[0]), s1 = cudaStreamSynchronize(st[1]);      // also on the error paths
    if (rc) return rc;
    if (e == cudaSuccess) e = s0 != cudaSuccess ? s0 : s1;
    if (e != cudaSuccess) return fail(MG_ECUDA, "mg_step_many_host: %s", cudaGetErrorString(e));
    return MG_OK;
}

int mg_poll_error(mg_engine* h, int64_t* env_index, int32_t* action)
{
    if (!h) return fail(MG_EINVAL, "mg_poll_error: null handle");
    MG_CUDA(cudaSetDevice(h->device));
    unsigned long long w = 0;
    MG_CUDA(cudaDeviceSynchronize());
    MG_CUDA(cudaMemcpy(&w, h->error_word, sizeof(w), cudaMemcpyDeviceToHost));
    if (!w) return MG_OK;
    MG_CUDA(cudaMemset(h->error_word, 0, sizeof(w)));
    if (env_index) *env_index = (int64_t)(w & 0xffffffffffull);
    if (w & 0x4000000000000000ull) return fail(MG_EINVAL, "mg_reset: layout id outside the pool");
    if (w & 0x2000000000000000ull) return fail(MG_ECUDA, "step kernel: a CTA waited 2 s for the previous launch of its envs (per-CTA launch chaining)");
    if (action) *action = (int32_t)((w >> 40) & 0xff);
    return fail(MG_EACTION, "Unknown action");
}

int mg_get_state(mg_engine* h, uint8_t* cells_host, mg_agent_state* agents_host)
{
    if (!h) return fail(MG_EINVAL, "mg_get_state: null handle");
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    if (cells_host) MG_CUDA(cudaMemcpy(cells_host, h->cells, (size_t)h->N * 4 * h->plane_stride, cudaMemcpyDeviceToHost));
    if (agents_host) MG_CUDA(cudaMemcpy(agents_host, h->agents, (size_t)h->N * sizeof(mg::AgentState), cudaMemcpyDeviceToHost));
    return MG_OK;
}

int mg_set_state(mg_engine* h, const uint8_t* cells_host, const mg_agent_state* agents_host)
{
    if (!h) return fail(MG_EINVAL, "mg_set_state: null handle");
    MG_CUDA(cudaSetDevice(h->device));
    MG_CUDA(cudaDeviceSynchronize());
    if (agents_host) {
        for (int i = 0; i < h->N; ++i) {
            const mg_agent_state& a = agents_host[i];
            if (a.class_id >= h->C || a.dir > 3 || a.x >= h->classes.c[a.class_id].W || a.y >= h->classes.c[a.class_id].H)
                return fail(MG_EINVAL, "mg_set_state: agent out of range");
            if (a.carry_type && !mg::code_representable(a.carry_type, a.carry_color, a.carry_state))
                return fail(MG_EINVAL, "mg_set_state: carried object not representable");
        }
        MG_CUDA(cudaMemcpy(h->agents, agents_host, (size_t)h->N * sizeof(mg::AgentState), cudaMemcpyHostToDevice));
    }
    if (cells_host) {
        // class ids come from the (possibly just updated) device agents
        mg_agent_state* tmp = (mg_agent_state*)malloc((size_t)h->N * sizeof(mg_agent_state));
        if (!tmp) return fail(MG_ENOMEM, "mg_set_state: out of host memory");
        cudaError_t e = cudaMemcpy(tmp, h->agents, (size_t)h->N * sizeof(mg::AgentState), cudaMemcpyDeviceToHost);
        int rc = MG_OK;
        for (int i = 0; i < h->N && e == cudaSuccess && rc == MG_OK; ++i) {
            const mg::ClassDev& cd = h->classes.c[tmp[i].class_id < h->C ? tmp[i].class_id : 0];
            rc = validate_cells(cells_host + (size_t)i * 4 * h->plane_stride, h->plane_stride, cd.W * cd.H);
        }
        free(tmp);
        if (e != cudaSuccess) return fail(MG_ECUDA, "mg_set_state: %s", cudaGetErrorString(e));
        if (rc) return rc;
        MG_CUDA(cudaMemcpy(h->cells, cells_host, (size_t)h->N * 4 * h->plane_stride, cudaMemcpyHostToDevice));
    }
    // rebuild the derived packed tiles
    mg::PackParams pk{h->cells, h->packed, reinterpret_cast<const uint8_t*>(h->agents) + 3, 16, h->N,
                      h->plane_stride, h->packed_stride, h->classes};
    MG_CUDA(mg::launch_pack(pk, nullptr));
    MG_CUDA(cudaDeviceSynchronize());
    h->launches++;
    return MG_OK;
}

}  // extern "C"
