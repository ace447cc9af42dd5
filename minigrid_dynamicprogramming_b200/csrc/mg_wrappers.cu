// mg_wrappers.cu — the reference's observation / reward wrappers (minigrid/wrappers.py) as kernels over the batch
// (SURVEY §8.6 N2).  Each one reads the engine's state (SoA planes + agent records) or the image a step produced and
// writes what the wrapper's observation() / step() would return for every env, in one launch.
//
//   mg_obs_fully      FullyObsWrapper          wrappers.py:384-423   uint8[W][H][3], agent cell = (10, 0, dir)
//   mg_obs_symbolic   SymbolicObsWrapper       wrappers.py:728-766   int64[W][H][3] = (x, y, type | -1), agent cell type 10
//   mg_obs_onehot     OneHotPartialObsWrapper  wrappers.py:220-287   uint8[...][20] one-hot of (type, color, state)
//   mg_obs_flat       FlatObsWrapper           wrappers.py:553-623   float32[147 + 28*96] image ++ one-hot mission
//   mg_bonus          ActionBonus / PositionBonus  wrappers.py:70-187  reward += 1/sqrt(visit count)
//   mg_obs_view       ViewSizeWrapper          wrappers.py:626-670   uint8[V][V][3] for any odd V in 3..15
//   mg_obs_goal_direction  DirectionObsWrapper wrappers.py:673-725   float64 slope (or arctan of it) towards the cached goal_position
//   mg_stochastic_actions  StochasticActionWrapper  wrappers.py:769-796  action kept with probability prob, else replaced
//
// All of them are streaming kernels bound by their HBM writes; none needs shared-memory tiles of the grid.
#include "../../include/mgb200.h"
#include "mg_engine.cuh"

namespace mg {

struct WrapParams {
    const uint8_t* cells;           // [N][4][plane_stride]
    const AgentState* agents;       // [N]
    int N, plane_stride;
    ClassTableDev classes;
};

// one thread per (env, output cell); output cell o = i*H + j (Grid.encode's [i][j] order, core/grid.py:252-266)
__global__ void __launch_bounds__(256) obs_fully_kernel(WrapParams p, uint8_t* out)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)p.N * p.plane_stride) return;
    const int e = (int)(idx / p.plane_stride), o = (int)(idx - (long long)e * p.plane_stride);
    const AgentState a = p.agents[e];
    const ClassDev cd = p.classes.c[a.cls];
    if (o >= cd.W * cd.H) return;
    const int i = o / cd.H, j = o - i * cd.H;
    const uint8_t* c = p.cells + (size_t)e * 4 * p.plane_stride + j * cd.W + i;
    uint8_t t = c[0], col = c[p.plane_stride], s = c[2 * p.plane_stride];
    if (i == a.x && j == a.y) { t = T_AGENT; col = 0; s = a.dir; }            // wrappers.py:419-421
    uint8_t* d = out + ((size_t)e * p.plane_stride + o) * 3;
    d[0] = t; d[1] = col; d[2] = s;
}

__global__ void __launch_bounds__(256) obs_symbolic_kernel(WrapParams p, long long* out)
{
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)p.N * p.plane_stride) return;
    const int e = (int)(idx / p.plane_stride), o = (int)(idx - (long long)e * p.plane_stride);
    const AgentState a = p.agents[e];
    const ClassDev cd = p.classes.c[a.cls];
    if (o >= cd.W * cd.H) return;
    const int i = o / cd.H, j = o - i * cd.H;
    const int t = p.cells[(size_t)e * 4 * p.plane_stride + j * cd.W + i];
    long long sym = t == T_EMPTY ? -1 : t;                                     // None -> -1 (wrappers.py:751-753)
    if (i == a.x && j == a.y) sym = T_AGENT;                                   // :764
    long long* d = out + ((size_t)e * p.plane_stride + o) * 3;
    d[0] = i; d[1] = j; d[2] = sym;
}

// 128 cells per CTA; each thread builds the 20 bytes of its cell as five words in shared memory (stride 5: conflict
// free), then the CTA streams the 2,560 bytes out with 16-byte stores.
__global__ void __launch_bounds__(128) obs_onehot_kernel(const uint8_t* image, long long ncells, uint8_t* out)
{
    __shared__ __align__(16) uint32_t stage[128 * 5];
    const long long cell0 = (long long)blockIdx.x * 128;
    const long long cell = cell0 + threadIdx.x;
    if (cell < ncells) {
        const uint8_t* c = image + cell * 3;
        const uint32_t k0 = c[0], k1 = 11u + c[1], k2 = 17u + c[2];          // len(OBJECT_TO_IDX) = 11, + len(COLOR_TO_IDX) = 17
#pragma unroll
        for (uint32_t w = 0; w < 5; ++w) {
            uint32_t v = 0;
            if ((k0 >> 2) == w) v |= 1u << (8 * (k0 & 3));
            if ((k1 >> 2) == w) v |= 1u << (8 * (k1 & 3));
            if ((k2 >> 2) == w) v |= 1u << (8 * (k2 & 3));
            stage[threadIdx.x * 5 + w] = v;
        }
    }
    __syncthreads();
    const long long left = ncells - cell0;
    const int nwords = (int)(left < 128 ? left : 128) * 5;
    uint32_t* o = reinterpret_cast<uint32_t*>(out + cell0 * 20);              // cell0*20 = blockIdx*2560: 16-byte aligned
    if (nwords == 640) {
        for (int q = threadIdx.x; q < 160; q += 128) __stcs(reinterpret_cast<uint4*>(o) + q, reinterpret_cast<const uint4*>(stage)[q]);
    } else {
        for (int q = threadIdx.x; q < nwords; q += 128) o[q] = stage[q];
    }
}

// out[e][k] = k < 147 ? float(image[e][k]) : mission_table[mission_of_layout[layout_id[e]]][k - 147]
__global__ void __launch_bounds__(256) obs_flat_kernel(const uint8_t* image, const AgentState* agents, const float* table,
                                                       const int32_t* mission_of_layout, int N, int mission_len, float* out)
{
    const int rec = OBS_BYTES + mission_len;
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)N * rec) return;
    const int e = (int)(idx / rec), k = (int)(idx - (long long)e * rec);
    float v;
    if (k < OBS_BYTES) v = (float)image[(size_t)e * OBS_BYTES + k];
    else v = __ldg(table + (size_t)__ldg(mission_of_layout + agents[e].layout_id) * mission_len + (k - OBS_BYTES));
    __stcs(out + idx, v);
}

// kind 0: key (x, y, dir, action) after the step (wrappers.py:103-104); kind 1: key (x, y) (:167-168)
__global__ void __launch_bounds__(256) bonus_kernel(WrapParams p, int kind, const uint8_t* actions, uint32_t* counts, double* reward)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.N) return;
    const AgentState a = p.agents[e];
    const ClassDev cd = p.classes.c[a.cls];
    const int cell = a.y * cd.W + a.x;
    size_t slot;
    if (kind == 0) slot = (size_t)e * p.plane_stride * 28 + ((size_t)cell * 4 + a.dir) * 7 + (actions[e] < 7 ? actions[e] : 6);
    else slot = (size_t)e * p.plane_stride + cell;
    const uint32_t n = counts[slot] + 1;
    counts[slot] = n;
    reward[e] = __dadd_rn(reward[e], __ddiv_rn(1.0, __dsqrt_rn((double)n)));   // reward += 1 / math.sqrt(new_count)
}

// ViewSizeWrapper (wrappers.py:626-670): gen_obs_grid(agent_view_size) + Grid.encode(vis_mask) for any odd view size
// 3..15, one thread per env, straight from the SoA planes.  View cell (i, j) is the world cell
// agent + f*(V-1-j) + r*(i - V/2)  (get_view_exts + rotate_left x (dir+1), minigrid_env.py:448-479, 606-607); outside the
// grid it is a wall (Grid.slice, core/grid.py:136-139).  process_vis (grid.py:291-328) runs on V-bit row masks exactly
// like the step kernel's 7-bit version: carry-chain floods both ways, 3-wide dilation into the row above.
__global__ void __launch_bounds__(128) obs_view_kernel(WrapParams p, int V, uint8_t* out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= p.N) return;
    const AgentState a = p.agents[e];
    const ClassDev cd = p.classes.c[a.cls];
    const uint8_t* c = p.cells + (size_t)e * 4 * p.plane_stride;
    const int fx = dir_dx(a.dir), fy = dir_dy(a.dir), rx = -fy, ry = fx;
    const int ox = (int)a.x + fx * (V - 1) - rx * (V / 2), oy = (int)a.y + fy * (V - 1) - ry * (V / 2);   // world cell of view (0, 0)
    const uint32_t full = (1u << V) - 1u;
    uint32_t vis[15];
    if (cd.see_through) {
        for (int j = 0; j < V; ++j) vis[j] = full;
    } else {
        uint32_t m = 1u << (V / 2);
        for (int j = V - 1; j >= 0; --j) {
            uint32_t t = 0;                                       // bit i = view cell (i, j) lets light through
            for (int i = 0; i < V; ++i) {
                const int wx = ox - fx * j + rx * i, wy = oy - fy * j + ry * i;
                bool opaque = true;                               // out of bounds = wall
                if ((unsigned)wx < (unsigned)cd.W && (unsigned)wy < (unsigned)cd.H) {
                    const int k = wy * cd.W + wx;
                    opaque = is_opaque(c[k], c[2 * p.plane_stride + k]);
                }
                t |= (opaque ? 0u : 1u) << i;
            }
            m = (m | ((t + (m & t)) ^ t)) & full;
            uint32_t mr = __brev(m) >> (32 - V);
            const uint32_t tr = __brev(t) >> (32 - V);
            mr |= ((tr + (mr & tr)) ^ tr);
            m = __brev(mr & full) >> (32 - V);
            vis[j] = m;
            const uint32_t lit = m & t;
            m = (lit | (lit << 1) | (lit >> 1)) & full;
        }
    }
    uint8_t* o = out + (size_t)e * V * V * 3;
    for (int i = 0; i < V; ++i)
        for (int j = 0; j < V; ++j) {
            uint8_t t = 0, col = 0, st = 0;
            if ((vis[j] >> i) & 1u) {
                if (i == V / 2 && j == V - 1) {                   // the agent's own cell shows what it carries (:618-625)
                    t = a.ct ? a.ct : T_EMPTY; col = a.ct ? a.cc : 0; st = a.ct ? a.cs : 0;
                } else {
                    const int wx = ox - fx * j + rx * i, wy = oy - fy * j + ry * i;
                    if ((unsigned)wx < (unsigned)cd.W && (unsigned)wy < (unsigned)cd.H) {
                        const int k = wy * cd.W + wx;
                        t = c[k]; col = c[p.plane_stride + k]; st = c[2 * p.plane_stride + k];
                    } else { t = T_WALL; col = 5; st = 0; }
                }
            }
            uint8_t* d = o + (i * V + j) * 3;
            d[0] = t; d[1] = col; d[2] = st;
        }
}

// DirectionObsWrapper.observation (wrappers.py:714-725): np.divide(goal_position[1] - agent_pos[1], goal_position[0] - agent_pos[0])
// on Python ints = one IEEE fp64 division (x/0 -> +-inf, 0/0 -> nan, as numpy returns them); type "angle" takes np.arctan.
__global__ void __launch_bounds__(256) goal_direction_kernel(const AgentState* agents, const int32_t* goal, int N, int angle, double* out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= N) return;
    const AgentState a = agents[e];
    const double slope = __ddiv_rn((double)(goal[2 * e + 1] - (int)a.y), (double)(goal[2 * e] - (int)a.x));
    out[e] = angle ? atan(slope) : slope;
}

// StochasticActionWrapper.action (wrappers.py:786-796): `uniform < prob` keeps the action; otherwise random_action, or (None)
// the caller's draw of np_random.integers(0, 6) for that env.
__global__ void __launch_bounds__(256) stochastic_actions_kernel(const uint8_t* actions, const double* u, const uint8_t* repl, int N,
                                                                 double prob, int random_action, uint8_t* out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= N) return;
    out[e] = u[e] < prob ? actions[e] : (random_action >= 0 ? (uint8_t)random_action : repl[e]);
}

__global__ void __launch_bounds__(256) layout_ids_kernel(const AgentState* agents, int N, int32_t* out)
{
    const int e = blockIdx.x * blockDim.x + threadIdx.x;
    if (e < N) out[e] = (int32_t)agents[e].layout_id;
}

}  // namespace mg

namespace {

#define MGW_CUDA(call)                                                                    \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess) return mg::set_error(MG_ECUDA, cudaGetErrorString(_e));    \
    } while (0)

mg::WrapParams wrap_params(mg_engine* h)
{
    mg::WrapParams p{};
    p.cells = h->cells; p.agents = h->agents; p.N = h->N; p.plane_stride = h->plane_stride; p.classes = h->classes;
    return p;
}

int ready(mg_engine* h, const void* out, const char* who)
{
    if (!h || !out) return mg::set_error(MG_EINVAL, who);
    if (h->P == 0) return mg::set_error(MG_ESTATE, "wrapper kernels: no layouts uploaded / envs never reset");
    return MG_OK;
}

}  // namespace

extern "C" {

int mg_obs_fully(mg_engine* h, uint8_t* image_dev, void* stream)
{
    int rc = ready(h, image_dev, "mg_obs_fully: null argument");
    if (rc) return rc;
    MGW_CUDA(cudaSetDevice(h->device));
    const long long n = (long long)h->N * h->plane_stride;
    mg::obs_fully_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(wrap_params(h), image_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_obs_symbolic(mg_engine* h, int64_t* image_dev, void* stream)
{
    int rc = ready(h, image_dev, "mg_obs_symbolic: null argument");
    if (rc) return rc;
    MGW_CUDA(cudaSetDevice(h->device));
    const long long n = (long long)h->N * h->plane_stride;
    mg::obs_symbolic_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(wrap_params(h), reinterpret_cast<long long*>(image_dev));
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_obs_onehot(mg_engine* h, const uint8_t* image_dev, int64_t num_images, uint8_t* out_dev, void* stream)
{
    if (!h || !image_dev || !out_dev || num_images <= 0) return mg::set_error(MG_EINVAL, "mg_obs_onehot: bad argument");
    if (((uintptr_t)out_dev & 15) != 0) return mg::set_error(MG_EINVAL, "mg_obs_onehot: out must be 16-byte aligned");
    MGW_CUDA(cudaSetDevice(h->device));
    const long long ncells = (long long)num_images * MG_VIEW * MG_VIEW;
    mg::obs_onehot_kernel<<<(unsigned)((ncells + 127) / 128), 128, 0, (cudaStream_t)stream>>>(image_dev, ncells, out_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_obs_flat(mg_engine* h, const uint8_t* image_dev, const float* mission_table_dev, const int32_t* mission_of_layout_dev,
                int32_t mission_len, float* out_dev, void* stream)
{
    int rc = ready(h, out_dev, "mg_obs_flat: null argument");
    if (rc) return rc;
    if (!image_dev || !mission_table_dev || !mission_of_layout_dev || mission_len <= 0) return mg::set_error(MG_EINVAL, "mg_obs_flat: bad argument");
    MGW_CUDA(cudaSetDevice(h->device));
    const long long n = (long long)h->N * (MG_OBS_BYTES + mission_len);
    mg::obs_flat_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(image_dev, h->agents, mission_table_dev,
                                                                                      mission_of_layout_dev, h->N, mission_len, out_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_obs_view(mg_engine* h, int32_t view_size, uint8_t* image_dev, void* stream)
{
    int rc = ready(h, image_dev, "mg_obs_view: null argument");
    if (rc) return rc;
    if (view_size < 3 || view_size > 15 || (view_size & 1) == 0) return mg::set_error(MG_EINVAL, "mg_obs_view: view size must be odd, 3..15");
    MGW_CUDA(cudaSetDevice(h->device));
    mg::obs_view_kernel<<<(h->N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(wrap_params(h), view_size, image_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_obs_goal_direction(mg_engine* h, const int32_t* goal_position_dev, int32_t angle, double* out_dev, void* stream)
{
    int rc = ready(h, out_dev, "mg_obs_goal_direction: null argument");
    if (rc) return rc;
    if (!goal_position_dev) return mg::set_error(MG_EINVAL, "mg_obs_goal_direction: null argument");
    MGW_CUDA(cudaSetDevice(h->device));
    mg::goal_direction_kernel<<<(h->N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(h->agents, goal_position_dev, h->N, angle, out_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_stochastic_actions(mg_engine* h, const uint8_t* actions_dev, const double* uniform_dev, const uint8_t* replacement_dev,
                          double prob, int32_t random_action, uint8_t* out_dev, void* stream)
{
    if (!h || !actions_dev || !uniform_dev || !out_dev || (random_action < 0 && !replacement_dev) || random_action > 255)
        return mg::set_error(MG_EINVAL, "mg_stochastic_actions: bad argument");
    MGW_CUDA(cudaSetDevice(h->device));
    mg::stochastic_actions_kernel<<<(h->N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(actions_dev, uniform_dev, replacement_dev, h->N,
                                                                                       prob, random_action, out_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_layout_ids(mg_engine* h, int32_t* ids_dev, void* stream)
{
    int rc = ready(h, ids_dev, "mg_layout_ids: null argument");
    if (rc) return rc;
    MGW_CUDA(cudaSetDevice(h->device));
    mg::layout_ids_kernel<<<(h->N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(h->agents, h->N, ids_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_bonus(mg_engine* h, int32_t kind, const uint8_t* actions_dev, uint32_t* counts_dev, double* reward_dev, void* stream)
{
    int rc = ready(h, reward_dev, "mg_bonus: null argument");
    if (rc) return rc;
    if (!counts_dev || (kind != MG_BONUS_ACTION && kind != MG_BONUS_POSITION) || (kind == MG_BONUS_ACTION && !actions_dev))
        return mg::set_error(MG_EINVAL, "mg_bonus: bad argument");
    MGW_CUDA(cudaSetDevice(h->device));
    mg::bonus_kernel<<<(h->N + 255) / 256, 256, 0, (cudaStream_t)stream>>>(wrap_params(h), kind, actions_dev, counts_dev, reward_dev);
    MGW_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

}  // extern "C"
