// mg_step.cu — fused step + partial-observation kernels (K1+K2+K3 of SURVEY §8.7).
//
// Replaces MiniGridEnv.step (minigrid/minigrid_env.py:520-590), gen_obs_grid / gen_obs (:592-645),
// Grid.slice / rotate_left / process_vis / encode (minigrid/core/grid.py:110-143, 244-328),
// _reward (:235-240), KeyCorridorEnv.step (minigrid/envs/keycorridor.py:129-137) and the
// state-copy part of reset (:119-157).
#include "mg_step.cuh"

namespace mg {

// ------------------------------------------------------------------------------------------------
// Visibility: Grid.process_vis (core/grid.py:291-328) in closed form on 7-bit row masks.
// `t[j]` = transparency bits of view row j (bit a = view column a).  Within a row a lit transparent
// cell lights its right neighbour during the left-to-right sweep and its left neighbour during the
// right-to-left sweep (opaque cells get lit but do not propagate); lit transparent cells also light
// the three cells above them (up, up-left, up-right) for the next row.
// The rightward flood is a carry chain: (t + (m & t)) ^ t marks each seed's run and the cell after it.
// The leftward flood is the same on bit-reversed masks.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void process_vis(const uint32_t t[VIEW], uint32_t vis[VIEW])
{
    uint32_t m = 1u << (VIEW / 2);                       // agent cell (3, 6), grid.py:294
#pragma unroll
    for (int j = VIEW - 1; j >= 0; --j) {
        const uint32_t tj = t[j];
        m |= ((tj + (m & tj)) ^ tj);                     // rightward sweep, grid.py:297-308
        m &= 0x7fu;
        uint32_t mr = __brev(m) >> 25, tr = __brev(tj) >> 25;
        mr |= ((tr + (mr & tr)) ^ tr);                   // leftward sweep, grid.py:310-321
        m = (__brev(mr) >> 25) & 0x7fu;
        vis[j] = m;
        const uint32_t e = m & tj;                       // lit transparent cells seed the row above
        m = (e | (e << 1) | (e >> 1)) & 0x7fu;
    }
}

struct EnvRegs {
    int x, y, dir, cls, steps;
    cell_t carry;
    uint32_t layout_id;
};

__device__ __forceinline__ cell_t load_cell(const uint8_t* cells, int ps, int k)
{
    return make_cell(cells[k], cells[ps + k], cells[2 * ps + k], cells[3 * ps + k]);
}
__device__ __forceinline__ void store_cell(uint8_t* cells, int ps, int k, cell_t c)
{
    cells[k] = (uint8_t)c_type(c); cells[ps + k] = (uint8_t)c_color(c);
    cells[2 * ps + k] = (uint8_t)c_state(c); cells[3 * ps + k] = (uint8_t)c_aux(c);
}

// gen_obs for one env, written as 147 bytes at `out` (shared memory), byte (a*7+b)*3+c.
// View cell (a,b) <- world (ax,ay) + f*(6-b) + r*(a-3), f = DIR_TO_VEC[dir], r = (-f.y, f.x);
// outside the grid -> Wall() (grid.py:136-139).  This is slice + rotate_left^(dir+1) in closed form.
__device__ __forceinline__ void gen_obs(const uint8_t* __restrict__ cells, int ps, const ClassDesc& cd,
                                        const EnvRegs& e, uint8_t* out)
{
    const int fx = dir_dx(e.dir), fy = dir_dy(e.dir);
    const int rx = -fy, ry = fx;
    uint32_t t[VIEW], vis[VIEW];
#pragma unroll
    for (int b = 0; b < VIEW; ++b) {
        uint32_t row = 0;
#pragma unroll
        for (int a = 0; a < VIEW; ++a) {
            const int x = e.x + fx * (VIEW - 1 - b) + rx * (a - VIEW / 2);
            const int y = e.y + fy * (VIEW - 1 - b) + ry * (a - VIEW / 2);
            bool opaque = true;
            if (x >= 0 && y >= 0 && x < cd.W && y < cd.H) {
                const int k = y * cd.W + x;
                opaque = is_opaque(cells[k], cells[2 * ps + k]);
            }
            row |= (opaque ? 0u : 1u) << a;
        }
        t[b] = row;
    }
    if (cd.see_through) {
#pragma unroll
        for (int b = 0; b < VIEW; ++b) vis[b] = 0x7fu;                       // minigrid_env.py:615-616
    } else {
        process_vis(t, vis);
    }
#pragma unroll
    for (int a = 0; a < VIEW; ++a) {
#pragma unroll
        for (int b = 0; b < VIEW; ++b) {
            uint8_t o0 = 0, o1 = 0, o2 = 0;                                   // invisible -> (0,0,0), grid.py:252,257
            if ((vis[b] >> a) & 1u) {
                if (a == VIEW / 2 && b == VIEW - 1) {                         // carried object shown at the agent cell, :621-625
                    const cell_t c = e.carry ? e.carry : CELL_NONE;
                    o0 = c_type(c); o1 = c_color(c); o2 = c_state(c);
                } else {
                    const int x = e.x + fx * (VIEW - 1 - b) + rx * (a - VIEW / 2);
                    const int y = e.y + fy * (VIEW - 1 - b) + ry * (a - VIEW / 2);
                    if (x >= 0 && y >= 0 && x < cd.W && y < cd.H) {
                        const int k = y * cd.W + x;
                        o0 = cells[k]; o1 = cells[ps + k]; o2 = cells[2 * ps + k];
                    } else { o0 = T_WALL; o1 = 5; o2 = 0; }
                }
            }
            uint8_t* o = out + (a * VIEW + b) * 3;
            o[0] = o0; o[1] = o1; o[2] = o2;
        }
    }
}

// Copy `nbytes` from shared memory to global memory with the whole CTA, 16 bytes per thread when aligned.
__device__ __forceinline__ void cta_copy_out(uint8_t* __restrict__ dst, const uint8_t* src, int nbytes)
{
    if ((((uintptr_t)dst) & 15) == 0) {
        const int n16 = nbytes >> 4;
        const uint4* s4 = reinterpret_cast<const uint4*>(src);
        uint4* d4 = reinterpret_cast<uint4*>(dst);
        for (int i = threadIdx.x; i < n16; i += blockDim.x) __stcs(d4 + i, s4[i]);
        for (int i = (n16 << 4) + threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    } else {
        for (int i = threadIdx.x; i < nbytes; i += blockDim.x) dst[i] = src[i];
    }
}

__device__ __forceinline__ void reset_env_from_pool(uint8_t* cells, int ps, const uint8_t* pool_cells,
                                                    const LayoutAgent* pool_agents, uint32_t lid, EnvRegs& e)
{
    const uint4* s = reinterpret_cast<const uint4*>(pool_cells + (size_t)lid * 4 * ps);
    uint4* d = reinterpret_cast<uint4*>(cells);
    for (int i = 0; i < (4 * ps) / 16; ++i) d[i] = s[i];
    const LayoutAgent la = pool_agents[lid];
    e.x = la.x; e.y = la.y; e.dir = la.dir; e.cls = la.cls; e.steps = 0; e.carry = 0; e.layout_id = lid;
}

// ------------------------------------------------------------------------------------------------
// v1 kernel: one thread per env, T steps per launch, cells read straight from the HBM/L2-resident
// SoA planes, obs staged in shared memory and written out coalesced.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(STEP_THREADS)
step_many_kernel(StepParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    const int env0 = blockIdx.x * STEP_THREADS;
    const int env = env0 + threadIdx.x;
    const int nvalid = min(STEP_THREADS, p.N - env0);
    const bool active = env < p.N;
    const int ps = p.plane_stride;
    uint8_t* cells = p.cells + (size_t)(active ? env : 0) * 4 * ps;
    EnvRegs e;
    {
        const AgentState a = p.agents[active ? env : 0];
        e.x = a.x; e.y = a.y; e.dir = a.dir; e.cls = a.cls; e.steps = a.steps;
        e.carry = make_cell(a.ct, a.cc, a.cs, a.ca); e.layout_id = a.layout_id;
        if (a.ct == 0) e.carry = 0;
    }
    for (int t = 0; t < p.T; ++t) {
        const size_t k = (size_t)t * p.N + env;
        if (active) {
            ClassDesc cd = p.classes.c[e.cls];
            double reward = 0.0; bool terminated = false, truncated = false;
            if (p.actions) {                                   // actions == NULL: observation only (reset path)
                const int action = p.actions[k];
                e.steps += 1;                                                   // minigrid_env.py:523
                const int fx = e.x + dir_dx(e.dir), fy = e.y + dir_dy(e.dir);  // front_pos :413-419
                const int fk = fy * cd.W + fx;
                const cell_t front = load_cell(cells, ps, fk);                  // :533
                const Transition tr = apply_action(action, e.x, e.y, e.dir, e.carry, fx, fy, front, cd.rule);
                if (tr.bad_action) {
                    atomicCAS(p.error_word, 0ull, 0x8000000000000000ull | ((unsigned long long)action << 40) | (unsigned long long)env);
                } else {
                    e.x = tr.x; e.y = tr.y; e.dir = tr.dir; e.carry = tr.carry;
                    if (tr.wrote_front) store_cell(cells, ps, fk, tr.front);
                    terminated = tr.terminated;
                    if (tr.success) reward = reward_fn(e.steps, cd.max_steps);
                    truncated = e.steps >= cd.max_steps;                        // :582-583
                    if (p.autoreset && (terminated || truncated)) {
                        reset_env_from_pool(cells, ps, p.pool_cells, p.pool_agents,
                                            (uint32_t)(((unsigned long long)e.layout_id + (unsigned)p.N) % (unsigned)p.P), e);
                        cd = p.classes.c[e.cls];
                    }
                }
                if (p.reward) p.reward[k] = reward;
                if (p.terminated) p.terminated[k] = terminated;
                if (p.truncated) p.truncated[k] = truncated;
            }
            if (p.direction) p.direction[k] = (uint8_t)e.dir;
            if (p.image) gen_obs(cells, ps, cd, e, smem + threadIdx.x * OBS_BYTES);
        }
        if (p.image) {
            __syncthreads();
            cta_copy_out(p.image + ((size_t)t * p.N + env0) * OBS_BYTES, smem, nvalid * OBS_BYTES);
            __syncthreads();
        }
    }
    if (active && p.actions) {
        AgentState a;
        a.x = e.x; a.y = e.y; a.dir = e.dir; a.cls = e.cls;
        a.ct = c_type(e.carry); a.cc = c_color(e.carry); a.cs = c_state(e.carry); a.ca = c_aux(e.carry);
        a.steps = (uint16_t)e.steps; a.pad = 0; a.layout_id = e.layout_id;
        p.agents[env] = a;
    }
}

// reset: env e <- pool[layout_ids[e]] (where mask says so)
__global__ void __launch_bounds__(256)
reset_kernel(ResetParams p)
{
    // one warp per env: 4*ps bytes copied as uint4
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= p.N) return;
    if (p.mask && !p.mask[warp]) return;
    const uint32_t lid = (uint32_t)p.layout_ids[warp];
    const uint4* s = reinterpret_cast<const uint4*>(p.pool_cells + (size_t)lid * 4 * p.plane_stride);
    uint4* d = reinterpret_cast<uint4*>(p.cells + (size_t)warp * 4 * p.plane_stride);
    for (int i = lane; i < (4 * p.plane_stride) / 16; i += 32) d[i] = s[i];
    if (lane == 0) {
        const LayoutAgent la = p.pool_agents[lid];
        AgentState a;
        a.x = la.x; a.y = la.y; a.dir = la.dir; a.cls = la.cls; a.ct = a.cc = a.cs = a.ca = 0;
        a.steps = 0; a.pad = 0; a.layout_id = lid;
        p.agents[warp] = a;
    }
}

cudaError_t launch_step_many(const StepParams& p, cudaStream_t stream)
{
    const int grid = (p.N + STEP_THREADS - 1) / STEP_THREADS;
    const size_t smem = (size_t)STEP_THREADS * OBS_BYTES;
    step_many_kernel<<<grid, STEP_THREADS, smem, stream>>>(p);
    return cudaGetLastError();
}

cudaError_t launch_reset(const ResetParams& p, cudaStream_t stream)
{
    const int threads = 256;
    const int grid = (int)(((size_t)p.N * 32 + threads - 1) / threads);
    reset_kernel<<<grid, threads, 0, stream>>>(p);
    return cudaGetLastError();
}

}  // namespace mg
