// mg_render.cu — RGB rendering of the batch: the tile atlas and the two frame kernels.
//
//   reference                                             here
//   Grid.render_tile + tile_cache  core/grid.py:145-198   atlas_kernel: every tile the engine can show, rasterised once per
//                                                         tile size on the device (mg_raster.cuh), kept in HBM / L2
//   get_pov_render  minigrid_env.py:647-661               render_pov_kernel:  uint8[N][7*ts][7*ts][3]
//   get_full_render minigrid_env.py:663-709               render_full_kernel: uint8[N][H*ts][W*ts][3]
//   Grid.render     core/grid.py:200-242                  the copy loop of both kernels
//
// A frame is a pure gather: every G-byte piece of the output lies inside one row of one tile, so a thread finds the
// tile id of its piece (one shared-memory byte), loads G bytes of the atlas through L1 and streams them out.  The
// kernels are bound by the HBM write of the frames (9,408 B per env for a 7x7 view at tile size 8).
//
// Atlas index: slot * 256 + (highlight << 7 | code7); slot 0 = no agent, 1..4 = agent facing 0..3; code7 is the packed
// cell code of mg_packed.cuh without its opaque bit.  A tile is ts*ts*3 bytes, padded to tile_pitch (multiple of 16).
#include "../../include/mgb200.h"
#include "mg_engine.cuh"
#include "mg_raster.cuh"

#include <cmath>

namespace mg {

#ifndef MG_FULL_MINB
#define MG_FULL_MINB 8
#endif
#ifndef MG_FULL_EPC
#define MG_FULL_EPC 32          // envs per CTA of the full-grid kernels (<= RENDER_ENVS)
#endif
#ifndef MG_POV_MINB
#define MG_POV_MINB 16
#endif
#ifndef MG_RENDER_ENVS
#define MG_RENDER_ENVS 32
#endif
constexpr int ATLAS_TILES = 5 * 256;
constexpr int RENDER_THREADS = 128;
constexpr int RENDER_ENVS = MG_RENDER_ENVS;          // envs per CTA (full render: fewer when the grid is large)

// ---- exact division of n < 2^31 by a launch-time constant: q = (n * mul) >> sh  (Granlund & Montgomery, thm 4.2) ----
struct FastDiv { uint32_t d, mul, sh; };
static FastDiv make_fastdiv(uint32_t d)
{
    uint32_t l = 0;
    while ((1ull << l) < d) ++l;                       // l = ceil(log2 d)
    FastDiv f;
    f.d = d; f.sh = 31 + l;
    f.mul = (uint32_t)((1ull << (31 + l)) / d + 1);
    return f;
}
__device__ __forceinline__ uint32_t qdiv(uint32_t n, const FastDiv& f) { return (uint32_t)(((uint64_t)n * f.mul) >> f.sh); }

struct RenderParams {
    const uint8_t* packed;          // [N][packed_stride]
    const AgentState* agents;       // [N]
    const uint8_t* atlas;           // [ATLAS_TILES][tile_pitch]
    uint8_t* out;                   // [N][env_stride]
    long long env_stride;           // bytes between consecutive env frames
    int N, packed_stride, ts, tile_pitch, highlight, cell_stride, envs_per_cta, strip_max;
    FastDiv by_ts, by_ppt, by_ppr, by_pieces;   // tile size, pieces per tile row / per pixel row / per frame (POV)
    FastDiv cls_ppr[MAX_CLASSES];               // full render: pieces per pixel row of each class
    FastDiv cls_w[MAX_CLASSES];                 // full render: grid width of each class
    FastDiv by_cells;                           // full render: cell_stride
    ClassTableDev classes;
};

// ---------------------------------------------------------------- atlas

__global__ void __launch_bounds__(128) atlas_kernel(uint8_t* atlas, int ts, int tile_pitch, raster::Rotations rot)
{
    const int tile = blockIdx.x, slot = tile >> 8, code = tile & 0x7f;
    const bool hl = (tile >> 7) & 1;
    const cell_t c = decode_code((uint32_t)code);
    int type = c_type(c), state = c_state(c);
    if (type < T_WALL || type > T_LAVA || state > S_LOCKED) { type = T_EMPTY; state = 0; }    // not a drawable object -> None
    uint8_t* dst = atlas + (size_t)tile * tile_pitch;
    for (int p = threadIdx.x; p < ts * ts; p += blockDim.x) {
        const int py = p / ts, px = p - py * ts;
        const uint32_t v = raster::tile_pixel(type, c_color(c), state, slot - 1, hl, rot, ts, px, py);
        dst[3 * p] = (uint8_t)v; dst[3 * p + 1] = (uint8_t)(v >> 8); dst[3 * p + 2] = (uint8_t)(v >> 16);
    }
}

// ---------------------------------------------------------------- piece copy

template <int G> struct Piece;
template <> struct Piece<1> { typedef uint8_t T; };
template <> struct Piece<4> { typedef uint32_t T; };
template <> struct Piece<8> { typedef uint2 T; };
template <> struct Piece<16> { typedef uint4 T; };

template <int G>
__device__ __forceinline__ void copy_piece(uint8_t* dst, const uint8_t* src)
{
    typedef typename Piece<G>::T T;
    __stcs(reinterpret_cast<T*>(dst), __ldg(reinterpret_cast<const T*>(src)));
}

// visibility of one env -> view codes; shared by both kernels
__device__ __forceinline__ void env_view(const RenderParams& p, int env, const AgentState& a, const ClassDev& cd,
                                         uint32_t lo[VIEW], uint32_t hi[VIEW], uint32_t& vlo, uint32_t& vhi)
{
    TileGeom g;
    g.W = cd.W; g.H = cd.H; g.PR = cd.PR; g.PC = cd.PC; g.offC = cd.offC; g.bytes = 0;
    const uint8_t* tile = p.packed + (size_t)env * p.packed_stride;
    view_codes(tile, g, a.x, a.y, a.dir, a.ct ? encode_code(a.ct, a.cc, a.cs) : CODE_EMPTY, cd.see_through != 0, lo, hi, vlo, vhi);
}

// ---------------------------------------------------------------- agent's point of view

// ids[e][a*8 + b] = highlight << 7 | code7 of view cell (a, b); cells outside vis_mask were set to None by
// Grid.process_vis (core/grid.py:323-326) and are not highlighted.
__device__ __forceinline__ void pov_ids(const RenderParams& p, int env, uint8_t* ids56)
{
    const AgentState a = p.agents[env];
    const ClassDev cd = p.classes.c[a.cls];
    uint32_t lo[VIEW], hi[VIEW], vlo, vhi;
    env_view(p, env, a, cd, lo, hi, vlo, vhi);
    uint32_t* w = reinterpret_cast<uint32_t*>(ids56);
#pragma unroll
    for (int v = 0; v < VIEW; ++v) {
        const uint32_t ml = ((vlo >> v) & 0x01010101u) * 0xFFu, mh = ((vhi >> v) & 0x01010101u) * 0xFFu;
        w[2 * v] = (((lo[v] & 0x7F7F7F7Fu) | 0x80808080u) & ml) | (0x01010101u & ~ml);
        w[2 * v + 1] = (((hi[v] & 0x7F7F7F7Fu) | 0x80808080u) & mh) | (0x01010101u & ~mh);
    }
}

// Direct path: every thread gathers one G-byte piece and stores it.  Used when the pieces are 16 bytes wide (tile
// sizes that are multiples of 16: the index arithmetic is amortised over a full-width store) and for odd tile sizes.
template <int G>
__global__ void __launch_bounds__(RENDER_THREADS) render_pov_kernel(const __grid_constant__ RenderParams p)
{
    __shared__ __align__(16) uint8_t ids[RENDER_ENVS][VIEW * 8];
    const int env0 = blockIdx.x * RENDER_ENVS;
    const int nvalid = min(RENDER_ENVS, p.N - env0);
    if ((int)threadIdx.x < nvalid) pov_ids(p, env0 + threadIdx.x, ids[threadIdx.x]);
    __syncthreads();
    const uint32_t trow = (uint32_t)p.ts * 3u;                     // bytes in one row of a tile
    const uint32_t total = (uint32_t)nvalid * p.by_pieces.d;
    uint8_t* out0 = p.out + (size_t)env0 * p.env_stride;
    for (uint32_t idx = threadIdx.x; idx < total; idx += RENDER_THREADS) {
        const uint32_t e = qdiv(idx, p.by_pieces), q = idx - e * p.by_pieces.d;
        const uint32_t y = qdiv(q, p.by_ppr), xb = q - y * p.by_ppr.d;          // pixel row, piece within the row
        const uint32_t b = qdiv(y, p.by_ts), py = y - b * p.by_ts.d;            // view row, pixel row within the tile
        const uint32_t a = qdiv(xb, p.by_ppt), pxb = xb - a * p.by_ppt.d;       // view column, piece within the tile row
        uint32_t t = ids[e][a * 8 + b];
        if (a == VIEW / 2 && b == VIEW - 1) t += 4u * 256u;                     // agent marker, facing up (minigrid_env.py:656-657)
        const uint8_t* src = p.atlas + (size_t)t * p.tile_pitch + py * trow + pxb * G;
        copy_piece<G>(out0 + (size_t)e * p.env_stride + (size_t)q * G, src);
    }
}

// Staged path (tile sizes 4, 8, 12, 20, 24: tile rows of 12..72 bytes).  The unit of work is a STRIP: one row of tiles
// of one frame = ts pixel rows = A*ts*ts*3 contiguous output bytes.  A warp assembles its strip in shared memory —
// one index computation per TILE ROW (3*ts bytes, copied as GP-byte pieces), destination offset = item * 3*ts because
// items are ordered [pixel row][tile] exactly like the output — then streams it out with 16-byte stores.  Warps work
// on their own strips: no CTA barrier after the ids are built.
// TS = compile-time tile size (0: read it from the launch parameters); with TS known every loop below unrolls.
template <int GP, int TS>
__device__ __forceinline__ void copy_tile_row(uint8_t* dst, const uint8_t* src, uint32_t ppt)
{
    typedef typename Piece<GP>::T T;
    if (TS) {
        T v[TS ? 3 * TS / GP : 1];
#pragma unroll
        for (int k = 0; k < 3 * TS / GP; ++k) v[k] = __ldg(reinterpret_cast<const T*>(src) + k);
#pragma unroll
        for (int k = 0; k < 3 * TS / GP; ++k) reinterpret_cast<T*>(dst)[k] = v[k];
    } else {
        for (uint32_t k = 0; k < ppt; ++k) reinterpret_cast<T*>(dst)[k] = __ldg(reinterpret_cast<const T*>(src) + k);
    }
}
__device__ __forceinline__ void strip_out(uint8_t* out, const uint8_t* buf, uint32_t bytes, uint32_t lane)
{
    for (uint32_t c = lane; c < bytes / 16; c += 32) __stcs(reinterpret_cast<uint4*>(out) + c, reinterpret_cast<const uint4*>(buf)[c]);
}

template <int GP, int TS>
__global__ void __launch_bounds__(RENDER_THREADS, MG_POV_MINB) render_pov_staged_kernel(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    uint8_t* ids = smem;                                            // [RENDER_ENVS][56]
    const int env0 = blockIdx.x * RENDER_ENVS;
    const int nvalid = min(RENDER_ENVS, p.N - env0);
    if ((int)threadIdx.x < nvalid) pov_ids(p, env0 + threadIdx.x, ids + threadIdx.x * (VIEW * 8));
    __syncthreads();
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t ts = TS ? TS : p.ts, trow = 3u * ts, items = VIEW * ts, strip_bytes = items * trow, ppt = trow / GP;
    uint8_t* buf = smem + RENDER_ENVS * VIEW * 8 + warp * strip_bytes;
    uint8_t* out0 = p.out + (size_t)env0 * p.env_stride;
    for (uint32_t s = warp; s < (uint32_t)nvalid * VIEW; s += RENDER_THREADS / 32) {
        const uint32_t e = s / VIEW, b = s - e * VIEW;
        const uint8_t* eid = ids + e * (VIEW * 8) + b;
#pragma unroll
        for (uint32_t it = lane; it < items; it += 32) {
            const uint32_t py = it / VIEW, a = it - py * VIEW;
            uint32_t t = eid[a * 8];
            if (a == VIEW / 2 && b == VIEW - 1) t += 4u * 256u;                 // agent marker, facing up
            copy_tile_row<GP, TS>(buf + it * trow, p.atlas + (size_t)t * p.tile_pitch + py * trow, ppt);
        }
        __syncwarp();
        strip_out(out0 + (size_t)e * p.env_stride + (size_t)b * strip_bytes, buf, strip_bytes, lane);
        __syncwarp();
    }
}

// ---------------------------------------------------------------- full grid

struct FullHdr { uint32_t vlo, vhi; uint8_t x, y, dir, cls; };

// phase 1 of both full-grid kernels: per-env header, then highlight << 7 | code7 for every cell at ids[e][j*W + i]
__device__ __forceinline__ void full_ids(const RenderParams& p, int env0, int nvalid, FullHdr* hdr, uint8_t* ids)
{
    if ((int)threadIdx.x < nvalid) {
        const int env = env0 + threadIdx.x;
        const AgentState a = p.agents[env];
        const ClassDev cd = p.classes.c[a.cls];
        uint32_t lo[VIEW], hi[VIEW];
        FullHdr h;
        env_view(p, env, a, cd, lo, hi, h.vlo, h.vhi);
        if (!p.highlight) { h.vlo = 0; h.vhi = 0; }
        h.x = a.x; h.y = a.y; h.dir = a.dir; h.cls = a.cls;
        hdr[threadIdx.x] = h;
    }
    __syncthreads();
    // the highlight mask is vis_mask carried back to world coordinates (minigrid_env.py:671-699):
    // world cell = agent + f*(6-b) + r*(a-3)  <=>  b = 6 - (rel . f), a = 3 + (rel . r)
    // flat over (env, cell slot) so that small grids keep every thread busy
    for (uint32_t idx = threadIdx.x; idx < (uint32_t)nvalid * p.by_cells.d; idx += RENDER_THREADS) {
        const uint32_t e = qdiv(idx, p.by_cells), k = idx - e * p.by_cells.d;
        const FullHdr h = hdr[e];
        const int W = p.classes.c[h.cls].W, PR = p.classes.c[h.cls].PR;
        if (k >= (uint32_t)(W * p.classes.c[h.cls].H)) continue;
        const uint8_t* R = p.packed + (size_t)(env0 + e) * p.packed_stride + PAD;
        const int fx = dir_dx(h.dir), fy = dir_dy(h.dir);
        const unsigned long long vis = (unsigned long long)h.vlo | ((unsigned long long)h.vhi << 32);
        const int j = (int)qdiv(k, p.cls_w[h.cls]), i = (int)k - j * W;
        const int rx = i - (int)h.x, ry = j - (int)h.y;
        const int vb = (VIEW - 1) - (rx * fx + ry * fy), va = VIEW / 2 + (-rx * fy + ry * fx);
        const bool in_view = (unsigned)va < (unsigned)VIEW && (unsigned)vb < (unsigned)VIEW;
        const uint32_t lit = in_view ? (uint32_t)(vis >> (vb * 8 + va)) & 1u : 0u;
        ids[(size_t)e * p.cell_stride + k] = (uint8_t)((__ldg(R + j * PR + i) & 0x7Fu) | (lit << 7));
    }
    __syncthreads();
}

template <int G>
__global__ void __launch_bounds__(RENDER_THREADS) render_full_kernel(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    FullHdr* hdr = reinterpret_cast<FullHdr*>(smem);
    uint8_t* ids = smem + sizeof(FullHdr) * RENDER_ENVS;           // [envs_per_cta][cell_stride]
    const int env0 = blockIdx.x * p.envs_per_cta;
    const int nvalid = min(p.envs_per_cta, p.N - env0);
    full_ids(p, env0, nvalid, hdr, ids);
    const uint32_t trow = (uint32_t)p.ts * 3u;
    for (int e = 0; e < nvalid; ++e) {
        const FullHdr h = hdr[e];
        const ClassDev cd = p.classes.c[h.cls];
        const uint32_t ppr = (uint32_t)cd.W * p.by_ppt.d;                       // pieces per pixel row
        const uint32_t rows = (uint32_t)cd.H * (uint32_t)p.ts;
        const uint8_t* eid = ids + (size_t)e * p.cell_stride;
        uint8_t* out = p.out + (size_t)(env0 + e) * p.env_stride;
        const uint32_t agent_cell = (uint32_t)h.y * cd.W + h.x, agent_slot = ((uint32_t)h.dir + 1u) * 256u;
        const FastDiv by_ppr = p.cls_ppr[h.cls];
        for (uint32_t q = threadIdx.x; q < rows * ppr; q += RENDER_THREADS) {
            const uint32_t y = qdiv(q, by_ppr), xb = q - y * ppr;
            const uint32_t j = qdiv(y, p.by_ts), py = y - j * p.by_ts.d;
            const uint32_t i = qdiv(xb, p.by_ppt), pxb = xb - i * p.by_ppt.d;
            const uint32_t cell = j * cd.W + i;
            uint32_t t = eid[cell];
            if (cell == agent_cell) t += agent_slot;                            // grid.py:229-235
            copy_piece<G>(out + (size_t)q * G, p.atlas + (size_t)t * p.tile_pitch + py * trow + pxb * G);
        }
    }
}

// staged variant: strip = grid row j of one env = W tiles x ts pixel rows (see render_pov_staged_kernel)
template <int GP, int TS>
__global__ void __launch_bounds__(RENDER_THREADS, TS ? MG_FULL_MINB : 8) render_full_staged_kernel(const __grid_constant__ RenderParams p)
{
    extern __shared__ __align__(16) uint8_t smem[];
    FullHdr* hdr = reinterpret_cast<FullHdr*>(smem);
    uint8_t* ids = smem + sizeof(FullHdr) * RENDER_ENVS;
    const int env0 = blockIdx.x * p.envs_per_cta;
    const int nvalid = min(p.envs_per_cta, p.N - env0);
    full_ids(p, env0, nvalid, hdr, ids);
    const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t ts = TS ? TS : p.ts, trow = 3u * ts, ppt = trow / GP;
    uint8_t* buf = ids + (((size_t)p.envs_per_cta * p.cell_stride + 15) & ~(size_t)15) + (size_t)warp * p.strip_max;
    for (int e = 0; e < nvalid; ++e) {
        const FullHdr h = hdr[e];
        const ClassDev cd = p.classes.c[h.cls];
        const FastDiv by_w = p.cls_w[h.cls];
        const uint32_t items = (uint32_t)cd.W * ts, strip_bytes = items * trow;
        const uint8_t* eid = ids + (size_t)e * p.cell_stride;
        uint8_t* out = p.out + (size_t)(env0 + e) * p.env_stride;
        const uint32_t agent_slot = ((uint32_t)h.dir + 1u) * 256u;
        for (uint32_t j = warp; j < (uint32_t)cd.H; j += RENDER_THREADS / 32) {
            const uint8_t* rid = eid + j * cd.W;
            const uint32_t agent_i = (j == h.y) ? (uint32_t)h.x : 0xFFFFFFFFu;
            for (uint32_t it = lane; it < items; it += 32) {
                const uint32_t py = qdiv(it, by_w), i = it - py * cd.W;
                uint32_t t = rid[i];
                if (i == agent_i) t += agent_slot;                              // grid.py:229-235
                copy_tile_row<GP, TS>(buf + it * trow, p.atlas + (size_t)t * p.tile_pitch + py * trow, ppt);
            }
            __syncwarp();
            strip_out(out + (size_t)j * strip_bytes, buf, strip_bytes, lane);
            __syncwarp();
        }
    }
}

}  // namespace mg

// ---------------------------------------------------------------- C ABI

namespace {

#define MGR_CUDA(call)                                                                    \
    do {                                                                                  \
        cudaError_t _e = (call);                                                          \
        if (_e != cudaSuccess) return mg::set_error(MG_ECUDA, cudaGetErrorString(_e));    \
    } while (0)

int piece_bytes(int ts, const void* out, long long env_stride)
{
    const int trow = ts * 3;
    const uintptr_t a = (uintptr_t)out;
    for (int g = 16; g >= 4; g >>= 1)
        if (trow % g == 0 && a % g == 0 && env_stride % g == 0) return g;
    return 1;
}

// The staged kernels need 16-byte aligned strips (tile size a multiple of 4) that fit a warp's shared-memory buffer.
bool staged_ok(int ts, size_t strip_bytes, const void* out, long long env_stride)
{
    return ts % 4 == 0 && strip_bytes <= 12 * 1024 && (uintptr_t)out % 16 == 0 && env_stride % 16 == 0;
}

// the atlas of one tile size, built on first use (Grid.tile_cache, core/grid.py:159-163, filled eagerly)
int get_atlas(mg_engine* h, int ts, cudaStream_t stream, const mg_atlas** out)
{
    for (int i = 0; i < h->num_atlases; ++i)
        if (h->atlases[i].tile_size == ts) { *out = &h->atlases[i]; return MG_OK; }
    if (h->num_atlases == 8) return mg::set_error(MG_EINVAL, "render: at most 8 different tile sizes per handle");
    mg_atlas a;
    a.tile_size = ts;
    a.tile_pitch = (ts * ts * 3 + 15) & ~15;
    a.dev = nullptr;
    MGR_CUDA(cudaMalloc(&a.dev, (size_t)mg::ATLAS_TILES * a.tile_pitch));
    mg::raster::Rotations rot;
    for (int k = 0; k < 4; ++k) {
        const double theta = 0.5 * M_PI * k;                 // grid.py:184; host libm = what CPython's math module calls
        rot.cs[k] = std::cos(-theta); rot.sn[k] = std::sin(-theta);
    }
    mg::atlas_kernel<<<mg::ATLAS_TILES, 128, 0, stream>>>(a.dev, ts, a.tile_pitch, rot);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) { cudaFree(a.dev); return mg::set_error(MG_ECUDA, cudaGetErrorString(e)); }
    h->launches++;
    h->atlases[h->num_atlases] = a;
    *out = &h->atlases[h->num_atlases++];
    return MG_OK;
}

mg::RenderParams base_params(mg_engine* h, const mg_atlas* a, uint8_t* out, long long env_stride)
{
    mg::RenderParams p{};
    p.packed = h->packed; p.agents = h->agents; p.atlas = a->dev; p.out = out; p.env_stride = env_stride;
    p.N = h->N; p.packed_stride = h->packed_stride; p.ts = a->tile_size; p.tile_pitch = a->tile_pitch;
    p.cell_stride = h->plane_stride; p.envs_per_cta = mg::RENDER_ENVS;
    p.classes = h->classes;
    return p;
}

}  // namespace

extern "C" {

int mg_render_pov(mg_engine* h, int32_t tile_size, uint8_t* image_dev, void* stream)
{
    if (!h || !image_dev || tile_size < 1 || tile_size > 64) return mg::set_error(MG_EINVAL, "mg_render_pov: bad argument (tile_size 1..64)");
    if (h->P == 0) return mg::set_error(MG_ESTATE, "mg_render_pov: no layouts uploaded / envs never reset");
    MGR_CUDA(cudaSetDevice(h->device));
    const cudaStream_t s = (cudaStream_t)stream;
    const mg_atlas* a = nullptr;
    int rc = get_atlas(h, tile_size, s, &a);
    if (rc) return rc;
    const int ts = tile_size;
    const long long frame = (long long)MG_VIEW * ts * MG_VIEW * ts * 3;
    const int G = piece_bytes(ts, image_dev, frame);
    mg::RenderParams p = base_params(h, a, image_dev, frame);
    p.by_ts = mg::make_fastdiv(ts);
    p.by_ppt = mg::make_fastdiv(ts * 3 / G);
    p.by_ppr = mg::make_fastdiv(MG_VIEW * ts * 3 / G);
    p.by_pieces = mg::make_fastdiv((uint32_t)(frame / G));
    const int grid = (h->N + mg::RENDER_ENVS - 1) / mg::RENDER_ENVS;
    const size_t strip = (size_t)MG_VIEW * ts * ts * 3;
    if (G != 16 && staged_ok(ts, strip, image_dev, frame)) {
        const size_t smem = (size_t)mg::RENDER_ENVS * MG_VIEW * 8 + (mg::RENDER_THREADS / 32) * strip;
#define MG_LAUNCH_POV(KERNEL)                                                                                     \
    do {                                                                                                         \
        MGR_CUDA(cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));          \
        KERNEL<<<grid, mg::RENDER_THREADS, smem, s>>>(p);                                                        \
    } while (0)
        if (ts == 8) MG_LAUNCH_POV((mg::render_pov_staged_kernel<8, 8>));      // the wrappers' default tile size
        else if (ts % 8 == 0) MG_LAUNCH_POV((mg::render_pov_staged_kernel<8, 0>));
        else MG_LAUNCH_POV((mg::render_pov_staged_kernel<4, 0>));
#undef MG_LAUNCH_POV
    } else {
        switch (G) {
        case 16: mg::render_pov_kernel<16><<<grid, mg::RENDER_THREADS, 0, s>>>(p); break;
        case 8: mg::render_pov_kernel<8><<<grid, mg::RENDER_THREADS, 0, s>>>(p); break;
        case 4: mg::render_pov_kernel<4><<<grid, mg::RENDER_THREADS, 0, s>>>(p); break;
        default: mg::render_pov_kernel<1><<<grid, mg::RENDER_THREADS, 0, s>>>(p); break;
        }
    }
    MGR_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int64_t mg_render_full_stride(const mg_engine* h, int32_t tile_size)
{
    if (!h || tile_size < 1) return 0;
    int64_t m = 0;
    for (int i = 0; i < h->C; ++i) {
        const int64_t b = (int64_t)h->classes.c[i].W * h->classes.c[i].H * tile_size * tile_size * 3;
        if (b > m) m = b;
    }
    return (m + 15) & ~(int64_t)15;
}

int mg_render_full(mg_engine* h, int32_t tile_size, int32_t highlight, uint8_t* image_dev, void* stream)
{
    if (!h || !image_dev || tile_size < 1 || tile_size > 64) return mg::set_error(MG_EINVAL, "mg_render_full: bad argument (tile_size 1..64)");
    if (h->P == 0) return mg::set_error(MG_ESTATE, "mg_render_full: no layouts uploaded / envs never reset");
    MGR_CUDA(cudaSetDevice(h->device));
    const cudaStream_t s = (cudaStream_t)stream;
    const mg_atlas* a = nullptr;
    int rc = get_atlas(h, tile_size, s, &a);
    if (rc) return rc;
    const int ts = tile_size;
    const long long stride = mg_render_full_stride(h, ts);
    const int G = piece_bytes(ts, image_dev, stride);
    mg::RenderParams p = base_params(h, a, image_dev, stride);
    p.highlight = highlight != 0;
    p.by_ts = mg::make_fastdiv(ts);
    p.by_ppt = mg::make_fastdiv(ts * 3 / G);
    for (int i = 0; i < h->C; ++i) p.cls_ppr[i] = mg::make_fastdiv((uint32_t)h->classes.c[i].W * (ts * 3 / G));
    for (int i = 0; i < h->C; ++i) p.cls_w[i] = mg::make_fastdiv((uint32_t)h->classes.c[i].W);
    p.by_cells = mg::make_fastdiv((uint32_t)h->plane_stride);
    int maxw = 0;
    for (int i = 0; i < h->C; ++i) maxw = h->classes.c[i].W > maxw ? h->classes.c[i].W : maxw;
    const size_t strip = (size_t)maxw * ts * ts * 3;
    const bool staged = G != 16 && staged_ok(ts, strip, image_dev, stride);
    const size_t strips = staged ? (mg::RENDER_THREADS / 32) * strip + 16 : 0;
    p.strip_max = (int)strip;
    // shared memory: headers + one id byte per cell per env (+ the strip buffers); fewer envs per CTA when the grid is large
    const size_t budget = 64 * 1024;
    int epc = (int)((budget - sizeof(mg::FullHdr) * mg::RENDER_ENVS) / (size_t)h->plane_stride);
    if (epc > MG_FULL_EPC) epc = MG_FULL_EPC;
    if (epc < 1) epc = 1;
    p.envs_per_cta = epc;
    const size_t smem = sizeof(mg::FullHdr) * mg::RENDER_ENVS + (size_t)epc * h->plane_stride + strips;
    const int grid = (h->N + epc - 1) / epc;
#define MG_LAUNCH_FULL(KERNEL)                                                                                   \
    do {                                                                                                         \
        MGR_CUDA(cudaFuncSetAttribute(KERNEL, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));          \
        KERNEL<<<grid, mg::RENDER_THREADS, smem, s>>>(p);                                                        \
    } while (0)
    if (staged) {
        if (ts == 8) MG_LAUNCH_FULL((mg::render_full_staged_kernel<8, 8>));
        else if (ts % 8 == 0) MG_LAUNCH_FULL((mg::render_full_staged_kernel<8, 0>));
        else MG_LAUNCH_FULL((mg::render_full_staged_kernel<4, 0>));
    } else {
        switch (G) {
        case 16: MG_LAUNCH_FULL(mg::render_full_kernel<16>); break;
        case 8: MG_LAUNCH_FULL(mg::render_full_kernel<8>); break;
        case 4: MG_LAUNCH_FULL(mg::render_full_kernel<4>); break;
        default: MG_LAUNCH_FULL(mg::render_full_kernel<1>); break;
        }
    }
#undef MG_LAUNCH_FULL
    MGR_CUDA(cudaGetLastError());
    h->launches++;
    return MG_OK;
}

int mg_render_atlas(mg_engine* h, int32_t tile_size, uint8_t* tiles_host)
{
    if (!h || !tiles_host || tile_size < 1 || tile_size > 64) return mg::set_error(MG_EINVAL, "mg_render_atlas: bad argument");
    MGR_CUDA(cudaSetDevice(h->device));
    const mg_atlas* a = nullptr;
    int rc = get_atlas(h, tile_size, nullptr, &a);
    if (rc) return rc;
    MGR_CUDA(cudaMemcpy2D(tiles_host, (size_t)tile_size * tile_size * 3, a->dev, a->tile_pitch,
                          (size_t)tile_size * tile_size * 3, mg::ATLAS_TILES, cudaMemcpyDeviceToHost));
    return MG_OK;
}

}  // extern "C"
