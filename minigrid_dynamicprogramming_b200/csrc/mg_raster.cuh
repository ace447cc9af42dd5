// mg_raster.cuh — per-pixel rasteriser for the tile atlas (Grid.render_tile, core/grid.py:145-198).
//
// The reference paints a (3*ts)x(3*ts) uint8 image fill by fill (utils/rendering.py:25-37: every subpixel is
// tested against every shape predicate), blends the highlight, then box-filters 3x3 down to ts x ts.  All of
// that is independent per output pixel, so here one thread owns one output pixel: it walks the tile's display
// list in drawing order for its nine subpixels (the last shape covering a subpixel wins), blends, and averages.
//
// The predicates are floating point.  numpy's promotion rules decide which operands are float32 (the np.array(...,
// dtype=np.float32) vertices of point_in_line / point_in_triangle, rendering.py:54-55,101-103) and which are float64
// (the subpixel coordinates); every product and sum below is written as a separately rounded operation in that
// type (no FMA contraction), so the in/out decisions — and therefore the bytes — match the reference's.
//
// __host__ __device__: tests/host builds the same code for the CPU and compares it with the reference's tiles.
#pragma once
#include "mg_common.cuh"

namespace mg {
namespace raster {

#ifdef __CUDA_ARCH__
__device__ __forceinline__ double dmul(double a, double b) { return __dmul_rn(a, b); }
__device__ __forceinline__ double dadd(double a, double b) { return __dadd_rn(a, b); }
__device__ __forceinline__ double dsub(double a, double b) { return __dsub_rn(a, b); }
__device__ __forceinline__ double ddiv(double a, double b) { return __ddiv_rn(a, b); }
__device__ __forceinline__ double dsqrt(double a) { return __dsqrt_rn(a); }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float fsqrt(float a) { return __fsqrt_rn(a); }
#else   // host twins: compile with -ffp-contract=off
inline double dmul(double a, double b) { volatile double r = a * b; return r; }
inline double dadd(double a, double b) { volatile double r = a + b; return r; }
inline double dsub(double a, double b) { volatile double r = a - b; return r; }
inline double ddiv(double a, double b) { volatile double r = a / b; return r; }
inline double dsqrt(double a) { return __builtin_sqrt(a); }
inline float fmul(float a, float b) { volatile float r = a * b; return r; }
inline float fadd(float a, float b) { volatile float r = a + b; return r; }
inline float fsub(float a, float b) { volatile float r = a - b; return r; }
inline float fdiv(float a, float b) { volatile float r = a / b; return r; }
inline float fsqrt(float a) { return __builtin_sqrtf(a); }
#endif

// cos(-theta), sin(-theta) for theta = 0.5*pi*agent_dir (grid.py:184), evaluated by the HOST libm — the function
// CPython's math.cos/math.sin call — and handed to the kernel, so no device libm rounding enters the picture.
struct Rotations { double cs[4], sn[4]; };

// core/constants.py:8-15 in COLOR_TO_IDX order (:20); r | g<<8 | b<<16
__host__ __device__ __forceinline__ uint32_t color_rgb(int c)
{
    switch (c) {
    case 0: return 255u;
    case 1: return 255u << 8;
    case 2: return 255u << 16;
    case 3: return 112u | (39u << 8) | (195u << 16);
    case 4: return 255u | (255u << 8);
    default: return 100u | (100u << 8) | (100u << 16);
    }
}
// per-channel (uint8)(channel * k) as numpy casts a float colour on assignment (Floor: /2 -> k = 0.5, world_object.py:134;
// locked door: k = 0.45, :226)
__host__ __device__ __forceinline__ uint32_t scale_rgb(uint32_t rgb, double k)
{
    uint32_t o = 0;
    for (int ch = 0; ch < 3; ++ch) {
        const double v = dmul(k, (double)((rgb >> (8 * ch)) & 0xffu));
        o |= (uint32_t)(int)v << (8 * ch);
    }
    return o;
}

__host__ __device__ __forceinline__ bool in_rect(double x, double y, double xmin, double xmax, double ymin, double ymax)
{   // rendering.py:93-97
    return x >= xmin && x <= xmax && y >= ymin && y <= ymax;
}
__host__ __device__ __forceinline__ bool in_circle(double x, double y, double cx, double cy, double r)
{   // rendering.py:86-90
    const double dx = dsub(x, cx), dy = dsub(y, cy);
    return dadd(dmul(dx, dx), dmul(dy, dy)) <= dmul(r, r);
}
__host__ __device__ inline bool in_line(double x, double y, double x0, double y0, double x1, double y1, double r)
{   // rendering.py:53-83
    const float p0x = (float)x0, p0y = (float)y0;
    float dx = fsub((float)x1, p0x), dy = fsub((float)y1, p0y);
    const float dist = fsqrt(fadd(fmul(dx, dx), fmul(dy, dy)));            // np.linalg.norm of a float32 vector
    dx = fdiv(dx, dist); dy = fdiv(dy, dist);
    const double xmin = dsub(x0 < x1 ? x0 : x1, r), xmax = dadd(x0 > x1 ? x0 : x1, r);
    const double ymin = dsub(y0 < y1 ? y0 : y1, r), ymax = dadd(y0 > y1 ? y0 : y1, r);
    if (x < xmin || x > xmax || y < ymin || y > ymax) return false;
    const double pqx = dsub(x, (double)p0x), pqy = dsub(y, (double)p0y);
    double a = dadd(dmul(pqx, (double)dx), dmul(pqy, (double)dy));
    const double hi = (double)dist;
    a = a < 0.0 ? 0.0 : (a > hi ? hi : a);                                  // np.clip(a, 0, dist)
    const double px = dadd((double)p0x, dmul(a, (double)dx)), py = dadd((double)p0y, dmul(a, (double)dy));
    const double ex = dsub(x, px), ey = dsub(y, py);
    return dsqrt(dadd(dmul(ex, ex), dmul(ey, ey))) <= r;
}
// point_in_triangle((0.12,0.19),(0.87,0.50),(0.12,0.81)) — the agent marker (grid.py:176-180), rendering.py:100-124
__host__ __device__ inline bool in_agent_triangle(double x, double y)
{
    const float ax = (float)0.12, ay = (float)0.19, bx = (float)0.87, by = (float)0.50, cx = (float)0.12, cy = (float)0.81;
    const float v0x = fsub(cx, ax), v0y = fsub(cy, ay), v1x = fsub(bx, ax), v1y = fsub(by, ay);
    const float dot00 = fadd(fmul(v0x, v0x), fmul(v0y, v0y));
    const float dot01 = fadd(fmul(v0x, v1x), fmul(v0y, v1y));
    const float dot11 = fadd(fmul(v1x, v1x), fmul(v1y, v1y));
    const float inv_denom = fdiv(1.0f, fsub(fmul(dot00, dot11), fmul(dot01, dot01)));
    const double v2x = dsub(x, (double)ax), v2y = dsub(y, (double)ay);
    const double dot02 = dadd(dmul((double)v0x, v2x), dmul((double)v0y, v2y));
    const double dot12 = dadd(dmul((double)v1x, v2x), dmul((double)v1y, v2y));
    const double u = dmul(dsub(dmul((double)dot11, dot02), dmul((double)dot01, dot12)), (double)inv_denom);
    const double v = dmul(dsub(dmul((double)dot00, dot12), dmul((double)dot01, dot02)), (double)inv_denom);
    return (u >= 0) && (v >= 0) && dadd(u, v) < 1;
}

// Colour of one subpixel of the supersampled tile before highlighting: the display list of render_tile.
// type == T_EMPTY draws no object; agent_dir < 0 draws no agent.
__host__ __device__ inline uint32_t subpixel(int type, int color, int state, int agent_dir, const Rotations& rot,
                                             double x, double y)
{
    uint32_t px = 0;
    const uint32_t GREY = 100u | (100u << 8) | (100u << 16);
    if (in_rect(x, y, 0, 0.031, 0, 1)) px = GREY;                          // grid lines, grid.py:169-171
    if (in_rect(x, y, 0, 1, 0, 0.031)) px = GREY;
    const uint32_t c = color_rgb(color);
    switch (type) {
    case T_GOAL: case T_WALL:                                              // world_object.py:117-118, 168-169
        if (in_rect(x, y, 0, 1, 0, 1)) px = c;
        break;
    case T_FLOOR:                                                          // :132-135
        if (in_rect(x, y, 0.031, 1, 0.031, 1)) px = scale_rgb(c, 0.5);
        break;
    case T_LAVA:                                                           // :145-158
        if (in_rect(x, y, 0, 1, 0, 1)) px = 255u | (128u << 8);
        for (int i = 0; i < 3; ++i) {
            const double ylo = dadd(0.3, dmul(0.2, (double)i)), yhi = dadd(0.4, dmul(0.2, (double)i));
            if (in_line(x, y, 0.1, ylo, 0.3, yhi, 0.03)) px = 0;
            if (in_line(x, y, 0.3, yhi, 0.5, ylo, 0.03)) px = 0;
            if (in_line(x, y, 0.5, ylo, 0.7, yhi, 0.03)) px = 0;
            if (in_line(x, y, 0.7, yhi, 0.9, ylo, 0.03)) px = 0;
        }
        break;
    case T_DOOR:                                                           // :215-237
        if (state == S_OPEN) {
            if (in_rect(x, y, 0.88, 1.00, 0.00, 1.00)) px = c;
            if (in_rect(x, y, 0.92, 0.96, 0.04, 0.96)) px = 0;
        } else if (state == S_LOCKED) {
            if (in_rect(x, y, 0.00, 1.00, 0.00, 1.00)) px = c;
            if (in_rect(x, y, 0.06, 0.94, 0.06, 0.94)) px = scale_rgb(c, 0.45);
            if (in_rect(x, y, 0.52, 0.75, 0.50, 0.56)) px = c;
        } else {
            if (in_rect(x, y, 0.00, 1.00, 0.00, 1.00)) px = c;
            if (in_rect(x, y, 0.04, 0.96, 0.04, 0.96)) px = 0;
            if (in_rect(x, y, 0.08, 0.92, 0.08, 0.92)) px = c;
            if (in_rect(x, y, 0.12, 0.88, 0.12, 0.88)) px = 0;
            if (in_circle(x, y, 0.75, 0.50, 0.08)) px = c;
        }
        break;
    case T_KEY:                                                            // :247-259
        if (in_rect(x, y, 0.50, 0.63, 0.31, 0.88)) px = c;
        if (in_rect(x, y, 0.38, 0.50, 0.59, 0.66)) px = c;
        if (in_rect(x, y, 0.38, 0.50, 0.81, 0.88)) px = c;
        if (in_circle(x, y, 0.56, 0.28, 0.190)) px = c;
        if (in_circle(x, y, 0.56, 0.28, 0.064)) px = 0;
        break;
    case T_BALL:                                                           // :267-268
        if (in_circle(x, y, 0.5, 0.5, 0.31)) px = c;
        break;
    case T_BOX:                                                            // :280-289
        if (in_rect(x, y, 0.12, 0.88, 0.12, 0.88)) px = c;
        if (in_rect(x, y, 0.18, 0.82, 0.18, 0.82)) px = 0;
        if (in_rect(x, y, 0.16, 0.84, 0.47, 0.53)) px = c;
        break;
    default: break;
    }
    if (agent_dir >= 0) {                                                  // grid.py:176-185 + rotate_fn, rendering.py:40-50
        const double cs = rot.cs[agent_dir], sn = rot.sn[agent_dir];
        const double xr = dsub(x, 0.5), yr = dsub(y, 0.5);
        const double x2 = dsub(dadd(0.5, dmul(xr, cs)), dmul(yr, sn));
        const double y2 = dadd(dadd(0.5, dmul(yr, cs)), dmul(xr, sn));
        if (in_agent_triangle(x2, y2)) px = 255u;
    }
    return px;
}

// One output pixel (px, py) of a ts x ts tile: r | g<<8 | b<<16.
__host__ __device__ inline uint32_t tile_pixel(int type, int color, int state, int agent_dir, bool highlight,
                                               const Rotations& rot, int ts, int px, int py)
{
    const double n = (double)(3 * ts);                                     // subdivs = 3, grid.py:151
    double rows[3][3];                                                     // [channel][sub-row] mean over the 3 x-subpixels
    for (int sy = 0; sy < 3; ++sy) {
        int acc[3] = {0, 0, 0};
        for (int sx = 0; sx < 3; ++sx) {
            const double yf = ddiv(dadd((double)(3 * py + sy), 0.5), n);   // rendering.py:32-33
            const double xf = ddiv(dadd((double)(3 * px + sx), 0.5), n);
            const uint32_t s = subpixel(type, color, state, agent_dir, rot, xf, yf);
            for (int ch = 0; ch < 3; ++ch) {
                int v = (int)((s >> (8 * ch)) & 0xffu);
                if (highlight) {                                           // highlight_img, rendering.py:127-135
                    double b = dadd((double)v, dmul(0.30, (double)(255 - v)));
                    b = b < 0.0 ? 0.0 : (b > 255.0 ? 255.0 : b);
                    v = (int)b;
                }
                acc[ch] += v;
            }
        }
        for (int ch = 0; ch < 3; ++ch) rows[ch][sy] = ddiv((double)acc[ch], 3.0);   // downsample, rendering.py:19-20
    }
    uint32_t o = 0;
    for (int ch = 0; ch < 3; ++ch) {
        const double m = ddiv(dadd(dadd(rows[ch][0], rows[ch][1]), rows[ch][2]), 3.0);
        o |= (uint32_t)(int)m << (8 * ch);                                 // uint8 cast of Grid.render's assignment, grid.py:240
    }
    return o;
}

}  // namespace raster
}  // namespace mg
