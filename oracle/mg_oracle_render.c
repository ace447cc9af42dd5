/*
 * mg_oracle_render.c — CPU restatement of the reference's RGB rendering.  TEST INFRASTRUCTURE ONLY
 * (same rules as mg_oracle.c: only tests/, smoke() and bench.py's CPU legs may load it).
 *
 * Follows the reference LITERALLY: a (3*ts)x(3*ts) uint8 supersampled image per tile, the shape
 * predicates of minigrid/utils/rendering.py evaluated at every subpixel for every fill in drawing
 * order, the highlight blend, then the 3x3 box downsample — with numpy's dtype promotion written
 * out (which operands are float32, which are float64) because the in/out tests are floating point.
 *
 * Parity pin: tests/golden/atlas.npz (every tile of Grid.render_tile at four tile sizes) and
 * tests/golden/frames.npz (get_frame of real episodes), recorded by oracle/gen_golden_render.py
 * from the unmodified reference; tests/test_oracle_golden.py compares.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { T_EMPTY = 1, T_WALL = 2, T_FLOOR = 3, T_DOOR = 4, T_KEY = 5, T_BALL = 6, T_BOX = 7, T_GOAL = 8, T_LAVA = 9 };
enum { CFG_W = 0, CFG_H = 1, CFG_MAX_STEPS = 2, CFG_SEE_THROUGH = 3, CFG_RULE = 4, CFG_VIEW = 5 };
enum { ST_X = 0, ST_Y = 1, ST_DIR = 2, ST_STEPS = 3, ST_CT = 4, ST_CC = 5, ST_CS = 6, ST_CA = 7 };

void mgo_gen_obs_grid(const uint8_t* cells, const int32_t* cfg, const int32_t* st, uint8_t* g4, uint8_t* vis);

/* core/constants.py:8-15, indexed by COLOR_TO_IDX (:20) */
static const double COLORS[6][3] = { {255, 0, 0}, {0, 255, 0}, {0, 0, 255}, {112, 39, 195}, {255, 255, 0}, {100, 100, 100} };

typedef struct { uint8_t* px; int n; } img_t;           /* n x n x 3 */

/* ---- shape predicates, utils/rendering.py ---- */
typedef struct shape shape_t;
struct shape {
    int kind;                       /* 0 rect, 1 circle, 2 line, 3 triangle, 4 rotated(triangle) */
    double xmin, xmax, ymin, ymax;  /* rect; line's bounding box */
    double cx, cy, r;               /* circle; line radius; rotation centre */
    float p0[2], dir[2], dist;      /* line (float32 arrays, rendering.py:54-58) */
    float a[2], v0[2], v1[2], dot00, dot01, dot11, inv_denom;   /* triangle */
    double cs, sn;                  /* cos(-theta), sin(-theta) */
};

static shape_t rect(double xmin, double xmax, double ymin, double ymax)
{   /* point_in_rect, rendering.py:93-97 */
    shape_t s; memset(&s, 0, sizeof s); s.kind = 0; s.xmin = xmin; s.xmax = xmax; s.ymin = ymin; s.ymax = ymax; return s;
}
static shape_t circle(double cx, double cy, double r)
{   /* point_in_circle, rendering.py:86-90 */
    shape_t s; memset(&s, 0, sizeof s); s.kind = 1; s.cx = cx; s.cy = cy; s.r = r; return s;
}
static shape_t line(double x0, double y0, double x1, double y1, double r)
{   /* point_in_line, rendering.py:53-64 */
    shape_t s; memset(&s, 0, sizeof s); s.kind = 2;
    s.p0[0] = (float)x0; s.p0[1] = (float)y0;
    const float p1[2] = { (float)x1, (float)y1 };
    float d[2] = { p1[0] - s.p0[0], p1[1] - s.p0[1] };
    volatile float q0 = d[0] * d[0], q1 = d[1] * d[1];
    s.dist = sqrtf(q0 + q1);                                  /* np.linalg.norm on float32 */
    s.dir[0] = d[0] / s.dist; s.dir[1] = d[1] / s.dist;
    s.xmin = (x0 < x1 ? x0 : x1) - r; s.xmax = (x0 > x1 ? x0 : x1) + r;
    s.ymin = (y0 < y1 ? y0 : y1) - r; s.ymax = (y0 > y1 ? y0 : y1) + r;
    s.r = r;
    return s;
}
static shape_t triangle(double ax, double ay, double bx, double by, double cx, double cy)
{   /* point_in_triangle, rendering.py:100-104 + the point-independent part of fn */
    shape_t s; memset(&s, 0, sizeof s); s.kind = 3;
    const float a[2] = { (float)ax, (float)ay }, b[2] = { (float)bx, (float)by }, c[2] = { (float)cx, (float)cy };
    s.a[0] = a[0]; s.a[1] = a[1];
    s.v0[0] = c[0] - a[0]; s.v0[1] = c[1] - a[1];
    s.v1[0] = b[0] - a[0]; s.v1[1] = b[1] - a[1];
    volatile float m0, m1;
    m0 = s.v0[0] * s.v0[0]; m1 = s.v0[1] * s.v0[1]; s.dot00 = m0 + m1;
    m0 = s.v0[0] * s.v1[0]; m1 = s.v0[1] * s.v1[1]; s.dot01 = m0 + m1;
    m0 = s.v1[0] * s.v1[0]; m1 = s.v1[1] * s.v1[1]; s.dot11 = m0 + m1;
    m0 = s.dot00 * s.dot11; m1 = s.dot01 * s.dot01;
    volatile float den = m0 - m1;
    s.inv_denom = 1.0f / den;                                 /* python int / np.float32 -> float32 */
    return s;
}

static int inside(const shape_t* s, double x, double y)
{
    switch (s->kind) {
    case 0: return x >= s->xmin && x <= s->xmax && y >= s->ymin && y <= s->ymax;
    case 1: {
        volatile double dx = x - s->cx, dy = y - s->cy;
        volatile double a = dx * dx, b = dy * dy, rr = s->r * s->r;
        return a + b <= rr;
    }
    case 2: {   /* rendering.py:66-81 */
        if (x < s->xmin || x > s->xmax || y < s->ymin || y > s->ymax) return 0;
        const double pq0 = x - (double)s->p0[0], pq1 = y - (double)s->p0[1];
        volatile double t0 = pq0 * (double)s->dir[0], t1 = pq1 * (double)s->dir[1];
        double a = t0 + t1;
        const double hi = (double)s->dist;
        a = a < 0.0 ? 0.0 : (a > hi ? hi : a);                                /* np.clip(a, 0, dist) */
        volatile double m0 = a * (double)s->dir[0], m1 = a * (double)s->dir[1];
        const double p0 = (double)s->p0[0] + m0, p1 = (double)s->p0[1] + m1;
        const double e0 = x - p0, e1 = y - p1;
        volatile double s0 = e0 * e0, s1 = e1 * e1;
        return sqrt(s0 + s1) <= s->r;
    }
    case 3: {   /* rendering.py:105-124 */
        const double v20 = x - (double)s->a[0], v21 = y - (double)s->a[1];
        volatile double t0, t1;
        t0 = (double)s->v0[0] * v20; t1 = (double)s->v0[1] * v21; const double dot02 = t0 + t1;
        t0 = (double)s->v1[0] * v20; t1 = (double)s->v1[1] * v21; const double dot12 = t0 + t1;
        t0 = (double)s->dot11 * dot02; t1 = (double)s->dot01 * dot12;
        const double u = (t0 - t1) * (double)s->inv_denom;
        t0 = (double)s->dot00 * dot12; t1 = (double)s->dot01 * dot02;
        const double v = (t0 - t1) * (double)s->inv_denom;
        return (u >= 0) && (v >= 0) && (u + v) < 1;
    }
    default: return 0;
    }
}

/* rotate_fn, rendering.py:40-50, around a triangle */
static int inside_rotated(const shape_t* tri, double cx, double cy, double cs, double sn, double x, double y)
{
    x = x - cx; y = y - cy;
    volatile double xc = x * cs, ys = y * sn, yc = y * cs, xs = x * sn;
    const double x2 = cx + xc - ys;
    const double y2 = cy + yc + xs;
    return inside(tri, x2, y2);
}

/* fill_coords, rendering.py:25-37; the colour is cast to uint8 on assignment */
static void fill(img_t* im, const shape_t* s, const double col[3])
{
    for (int y = 0; y < im->n; y++)
        for (int x = 0; x < im->n; x++) {
            const double yf = (y + 0.5) / im->n, xf = (x + 0.5) / im->n;
            if (inside(s, xf, yf))
                for (int k = 0; k < 3; k++) im->px[(y * im->n + x) * 3 + k] = (uint8_t)col[k];
        }
}
static void fill3(img_t* im, shape_t s, double r, double g, double b)
{
    const double c[3] = { r, g, b };
    fill(im, &s, c);
}
static void fillc(img_t* im, shape_t s, const double c[3]) { fill(im, &s, c); }

/* WorldObj.render of each class, core/world_object.py */
static void render_obj(img_t* im, int type, int color, int state)
{
    const double* c = COLORS[color < 6 ? color : 5];
    switch (type) {
    case T_GOAL: fillc(im, rect(0, 1, 0, 1), c); break;                                   /* :117-118 */
    case T_FLOOR: {                                                                       /* :132-135 */
        const double h[3] = { c[0] / 2, c[1] / 2, c[2] / 2 };
        fillc(im, rect(0.031, 1, 0.031, 1), h);
        break;
    }
    case T_LAVA:                                                                          /* :145-158 */
        fill3(im, rect(0, 1, 0, 1), 255, 128, 0);
        for (int i = 0; i < 3; i++) {
            const double ylo = 0.3 + 0.2 * i, yhi = 0.4 + 0.2 * i;
            fill3(im, line(0.1, ylo, 0.3, yhi, 0.03), 0, 0, 0);
            fill3(im, line(0.3, yhi, 0.5, ylo, 0.03), 0, 0, 0);
            fill3(im, line(0.5, ylo, 0.7, yhi, 0.03), 0, 0, 0);
            fill3(im, line(0.7, yhi, 0.9, ylo, 0.03), 0, 0, 0);
        }
        break;
    case T_WALL: fillc(im, rect(0, 1, 0, 1), c); break;                                   /* :168-169 */
    case T_DOOR:                                                                          /* :215-237 */
        if (state == 0) {
            fillc(im, rect(0.88, 1.00, 0.00, 1.00), c);
            fill3(im, rect(0.92, 0.96, 0.04, 0.96), 0, 0, 0);
        } else if (state == 2) {
            const double d[3] = { 0.45 * c[0], 0.45 * c[1], 0.45 * c[2] };
            fillc(im, rect(0.00, 1.00, 0.00, 1.00), c);
            fillc(im, rect(0.06, 0.94, 0.06, 0.94), d);
            fillc(im, rect(0.52, 0.75, 0.50, 0.56), c);
        } else {
            fillc(im, rect(0.00, 1.00, 0.00, 1.00), c);
            fill3(im, rect(0.04, 0.96, 0.04, 0.96), 0, 0, 0);
            fillc(im, rect(0.08, 0.92, 0.08, 0.92), c);
            fill3(im, rect(0.12, 0.88, 0.12, 0.88), 0, 0, 0);
            fillc(im, circle(0.75, 0.50, 0.08), c);
        }
        break;
    case T_KEY:                                                                           /* :247-259 */
        fillc(im, rect(0.50, 0.63, 0.31, 0.88), c);
        fillc(im, rect(0.38, 0.50, 0.59, 0.66), c);
        fillc(im, rect(0.38, 0.50, 0.81, 0.88), c);
        fillc(im, circle(0.56, 0.28, 0.190), c);
        fill3(im, circle(0.56, 0.28, 0.064), 0, 0, 0);
        break;
    case T_BALL: fillc(im, circle(0.5, 0.5, 0.31), c); break;                             /* :267-268 */
    case T_BOX:                                                                           /* :280-289 */
        fillc(im, rect(0.12, 0.88, 0.12, 0.88), c);
        fill3(im, rect(0.18, 0.82, 0.18, 0.82), 0, 0, 0);
        fillc(im, rect(0.16, 0.84, 0.47, 0.53), c);
        break;
    default: break;
    }
}

/* Grid.render_tile, core/grid.py:145-198 (+ the uint8 cast of Grid.render's assignment, :240).
 * type == T_EMPTY means obj is None; agent_dir < 0 means None.  out: uint8[ts][ts][3]. */
void mgo_render_tile(int type, int color, int state, int agent_dir, int highlight, int ts, uint8_t* out)
{
    const int n = ts * 3;                                   /* subdivs = 3 */
    img_t im = { (uint8_t*)calloc((size_t)n * n * 3, 1), n };
    fill3(&im, rect(0, 0.031, 0, 1), 100, 100, 100);
    fill3(&im, rect(0, 1, 0, 0.031), 100, 100, 100);
    if (type != T_EMPTY) render_obj(&im, type, color, state);
    if (agent_dir >= 0) {
        const shape_t tri = triangle(0.12, 0.19, 0.87, 0.50, 0.12, 0.81);
        const double theta = 0.5 * M_PI * agent_dir;
        const double cs = cos(-theta), sn = sin(-theta);
        for (int y = 0; y < n; y++)
            for (int x = 0; x < n; x++) {
                const double yf = (y + 0.5) / n, xf = (x + 0.5) / n;
                if (inside_rotated(&tri, 0.5, 0.5, cs, sn, xf, yf)) {
                    uint8_t* p = im.px + (y * n + x) * 3;
                    p[0] = 255; p[1] = 0; p[2] = 0;
                }
            }
    }
    if (highlight)                                          /* highlight_img, rendering.py:127-135 */
        for (int k = 0; k < n * n * 3; k++) {
            const uint8_t v = im.px[k];
            volatile double d = 0.30 * (double)(uint8_t)(255 - v);
            double b = (double)v + d;
            b = b < 0 ? 0 : (b > 255 ? 255 : b);
            im.px[k] = (uint8_t)b;
        }
    /* downsample, rendering.py:8-22: mean over the x-subpixels, then over the y-subpixels, in float64 */
    for (int y = 0; y < ts; y++)
        for (int x = 0; x < ts; x++)
            for (int k = 0; k < 3; k++) {
                double rows[3];
                for (int sy = 0; sy < 3; sy++) {
                    double acc = 0;
                    for (int sx = 0; sx < 3; sx++) acc += im.px[((3 * y + sy) * n + 3 * x + sx) * 3 + k];
                    rows[sy] = acc / 3.0;
                }
                const double m = (rows[0] + rows[1] + rows[2]) / 3.0;
                out[(y * ts + x) * 3 + k] = (uint8_t)m;
            }
    free(im.px);
}

/* Grid.render, core/grid.py:200-242, on a w x h grid given as g4[(j*w+i)*4] with per-cell highlight
 * hl[i*h + j] (may be NULL); agent at (ax, ay) with direction adir (adir < 0: no agent). */
static void render_grid(const uint8_t* g4, int w, int h, const uint8_t* hl, int ax, int ay, int adir, int ts, uint8_t* out)
{
    uint8_t* tile = (uint8_t*)malloc((size_t)ts * ts * 3);
    for (int j = 0; j < h; j++)
        for (int i = 0; i < w; i++) {
            const uint8_t* c = g4 + (j * w + i) * 4;
            const int here = (i == ax && j == ay);
            mgo_render_tile(c[0], c[1], c[2], here ? adir : -1, hl ? hl[i * h + j] : 0, ts, tile);
            for (int y = 0; y < ts; y++)
                memcpy(out + (((size_t)(j * ts + y) * w + i) * ts) * 3, tile + (size_t)y * ts * 3, (size_t)ts * 3);
        }
    free(tile);
}

/* get_pov_render, minigrid_env.py:647-661.  out: uint8[V*ts][V*ts][3] */
void mgo_render_pov(const uint8_t* cells, const int32_t* cfg, const int32_t* st, int ts, uint8_t* out)
{
    const int V = cfg[CFG_VIEW];
    uint8_t g4[15 * 15 * 4], vis[15 * 15];
    mgo_gen_obs_grid(cells, cfg, st, g4, vis);
    render_grid(g4, V, V, vis, V / 2, V - 1, 3, ts, out);
}

/* get_full_render, minigrid_env.py:663-709.  out: uint8[H*ts][W*ts][3] */
void mgo_render_full(const uint8_t* cells, const int32_t* cfg, const int32_t* st, int highlight, int ts, uint8_t* out)
{
    const int W = cfg[CFG_W], H = cfg[CFG_H], V = cfg[CFG_VIEW], n = W * H;
    static const int DIR_TO_VEC[4][2] = { {1, 0}, {0, 1}, {-1, 0}, {0, -1} };
    uint8_t g4v[15 * 15 * 4], vis[15 * 15];
    mgo_gen_obs_grid(cells, cfg, st, g4v, vis);
    const int dir = st[ST_DIR];
    const int fx = DIR_TO_VEC[dir][0], fy = DIR_TO_VEC[dir][1];
    const int rx = -fy, ry = fx;                                        /* right_vec, minigrid_env.py:397-403 */
    const int tlx = st[ST_X] + fx * (V - 1) - rx * (V / 2), tly = st[ST_Y] + fy * (V - 1) - ry * (V / 2);
    uint8_t* hl = (uint8_t*)calloc((size_t)n, 1);
    for (int vj = 0; vj < V; vj++)
        for (int vi = 0; vi < V; vi++) {
            if (!vis[vi * V + vj]) continue;
            const int ai = tlx - fx * vj + rx * vi, aj = tly - fy * vj + ry * vi;
            if (ai < 0 || ai >= W || aj < 0 || aj >= H) continue;
            hl[ai * H + aj] = 1;
        }
    uint8_t* g4 = (uint8_t*)malloc((size_t)n * 4);
    for (int k = 0; k < n; k++)
        for (int p = 0; p < 4; p++) g4[k * 4 + p] = cells[p * n + k];
    render_grid(g4, W, H, highlight ? hl : NULL, st[ST_X], st[ST_Y], dir, ts, out);
    free(g4); free(hl);
}
