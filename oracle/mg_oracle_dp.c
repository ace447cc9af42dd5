/*
 * mg_oracle_dp.c — CPU oracle for the DP path.  TEST INFRASTRUCTURE ONLY (see mg_oracle.c header).
 *
 * The reference contains NO dynamic-programming code: the MDP is defined only by
 * MiniGridEnv.step (minigrid/minigrid_env.py:520-590) + KeyCorridorEnv.step (envs/keycorridor.py:129-137).
 * This file therefore (1) enumerates the reachable state space by serial FIFO breadth-first search
 * THROUGH the oracle's literal step restatement (mg_oracle.c: mgo_step), and (2) runs plain fp32
 * Jacobi value iteration / policy evaluation on the resulting successor table.
 * DP parity inside the reference is UNPINNED (nothing to compare with); what IS pinned is the
 * transition function (golden trajectories) and, via oracle/gen_golden.py, the successor tables of
 * small MDPs enumerated through the reference's own step() in the build container.
 *
 * Conventions (SURVEY §8.3 "DP conventions"):
 *   state   = everything step() reads except step_count: agent (x,y,dir), carrying, mutable cells.
 *   ids     = serial FIFO BFS discovery order from s0 = 0, parents in id order, actions in acts[] order.
 *   succ    = int32[S][A]; terminal transitions carry sentinels: MGO_WIN (-1) reward 1, MGO_LOSE (-2) reward 0.
 *   reward  = 1.0 on the transition that triggers _reward() (stationary: _reward() at step_count 0), else 0.
 *   VI      = fp32 Jacobi from V=0:  V'[s] = max_a q(s,a),  q = 1 (WIN) | 0 (LOSE) | gamma*V[s'];
 *             stop after the first sweep whose max|V'-V| < tol; pi = argmax of that last sweep, lowest action on ties.
 */
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define MGO_WIN  (-1)
#define MGO_LOSE (-2)
#define MAXD 24

enum { CFG_W = 0, CFG_H = 1, CFG_MAX_STEPS = 2, CFG_SEE_THROUGH = 3, CFG_RULE = 4, CFG_VIEW = 5, CFG_N = 6 };
enum { ST_X = 0, ST_Y = 1, ST_DIR = 2, ST_STEPS = 3, ST_CT = 4, ST_CC = 5, ST_CS = 6, ST_CA = 7, ST_N = 8 };

int mgo_find_target(const uint8_t* cells, int n, int rule);
int mgo_step_t(uint8_t* cells, const int32_t* cfg, int32_t* st, int action, int target,
               uint8_t* image, double* reward, uint8_t* terminated, uint8_t* truncated);
int mgo_step(uint8_t* cells, const int32_t* cfg, int32_t* st, int action,
             uint8_t* image, double* reward, uint8_t* terminated, uint8_t* truncated);

/* exact canonical state record: agent + carry + sorted list of cells that differ from the initial layout */
typedef struct {
    uint8_t x, y, dir, ct, cc, cs, ca, nd;
    uint16_t idx[MAXD];
    uint8_t  val[MAXD][4];
} rec_t;

static uint64_t rec_hash(const rec_t* r)
{
    uint64_t h = 1469598103934665603ull;
    const uint8_t* p = (const uint8_t*)r;
    for (size_t i = 0; i < sizeof(rec_t); i++) { h ^= p[i]; h *= 1099511628211ull; }
    return h;
}

static void rec_apply(const rec_t* r, const uint8_t* base, uint8_t* cells, int ncell, int32_t* st)
{
    memcpy(cells, base, (size_t)ncell * 4);
    for (int d = 0; d < r->nd; d++)
        for (int p = 0; p < 4; p++) cells[p * ncell + r->idx[d]] = r->val[d][p];
    st[ST_X] = r->x; st[ST_Y] = r->y; st[ST_DIR] = r->dir; st[ST_STEPS] = 0;
    st[ST_CT] = r->ct; st[ST_CC] = r->cc; st[ST_CS] = r->cs; st[ST_CA] = r->ca;
}

/* rebuild the diff list by scanning the grid (simple and obviously right; grids are <= 19x19) */
static int rec_from(rec_t* r, const uint8_t* base, const uint8_t* cells, int ncell, const int32_t* st)
{
    memset(r, 0, sizeof(*r));
    r->x = (uint8_t)st[ST_X]; r->y = (uint8_t)st[ST_Y]; r->dir = (uint8_t)st[ST_DIR];
    r->ct = (uint8_t)st[ST_CT]; r->cc = (uint8_t)st[ST_CC]; r->cs = (uint8_t)st[ST_CS]; r->ca = (uint8_t)st[ST_CA];
    int nd = 0;
    for (int k = 0; k < ncell; k++) {
        if (cells[k] != base[k] || cells[ncell + k] != base[ncell + k] ||
            cells[2 * ncell + k] != base[2 * ncell + k] || cells[3 * ncell + k] != base[3 * ncell + k]) {
            if (nd >= MAXD) return -1;
            r->idx[nd] = (uint16_t)k;
            for (int p = 0; p < 4; p++) r->val[nd][p] = cells[p * ncell + k];
            nd++;
        }
    }
    r->nd = (uint8_t)nd;
    return 0;
}

/*
 * Serial BFS.  Returns the number of states S (>0), -1 if max_states is exceeded, -2 if a state has
 * more than MAXD modified cells.  succ is int32[max_states][A]; info is uint8[max_states][8] =
 * (x, y, dir, carry type, carry color, carry state, carry aux, number of modified cells).
 */
/* a finished enumeration kept alive (mgo_space_build): state records + hash table, so that callers can map a full state
 * (cells + agent) to its BFS id and back — used by oracle/gen_golden_big.py to replay sampled rows of a 16.8 M-state table
 * through the reference's own step() */
typedef struct {
    long S, cap;
    int ncell;
    int32_t* table;
    rec_t* recs;
    uint8_t* base;
} mgo_space;

static long enumerate_impl(const uint8_t* cells0, const int32_t* cfg_in, const int32_t* st0,
                           const int32_t* acts, int A, long max_states, int32_t* succ, uint8_t* info, mgo_space** keep);

long mgo_enumerate(const uint8_t* cells0, const int32_t* cfg_in, const int32_t* st0,
                   const int32_t* acts, int A, long max_states, int32_t* succ, uint8_t* info)
{
    return enumerate_impl(cells0, cfg_in, st0, acts, A, max_states, succ, info, NULL);
}

/* same BFS; on success *out owns the records (free with mgo_space_free) */
long mgo_space_build(const uint8_t* cells0, const int32_t* cfg_in, const int32_t* st0,
                     const int32_t* acts, int A, long max_states, int32_t* succ, uint8_t* info, mgo_space** out)
{
    return enumerate_impl(cells0, cfg_in, st0, acts, A, max_states, succ, info, out);
}

/* BFS id of the state (cells uint8[4][ncell], st), or -1 if it was never reached, -2 if it has too many modified cells */
long mgo_space_lookup(const mgo_space* sp, const uint8_t* cells, const int32_t* st)
{
    rec_t r;
    if (rec_from(&r, sp->base, cells, sp->ncell, st)) return -2;
    uint64_t h = rec_hash(&r) & (uint64_t)(sp->cap - 1);
    for (;;) {
        const int32_t id = sp->table[h];
        if (id < 0) return -1;
        if (memcmp(&sp->recs[id], &r, sizeof(rec_t)) == 0) return id;
        h = (h + 1) & (uint64_t)(sp->cap - 1);
    }
}

/* the full state of BFS id `id`: cells uint8[4][ncell], st int32[ST_N] (step_count 0) */
int mgo_space_state(const mgo_space* sp, long id, uint8_t* cells, int32_t* st)
{
    if (id < 0 || id >= sp->S) return -1;
    rec_apply(&sp->recs[id], sp->base, cells, sp->ncell, st);
    return 0;
}

void mgo_space_free(mgo_space* sp)
{
    if (!sp) return;
    free(sp->table); free(sp->recs); free(sp->base); free(sp);
}

static long enumerate_impl(const uint8_t* cells0, const int32_t* cfg_in, const int32_t* st0,
                           const int32_t* acts, int A, long max_states, int32_t* succ, uint8_t* info, mgo_space** keep)
{
    const int W = cfg_in[CFG_W], H = cfg_in[CFG_H], ncell = W * H;
    int32_t cfg[CFG_N];
    memcpy(cfg, cfg_in, sizeof(cfg));
    cfg[CFG_MAX_STEPS] = 0x7fffffff;                 /* no truncation inside the DP */

    long cap = 1; while (cap < max_states * 2 + 16) cap <<= 1;
    int32_t* table = (int32_t*)malloc(sizeof(int32_t) * cap);
    rec_t* recs = (rec_t*)malloc(sizeof(rec_t) * (size_t)max_states);
    uint8_t* cells = (uint8_t*)malloc((size_t)ncell * 4);
    uint8_t image[15 * 15 * 3];
    if (!table || !recs || !cells) { free(table); free(recs); free(cells); return -3; }
    for (long i = 0; i < cap; i++) table[i] = -1;

    long S = 0, rc = 0;
    const int target0 = mgo_find_target(cells0, ncell, cfg[4]);     /* target_pos is fixed at generation time */
    int32_t st[ST_N];
    rec_t r;
    memcpy(st, st0, sizeof(st));
    if (rec_from(&r, cells0, cells0, ncell, st)) { rc = -2; goto out; }
    recs[S] = r;
    { uint64_t h = rec_hash(&r) & (cap - 1); table[h] = 0; }
    S = 1;

    for (long s = 0; s < S; s++) {
        for (int ai = 0; ai < A; ai++) {
            rec_apply(&recs[s], cells0, cells, ncell, st);
            double rew; uint8_t term, trunc;
            mgo_step_t(cells, cfg, st, acts[ai], target0, image, &rew, &term, &trunc);
            int32_t out;
            if (term) {
                out = (rew > 0.0) ? MGO_WIN : MGO_LOSE;
            } else {
                if (rec_from(&r, cells0, cells, ncell, st)) { rc = -2; goto out; }
                uint64_t h = rec_hash(&r) & (cap - 1);
                for (;;) {
                    int32_t id = table[h];
                    if (id < 0) {
                        if (S >= max_states) { rc = -1; goto out; }
                        table[h] = (int32_t)S; recs[S] = r; id = (int32_t)S; S++;
                        out = id; break;
                    }
                    if (memcmp(&recs[id], &r, sizeof(rec_t)) == 0) { out = id; break; }
                    h = (h + 1) & (cap - 1);
                }
            }
            succ[s * A + ai] = out;
        }
    }
    if (info)
        for (long s = 0; s < S; s++) {
            uint8_t* o = info + s * 8;
            o[0] = recs[s].x; o[1] = recs[s].y; o[2] = recs[s].dir; o[3] = recs[s].ct;
            o[4] = recs[s].cc; o[5] = recs[s].cs; o[6] = recs[s].ca; o[7] = recs[s].nd;
        }
    rc = S;
    if (keep) {
        mgo_space* sp = (mgo_space*)malloc(sizeof(mgo_space));
        uint8_t* base = (uint8_t*)malloc((size_t)ncell * 4);
        if (sp && base) {
            memcpy(base, cells0, (size_t)ncell * 4);
            sp->S = S; sp->cap = cap; sp->ncell = ncell; sp->table = table; sp->recs = recs; sp->base = base;
            *keep = sp;
            free(cells);
            return rc;
        }
        free(sp); free(base);
        rc = -3;
    }
out:
    free(table); free(recs); free(cells);
    return rc;
}

static inline float q_value(int32_t sp, const float* V, float gamma)
{
    if (sp == MGO_WIN) return 1.0f;
    if (sp == MGO_LOSE) return 0.0f;
    return gamma * V[sp];
}

/*
 * fp32 Jacobi value iteration (pi_fixed == NULL) or policy evaluation (pi_fixed != NULL: q(s, pi_fixed[s])).
 * V (in/out, length S) holds the start values on entry (callers pass zeros) and the result on exit.
 * Returns the number of sweeps performed; resid_hist[k] (if not NULL, length max_sweeps) = max|V'-V| of sweep k.
 */
int mgo_vi_sweeps(const int32_t* succ, long S, int A, float gamma, float tol, int max_sweeps,
                  float* V, uint8_t* pi, const uint8_t* pi_fixed, float* resid_hist)
{
    float* Vn = (float*)malloc(sizeof(float) * (size_t)S);
    int k = 0;
    for (; k < max_sweeps; ) {
        float resid = 0.0f;
        for (long s = 0; s < S; s++) {
            float best; int ba;
            if (pi_fixed) {
                ba = pi_fixed[s];
                best = q_value(succ[s * A + ba], V, gamma);
            } else {
                best = q_value(succ[s * A], V, gamma); ba = 0;
                for (int a = 1; a < A; a++) {
                    float q = q_value(succ[s * A + a], V, gamma);
                    if (q > best) { best = q; ba = a; }
                }
            }
            Vn[s] = best;
            if (pi) pi[s] = (uint8_t)ba;
            float d = best - V[s]; if (d < 0) d = -d;
            if (d > resid) resid = d;
        }
        memcpy(V, Vn, sizeof(float) * (size_t)S);
        if (resid_hist) resid_hist[k] = resid;
        k++;
        if (resid < tol) break;
    }
    free(Vn);
    return k;
}
