// mg_enum.cu — K3b: reachable-state enumeration of ONE layout or of a whole BATCH of layouts on the GPU.
//
// The reference has no such code.  The MDP is the one MiniGridEnv.step defines (minigrid/minigrid_env.py:520-590,
// envs/keycorridor.py:129-137 and the other step overrides): a thread expands one (state, action) pair with apply_action() —
// the transition function of the step kernel (mg_common.cuh) — plus the per-class terminal rules of the EXT step kernel.
//
// State = layout id + agent cell + dir + carried movable + position of every movable (key / ball / box) + state of every
// door + "this box has been opened" bits, packed into a key of 1..4 64-bit words.  Field widths follow the batch (ceil(log2)
// of the cell count, of the number of movables ...) and no field straddles a word, so DoorKey-16x16 needs 22 bits and
// ObstructedMaze-Full (17 movables, 12 doors) 3 words.  The static part of every layout (walls, goal, lava, floor) is a byte
// grid of packed codes.
//
// Level-synchronous BFS over ALL layouts at once, five launches per level whatever the batch size:
//   expand   one thread per (frontier state, action): successor key -> candidate array; claim it in an open-addressing table
//            whose 64-bit slots hold the claimer's rank (parent id * A + action index; atomicMin keeps the lowest) or, once
//            assigned, the state id — the key of a slot's owner is read back through that index, so keys of any width need
//            no multi-word atomics and no locks
//   mark     first discoverer (lowest rank) of a state that is new at this level
//   scan     exclusive sum of the marks (CUB, plumbing)
//   assign   ids in rank order = exactly the order a serial FIFO BFS hands them out; keys, info records
//   resolve  successor ids of the frontier's rows
// The frontier of a level is layout-major (layout 0's states in BFS order, then layout 1's ...) and children inherit the
// layout of their parent, so by induction every level is; a stable radix sort by layout id at the end (CUB) therefore yields
// every layout's states contiguous AND in that layout's own serial BFS order: tables of a batch are compared `==` with the
// single-layout enumerations, which are compared `==` with the oracle's BFS and with tables recorded through the reference's
// step().
#include "../../include/mgb200.h"
#include "../../include/mgb200_dp.h"
#include "mg_packed.cuh"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

namespace mg { int set_error(int code, const char* msg); }       // mg_capi.cu (thread-local mg_last_error text)

namespace mgdp {

int fail(int code, const char* msg) { return mg::set_error(code, msg); }
extern int64_t g_launches;                                       // mg_dp.cu

constexpr int MAXA = MGDP_MAX_ACTIONS;
constexpr int MAXM = 20, MAXD = 32, MAXKW = 4;
constexpr uint64_t EMPTY_SLOT = ~0ull;
constexpr uint64_t PROV = 1ull << 63;

// where the fields of a key live (bit offsets; computed on the host for the batch, passed by value)
struct KeyFmt {
    int16_t words, M, D;               // key words; movable / door slots (max over the batch)
    int16_t nb_layout, nb_cell, nb_carry, nb_pos;
    int16_t off_layout, off_cell, off_dir, off_carry, off_open;
    int16_t off_pos[MAXM], off_door[MAXD];
    uint32_t pos_carried, pos_gone;    // special positions: in the agent's hands / an EMPTY box that was opened (Box.toggle -> None)
};

template <int KW> struct Key { uint64_t w[KW]; };

template <int KW> __host__ __device__ __forceinline__ uint32_t kget(const Key<KW>& k, int off, int nb)
{
    const int i = off >> 6, sh = off & 63;
    uint64_t v = k.w[0];
#pragma unroll
    for (int j = 1; j < KW; ++j) if (j == i) v = k.w[j];
    return (uint32_t)((v >> sh) & ((1ull << nb) - 1ull));
}
template <int KW> __host__ __device__ __forceinline__ void kset(Key<KW>& k, int off, int nb, uint32_t val)
{
    const int i = off >> 6, sh = off & 63;
    const uint64_t m = ((1ull << nb) - 1ull) << sh, v = (uint64_t)val << sh;
#pragma unroll
    for (int j = 0; j < KW; ++j) if (KW == 1 || j == i) k.w[j] = (k.w[j] & ~m) | v;
}
template <int KW> __device__ __forceinline__ bool keq(const Key<KW>& a, const Key<KW>& b)
{
    bool e = true;
#pragma unroll
    for (int j = 0; j < KW; ++j) e &= a.w[j] == b.w[j];
    return e;
}
__device__ __forceinline__ uint64_t hash64(uint64_t x)
{
    x ^= x >> 30; x *= 0xbf58476d1ce4e5b9ull; x ^= x >> 27; x *= 0x94d049bb133111ebull; x ^= x >> 31;
    return x;
}
template <int KW> __device__ __forceinline__ uint64_t khash(const Key<KW>& k)
{
    uint64_t h = hash64(k.w[0]);
#pragma unroll
    for (int j = 1; j < KW; ++j) h = hash64(h ^ (k.w[j] + 0x9e3779b97f4a7c15ull * (uint64_t)j));
    return h;
}
// keys cross between threads through L2 only (ld.cg / st.cg): a line of the candidate array must never be served from an L1
// that cached it before a neighbouring candidate was written
template <int KW> __device__ __forceinline__ Key<KW> kload(const uint64_t* p)
{
    Key<KW> k;
#pragma unroll
    for (int j = 0; j < KW; ++j) k.w[j] = __ldcg(p + j);
    return k;
}
template <int KW> __device__ __forceinline__ void kstore(uint64_t* p, const Key<KW>& k)
{
#pragma unroll
    for (int j = 0; j < KW; ++j) __stcg(p + j, k.w[j]);
}

// per-layout description of everything that is not in the key (device array [B])
struct LayoutCtx {
    int16_t M, D;
    int16_t target_door;       // index of the door with the target bit (RULE_UNLOCK; the RED door of RULE_REDBLUE), -1 = none
    int16_t blue_door;         // RULE_REDBLUE: index of the flagged blue door, -1 = none
    int16_t tgt_a, tgt_b;      // cell (y*W+x) of the mission's target position(s) in the INITIAL layout, -1 = none:
                               // RULE_GOTO target_pos; RULE_MEMORY success_pos / failure_pos (aux 0x80 / 0x81)
    uint8_t mov_type[MAXM], mov_color[MAXM], mov_aux[MAXM];
    uint16_t mov_init[MAXM];
    uint16_t door_cell[MAXD];
    uint8_t door_color[MAXD], door_init[MAXD];
};

struct EnumCtx {
    int W, H, rule, A;
    int acts[MAXA];
    KeyFmt f;
    const LayoutCtx* lay;      // dev [B]
    const uint8_t* base;       // dev uint8[B][W*H] packed codes of the static layer (movables removed, doors as placed)
    const uint8_t* door_idx;   // dev uint8[B][W*H] door number or 0xFF
};

// what movable j currently is: itself, or — a box that has been opened (Box.toggle, core/world_object.py:291-294) — its
// contents (aux bits 0-3 type, 4-6 colour)
__device__ __forceinline__ mg::cell_t movable_cell(const LayoutCtx& c, int j, uint32_t opened)
{
    if ((opened >> j) & 1u) return mg::make_cell(c.mov_aux[j] & 0x0f, (c.mov_aux[j] >> 4) & 7, 0, 0);
    return mg::make_cell(c.mov_type[j], c.mov_color[j], 0, c.mov_aux[j]);
}

// Slot words and keys travel through L2 only (relaxed gpu-scope accesses / ld.cg / st.cg).  A claimer stores its candidate
// key, fences, and only then publishes its rank with the CAS; a reader that got a rank out of a slot fetches that candidate
// with a load whose address depends on the rank it read — the key it sees is the one stored before the publication.
__device__ __forceinline__ uint64_t ld_slot(const uint64_t* p)
{
    uint64_t v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t cas_slot(uint64_t* p, uint64_t cmp, uint64_t val)
{
    uint64_t old;
    asm volatile("atom.relaxed.gpu.global.cas.b64 %0, [%1], %2, %3;" : "=l"(old) : "l"(p), "l"(cmp), "l"(val) : "memory");
    return old;
}

// Claim `key` for the candidate of rank `rank` (its key is already in the candidate array and fenced).  A slot holds EMPTY, PROV | rank of its lowest-ranked claimer at this level, or a final state id.
// Returns the slot, or -3 when the table is full (the host turns that into MG_ENOMEM).
template <int KW>
__device__ __forceinline__ int64_t table_claim(uint64_t* slots, uint64_t mask, const Key<KW>& key, uint64_t rank,
                                               const uint64_t* cand, uint64_t rank_base, const uint64_t* state_keys, const int* overflow)
{
    uint64_t slot = khash(key) & mask;
    for (uint64_t probes = 0; probes <= mask; ++probes) {
        if ((probes & 63) == 63 && *reinterpret_cast<const volatile int*>(overflow)) return -3;      // somebody found the table full: stop probing
        uint64_t v = ld_slot(slots + slot);
        if (v == EMPTY_SLOT) {
            v = cas_slot(slots + slot, EMPTY_SLOT, PROV | rank);
            if (v == EMPTY_SLOT) return (int64_t)slot;
        }
        const uint64_t* kp = (v & PROV) ? cand + ((v & ~PROV) - rank_base) * KW : state_keys + v * KW;
        if (keq(kload<KW>(kp), key)) {
            if ((v & PROV) && (PROV | rank) < v) atomicMin(reinterpret_cast<unsigned long long*>(slots + slot), (unsigned long long)(PROV | rank));
            return (int64_t)slot;
        }
        slot = (slot + 1) & mask;
    }
    return -3;
}

template <int KW>
__global__ void enum_init_kernel(EnumCtx c, int B, uint64_t* slots, uint64_t mask, const uint64_t* __restrict__ keys0, uint64_t* state_keys, int* overflow)
{
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= B) return;
    const Key<KW> k = kload<KW>(keys0 + (size_t)l * KW);
    kstore<KW>(state_keys + (size_t)l * KW, k);
    // distinct layouts never collide (the layout id is part of the key): plain insertion of final ids
    uint64_t slot = khash(k) & mask;
    for (uint64_t probes = 0; probes <= mask; ++probes) {
        if (cas_slot(slots + slot, EMPTY_SLOT, (uint64_t)l) == EMPTY_SLOT) return;
        slot = (slot + 1) & mask;
    }
    *overflow = 1;
}

// one thread per (frontier state, action index)
template <int KW>
__global__ void __launch_bounds__(256)
enum_expand_kernel(EnumCtx c, const uint64_t* __restrict__ state_keys, int64_t lo, int64_t hi,
                   uint64_t* slots, uint64_t mask, uint64_t* cand, int64_t* __restrict__ slot_of, int* overflow)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t F = hi - lo;
    if (i >= F * c.A) return;
    const int64_t s = lo + i / c.A;
    const int ai = (int)(i % c.A);
    const KeyFmt& f = c.f;
    Key<KW> k = kload<KW>(state_keys + (size_t)s * KW);
    const int L = f.nb_layout ? (int)kget(k, f.off_layout, f.nb_layout) : 0;
    const LayoutCtx& lc = c.lay[L];
    const int ncell = c.W * c.H;
    MG_ASSERT((int)kget(k, f.off_cell, f.nb_cell) < ncell && lc.M <= MAXM && lc.D <= MAXD);
    const uint8_t* base = c.base + (size_t)L * ncell;
    const uint8_t* didx = c.door_idx + (size_t)L * ncell;
    const int cell = (int)kget(k, f.off_cell, f.nb_cell), dir = (int)kget(k, f.off_dir, 2);
    const int carry_i = f.nb_carry ? (int)kget(k, f.off_carry, f.nb_carry) : 0;
    uint32_t opened = f.M ? kget(k, f.off_open, f.M) : 0u;
    const int x = cell % c.W, y = cell / c.W;
    const int fx = x + mg::dir_dx(dir), fy = y + mg::dir_dy(dir);
    mg::cell_t front = mg::CELL_WALL;
    int fk = -1, fmov = -1, fdoor = -1;
    if (fx >= 0 && fy >= 0 && fx < c.W && fy < c.H) {
        fk = fy * c.W + fx;
        for (int j = 0; j < lc.M; ++j) if (kget(k, f.off_pos[j], f.nb_pos) == (uint32_t)fk) fmov = j;
        if (fmov >= 0) front = movable_cell(lc, fmov, opened);
        else if (didx[fk] != 0xFF) { fdoor = didx[fk]; front = mg::make_cell(mg::T_DOOR, lc.door_color[fdoor], (int)kget(k, f.off_door[fdoor], 2), 0); }
        else front = mg::decode_code(base[fk]);
    }
    const mg::cell_t carry = carry_i ? movable_cell(lc, carry_i - 1, opened) : 0;
    const bool pre_carrying = carry_i != 0;                                                      // putnear.py:177
    int action = c.acts[ai];
    if (c.rule == mg::RULE_MEMORY && action == mg::A_PICKUP) action = mg::A_TOGGLE;              // memory.py:154-155
    const mg::Transition tr = mg::apply_action(action, x, y, dir, carry, fx, fy, front, c.rule);
    if (tr.terminated) { slot_of[i] = tr.success ? -1 : -2; return; }
    auto door_open = [&](int d) { return d >= 0 && kget(k, f.off_door[d], 2) == (uint32_t)mg::S_OPEN; };
    const bool red_before = c.rule == mg::RULE_REDBLUE && door_open(lc.target_door);
    const bool blue_before = c.rule == mg::RULE_REDBLUE && door_open(lc.blue_door);
    const int ncell_new = tr.y * c.W + tr.x;
    kset(k, f.off_cell, f.nb_cell, (uint32_t)ncell_new);
    kset(k, f.off_dir, 2, (uint32_t)tr.dir);
    int carry_new = carry_i;
    if (tr.wrote_front) {
        if (fmov >= 0 && tr.carry != 0 && carry == 0) { kset(k, f.off_pos[fmov], f.nb_pos, f.pos_carried); carry_new = fmov + 1; }   // pickup
        else if (carry != 0 && tr.carry == 0) { kset(k, f.off_pos[carry_i - 1], f.nb_pos, (uint32_t)fk); carry_new = 0; }            // drop
        else if (fdoor >= 0) kset(k, f.off_door[fdoor], 2, (uint32_t)mg::c_state(tr.front));                                          // toggle
        else if (fmov >= 0 && mg::c_type(front) == mg::T_BOX) {                                                                       // Box.toggle
            opened |= 1u << fmov;
            kset(k, f.off_open, f.M, opened);
            if (mg::c_type(tr.front) <= mg::T_EMPTY) kset(k, f.off_pos[fmov], f.nb_pos, f.pos_gone);                                 // it was empty
        }
        if (f.nb_carry) kset(k, f.off_carry, f.nb_carry, (uint32_t)carry_new);
    }
    // UnlockEnv.step (envs/unlock.py:92-95): any toggle while the mission's door is open pays and terminates
    if (c.rule == mg::RULE_UNLOCK && c.acts[ai] == mg::A_TOGGLE && door_open(lc.target_door)) { slot_of[i] = -1; return; }
    // the same per-class terminal rules the EXT step kernel applies after the base transition (mg_step.cu), on the successor
    if (c.rule == mg::RULE_FETCH && carry_new != 0) {                                           // fetch.py:165-174
        slot_of[i] = (mg::c_aux(movable_cell(lc, carry_new - 1, opened)) & 0x80) ? -1 : -2;
        return;
    }
    if (c.rule == mg::RULE_PUTNEAR) {                                                           // putnear.py:181-199
        if (action == mg::A_PICKUP && carry_new != 0) {
            const int aux = mg::c_aux(movable_cell(lc, carry_new - 1, opened));
            if (!((aux & 0x80) && (aux & 0x0f) != 1)) { slot_of[i] = -2; return; }               // not the object to move
        }
        if (action == mg::A_DROP && pre_carrying) {
            bool near = false;
            if (carry_new == 0 && lc.tgt_b >= 0) {                                              // the drop happened: near target_pos?
                const int tx = lc.tgt_b % c.W, ty = lc.tgt_b / c.W;
                near = abs(fx - tx) <= 1 && abs(fy - ty) <= 1;
            }
            slot_of[i] = near ? -1 : -2;
            return;
        }
    }
    if (c.rule == mg::RULE_GOTO && (action == mg::A_TOGGLE || action == mg::A_DONE)) {          // gotodoor.py:136-148, gotoobject.py:148-160
        bool near = false;
        if (action == mg::A_DONE && lc.tgt_a >= 0) {
            const int tx = lc.tgt_a % c.W, ty = lc.tgt_a / c.W;
            near = (tr.x == tx && abs(tr.y - ty) == 1) || (tr.y == ty && abs(tr.x - tx) == 1);
        }
        slot_of[i] = near ? -1 : -2;
        return;
    }
    if (c.rule == mg::RULE_MEMORY && (ncell_new == lc.tgt_a || ncell_new == lc.tgt_b)) {         // memory.py:158-163 (failure_pos wins a tie)
        slot_of[i] = ncell_new == lc.tgt_b ? -2 : -1;
        return;
    }
    if (c.rule == mg::RULE_REDBLUE) {                                                           // redbluedoors.py:114-131
        const bool red_after = door_open(lc.target_door), blue_after = door_open(lc.blue_door);
        if (blue_after) { slot_of[i] = red_before ? -1 : -2; return; }
        if (red_after && blue_before) { slot_of[i] = -2; return; }
    }
    kstore<KW>(cand + (size_t)i * KW, k);
    __threadfence();                                               // the candidate key is in L2 before its rank can be seen in a slot
    const uint64_t rank_base = (uint64_t)lo * (uint64_t)c.A;
    const int64_t slot = table_claim<KW>(slots, mask, k, rank_base + (uint64_t)i, cand, rank_base, state_keys, overflow);
    if (slot == -3) *overflow = 1;
    slot_of[i] = slot;
}

// flag[i] = 1 iff (state, action) i is the FIRST discoverer (lowest rank) of a state that is new at this level
__global__ void __launch_bounds__(256)
enum_mark_kernel(int A, int64_t lo, int64_t n, const int64_t* __restrict__ slot_of, const uint64_t* __restrict__ slots, int32_t* __restrict__ flag)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t slot = slot_of[i];
    const uint64_t rank = (uint64_t)(lo * A + i);
    flag[i] = (slot >= 0 && slots[slot] == (PROV | rank)) ? 1 : 0;
}

template <int KW>
__global__ void __launch_bounds__(256)
enum_assign_kernel(EnumCtx c, int64_t n, int64_t hi, const int64_t* __restrict__ slot_of, const int32_t* __restrict__ flag,
                   const int32_t* __restrict__ offs, const uint64_t* __restrict__ cand, uint64_t* slots,
                   uint64_t* __restrict__ state_keys, uint8_t* __restrict__ info, uint32_t* __restrict__ layout_of,
                   int32_t* __restrict__ layout_count, int64_t max_states)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !flag[i]) return;
    const int64_t id = hi + offs[i];
    if (id >= max_states) return;                 // the host notices the overflow from the total
    MG_ASSERT(slot_of[i] >= 0 && id >= hi);
    const KeyFmt& f = c.f;
    const Key<KW> k = kload<KW>(cand + (size_t)i * KW);
    slots[slot_of[i]] = (uint64_t)id;
    kstore<KW>(state_keys + (size_t)id * KW, k);
    const int L = f.nb_layout ? (int)kget(k, f.off_layout, f.nb_layout) : 0;
    const LayoutCtx& lc = c.lay[L];
    if (layout_of) { layout_of[id] = (uint32_t)L; atomicAdd(layout_count + L, 1); }
    const uint32_t opened = f.M ? kget(k, f.off_open, f.M) : 0u;
    int nd = 0;
    for (int j = 0; j < lc.M; ++j) {
        const uint32_t p = kget(k, f.off_pos[j], f.nb_pos);
        if (p != lc.mov_init[j]) nd += (p == f.pos_carried || p == f.pos_gone) ? 1 : 2;
        else if ((opened >> j) & 1u) nd += 1;                          // opened in place: the cell shows the contents
    }
    for (int d = 0; d < lc.D; ++d) if (kget(k, f.off_door[d], 2) != lc.door_init[d]) nd += 1;
    const int cell = (int)kget(k, f.off_cell, f.nb_cell);
    const int carry_i = f.nb_carry ? (int)kget(k, f.off_carry, f.nb_carry) : 0;
    uint8_t* o = info + id * 8;
    o[0] = (uint8_t)(cell % c.W); o[1] = (uint8_t)(cell / c.W); o[2] = (uint8_t)kget(k, f.off_dir, 2);
    const mg::cell_t cc = carry_i ? movable_cell(lc, carry_i - 1, opened) : 0;
    o[3] = (uint8_t)mg::c_type(cc); o[4] = (uint8_t)mg::c_color(cc); o[5] = 0;
    o[6] = (uint8_t)mg::c_aux(cc); o[7] = (uint8_t)min(nd, 255);
}

// succ_tmp is action-major with stride max_states
__global__ void __launch_bounds__(256)
enum_resolve_kernel(int A, int64_t lo, int64_t n, const int64_t* __restrict__ slot_of,
                    const uint64_t* __restrict__ slots, int32_t* __restrict__ succ_tmp, int64_t stride)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const int64_t s = lo + i / A;
    const int ai = (int)(i % A);
    const int64_t slot = slot_of[i];
    succ_tmp[(int64_t)ai * stride + s] = slot >= 0 ? (int32_t)slots[slot] : (int32_t)slot;
}

__global__ void fill_u64_kernel(uint64_t* p, uint64_t v, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}
__global__ void iota_kernel(uint32_t* p, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = (uint32_t)i;
}
__global__ void invert_kernel(const uint32_t* __restrict__ perm, int32_t* __restrict__ inv, int64_t n)
{
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) inv[perm[i]] = (int32_t)i;
}
// final tables in layout-major order: state f of the output is temp state perm[f]; successor ids go through inv
__global__ void __launch_bounds__(256)
permute_kernel(int A, int64_t S, int64_t S_pad, int64_t stride, const uint32_t* __restrict__ perm, const int32_t* __restrict__ inv,
               const int32_t* __restrict__ succ_tmp, const uint8_t* __restrict__ info_tmp, int32_t* __restrict__ succ, uint8_t* __restrict__ info)
{
    const int64_t fidx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (fidx >= S) return;
    const int64_t t = perm ? (int64_t)perm[fidx] : fidx;
    for (int a = 0; a < A; ++a) {
        const int32_t sp = succ_tmp[(int64_t)a * stride + t];
        succ[(int64_t)a * S_pad + fidx] = (sp >= 0 && inv) ? inv[sp] : sp;
    }
    if (perm) reinterpret_cast<uint2*>(info)[fidx] = reinterpret_cast<const uint2*>(info_tmp)[t];
}

struct Scratch {                   // freed on every exit path
    std::vector<void*> ptrs;
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t alloc(T** p, size_t n)
    {
        cudaError_t e = cudaMalloc(p, (n ? n : 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
    void release(void* p) { for (size_t i = 0; i < ptrs.size(); ++i) if (ptrs[i] == p) { ptrs.erase(ptrs.begin() + i); return; } }
};

#define DP_CUDA(call) do { cudaError_t _e = (call); if (_e != cudaSuccess) return mgdp::fail(MG_ECUDA, cudaGetErrorString(_e)); } while (0)

static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }
static inline int bits_for(int64_t values) { int b = 0; while ((1ll << b) < values) ++b; return b; }     // ceil(log2(values))

}  // namespace mgdp

struct mgdp_mdp {
    int device = 0, A = 0, levels = 0, B = 1, key_words = 1;
    int64_t S = 0, S_pad = 0;
    int32_t* succ = nullptr;       // dev [A][S_pad]
    uint8_t* info = nullptr;       // dev [S][8]
    std::vector<int64_t> offsets;  // [B + 1] first state of every layout (+ S)
};

namespace mgdp {

template <int KW>
static int enumerate_impl(mgdp_mdp** out, int device, int B, int W, int H, int rule, const int32_t* actions, int A,
                          int64_t max_states, const KeyFmt& fmt, const std::vector<LayoutCtx>& lays,
                          const std::vector<uint8_t>& base, const std::vector<uint8_t>& didx, const std::vector<uint64_t>& keys0)
{
    const int n = W * H;
    DP_CUDA(cudaSetDevice(device));
    Scratch sc;
    EnumCtx c{};
    c.W = W; c.H = H; c.rule = rule; c.A = A; c.f = fmt;
    for (int i = 0; i < A; ++i) c.acts[i] = actions[i];
    uint8_t *d_base = nullptr, *d_didx = nullptr; LayoutCtx* d_lay = nullptr; uint64_t* d_keys0 = nullptr;
    DP_CUDA(sc.alloc(&d_base, (size_t)B * n)); DP_CUDA(sc.alloc(&d_didx, (size_t)B * n));
    DP_CUDA(sc.alloc(&d_lay, (size_t)B)); DP_CUDA(sc.alloc(&d_keys0, (size_t)B * KW));
    DP_CUDA(cudaMemcpy(d_base, base.data(), (size_t)B * n, cudaMemcpyHostToDevice));
    DP_CUDA(cudaMemcpy(d_didx, didx.data(), (size_t)B * n, cudaMemcpyHostToDevice));
    DP_CUDA(cudaMemcpy(d_lay, lays.data(), (size_t)B * sizeof(LayoutCtx), cudaMemcpyHostToDevice));
    DP_CUDA(cudaMemcpy(d_keys0, keys0.data(), (size_t)B * KW * 8, cudaMemcpyHostToDevice));
    c.base = d_base; c.door_idx = d_didx; c.lay = d_lay;

    uint64_t cap = 1024; while (cap < (uint64_t)max_states * 2) cap <<= 1;
    const uint64_t mask = cap - 1;
    uint64_t *slots = nullptr, *skeys = nullptr; int32_t* succ_tmp = nullptr; uint8_t* info_tmp = nullptr;
    uint32_t* layout_of = nullptr; int32_t* layout_count = nullptr; int* d_overflow = nullptr;
    DP_CUDA(sc.alloc(&slots, cap)); DP_CUDA(sc.alloc(&skeys, (size_t)max_states * KW));
    DP_CUDA(sc.alloc(&succ_tmp, (size_t)max_states * A));
    DP_CUDA(sc.alloc(&info_tmp, (size_t)max_states * 8));
    DP_CUDA(sc.alloc(&d_overflow, 1));
    DP_CUDA(cudaMemset(d_overflow, 0, sizeof(int)));
    if (B > 1) {
        DP_CUDA(sc.alloc(&layout_of, (size_t)max_states)); DP_CUDA(sc.alloc(&layout_count, (size_t)B));
        DP_CUDA(cudaMemset(layout_count, 0, (size_t)B * sizeof(int32_t)));
    }
    fill_u64_kernel<<<1024, 256>>>(slots, EMPTY_SLOT, (int64_t)cap);
    enum_init_kernel<KW><<<nblk(B), 256>>>(c, B, slots, mask, d_keys0, skeys, d_overflow);
    {
        std::vector<uint8_t> i0((size_t)B * 8, 0);
        std::vector<uint32_t> l0(B);
        std::vector<int32_t> c0(B, 1);
        for (int l = 0; l < B; ++l) {
            Key<KW> k; for (int j = 0; j < KW; ++j) k.w[j] = keys0[(size_t)l * KW + j];
            const int cell = (int)kget(k, fmt.off_cell, fmt.nb_cell);
            i0[(size_t)l * 8] = (uint8_t)(cell % W); i0[(size_t)l * 8 + 1] = (uint8_t)(cell / W); i0[(size_t)l * 8 + 2] = (uint8_t)kget(k, fmt.off_dir, 2);
            l0[l] = (uint32_t)l;
        }
        DP_CUDA(cudaMemcpy(info_tmp, i0.data(), i0.size(), cudaMemcpyHostToDevice));
        if (B > 1) {
            DP_CUDA(cudaMemcpy(layout_of, l0.data(), (size_t)B * 4, cudaMemcpyHostToDevice));
            DP_CUDA(cudaMemcpy(layout_count, c0.data(), (size_t)B * 4, cudaMemcpyHostToDevice));
        }
    }
    g_launches += 2;

    // per-level scratch grows with the frontier
    int64_t cap_items = 0; int64_t* slot_of = nullptr; int32_t *flag = nullptr, *offs = nullptr; uint64_t* cand = nullptr;
    void* cub_tmp = nullptr; size_t cub_bytes = 0;
    int64_t lo = 0, hi = B; int levels = 0;
    while (lo < hi) {
        const int64_t items = (hi - lo) * A;
        if (items > 0x7fffffffll) return fail(MG_EINVAL, "mgdp_enumerate: frontier too large");
        if (items > cap_items) {
            cap_items = items * 2;
            DP_CUDA(sc.alloc(&slot_of, (size_t)cap_items)); DP_CUDA(sc.alloc(&flag, (size_t)cap_items)); DP_CUDA(sc.alloc(&offs, (size_t)cap_items));
            DP_CUDA(sc.alloc(&cand, (size_t)cap_items * KW));
            size_t need = 0;
            cub::DeviceScan::ExclusiveSum(nullptr, need, flag, offs, (int)cap_items);
            uint8_t* t8 = nullptr; DP_CUDA(sc.alloc(&t8, need)); cub_tmp = t8; cub_bytes = need;
        }
        enum_expand_kernel<KW><<<nblk(items), 256>>>(c, skeys, lo, hi, slots, mask, cand, slot_of, d_overflow);
        enum_mark_kernel<<<nblk(items), 256>>>(A, lo, items, slot_of, slots, flag);
        size_t tb = cub_bytes;
        DP_CUDA(cub::DeviceScan::ExclusiveSum(cub_tmp, tb, flag, offs, (int)items));
        int32_t last_off = 0, last_flag = 0; int overflow = 0;
        DP_CUDA(cudaMemcpy(&last_off, offs + items - 1, 4, cudaMemcpyDeviceToHost));
        DP_CUDA(cudaMemcpy(&last_flag, flag + items - 1, 4, cudaMemcpyDeviceToHost));
        DP_CUDA(cudaMemcpy(&overflow, d_overflow, 4, cudaMemcpyDeviceToHost));
        const int64_t fresh = (int64_t)last_off + last_flag;
        if (overflow || hi + fresh > max_states) return fail(MG_ENOMEM, "mgdp_enumerate: more reachable states than max_states");
        enum_assign_kernel<KW><<<nblk(items), 256>>>(c, items, hi, slot_of, flag, offs, cand, slots, skeys, info_tmp, layout_of, layout_count, max_states);
        enum_resolve_kernel<<<nblk(items), 256>>>(A, lo, items, slot_of, slots, succ_tmp, max_states);
        g_launches += 5;
        lo = hi; hi += fresh; levels++;
    }
    DP_CUDA(cudaDeviceSynchronize());
    const int64_t S = hi, S_pad = (S + 31) & ~31ll;
    mgdp_mdp* m = new (std::nothrow) mgdp_mdp();
    if (!m) return fail(MG_ENOMEM, "mgdp_enumerate: out of host memory");
    m->device = device; m->A = A; m->levels = levels; m->S = S; m->S_pad = S_pad; m->B = B; m->key_words = KW;
    m->offsets.assign((size_t)B + 1, 0);
    cudaError_t e = cudaMalloc(&m->succ, (size_t)S_pad * A * sizeof(int32_t));
    if (e == cudaSuccess) e = cudaMemset(m->succ, 0xFF, (size_t)S_pad * A * sizeof(int32_t));     // padding rows read as WIN; never used
    if (e != cudaSuccess) { delete m; return fail(MG_ECUDA, cudaGetErrorString(e)); }
    uint32_t* perm = nullptr; int32_t* inv = nullptr;
    auto bail = [&](cudaError_t err) { cudaFree(m->succ); cudaFree(m->info); delete m; return fail(MG_ECUDA, cudaGetErrorString(err)); };
    if (B > 1) {
        // layout-major final order: stable radix sort of the (level-major) temp ids by layout id
        uint32_t *ids_in = nullptr, *lay_out = nullptr;
        if ((e = sc.alloc(&ids_in, (size_t)S)) != cudaSuccess || (e = sc.alloc(&perm, (size_t)S)) != cudaSuccess ||
            (e = sc.alloc(&lay_out, (size_t)S)) != cudaSuccess || (e = sc.alloc(&inv, (size_t)S)) != cudaSuccess) return bail(e);
        iota_kernel<<<1024, 256>>>(ids_in, S);
        size_t need = 0;
        cub::DeviceRadixSort::SortPairs(nullptr, need, layout_of, lay_out, ids_in, perm, (int)S, 0, fmt.nb_layout);
        uint8_t* tmp = nullptr;
        if ((e = sc.alloc(&tmp, need)) != cudaSuccess) return bail(e);
        if ((e = cub::DeviceRadixSort::SortPairs(tmp, need, layout_of, lay_out, ids_in, perm, (int)S, 0, fmt.nb_layout)) != cudaSuccess) return bail(e);
        invert_kernel<<<1024, 256>>>(perm, inv, S);
        if ((e = cudaMalloc(&m->info, (size_t)S * 8)) != cudaSuccess) return bail(e);
        std::vector<int32_t> cnt(B);
        if ((e = cudaMemcpy(cnt.data(), layout_count, (size_t)B * 4, cudaMemcpyDeviceToHost)) != cudaSuccess) return bail(e);
        for (int l = 0; l < B; ++l) m->offsets[l + 1] = m->offsets[l] + cnt[l];
        g_launches += 3;
    } else {
        m->info = info_tmp;                                 // ownership moves to the handle
        sc.release(info_tmp);
        m->offsets[1] = S;
    }
    permute_kernel<<<nblk(S), 256>>>(A, S, S_pad, max_states, perm, inv, succ_tmp, info_tmp, m->succ, m->info);
    g_launches += 1;
    if ((e = cudaDeviceSynchronize()) != cudaSuccess) return bail(e);
    if (m->offsets[B] != S) { cudaFree(m->succ); cudaFree(m->info); delete m; return fail(MG_ECUDA, "mgdp_enumerate: layout counts do not add up"); }
    *out = m;
    return MG_OK;
}

}  // namespace mgdp

extern "C" {

int mgdp_enumerate_batch(mgdp_mdp** out, int32_t device, int32_t B, const uint8_t* cells, int32_t W, int32_t H,
                         const int32_t* agents, int32_t rule, const int32_t* actions, int32_t A, int64_t max_states)
{
    using namespace mgdp;
    if (!out || !cells || !agents || !actions || B < 1 || A < 1 || A > MAXA || W < 3 || H < 3 || W * H > 4094 || max_states < B)
        return fail(MG_EINVAL, "mgdp_enumerate: bad argument (grid up to 4094 cells, 1..8 actions, max_states >= number of layouts)");
    if (max_states > (1ll << 31) - 16) return fail(MG_EINVAL, "mgdp_enumerate: max_states must fit int32 ids");
    if (rule < mg::RULE_NONE || rule > mg::RULE_PUTNEAR)
        return fail(MG_EINVAL, "mgdp_enumerate: every rule but MG_RULE_DYNOBS is understood by the enumeration (DynamicObstacles is not a "
                               "deterministic MDP)");
    const int n = W * H;
    std::vector<LayoutCtx> lays((size_t)B);
    std::vector<uint8_t> base((size_t)B * n), didx((size_t)B * n, 0xFF);
    int Mmax = 0, Dmax = 0;
    for (int l = 0; l < B; ++l) {
        const uint8_t* cl = cells + (size_t)l * 4 * n;
        LayoutCtx& c = lays[l];
        memset(&c, 0, sizeof(c));
        c.target_door = -1; c.blue_door = -1; c.tgt_a = -1; c.tgt_b = -1;
        const int ax = agents[3 * l], ay = agents[3 * l + 1], adir = agents[3 * l + 2];
        if (ax < 0 || ay < 0 || ax >= W || ay >= H || adir < 0 || adir > 3) return fail(MG_EINVAL, "mgdp_enumerate: agent outside the grid / bad direction");
        for (int k = 0; k < n; ++k) {
            const int t = cl[k], col = cl[n + k], s = cl[2 * n + k], aux = cl[3 * n + k];
            uint8_t& b = base[(size_t)l * n + k];
            if (!mg::code_representable(t, col, s)) return fail(MG_EINVAL, "mgdp_enumerate: cell not representable");
            if (t == mg::T_BOX && ((aux & 0x0f) == mg::T_BOX || (aux & 0x0f) == mg::T_DOOR))
                return fail(MG_EINVAL, "mgdp_enumerate: a box may hide a key or a ball (or nothing)");
            if (t == mg::T_KEY || t == mg::T_BALL || t == mg::T_BOX) {
                if (c.M >= MAXM) return fail(MG_EINVAL, "mgdp_enumerate: more than 20 movable objects");
                c.mov_type[c.M] = (uint8_t)t; c.mov_color[c.M] = (uint8_t)col; c.mov_aux[c.M] = (uint8_t)aux; c.mov_init[c.M] = (uint16_t)k;
                c.M++;
                b = (uint8_t)mg::CODE_EMPTY;
            } else if (t == mg::T_DOOR) {
                if (c.D >= MAXD) return fail(MG_EINVAL, "mgdp_enumerate: more than 32 doors");
                c.door_cell[c.D] = (uint16_t)k; c.door_color[c.D] = (uint8_t)col; c.door_init[c.D] = (uint8_t)s;
                didx[(size_t)l * n + k] = (uint8_t)c.D;
                if (rule == mg::RULE_REDBLUE) {                       // both doors are flagged: red (colour 0) and blue (colour 2)
                    if ((aux & 0x80) && col == 0 && c.target_door < 0) c.target_door = c.D;
                    if ((aux & 0x80) && col == 2 && c.blue_door < 0) c.blue_door = c.D;
                } else if ((aux & 0x80) && c.target_door < 0) c.target_door = c.D;
                c.D++;
                b = (uint8_t)mg::encode_code(t, col, s);
            } else {
                b = (uint8_t)mg::encode_code(t, col, s);
            }
        }
        for (int k = 0; k < n; ++k) {                                // the mission's position targets (the oracle's find_target)
            const int aux = cl[3 * n + k];
            if (!(aux & 0x80)) continue;
            if ((rule == mg::RULE_MEMORY || rule == mg::RULE_PUTNEAR) && (aux & 0x0f) == 1) { if (c.tgt_b < 0) c.tgt_b = (int16_t)k; }
            else if (c.tgt_a < 0) c.tgt_a = (int16_t)k;
        }
        if (c.M > Mmax) Mmax = c.M;
        if (c.D > Dmax) Dmax = c.D;
    }
    // key format: layout | cell | dir | carry | pos[M] | door[D] | opened; a field that would straddle a word starts the next one
    KeyFmt f{};
    f.M = (int16_t)Mmax; f.D = (int16_t)Dmax;
    f.nb_layout = (int16_t)bits_for(B); f.nb_cell = (int16_t)bits_for(n); f.nb_carry = (int16_t)bits_for(Mmax + 1);
    f.nb_pos = (int16_t)bits_for(n + 2);
    f.pos_carried = (1u << f.nb_pos) - 1u; f.pos_gone = f.pos_carried - 1u;
    int next = 0;
    auto place = [&](int nb) { if (nb > 0 && (next & 63) + nb > 64) next = (next + 63) & ~63; const int o = next; next += nb; return (int16_t)o; };
    f.off_layout = place(f.nb_layout); f.off_cell = place(f.nb_cell); f.off_dir = place(2); f.off_carry = place(f.nb_carry);
    for (int j = 0; j < Mmax; ++j) f.off_pos[j] = place(f.nb_pos);
    for (int d = 0; d < Dmax; ++d) f.off_door[d] = place(2);
    f.off_open = place(Mmax);
    const int KW = next == 0 ? 1 : (next + 63) / 64;
    if (KW > MAXKW) return fail(MG_EINVAL, "mgdp_enumerate: state does not fit a 256-bit key");
    f.words = (int16_t)KW;
    std::vector<uint64_t> keys0((size_t)B * KW, 0);
    auto hset = [&](uint64_t* w, int off, int nb, uint32_t v) { if (nb > 0) w[off >> 6] |= ((uint64_t)v & ((1ull << nb) - 1ull)) << (off & 63); };
    for (int l = 0; l < B; ++l) {
        uint64_t* w = keys0.data() + (size_t)l * KW;
        const LayoutCtx& c = lays[l];
        hset(w, f.off_layout, f.nb_layout, (uint32_t)l);
        hset(w, f.off_cell, f.nb_cell, (uint32_t)(agents[3 * l + 1] * W + agents[3 * l]));
        hset(w, f.off_dir, 2, (uint32_t)agents[3 * l + 2]);
        for (int j = 0; j < Mmax; ++j) hset(w, f.off_pos[j], f.nb_pos, j < c.M ? c.mov_init[j] : f.pos_gone);
        for (int d = 0; d < Dmax; ++d) hset(w, f.off_door[d], 2, d < c.D ? c.door_init[d] : 0);
    }
    switch (KW) {
    case 1: return enumerate_impl<1>(out, device, B, W, H, rule, actions, A, max_states, f, lays, base, didx, keys0);
    case 2: return enumerate_impl<2>(out, device, B, W, H, rule, actions, A, max_states, f, lays, base, didx, keys0);
    case 3: return enumerate_impl<3>(out, device, B, W, H, rule, actions, A, max_states, f, lays, base, didx, keys0);
    default: return enumerate_impl<4>(out, device, B, W, H, rule, actions, A, max_states, f, lays, base, didx, keys0);
    }
}

int mgdp_enumerate(mgdp_mdp** out, int32_t device, const uint8_t* cells, int32_t W, int32_t H,
                   int32_t ax, int32_t ay, int32_t adir, int32_t rule,
                   const int32_t* actions, int32_t A, int64_t max_states)
{
    const int32_t agent[3] = {ax, ay, adir};
    return mgdp_enumerate_batch(out, device, 1, cells, W, H, agent, rule, actions, A, max_states);
}

void mgdp_free(mgdp_mdp* m)
{
    if (!m) return;
    cudaSetDevice(m->device);
    cudaFree(m->succ); cudaFree(m->info);
    delete m;
}
int64_t mgdp_num_states(const mgdp_mdp* m) { return m ? m->S : 0; }
int64_t mgdp_padded_states(const mgdp_mdp* m) { return m ? m->S_pad : 0; }
int32_t mgdp_num_levels(const mgdp_mdp* m) { return m ? m->levels : 0; }
int32_t mgdp_num_layouts(const mgdp_mdp* m) { return m ? m->B : 0; }
int32_t mgdp_key_words(const mgdp_mdp* m) { return m ? m->key_words : 0; }
const int32_t* mgdp_succ_dev(const mgdp_mdp* m) { return m ? m->succ : nullptr; }
const uint8_t* mgdp_info_dev(const mgdp_mdp* m) { return m ? m->info : nullptr; }

int mgdp_layout_offsets(const mgdp_mdp* m, int64_t* offsets_host)
{
    if (!m || !offsets_host) return mgdp::fail(MG_EINVAL, "mgdp_layout_offsets: null argument");
    for (size_t i = 0; i < m->offsets.size(); ++i) offsets_host[i] = m->offsets[i];
    return MG_OK;
}

int mgdp_copy_out(const mgdp_mdp* m, int32_t* succ_dst_dev, uint8_t* info_dst_dev, void* stream)
{
    if (!m) return mgdp::fail(MG_EINVAL, "mgdp_copy_out: null handle");
    cudaStream_t st = (cudaStream_t)stream;
    if (succ_dst_dev) DP_CUDA(cudaMemcpyAsync(succ_dst_dev, m->succ, (size_t)m->S_pad * m->A * sizeof(int32_t), cudaMemcpyDeviceToDevice, st));
    if (info_dst_dev) DP_CUDA(cudaMemcpyAsync(info_dst_dev, m->info, (size_t)m->S * 8, cudaMemcpyDeviceToDevice, st));
    DP_CUDA(cudaStreamSynchronize(st));
    return MG_OK;
}

}  // extern "C"
