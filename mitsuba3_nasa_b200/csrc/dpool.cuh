// dpool.cuh -- k_forward_pool: the forward walk as a persistent, CTA-resident path pool.
//
// One CTA per SM keeps P paths ("slots") in shared memory (structure of arrays, one 32-bit word
// per field and slot).  Every slot sits in exactly one stage queue (a ring of slot ids in shared
// memory).  A warp repeatedly
//     1. claims up to 32 slot ids from ONE queue (warp-wide scan of the queue heads, one CAS),
//     2. loads the fields that stage needs into registers,
//     3. runs the stage (dstage.cuh) -- all 32 lanes execute the SAME stage,
//     4. stores the changed fields and appends the slot id to the queue of its next stage.
// All warps of the CTA serve the same group of stages until it runs out of full batches and then
// move on together (pool_claim): the twelve warps share the instruction-cache lines of one stage.
// There is no CTA-wide barrier in the steady state; P is 2-3x the thread count so that a full
// batch of 32 is (almost) always available.  A finished path goes back to START, which adds its
// radiance to the film / render_1 sums and starts the next lane of the CTA in the same slot.
//
// Lanes are dealt to CTAs in interleaved chunks of kPoolChunk lanes (chunk c belongs to CTA
// c mod gridDim), so every CTA samples the whole image and the load is balanced without a global
// work counter.  Per-lane seeds depend only on the global lane index (integrator.cpp:323-340), so
// the result is independent of which slot / warp / CTA walked a lane.
//
// Why this shape: the one-thread-one-path kernel (k_forward in kernels.cuh) measured 6-10 of 32
// lanes active per warp instruction and `no_instruction` as the top stall (profiles/r01_*): lanes of
// a warp sit in different branches of a 200 KB loop body.  Here a warp fetches one stage's code once
// for 32 paths.
#pragma once
#include <cstdio>
#include "dstage.cuh"

#ifndef MIB_POOL_NT
#define MIB_POOL_NT 384        // max threads per CTA (register cap 65536 / 384 = 168)
#endif

namespace mib {

#ifndef MIB_POOL_GROUPS
#define MIB_POOL_GROUPS 5
#endif
#ifndef MIB_POOL_RINGS
#define MIB_POOL_RINGS 8
#endif
constexpr int kPoolRings = MIB_POOL_RINGS;   // stage rings laid out in shared memory (forward: 6 classes, adjoint: 7)
// Ring capacity = P + kPoolRingSlack cells.  A cell is reused kPoolRingCap positions later; a queue never holds more than P
// entries, so the cell being written was claimed at least kPoolRingSlack positions ago -- 16 batches, while a consumer takes
// its entry out of the ring within a few hundred cycles of claiming it.  (With capacity P exactly, the push of the last
// batch of a group could land on the cells a warp had just claimed when all other slots were queued in one class: lost
// entries.)  pool_push checks that the cell it writes is empty and raises error 3 otherwise.
constexpr int kPoolRingSlack = 512;
#ifndef MIB_POOL_PATIENCE
#define MIB_POOL_PATIENCE 0
#endif
constexpr int kPoolPatience = MIB_POOL_PATIENCE;   // polls (100 ns each) a warp waits for a full batch before it takes a partial one
template <int P> __host__ __device__ constexpr int pool_ring_cap() { return P + kPoolRingSlack; }
constexpr int kPoolChunk = 256;            // lanes per interleaved chunk
#ifndef MIB_POOL_MAIN_ITERS
#define MIB_POOL_MAIN_ITERS 2
#endif
#ifndef MIB_POOL_NEE_ITERS
#define MIB_POOL_NEE_ITERS 4
#endif
constexpr int kPoolMainIters = MIB_POOL_MAIN_ITERS;   // MAIN: null collisions handled per batch before the slot is re-queued
constexpr int kPoolNeeIters = MIB_POOL_NEE_ITERS;     // NEE: shadow steps per batch
// The adjoint pool wants shorter batches (profiles/r02_ab_pool.md run 10: C4 +19 % with 2 / 1): its steps are heavier (grid
// look-up + 8 gradient atomics), so re-queueing a slot costs relatively less than the lanes idling while the longest walk of
// the batch finishes; the forward kernels lose 4-10 % with shorter batches.
#ifndef MIB_APOOL_MAIN_ITERS
#define MIB_APOOL_MAIN_ITERS 1
#endif
#ifndef MIB_APOOL_NEE_ITERS
#define MIB_APOOL_NEE_ITERS 2
#endif
constexpr int kAdjMainIters = MIB_APOOL_MAIN_ITERS, kAdjNeeIters = MIB_APOOL_NEE_ITERS;
constexpr unsigned kPoolSpinLimit = 1u << 22;   // watchdog of the two wait loops (consecutive idle polls x 100 ns; ring entry polls x ~50 ns)
constexpr uint32_t kNoLane = 0xffffffffu;
constexpr unsigned kRingEmpty = 0xffffu;

// Slot bookkeeping (MIB_POOL_BOS):
//   1 (shipped)  bank-owned slots: slot s belongs to shared-memory bank s % 32 and is only ever touched by lane s % 32 of some
//                warp, so every field access of a warp (pw[f * P + s], P % 32 == 0) hits 32 distinct banks by construction.
//                Per (class, bank) one 32-bit mask of waiting slots; claim = the lane clears one bit of ITS bank's mask
//                (atomicAnd), hand-over = atomicOr.  No rings, no head CAS, no waiting for a producer.
//   0            one ring of slot ids per class (round 1): a warp's 32 slot ids are arbitrary -> 49 % of the shared-memory
//                wavefronts were bank conflicts (profiles/r01_ncu_l_c2_kforward_pool.txt).
#ifndef MIB_POOL_BOS
#define MIB_POOL_BOS 1
#endif
#if MIB_POOL_BOS
constexpr int kPoolHeaderBytes = 64 + 8 * 32 * 4;      // PoolShared + masks[8][32]
#ifdef MIB_POOL_DEBUG
constexpr int kPoolDebugBytes = 2048;                  // MIB_POOL_DEBUG: one state byte + one class byte per slot, behind the state words
#else
constexpr int kPoolDebugBytes = 0;
#endif
__host__ __device__ constexpr int pool_slots_for(int f_count) {
    int p = ((232448 - kPoolHeaderBytes - kPoolDebugBytes) / (f_count * 4)) / 32 * 32;
    return p > 1024 ? 1024 : p;                        // at most 32 slots per bank (one mask word)
}
#else
__host__ __device__ constexpr int pool_slots_for(int f_count) {
    return ((232448 - 256 - 2 * kPoolRings * kPoolRingSlack) / (f_count * 4 + 2 * kPoolRings)) / 32 * 32;
}
#endif

enum { C_START = 0, C_MAIN = 1, C_SCAT = 2, C_NEE = 3, C_PHASE = 4, C_BSDF = 5, C_NEEG = 6, C_MAINB = 7, C_COUNT = 6 };   // C_NEEG, C_MAINB: adjoint only (apool.cuh)
// Group of every class, 4 bits per class (pool_claim): all warps of a CTA serve ONE group at a time.  The finer the groups
// the better (B200, Msamples/s, c3 / c3 polarized / c2 / c1 / c4 adjoint; profiles/r01_ab_variants.md):
//   MIB_POOL_GROUPS = 1  {MAIN, NEE, NEEG, MAINB} | rest                         275 / 173 / 313 /  910 / 37.3
//                     2  ... | {SCATTER, PHASE} | {START, BSDF}                  293 / 180 / 313 /  916 / 38.4
//                     3  {MAIN, MAINB} | {NEE, NEEG} | {SCATTER, PHASE} | {START, BSDF}   323 / 193 / 346 / 1050 / 40.3
//                     5  {MAIN, MAINB} | {NEE, NEEG} | SCATTER | PHASE | START | BSDF     347 / 210 / 393 / 1127 / 41.7  (shipped)
//                     4  every class alone                                        345 / 206 / 387 / 1127 / 42.4
#if MIB_POOL_GROUPS == 4   // every class its own group: the CTA serves one stage at a time
constexpr unsigned kFwdGroupMap = 0x76543210u, kAdjGroupMap = 0x76543210u;
#elif MIB_POOL_GROUPS == 5 // {MAIN, MAINB} | {NEE, NEEG} | {SCATTER} | {PHASE} | {START} | {BSDF}
constexpr unsigned kFwdGroupMap = 0x01531204u, kAdjGroupMap = 0x01531204u;
#elif MIB_POOL_GROUPS == 3   // {MAIN, MAINB} | {NEE, NEEG} | {SCATTER, PHASE} | {START, BSDF}
constexpr unsigned kFwdGroupMap = 0x01321203u, kAdjGroupMap = 0x01321203u;
#elif MIB_POOL_GROUPS == 2
constexpr unsigned kFwdGroupMap = 0x00210102u, kAdjGroupMap = 0x00210102u;
#else
constexpr unsigned kFwdGroupMap = 0x00110101u, kAdjGroupMap = 0x00110101u;
#endif

// Field map of one slot (units: 32-bit words; word f of slot s lives at pw[f * P + s]).
template <int SPEC> struct PoolL {
    static constexpr int NC = ST<SPEC>::NC, NT = ST<SPEC>::NT, NR = ST<SPEC>::NR;
    static constexpr int F_RNG = 0;                    // state lo, state hi, inc lo, inc hi
    static constexpr int F_LANE = 4;                   // CTA-local lane index of the path in this slot (kNoLane: none)
    static constexpr int F_DEPTH = 5;
    static constexpr int F_FLAGS = 6;
    static constexpr int F_ETA = 7, F_LASTPDF = 8;
    static constexpr int F_RAY = 9;                    // o(3) d(3) maxt
    static constexpr int F_SI = 16;                    // t, p(3), shape | face << 16
    static constexpr int F_SCP = 21;                   // scatter_p(3)
    static constexpr int F_T = 24;
    static constexpr int F_R = F_T + NT;
    static constexpr int F_NRAY = F_R + NR;            // shadow ray o(3) d(3)
    static constexpr int F_NSI = F_NRAY + 6;           // t, p(3), shape | face << 16
    static constexpr int F_NED = F_NSI + 5;            // emitter_d(3)
    static constexpr int F_NTR = F_NED + 3;
    static constexpr int F_NEV = F_NTR + NC;           // emitter value; between MAIN and SCATTER: the albedo factor u
    static constexpr int F_NDIST = F_NEV + NC;         // total_dist, max_dist, ds_dist
    static constexpr int F_NA = F_NDIST + 3;
    static constexpr int F_NW = F_NA + NR;
    static constexpr int F_COUNT = F_NW + 1;
    // The position of the last scattering event (MIS pdf of an area light that is hit) shares F_SCP: it IS scatter_p after a
    // real scatter, and a non-null BSDF event overwrites it with si.p when scatter_p is dead (used with area emitters only).
    static constexpr int F_LSP = F_SCP;
    // slots per CTA: as many as fit into the 227 KB of opt-in shared memory (state words + one 16-bit entry per stage ring),
    // a compile-time constant so that every field access is one LDS / STS with an immediate offset
    static constexpr int P = pool_slots_for(F_COUNT);
};

// flags word
constexpr uint32_t FL_CHANNEL = 0x3, FL_SPEC_CHAIN = 0x4, FL_NEEDS_INT = 0x8, FL_NEE_ACTIVE = 0x10, FL_NEE_NEEDS_INT = 0x20,
                   FL_NEE_MULW = 0x40, FL_NEE_SURFACE = 0x80;
MIB_D uint32_t fl_medium(uint32_t f) { return (f >> 8) & 0xff; }         // medium + 1
MIB_D uint32_t fl_nee_medium(uint32_t f) { return (f >> 16) & 0xff; }    // nee.medium + 1

#if MIB_POOL_BOS
struct PoolShared {
    unsigned retired, error, group, pad[13];
};
static_assert(sizeof(PoolShared) == 64, "PoolShared");
__host__ __device__ constexpr size_t pool_ring_offset() { return sizeof(PoolShared); }                       // the masks
__host__ __device__ constexpr size_t pool_state_offset(int) { return (size_t) kPoolHeaderBytes; }
#else
struct PoolShared {
    unsigned head[8], tail[8];
    unsigned next_local, total_local, retired, error;
    unsigned group, pad[3];
};

__host__ __device__ constexpr size_t pool_ring_offset() { return (sizeof(PoolShared) + 15) & ~(size_t) 15; }
__host__ __device__ constexpr size_t pool_state_offset(int P) { return (pool_ring_offset() + (size_t) kPoolRings * (P + kPoolRingSlack) * sizeof(uint16_t) + 15) & ~(size_t) 15; }
#endif
template <int SPEC> __host__ __device__ constexpr size_t pool_smem_bytes() {
    return pool_state_offset(PoolL<SPEC>::P) + (size_t) PoolL<SPEC>::F_COUNT * PoolL<SPEC>::P * sizeof(uint32_t) + kPoolDebugBytes;
}

// slot accessors
template <int P> struct PoolMem {
    uint32_t *pw;
#ifdef MIB_POOL_DEBUG
    int n_fields;
    MIB_D void chk(int field, int slot) const {
        if (field < 0 || field >= n_fields || slot < 0 || slot >= P) { printf("POOL OOB field %d slot %d (fields %d, P %d) cta %d thread %d\n", field, slot, n_fields, P, blockIdx.x, threadIdx.x); __trap(); }
    }
#else
    MIB_D void chk(int, int) const { }
#endif
    MIB_D float f(int field, int slot) const { chk(field, slot); return __uint_as_float(pw[field * P + slot]); }
    MIB_D uint32_t u(int field, int slot) const { chk(field, slot); return pw[field * P + slot]; }
    MIB_D void sf(int field, int slot, float v) const { chk(field, slot); pw[field * P + slot] = __float_as_uint(v); }
    MIB_D void su(int field, int slot, uint32_t v) const { chk(field, slot); pw[field * P + slot] = v; }
    MIB_D V3 v3f(int field, int slot) const { return v3(f(field, slot), f(field + 1, slot), f(field + 2, slot)); }
    MIB_D void sv3(int field, int slot, V3 v) const { sf(field, slot, v.x); sf(field + 1, slot, v.y); sf(field + 2, slot, v.z); }
    MIB_D void load_rng(int slot, Pcg32 &r) const {
        r.state = (uint64_t) u(0, slot) | ((uint64_t) u(1, slot) << 32);
        r.inc = (uint64_t) u(2, slot) | ((uint64_t) u(3, slot) << 32);
    }
    MIB_D void store_rng_state(int slot, const Pcg32 &r) const { su(0, slot, (uint32_t) r.state); su(1, slot, (uint32_t) (r.state >> 32)); }
    MIB_D void store_rng(int slot, const Pcg32 &r) const {
        store_rng_state(slot, r);
        su(2, slot, (uint32_t) r.inc); su(3, slot, (uint32_t) (r.inc >> 32));
    }
    MIB_D void load_si(int field, int slot, SurfIt &si) const {
        si.t = f(field, slot);
        si.p = v3f(field + 1, slot);
        uint32_t sf_ = u(field + 4, slot);
        si.shape = (int) (sf_ & 0xffffu) - 1;
        si.face = (int) (sf_ >> 16);
    }
    MIB_D void store_si(int field, int slot, const SurfIt &si) const {
        sf(field, slot, si.t);
        sv3(field + 1, slot, si.p);
        su(field + 4, slot, (uint32_t) (si.shape + 1) | ((uint32_t) si.face << 16));
    }
};

#if MIB_POOL_BOS
// ---------------------------------------------------------------------------------------------
// Slot bookkeeping shared by the forward and the adjoint pool kernels: bank-owned slots.
//
// masks[c * 32 + b], bit j set  <=>  slot j * 32 + b waits in class c.  A slot waits in at most one class; while a lane works
// on it, its bit is clear everywhere.  Lane b of ANY warp of the CTA serves bank b:
//   claim  every lane reads the 8 masks of its bank, 8 ballots give the number of banks with a waiting slot per class, the
//          warp picks ONE class (group policy below), and each lane clears one bit of masks[class][its bank] with atomicAnd.
//          Lanes whose bank has nothing in that class sit the batch out.
//   push   atomicOr of the slot's bit into masks[next class][its bank] -- one instruction for the warp, 32 distinct words.
// __threadfence_block() orders the state stores before the atomicOr and the atomicAnd before the state loads.
//
// Group policy (MIB_POOL_GROUPS, as in round 1): all warps of the CTA serve the classes of ONE group until fewer than
// kPoolFull banks have a waiting slot there, and then switch together, so that the code live in the instruction cache is one
// group's and the warps share its lines.
//
// Lanes: the warps draw their next lanes from ONE counter in global memory (lp.counters[6], one atomicAdd per START batch),
// so a CTA that happens to get short paths simply takes more lanes (no static deal, no per-CTA tail), and at any time all
// CTAs work on neighbouring pixels (grid and gradient lines are shared through the L2).  Per-lane seeds depend only on the
// lane index (integrator.cpp:323-340), so the result does not depend on which warp walked a lane.
// ---------------------------------------------------------------------------------------------
#ifndef MIB_POOL_DEBUG_SPINS
#define MIB_POOL_DEBUG_SPINS 30000000u     // MIB_POOL_DEBUG: claim-loop passes of one warp before the kernel prints its state and traps
#endif
#ifndef MIB_POOL_DEBUG_ITERS
#define MIB_POOL_DEBUG_ITERS 3000000u      //                 batches of one warp
#endif
#ifndef MIB_POOL_FULL
#define MIB_POOL_FULL 32
#endif
#ifndef MIB_POOL_STAY
#define MIB_POOL_STAY 24       // lanes a batch of the current stage group must still fill for the CTA to stay with that group (0: round-1 rule)
#endif
#ifndef MIB_POOL_STEAL
#define MIB_POOL_STEAL 0       // banks a lane may look into beyond its own when that has nothing in the chosen class (0: none)
#endif
constexpr unsigned kPoolFull = MIB_POOL_FULL;     // banks with a waiting slot that count as a full batch
template <int P, unsigned GROUPMAP> struct PoolQ {
    static constexpr unsigned FULL = 0xffffffffu;
    PoolShared *ps;
    unsigned *masks;
    unsigned lane_id, idle_spins, rot;
#ifdef MIB_POOL_DEBUG
    unsigned dbg_iters = 0, dbg_spins = 0;
#endif
    uint64_t n_lanes;                                  // lanes of the launch
    unsigned long long *lane_counter;                  // next lane of the launch (global memory, zeroed by the host)

#ifdef MIB_POOL_DEBUG
    volatile unsigned char *dbg_state, *dbg_class;     // per slot: 0 waiting / 1 in flight / 2 retired; class it waits in / runs
    MIB_D void dbg_init(unsigned char *end_of_state, int start_class) {
        dbg_state = end_of_state; dbg_class = end_of_state + 1024;
        for (int i = threadIdx.x; i < P; i += blockDim.x) { dbg_state[i] = 0; dbg_class[i] = (unsigned char) start_class; }
    }
    MIB_D void dbg_fail(const char *what, int slot, int cls) {
        printf("POOL %s: slot %d class %d (recorded state %d class %d) cta %d warp %d lane %d iters %u\n", what, slot, cls,
               (int) dbg_state[slot], (int) dbg_class[slot], blockIdx.x, threadIdx.x >> 5, threadIdx.x & 31, dbg_iters);
        __trap();
    }
#endif
    MIB_D void init(unsigned char *smem_raw, uint64_t n, unsigned long long *counter, int start_class) {
        ps = reinterpret_cast<PoolShared *>(smem_raw);
        masks = reinterpret_cast<unsigned *>(smem_raw + pool_ring_offset());
        lane_id = threadIdx.x & 31; idle_spins = 0; rot = threadIdx.x >> 5;
        n_lanes = n; lane_counter = counter;
        if (threadIdx.x == 0) {
            ps->retired = 0; ps->error = 0;
            ps->group = (GROUPMAP >> (4 * start_class)) & 15u;
        }
        constexpr unsigned per_bank = P / 32;
        for (int i = threadIdx.x; i < 8 * 32; i += blockDim.x)
            masks[i] = (i >> 5) == start_class ? (per_bank >= 32u ? FULL : ((1u << per_bank) - 1u)) : 0u;
    }

    // Returns false when every slot of the CTA has retired (or the watchdog fired).  has: this lane holds a slot.
    MIB_D bool claim(int &cls, bool &has, int &slot) {
        volatile unsigned *vm = masks;
        while (true) {
#ifdef MIB_POOL_DEBUG
            if (++dbg_spins > MIB_POOL_DEBUG_SPINS) {
                if (lane_id == 0) {
                    printf("POOL claim spins cta %d warp %d: retired %u of %d, error %u, group %u, iters %u, waiting per class:", blockIdx.x, threadIdx.x >> 5, ps->retired, P, ps->error, ps->group, dbg_iters);
                    for (int c = 0; c < 8; ++c) { unsigned n = 0; for (int b = 0; b < 32; ++b) n += __popc(masks[c * 32 + b]); printf(" %u", n); }
                    printf("\n");
                    int shown = 0;
                    for (int i = 0; i < P && shown < 12; ++i)
                        if (dbg_state[i] == 1) { printf("  slot %d in flight since class %d\n", i, (int) dbg_class[i]); ++shown; }
                }
                __trap();
            }
#endif
            unsigned kin = 0, kout = 0;
#if MIB_POOL_GROUPS
            // lane 0's view of the group for the whole warp: other warps change ps->group while this one reads it, and the
            // class chosen below must be the same for all 32 lanes
            const unsigned g = __shfl_sync(FULL, *(volatile unsigned *) &ps->group, 0);
#else
            const unsigned g = 0;
#endif
#pragma unroll
            for (unsigned c = 0; c < 8; ++c) {
#if MIB_POOL_STEAL
                const unsigned n = min(32u, __reduce_add_sync(FULL, (unsigned) __popc(vm[c * 32 + lane_id])));   // waiting slots
#else
                const unsigned n = (unsigned) __popc(__ballot_sync(FULL, vm[c * 32 + lane_id] != 0u));          // banks with one
#endif
                const unsigned key = n ? ((n << 3) | c) : 0u;
                if (((GROUPMAP >> (4u * c)) & 15u) == g) kin = max(kin, key); else kout = max(kout, key);
            }
            const unsigned n_in = kin >> 3, n_out = kout >> 3;
            int best = -1;
            bool flip = false;
#if MIB_POOL_STAY
            // stay with the current group while it still fills kPoolStay lanes, or nothing outside would fill more; otherwise
            // move the CTA on to the fullest class (round 1's rule drained a group to its last slot: half-empty batches)
            if (n_in >= (unsigned) MIB_POOL_STAY || (n_in > 0u && n_in >= n_out)) best = (int) (kin & 7u);
            else if (n_out > 0u) { best = (int) (kout & 7u); flip = true; }
#else
            if (n_in >= kPoolFull || (n_in > 0u && n_out < kPoolFull)) best = (int) (kin & 7u);
            else if (n_out > 0u) { best = (int) (kout & 7u); flip = true; }
#endif
            if (best >= 0) {
                unsigned m = vm[best * 32 + lane_id];
                bool ok = false;
                int s = 0;
                while (m) {
                    // start the search at a rotating bit: no slot of a bank waits for ever behind busier ones
                    const unsigned r = rot & 31u, mr = __funnelshift_r(m, m, r);
                    const unsigned j = ((unsigned) __ffs(mr) - 1u + r) & 31u;
                    const unsigned old = atomicAnd(&masks[best * 32 + lane_id], ~(1u << j));
                    if ((old >> j) & 1u) { ok = true; s = (int) (j * 32u + lane_id); break; }
                    m = old & ~(1u << j);
                }
#if MIB_POOL_STEAL
                // Lanes whose own bank has nothing waiting in this class take a slot of one of the next banks instead of
                // sitting the batch out: their field accesses then share a bank with that bank's own lane (2-way conflict for
                // these lanes only), which costs far less than an idle lane costs in issue slots.
                if (!ok) {
#pragma unroll 1
                    for (unsigned d = 1; d <= (unsigned) MIB_POOL_STEAL && !ok; ++d) {
                        const unsigned b = (lane_id + d) & 31u;
                        m = vm[best * 32 + b];
                        while (m) {
                            const unsigned j = (unsigned) __ffs(m) - 1u;
                            const unsigned old = atomicAnd(&masks[best * 32 + b], ~(1u << j));
                            if ((old >> j) & 1u) { ok = true; s = (int) (j * 32u + b); break; }
                            m = old & ~(1u << j);
                        }
                    }
                }
#endif
#ifdef MIB_POOL_DEBUG
                if (ok) {
                    if (dbg_state[s] != 0) dbg_fail("claimed a slot that is not waiting", s, best);
                    if (dbg_class[s] != best) dbg_fail("claimed a slot in the wrong class", s, best);
                    dbg_state[s] = 1;
                }
#endif
                if (__any_sync(FULL, ok)) {
#ifdef MIB_POOL_STATS
                    // batches and claimed lanes per class behind the event counters (mib_api.cu prints them: MIB_POOL_STATS build only)
                    { const unsigned nb = (unsigned) __popc(__ballot_sync(FULL, ok));
                      if (lane_id == 0) { atomicAdd(lane_counter + 2 + 2 * best, 1ull); atomicAdd(lane_counter + 3 + 2 * best, (unsigned long long) nb); } }
#endif
                    if (flip && lane_id == 0) ps->group = (GROUPMAP >> (4u * (unsigned) best)) & 15u;
                    ++rot;
                    idle_spins = 0;                    // the idle watchdog counts CONSECUTIVE polls
                    cls = best; has = ok; slot = s;
                    __threadfence_block();
                    return true;
                }
                continue;                              // other warps took the slots between the look and the claim
            }
            int stop = 0;
            if (lane_id == 0) {
                stop = *(volatile unsigned *) &ps->retired >= (unsigned) P || *(volatile unsigned *) &ps->error;
                if (!stop) {
                    __nanosleep(100);
                    if (++idle_spins > kPoolSpinLimit) { ps->error = 1; stop = 1; }
                }
            }
            stop = __shfl_sync(FULL, stop, 0);
            if (stop) { cls = -1; has = false; slot = 0; return false; }
        }
    }

    // Hand the lane's slot to class `next` (< 0: the lane held no slot).
    MIB_D void push(int next, int slot) {
#ifdef MIB_POOL_DEBUG
        if (next >= 8 || slot < 0 || slot >= P) { printf("POOL bad push next %d slot %d\n", next, slot); __trap(); }
        if (next >= 0) for (int c = 0; c < 8; ++c) if (masks[c * 32 + (slot & 31)] & (1u << ((unsigned) slot >> 5))) { printf("POOL slot %d pushed to %d while waiting in %d\n", slot, next, c); __trap(); }
        if (++dbg_iters > MIB_POOL_DEBUG_ITERS && lane_id == 0) {
            printf("POOL stuck cta %d warp %d: retired %u of %d, group %u, waiting per class:", blockIdx.x, threadIdx.x >> 5, ps->retired, P, ps->group);
            for (int c = 0; c < 8; ++c) { unsigned n = 0; for (int b = 0; b < 32; ++b) n += __popc(masks[c * 32 + b]); printf(" %u", n); }
            printf("\n"); __trap();
        }
#endif
#ifdef MIB_POOL_DEBUG
        if (next >= 0) {
            if (dbg_state[slot] != 1) dbg_fail("pushed a slot that is not in flight", slot, next);
            dbg_class[slot] = (unsigned char) next; dbg_state[slot] = 0;
        }
#endif
        __threadfence_block();
        if (next >= 0) atomicOr(&masks[next * 32 + (slot & 31)], 1u << ((unsigned) slot >> 5));
    }
#ifdef MIB_POOL_DEBUG
    MIB_D void retire(int slot) {
        if (dbg_state[slot] != 1) dbg_fail("retired a slot that is not in flight", slot, -1);
        dbg_state[slot] = 2;
        atomicAdd(&ps->retired, 1u);
    }
#else
    MIB_D void retire(int) { atomicAdd(&ps->retired, 1u); }
#endif

    // START: the lanes with `need` take the next lanes of the launch (one atomic per warp on the launch's lane counter);
    // kNoLane when the launch has run out of lanes.  Consecutive lanes are samples of one pixel, so a warp's batch starts
    // out as 32 neighbouring primary rays.
    MIB_D uint32_t take_lanes(bool need) {
        const unsigned m = __ballot_sync(FULL, need);
        if (!m) return kNoLane;
        const unsigned want = (unsigned) __popc(m), rank = (unsigned) __popc(m & ((1u << lane_id) - 1u));
        unsigned long long v = 0;
        if (lane_id == 0) v = atomicAdd(lane_counter, (unsigned long long) want);
        v = __shfl_sync(FULL, v, 0) + rank;
        return (need && v < n_lanes) ? (uint32_t) v : kNoLane;
    }
    MIB_D uint64_t lane_of(uint32_t j) const { return j; }          // index within the launch -> (virtual) lane of the rank
};
#else
// ---------------------------------------------------------------------------------------------
// Queue machinery shared by the forward and the adjoint pool kernels.
// ---------------------------------------------------------------------------------------------
// All slots start in class `start_class` (ring position i holds slot i).
template <int P>
MIB_D void pool_init(PoolShared *ps, volatile uint16_t *ring, unsigned total_local, int start_class) {
    if (threadIdx.x == 0) {
        for (int c = 0; c < 8; ++c) { ps->head[c] = 0; ps->tail[c] = 0; }
        ps->tail[start_class] = (unsigned) P;
        ps->next_local = 0;
        ps->total_local = total_local;
        ps->retired = 0;
        ps->error = 0;
        ps->group = (kFwdGroupMap >> (4 * start_class)) & 15u;   // the group of the start class
    }
    constexpr int R = pool_ring_cap<P>();
    for (int i = threadIdx.x; i < kPoolRings * R; i += blockDim.x) {
        int c = i / R, k = i - c * R;
        ring[i] = (uint16_t) ((c == start_class && k < P) ? k : kRingEmpty);
    }
}

// Claim up to 32 slots of ONE class: the warp scans the queues (lane c reads class c), lane 0 moves one head with a CAS,
// and every lane then takes its slot id out of the ring (a producer that has reserved the position but not yet written it is waited for).
// Policy (MIB_POOL_GROUPS): the classes form groups (GROUPMAP) -- the loops of the walk (MAIN, NEE: a slot is executed
// several times in a row), the scatter stages, and START / BSDF.  All warps of the CTA serve the current group until it has no
// full batch left and then switch together, so that the code live in the instruction cache at any time is one
// group's, not the whole kernel's (two groups: +8 ... +29 %, three: another +3 ... +7 %, profiles/r01_ab_variants.md).
// Returns false when every slot of the CTA has retired (or the watchdog fired).
template <int P, unsigned GROUPMAP>
MIB_D bool pool_claim(PoolShared *ps, volatile uint16_t *ring, unsigned lane_id, unsigned &idle_spins, int &cls, unsigned &n, int &slot) {
    constexpr unsigned FULL = 0xffffffffu;
    int c_sel = -1;
    unsigned base = 0, take = 0, patience = 0;
    // The whole warp runs the scan: lane c reads head / tail of class c, three shuffle steps find the fullest queue
    // inside and outside the current group (a lane-0-only scan cost ~12 % of all issued instructions).
    while (true) {
        unsigned hv = 0, av = 0;
        if (lane_id < 8u) {
            hv = *(volatile unsigned *) &ps->head[lane_id];
            av = *(volatile unsigned *) &ps->tail[lane_id] - hv;
        }
#if MIB_POOL_GROUPS
        const unsigned g = *(volatile unsigned *) &ps->group;
        const bool in = lane_id < 8u && (((GROUPMAP >> (4u * lane_id)) & 15u) == g);
#else
        const unsigned g = 0;
        const bool in = lane_id < 8u;
#endif
        unsigned kin = in ? ((av << 3) | lane_id) : 0u, kout = (!in && lane_id < 8u) ? ((av << 3) | lane_id) : 0u;
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
            kin = max(kin, __shfl_xor_sync(FULL, kin, o));
            kout = max(kout, __shfl_xor_sync(FULL, kout, o));
        }
        kin = __shfl_sync(FULL, kin, 0);
        kout = __shfl_sync(FULL, kout, 0);
        const unsigned n_in = kin >> 3, n_out = kout >> 3;
        int best = -1;
        unsigned n_best = 0;
        bool flip = false;
        if (n_in >= 32u || (n_in > 0u && n_out < 32u)) { best = (int) (kin & 7u); n_best = n_in; }
        else if (n_out > 0u) { best = (int) (kout & 7u); n_best = n_out; flip = true; }
        if (best >= 0) {
            // a batch that is not full costs as much instruction fetch as a full one: give the warps in flight a moment
            // to refill the queues before settling for a partial batch
            if (n_best < 32u && patience < (unsigned) kPoolPatience) { ++patience; __nanosleep(100); continue; }
            const unsigned hh = __shfl_sync(FULL, hv, best);
            const unsigned t = n_best < 32u ? n_best : 32u;
            int ok = 0;
            if (lane_id == 0) {
                ok = atomicCAS(&ps->head[best], hh, hh + t) == hh;
                if (ok && flip) ps->group = (GROUPMAP >> (4u * (unsigned) best)) & 15u;
            }
            ok = __shfl_sync(FULL, ok, 0);
            if (ok) { c_sel = best; base = hh; take = t; idle_spins = 0; break; }      // the idle watchdog counts CONSECUTIVE polls
            continue;
        }
        int stop = 0;
        if (lane_id == 0) {
            stop = *(volatile unsigned *) &ps->retired >= (unsigned) P || *(volatile unsigned *) &ps->error;
            if (!stop) {
                __nanosleep(100);
                if (++idle_spins > kPoolSpinLimit) { ps->error = 1; stop = 1; }
            }
        }
        stop = __shfl_sync(FULL, stop, 0);
        if (stop) break;
    }
    cls = c_sel; n = take; slot = 0;
    if (c_sel < 0) return false;
    if (lane_id < take) {
        constexpr unsigned R = (unsigned) pool_ring_cap<P>();
        unsigned pos = (base + lane_id) % R;
        unsigned v, sp = 0;
        while ((v = ring[c_sel * R + pos]) == kRingEmpty) {
            if (++sp > kPoolSpinLimit) { ps->error = 2; v = 0; break; }
        }
        ring[c_sel * R + pos] = (uint16_t) kRingEmpty;
        slot = (int) v;
    }
    return true;
}

// Append each lane's slot to the queue of class `next` (< 0: nothing to append): one atomic per class and warp.
template <int P>
MIB_D void pool_push(PoolShared *ps, volatile uint16_t *ring, unsigned lane_id, int next, int slot) {
    constexpr unsigned FULL = 0xffffffffu;
    __threadfence_block();
#pragma unroll
    for (int c = 0; c < 8; ++c) {
        unsigned m = __ballot_sync(FULL, next == c);
        if (m) {
            int leader = __ffs(m) - 1;
            unsigned b0 = 0;
            if ((int) lane_id == leader) b0 = atomicAdd(&ps->tail[c], (unsigned) __popc(m));
            b0 = __shfl_sync(FULL, b0, leader);
            if (next == c) {
                constexpr unsigned R = (unsigned) pool_ring_cap<P>();
                unsigned pos = (b0 + (unsigned) __popc(m & ((1u << lane_id) - 1u))) % R;
                if (ring[c * R + pos] != kRingEmpty) ps->error = 3;
                ring[c * R + pos] = (uint16_t) slot;
            }
        }
    }
}

// Lanes of CTA `cta`: interleaved chunks of kPoolChunk lanes (chunk c belongs to CTA c mod n_cta)
MIB_D uint64_t pool_local_total(uint64_t N, uint32_t cta, uint32_t n_cta) {
    const uint64_t n_chunks = (N + kPoolChunk - 1) / kPoolChunk;
    uint64_t owned = n_chunks > cta ? (n_chunks - 1 - cta) / n_cta + 1 : 0;
    if (!owned) return 0;
    uint64_t last = cta + (owned - 1) * n_cta;
    uint64_t last_size = (last == n_chunks - 1) ? N - last * kPoolChunk : (uint64_t) kPoolChunk;
    return (owned - 1) * kPoolChunk + last_size;
}

// CTA-local lane index -> lane of the wavefront (interleaved chunks)
MIB_D uint64_t pool_lane(uint32_t i, uint32_t cta, uint32_t n_cta) {
    uint64_t chunk = (uint64_t) cta + (uint64_t) (i / kPoolChunk) * n_cta;
    return chunk * kPoolChunk + (i % kPoolChunk);
}

// the round-1 rings behind the interface of the bank-owned bookkeeping
template <int P, unsigned GROUPMAP> struct PoolQ {
    static constexpr unsigned FULL = 0xffffffffu;
    PoolShared *ps;
    volatile uint16_t *ring;
    unsigned lane_id, idle_spins;
    MIB_D void init(unsigned char *smem_raw, uint64_t n, unsigned long long *, int start_class) {
        ps = reinterpret_cast<PoolShared *>(smem_raw);
        ring = reinterpret_cast<volatile uint16_t *>(smem_raw + pool_ring_offset());
        lane_id = threadIdx.x & 31; idle_spins = 0;
        pool_init<P>(ps, ring, (unsigned) pool_local_total(n, blockIdx.x, gridDim.x), start_class);
    }
    MIB_D bool claim(int &cls, bool &has, int &slot) {
        unsigned n;
        const bool ok = pool_claim<P, GROUPMAP>(ps, ring, lane_id, idle_spins, cls, n, slot);
        has = ok && lane_id < n;
        __threadfence_block();
        return ok;
    }
    MIB_D void push(int next, int slot) { pool_push<P>(ps, ring, lane_id, next, slot); }
    MIB_D uint32_t take_lanes(bool need) {
        const unsigned m = __ballot_sync(FULL, need);
        unsigned b0 = 0;
        if (m && (int) lane_id == __ffs(m) - 1) b0 = atomicAdd(&ps->next_local, (unsigned) __popc(m));
        if (m) b0 = __shfl_sync(FULL, b0, __ffs(m) - 1);
        if (!need) return kNoLane;
        const uint32_t jn = b0 + (uint32_t) __popc(m & ((1u << lane_id) - 1u));
        return jn < ps->total_local ? jn : kNoLane;
    }
    MIB_D uint64_t lane_of(uint32_t j) const { return pool_lane(j, blockIdx.x, gridDim.x); }
    MIB_D void retire(int) { atomicAdd(&ps->retired, 1u); }
};
#endif

// V: id of the scene-specialised instantiation (inst.cu compiles one per MIB_FEAT set; 0 = every feature)
template <int SPEC, int MODE, int V>
__global__ void __launch_bounds__(MIB_POOL_NT, 1) k_forward_pool(const __grid_constant__ LaunchParams lp) {
    using L = PoolL<SPEC>;
    constexpr int C = RawChannels<SPEC>::C, NR = ST<SPEC>::NR, NC = ST<SPEC>::NC, NT = ST<SPEC>::NT;
    constexpr bool AOS = MODE == 1;
    constexpr unsigned FULL = 0xffffffffu;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int P = L::P;
    static_assert(pool_smem_bytes<SPEC>() <= 232448, "pool does not fit the shared memory of sm_100");
    PoolMem<P> pm;
    pm.pw = reinterpret_cast<uint32_t *>(smem_raw + pool_state_offset(P));
#ifdef MIB_POOL_DEBUG
    pm.n_fields = L::F_COUNT;
#endif
    const mib_sensor &sn = sc.d.sensor;
    const unsigned lane_id = threadIdx.x & 31;

    PoolQ<P, kFwdGroupMap> q;
    q.init(smem_raw, lp.lane_end - lp.lane_begin, lp.counters + 6, C_START);
#ifdef MIB_POOL_DEBUG
    q.dbg_init(smem_raw + pool_state_offset(P) + (size_t) L::F_COUNT * P * sizeof(uint32_t), C_START);
#endif
    PoolShared *ps = q.ps;
    for (int s = threadIdx.x; s < P; s += blockDim.x) pm.su(L::F_LANE, s, kNoLane);
    __syncthreads();

    Counters cnt = { 0, 0, 0, 0, 0 };
    unsigned long long n_samples = 0;
    double dacc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) dacc[c] = 0.0;
    const float scale_x = 1.f / (float) sn.crop_width, scale_y = 1.f / (float) sn.crop_height;
    const float offset_x = -(float) sn.crop_offset_x * scale_x, offset_y = -(float) sn.crop_offset_y * scale_y;

    // integrator.cpp:323-340 (lane -> pixel) + render_sample :1139-1170 (film position, sensor ray)
    auto start = [&](uint64_t lane, Pcg32 &rng, float &sx, float &sy) -> Ray {
        lane = stripe_map(lp, lane);
        rng.seed_lane(lp.seed_value, (uint32_t) lane);
        uint32_t idx = (uint32_t) (lane / lp.spp);
        uint32_t py = idx / (uint32_t) sn.crop_width, px = idx - (uint32_t) sn.crop_width * py;
        float posx = (float) ((int) px + sn.crop_offset_x), posy = (float) ((int) py + sn.crop_offset_y);
        float jx = rng.next(), jy = rng.next();
        sx = posx + jx; sy = posy + jy;
        float ax = fma_(sx, scale_x, offset_x), ay = fma_(sy, scale_y, offset_y);
        float apx = .5f, apy = .5f;
        if (sn.needs_aperture_sample) { apx = rng.next(); apy = rng.next(); }
        return sensor_ray(ax, ay, apx, apy);
    };
    // Film / render_1 accumulation of one finished sample (imageblock.cpp:206-232 box put)
    auto finish = [&](uint64_t lane, float *R) {
        n_samples++;
        Pcg32 tmp;
        float sx, sy;
        Ray ray0 = start(lane, tmp, sx, sy);           // film position and sensor ray are functions of the lane only
        float aov[C];
        if constexpr (SPEC == MIB_SPEC_RGB) { aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = 1.f; }
        else if constexpr (SPEC == MIB_SPEC_MONO) { aov[0] = R[0]; aov[1] = 1.f; }
        else { to_sensor_stokes(ray0, R); aov[0] = R[0]; aov[1] = R[1]; aov[2] = R[2]; aov[3] = R[3]; }
        if (lp.mode == 2) {
#pragma unroll
            for (int c = 0; c < C; ++c) lp.per_sample[(lane - lp.lane_begin) * C + c] = aov[c];
            return;
        }
        if (lp.mode == 0 && (MIB_FT(FT_FILTER) && sn.rfilter != MIB_RFILTER_BOX)) {
            filter_footprint(lp.spp, sx, sy, [&](size_t pixel, float w) {
#pragma unroll
                for (int c = 0; c < C; ++c) atomicAdd(lp.film + pixel * C + c, aov[c] * w);
            });
            return;
        }
        int ix = (int) floorf(sx) - sn.crop_offset_x, iy = (int) floorf(sy) - sn.crop_offset_y;
        if ((uint32_t) ix >= (uint32_t) sn.crop_width || (uint32_t) iy >= (uint32_t) sn.crop_height) return;
        if (lp.mode == 1) {
#pragma unroll
            for (int c = 0; c < C; ++c) dacc[c] += (double) aov[c];
        } else {
            size_t pixel = (size_t) iy * sn.crop_width + ix;
            if constexpr (C == 2) atomicAdd(reinterpret_cast<float2 *>(lp.film) + pixel, make_float2(aov[0], aov[1]));
            else atomicAdd(reinterpret_cast<float4 *>(lp.film) + pixel, make_float4(aov[0], aov[1], aov[2], aov[3]));
        }
    };

    while (true) {
        // ---- 1. claim a batch of up to 32 slots of one class ----
        int cls, slot;
        bool has;
        if (!q.claim(cls, has, slot)) break;

        // ---- 2.-3. run the stage ----
        int next = -1;
        switch (cls) {
        case C_START: {
            // finish the path that ended in this slot, then start the CTA's next lane in it
            if (has) {
                uint32_t j = pm.u(L::F_LANE, slot);
                if (j != kNoLane) {
                    float R[NR];
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] = pm.f(L::F_R + r, slot);
                    finish(lp.lane_begin + q.lane_of(j), R);
                }
            }
            const uint32_t j = q.take_lanes(has);
            if (has) {
                if (j != kNoLane) {
                    Pcg32 rng;
                    float sx, sy;
                    Ray ray0 = start(lp.lane_begin + q.lane_of(j), rng, sx, sy);
                    WalkState<SPEC> st;
                    walk_init<SPEC>(rng, ray0, st);
                    pm.store_rng(slot, rng);
                    pm.su(L::F_LANE, slot, j);
                    pm.su(L::F_DEPTH, slot, 0u);
                    pm.su(L::F_FLAGS, slot, st.channel | (st.specular_chain ? FL_SPEC_CHAIN : 0u) | FL_NEEDS_INT |
                                                ((uint32_t) (st.medium + 1) << 8));
                    pm.sf(L::F_ETA, slot, 1.f);
                    pm.sf(L::F_LASTPDF, slot, 1.f);
                    if ((MIB_FT(FT_AREA) && sc.has_area_emitter)) pm.sv3(L::F_LSP, slot, v3(0.f));
                    pm.sv3(L::F_RAY, slot, ray0.o); pm.sv3(L::F_RAY + 3, slot, ray0.d); pm.sf(L::F_RAY + 6, slot, ray0.maxt);
                    pm.store_si(L::F_SI, slot, st.si);
#pragma unroll
                    for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
#pragma unroll
                    for (int r = 0; r < NR; ++r) pm.sf(L::F_R + r, slot, 0.f);
                    next = C_MAIN;
                } else {
                    pm.su(L::F_LANE, slot, kNoLane);
                    q.retire(slot);
                }
            }
            break;
        }
        case C_MAIN: {
            if (has) {
                WalkState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.depth = pm.u(L::F_DEPTH, slot);
                st.channel = fl & FL_CHANNEL;
                st.specular_chain = fl & FL_SPEC_CHAIN;
                st.needs_intersection = fl & FL_NEEDS_INT;
                st.medium = (int) fl_medium(fl) - 1;
                st.eta = pm.f(L::F_ETA, slot);
                st.last_pdf = pm.f(L::F_LASTPDF, slot);
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                pm.load_si(L::F_SI, slot, st.si);
                if (!st.needs_intersection && st.si.shape >= 0) si_finalize(st.ray.d, st.si);
                st.last_p = (MIB_FT(FT_AREA) && sc.has_area_emitter) ? pm.v3f(L::F_LSP, slot) : v3(0.f);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
#pragma unroll
                for (int r = 0; r < NR; ++r) st.R[r] = pm.f(L::F_R + r, slot);
                float u[NC];
                bool T_dirty = false, R_dirty = false;
                int ev = EV_NULL;
                for (int it = 0; it < kPoolMainIters && ev == EV_NULL; ++it)
                    ev = stg_main<SPEC, AOS>(rng, cnt, st, u, T_dirty, R_dirty);
                pm.store_rng_state(slot, rng);
                if (T_dirty) {
#pragma unroll
                    for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                }
                if (R_dirty) {
#pragma unroll
                    for (int r = 0; r < NR; ++r) pm.sf(L::F_R + r, slot, st.R[r]);
                }
                if (ev == EV_DONE) next = C_START;
                else {
                    pm.su(L::F_DEPTH, slot, st.depth);
                    fl = (fl & ~(FL_NEEDS_INT | FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | FL_NEE_MULW | FL_NEE_SURFACE | 0xff0000u)) |
                         (st.needs_intersection ? FL_NEEDS_INT : 0u);
                    pm.sv3(L::F_RAY, slot, st.ray.o); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                    pm.store_si(L::F_SI, slot, st.si);
                    if (ev == EV_NULL) next = C_MAIN;
                    else if (ev == EV_SCATTER) {
                        pm.sv3(L::F_SCP, slot, st.scatter_p);
#pragma unroll
                        for (int c = 0; c < NC; ++c) pm.sf(L::F_NEV + c, slot, u[c]);
                        next = C_SCAT;
                    } else if (ev == EV_NEE) {
                        const NeeState<SPEC> &ns = st.nee;
                        pm.sv3(L::F_NRAY, slot, ns.ray.o); pm.sv3(L::F_NRAY + 3, slot, ns.ray.d);
                        pm.store_si(L::F_NSI, slot, ns.si);
                        pm.sv3(L::F_NED, slot, ns.emitter_d);
#pragma unroll
                        for (int c = 0; c < NC; ++c) { pm.sf(L::F_NTR + c, slot, ns.tr[c]); pm.sf(L::F_NEV + c, slot, ns.ev[c]); }
                        pm.sf(L::F_NDIST, slot, ns.total_dist); pm.sf(L::F_NDIST + 1, slot, ns.max_dist); pm.sf(L::F_NDIST + 2, slot, ns.ds_dist);
#pragma unroll
                        for (int r = 0; r < NR; ++r) pm.sf(L::F_NA + r, slot, ns.a[r]);
                        pm.sf(L::F_NW, slot, ns.w);
                        fl |= (ns.active ? FL_NEE_ACTIVE : 0u) | (ns.needs_intersection ? FL_NEE_NEEDS_INT : 0u) |
                              (ns.mul_w_after ? FL_NEE_MULW : 0u) | FL_NEE_SURFACE | ((uint32_t) (ns.medium + 1) << 16);
                        next = C_NEE;
                    } else next = C_BSDF;
                    pm.su(L::F_FLAGS, slot, fl);
                }
            }
            break;
        }
        case C_SCAT: {
            if (has) {
                WalkState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.channel = fl & FL_CHANNEL;
                st.ray.d = pm.v3f(L::F_RAY + 3, slot);
                st.scatter_p = pm.v3f(L::F_SCP, slot);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                float u[NC];
#pragma unroll
                for (int c = 0; c < NC; ++c) u[c] = pm.f(L::F_NEV + c, slot);
                bool nee = stg_scatter<SPEC, AOS>(rng, st, u);
#pragma unroll
                for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                fl = (fl & ~(FL_SPEC_CHAIN | FL_NEE_ACTIVE | FL_NEE_NEEDS_INT | FL_NEE_MULW | FL_NEE_SURFACE | 0xff0000u)) |
                     (st.specular_chain ? FL_SPEC_CHAIN : 0u);
                if (nee) {
                    const NeeState<SPEC> &ns = st.nee;
                    pm.store_rng_state(slot, rng);
                    pm.sv3(L::F_NRAY, slot, ns.ray.o); pm.sv3(L::F_NRAY + 3, slot, ns.ray.d);
                    pm.store_si(L::F_NSI, slot, ns.si);
                    pm.sv3(L::F_NED, slot, ns.emitter_d);
#pragma unroll
                    for (int c = 0; c < NC; ++c) { pm.sf(L::F_NTR + c, slot, ns.tr[c]); pm.sf(L::F_NEV + c, slot, ns.ev[c]); }
                    pm.sf(L::F_NDIST, slot, ns.total_dist); pm.sf(L::F_NDIST + 1, slot, ns.max_dist); pm.sf(L::F_NDIST + 2, slot, ns.ds_dist);
#pragma unroll
                    for (int r = 0; r < NR; ++r) pm.sf(L::F_NA + r, slot, ns.a[r]);
                    pm.sf(L::F_NW, slot, ns.w);
                    fl |= (ns.active ? FL_NEE_ACTIVE : 0u) | (ns.needs_intersection ? FL_NEE_NEEDS_INT : 0u) |
                          (ns.mul_w_after ? FL_NEE_MULW : 0u) | ((uint32_t) (ns.medium + 1) << 16);
                    next = C_NEE;
                } else next = C_PHASE;
                pm.su(L::F_FLAGS, slot, fl);
            }
            break;
        }
        case C_NEE: {
            if (has) {
                NeeState<SPEC> ns;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                const uint32_t channel = fl & FL_CHANNEL;
                ns.active = fl & FL_NEE_ACTIVE;
                ns.needs_intersection = fl & FL_NEE_NEEDS_INT;
                ns.medium = (int) fl_nee_medium(fl) - 1;
                ns.ray.o = pm.v3f(L::F_NRAY, slot); ns.ray.d = pm.v3f(L::F_NRAY + 3, slot); ns.ray.maxt = 0.f;
                pm.load_si(L::F_NSI, slot, ns.si);
                if (!ns.needs_intersection && ns.si.shape >= 0) si_finalize(ns.ray.d, ns.si);
                ns.emitter_d = pm.v3f(L::F_NED, slot);
#pragma unroll
                for (int c = 0; c < NC; ++c) ns.tr[c] = pm.f(L::F_NTR + c, slot);
                ns.total_dist = pm.f(L::F_NDIST, slot); ns.max_dist = pm.f(L::F_NDIST + 1, slot); ns.ds_dist = pm.f(L::F_NDIST + 2, slot);
                for (int it = 0; it < kPoolNeeIters && ns.active; ++it) nee_step<SPEC, AOS>(rng, cnt, channel, ns);
                pm.store_rng_state(slot, rng);
                if (ns.active) {
                    pm.sv3(L::F_NRAY, slot, ns.ray.o); pm.sv3(L::F_NRAY + 3, slot, ns.ray.d);
                    pm.store_si(L::F_NSI, slot, ns.si);
#pragma unroll
                    for (int c = 0; c < NC; ++c) pm.sf(L::F_NTR + c, slot, ns.tr[c]);
                    pm.sf(L::F_NDIST, slot, ns.total_dist);
                    fl = (fl & ~(FL_NEE_NEEDS_INT | 0xff0000u)) | (ns.needs_intersection ? FL_NEE_NEEDS_INT : 0u) |
                         ((uint32_t) (ns.medium + 1) << 16);
                    pm.su(L::F_FLAGS, slot, fl);
                    next = C_NEE;
                } else {
#pragma unroll
                    for (int c = 0; c < NC; ++c) ns.ev[c] = pm.f(L::F_NEV + c, slot);
#pragma unroll
                    for (int r = 0; r < NR; ++r) ns.a[r] = pm.f(L::F_NA + r, slot);
                    ns.w = pm.f(L::F_NW, slot);
                    ns.mul_w_after = fl & FL_NEE_MULW;
                    float R[NR];
#pragma unroll
                    for (int r = 0; r < NR; ++r) R[r] = pm.f(L::F_R + r, slot);
                    nee_finish<SPEC>(ns, R);
#pragma unroll
                    for (int r = 0; r < NR; ++r) pm.sf(L::F_R + r, slot, R[r]);
                    next = (fl & FL_NEE_SURFACE) ? C_BSDF : C_PHASE;
                }
            }
            break;
        }
        case C_PHASE: {
            if (has) {
                WalkState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.channel = fl & FL_CHANNEL;
                st.needs_intersection = fl & FL_NEEDS_INT;
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                st.scatter_p = pm.v3f(L::F_SCP, slot);
                st.last_pdf = pm.f(L::F_LASTPDF, slot);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                stg_phase<SPEC, AOS>(rng, st);
                pm.store_rng_state(slot, rng);
#pragma unroll
                for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                pm.sv3(L::F_RAY, slot, st.ray.o); pm.sv3(L::F_RAY + 3, slot, st.ray.d); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                pm.sf(L::F_LASTPDF, slot, st.last_pdf);
                pm.su(L::F_FLAGS, slot, (fl & ~FL_NEEDS_INT) | (st.needs_intersection ? FL_NEEDS_INT : 0u));
                next = C_MAIN;
            }
            break;
        }
        default: {   // C_BSDF
            if (has) {
                WalkState<SPEC> st;
                Pcg32 rng;
                pm.load_rng(slot, rng);
                uint32_t fl = pm.u(L::F_FLAGS, slot);
                st.medium = (int) fl_medium(fl) - 1;
                st.channel = fl & FL_CHANNEL;
                st.specular_chain = fl & FL_SPEC_CHAIN;
                st.depth = pm.u(L::F_DEPTH, slot);
                st.last_pdf = pm.f(L::F_LASTPDF, slot);
                st.last_p = (MIB_FT(FT_AREA) && sc.has_area_emitter) ? pm.v3f(L::F_LSP, slot) : v3(0.f);
                st.ray.o = pm.v3f(L::F_RAY, slot); st.ray.d = pm.v3f(L::F_RAY + 3, slot); st.ray.maxt = pm.f(L::F_RAY + 6, slot);
                pm.load_si(L::F_SI, slot, st.si);
                si_finalize(st.ray.d, st.si);
#pragma unroll
                for (int i = 0; i < NT; ++i) st.T[i] = pm.f(L::F_T + i, slot);
                stg_bsdf<SPEC, AOS>(rng, st);
                pm.store_rng_state(slot, rng);
#pragma unroll
                for (int i = 0; i < NT; ++i) pm.sf(L::F_T + i, slot, st.T[i]);
                pm.sv3(L::F_RAY, slot, st.ray.o); pm.sv3(L::F_RAY + 3, slot, st.ray.d); pm.sf(L::F_RAY + 6, slot, st.ray.maxt);
                pm.sf(L::F_LASTPDF, slot, st.last_pdf);
                if ((MIB_FT(FT_AREA) && sc.has_area_emitter)) pm.sv3(L::F_LSP, slot, st.last_p);
                pm.su(L::F_DEPTH, slot, st.depth);
                pm.su(L::F_FLAGS, slot, (fl & ~(FL_NEEDS_INT | FL_SPEC_CHAIN | 0xff00u)) | FL_NEEDS_INT |
                                            (st.specular_chain ? FL_SPEC_CHAIN : 0u) | ((uint32_t) (st.medium + 1) << 8));
                next = C_MAIN;
            }
            break;
        }
        }
        // ---- 4. hand the slots to their next classes ----
        q.push(next, slot);
    }

    if (lp.mode == 1) {
#pragma unroll
        for (int c = 0; c < C; ++c) {
            double v = dacc[c];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
            if (lane_id == 0) atomicAdd(lp.sums + c, v);
        }
    }
    unsigned long long cv[6] = { cnt.n_lookups, cnt.n_real, cnt.n_null, cnt.n_nee_steps, cnt.n_surface, n_samples };
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        unsigned long long v = cv[k];
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
        if (lane_id == 0 && v) atomicAdd(lp.counters + k, v);
    }
    if (threadIdx.x == 0 && ps->error) atomicExch((unsigned int *) (lp.counters + 7), ps->error);
}

template <int SPEC, int MODE, int V> void launch_k_forward_pool(const DScene &scene, const LaunchParams &lp, unsigned blocks, unsigned threads, cudaStream_t st);
template <int SPEC, int MODE> int pool_slots();   // paths resident per CTA (PoolL<SPEC>::P)

} // namespace mib
