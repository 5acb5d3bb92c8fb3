// 2-player Hanabi step kernels (sm_100a): action apply, deck draw, obs / state / legal-move-mask encoding, score, done,
// in-step reset.  Behaviour follows src/hanabi_env/sim.cpp:45-891 of the reference including its quirks (cited in
// hanabi_game.cuh, which holds the game logic); the structure does not: one CTA owns a tile of 32 consecutive worlds,
// lane = world, a world's packed record (12 words) lives in the REGISTERS of its lane of the game warp, only the deck
// (50 bytes, indexed at random) sits in shared memory as [word][lane] (conflict free), and a second warp turns the rows
// the game warp composes as bit streams into bytes in HBM.
//
// HBM layout:
//   game   u32 [GAME_WORDS][N]    packed per-world record (13 deck words, 12 record words), word-major so a warp's 32 worlds coalesce
//   obs    u8  [2][N][658]   state u8 [2][N][783]   mask i32 [2][N][20]   active i32 [2][N]
//   reward f32 [2][N]        done  i32 [N]          actions i32 [2][N]
// Only the NEW current player's obs/state/mask rows are rewritten each step (sim.cpp:794-810); the other player's
// rows keep their previous content, exactly like the reference's exported columns.  On a reset both are rewritten.
#include <algorithm>
#include <cmath>
#include <cstring>

#include <cstdlib>
// back-off between two polls of a reset count that has not been published yet: the game warps are the older warps of
// their SM's schedulers and a spinning poll loop took issue slots (and ~200 MB of L2 reads per step) from the writing
// warps — 21.0 -> 20.5 us per fused step; Cartpole / Balance Beam (latency-bound) keep polling flat out
#ifndef B200_POLL_SLEEP_NS
#define B200_POLL_SLEEP_NS 400
#endif
#include "common.cuh"
#include "hanabi_game.cuh"
#include "hanabi_host.h"
#include "random_policy_host.h"
#include "sm_rates_host.h"

namespace b200 {
namespace hanabi {

static_assert(GAME_WORDS == DECK_WORDS + REC_WORDS, "game record = deck + register record");

struct Rules : GameRules {
    uint32_t fresh[DECK_WORDS];   // the deck of a freshly reset game (identical for every reset), built on the host
};

typedef DeckT<128> Deck;   // [word][lane] shared tile, base = tile + lane * 4

// the record's words in HBM / shared-memory order
#define HB_REC_WORDS(g, F)                                                                                       \
    F(0, (g).board) F(1, (g).turn) F(2, (g).lm) F(3, (g).disc_lo) F(4, (g).disc_hi) F(5, (g).rng)                 \
    F(6, (g).cur.cards) F(7, (g).cur.know) F(8, (g).cur.plaus) F(9, (g).oth.cards) F(10, (g).oth.know) F(11, (g).oth.plaus)

__device__ __forceinline__ void rec_load(Rec &g, const uint32_t *src, size_t stride)   // src[w * stride]
{
#define F(i, f) f = src[(size_t)(i) * stride];
    HB_REC_WORDS(g, F)
#undef F
}
__device__ __forceinline__ void rec_store(const Rec &g, uint32_t *dst, size_t stride)
{
#define F(i, f) dst[(size_t)(i) * stride] = f;
    HB_REC_WORDS(g, F)
#undef F
}
__device__ __forceinline__ Rec rec_shfl(const Rec &g, int src_lane)
{
    Rec r;
#define F(i, f) f = __shfl_sync(0xffffffffu, f, src_lane);
    r = g;
    HB_REC_WORDS(r, F)
#undef F
    return r;
}

// 16 bits -> 16 bytes of 0 / 1
__device__ __forceinline__ uint4 expand16(uint32_t v)
{
    uint4 r;
    r.x = ((v & 0xFu) * 0x00204081u) & 0x01010101u;
    r.y = (((v >> 4) & 0xFu) * 0x00204081u) & 0x01010101u;
    r.z = (((v >> 8) & 0xFu) * 0x00204081u) & 0x01010101u;
    r.w = (((v >> 12) & 0xFu) * 0x00204081u) & 0x01010101u;
    return r;
}

// The same through a table in shared memory: 8 bits -> 8 bytes per 64-bit load (256 entries, 2 KB per CTA).  The
// arithmetic expansion costs 12 integer-pipe instructions per 16-byte window (shift, mask, multiply, mask per nibble) and
// the row writer is what bounds both Hanabi kernels; the table costs two loads and three.  The rows are sparse one-hot
// vectors, so most lanes of a load hit entry 0 (one broadcast) and bank conflicts stay rare.  Per-step kernel: 29.9 ->
// 28.6 us per step.  NOT in the fused rollout: there the extra shared-memory traffic slows the game warps' chain that
// all tiles wait for every step (21.8 against 17.9 us).
#ifndef HB_LUT
#define HB_LUT 1
#endif
__device__ __forceinline__ void expand_lut_fill(uint2 *lut, int tid, int nthreads)
{
    for (int i = tid; i < 256; i += nthreads)
        lut[i] = make_uint2(((i & 0xF) * 0x00204081u) & 0x01010101u, ((i >> 4) * 0x00204081u) & 0x01010101u);
}
__device__ __forceinline__ uint4 expand16_lut(const uint2 *lut, uint32_t v)
{
    const uint2 lo = lut[v & 0xFFu], hi = lut[(v >> 8) & 0xFFu];
    return make_uint4(lo.x, lo.y, hi.x, hi.y);
}

// The warp writes bits [first_bit, first_bit + nbytes) of one shared row as `nbytes` bytes at dst (any alignment):
// the 16-byte-aligned interior goes out as 128-bit stores, one chunk per lane and iteration; the < 16 bytes before and
// after it are written one byte per lane (lanes 0-15 the head, lanes 16-31 the tail) — a byte IS its bit.
__device__ __forceinline__ void write_row(uint8_t *dst, const uint32_t *row, int first_bit, int nbytes, int lane)
{
    const int mis = (int)(reinterpret_cast<uintptr_t>(dst) & 15u);
    const int head = min((16 - mis) & 15, nbytes);
    const int nfull = (nbytes - head) >> 4;
    const int tail = nbytes - head - 16 * nfull;
    uint4 *d4 = reinterpret_cast<uint4 *>(dst + head);
    for (int j = lane; j < nfull; j += 32) {
        const int b = first_bit + head + 16 * j;
        d4[j] = expand16(__funnelshift_r(row[b >> 5], row[(b >> 5) + 1], b & 31) & 0xFFFFu);
    }
    const int i = lane & 15;
    const bool is_tail = lane >= 16;
    if (i < (is_tail ? tail : head)) {
        const int byte = is_tail ? head + 16 * nfull + i : i;
        const int b = first_bit + byte;
        dst[byte] = (uint8_t)((row[b >> 5] >> (b & 31)) & 1u);
    }
}

// Policy-ready rows (SURVEY 8(f) row 3 for Hanabi: every consumer starts with `obs.float()`,
// pantheonrl_extension/hanabi_agent.py:257,292): the same bits written as bfloat16 (E = uint16_t, 1.0 = 0x3F80) or float
// (E = float) ELEMENTS instead of bytes.  One 16-byte chunk = 8 / 4 elements from the low bits of v.
template <class E> __device__ __forceinline__ uint4 expand_chunk(uint32_t v);
template <> __device__ __forceinline__ uint4 expand_chunk<uint16_t>(uint32_t v)
{
    uint4 r;
    r.x = ((v & 1u) * 0x3F80u) | (((v >> 1) & 1u) * 0x3F800000u);
    r.y = (((v >> 2) & 1u) * 0x3F80u) | (((v >> 3) & 1u) * 0x3F800000u);
    r.z = (((v >> 4) & 1u) * 0x3F80u) | (((v >> 5) & 1u) * 0x3F800000u);
    r.w = (((v >> 6) & 1u) * 0x3F80u) | (((v >> 7) & 1u) * 0x3F800000u);
    return r;
}
template <> __device__ __forceinline__ uint4 expand_chunk<float>(uint32_t v)
{
    return make_uint4((v & 1u) * 0x3F800000u, ((v >> 1) & 1u) * 0x3F800000u, ((v >> 2) & 1u) * 0x3F800000u,
                      ((v >> 3) & 1u) * 0x3F800000u);
}
__device__ __forceinline__ void store_element(uint16_t *dst, uint32_t bit) { *dst = (uint16_t)(bit * 0x3F80u); }
__device__ __forceinline__ void store_element(float *dst, uint32_t bit) { *dst = bit ? 1.0f : 0.0f; }

// The warp writes bits [first_bit, first_bit + n) of one shared row as n elements at dst (element aligned): the
// 16-byte-aligned interior as 128-bit streaming stores, one chunk per lane and iteration (512 contiguous bytes per
// instruction), the < 16 bytes before and after it one element per lane.
template <class E>
__device__ __forceinline__ void write_row_typed(E *dst, const uint32_t *row, int first_bit, int n, int lane)
{
    constexpr int PER = 16 / (int)sizeof(E);
    const int mis = (int)((reinterpret_cast<uintptr_t>(dst) & 15u) / sizeof(E));
    const int head = min((PER - mis) % PER, n);
    const int nfull = (n - head) / PER;
    const int tail = n - head - PER * nfull;
    uint4 *d4 = reinterpret_cast<uint4 *>(dst + head);
    // a row has at most MAXIT chunks per lane: unrolled with predicates, so that the shared loads of all of them are in
    // flight together (a loop with a run-time trip count exposed one load latency per chunk)
    constexpr int MAXIT = (STATE * (int)sizeof(E) / 16 + 31) / 32 + 1;
#pragma unroll
    for (int it = 0; it < MAXIT; it++) {
        const int j = lane + 32 * it;
        if (j < nfull) {
            const int b = first_bit + head + PER * j;
            __stcs(d4 + j, expand_chunk<E>(__funnelshift_r(row[b >> 5], row[(b >> 5) + 1], b & 31)));
        }
    }
    const int i = lane & 15;
    const bool is_tail = lane >= 16;
    if (i < (is_tail ? tail : head)) {
        const int e = is_tail ? head + PER * nfull + i : i;
        const int b = first_bit + e;
        store_element(dst + e, (row[b >> 5] >> (b & 31)) & 1u);
    }
}

// n (1, 2, 4 or 8) bits -> n bytes of 0 / 1, stored with one n-byte access at an n-aligned address
__device__ __forceinline__ void store_bits(uint8_t *dst, uint32_t v, int n, uint64_t keep)
{
    // `keep` = an L2 evict_last policy: the partial sectors at the ends of a row are completed by the neighbouring
    // world's row a step or two later (players alternate); kept in L2 until then they never reach DRAM half-written
    // (a read-modify-write each, i.e. reads sprinkled into the write stream)
    const uint32_t lo = ((v & 0xFu) * 0x00204081u) & 0x01010101u;
    if (n == 1) asm volatile("st.global.L2::cache_hint.u8 [%0], %1, %2;" ::"l"(dst), "r"(v & 1u), "l"(keep) : "memory");
    else if (n == 2)
        asm volatile("st.global.L2::cache_hint.u16 [%0], %1, %2;" ::"l"(dst), "r"((v & 1u) | ((v & 2u) << 7)), "l"(keep) : "memory");
    else if (n == 4) asm volatile("st.global.L2::cache_hint.u32 [%0], %1, %2;" ::"l"(dst), "r"(lo), "l"(keep) : "memory");
    else
        asm volatile("st.global.L2::cache_hint.v2.u32 [%0], {%1, %2}, %3;" ::"l"(dst), "r"(lo),
                     "r"((((v >> 4) & 0xFu) * 0x00204081u) & 0x01010101u), "l"(keep) : "memory");
    // (as STREAMING stores they were evicted half-written and lost the whole gain of the streaming windows: 31.8 vs
    // 27.2 us per fused step; plain stores 27.2 / 39.3 per-step; evict_last 27.0 / 37.7)
}

// A row to write: world within the tile | player << 8 | valid << 16 (jobs of one batch sit in shared memory)
constexpr uint32_t JOB_VALID = 1u << 16;
__device__ __forceinline__ size_t job_row(uint32_t job, size_t N, uint32_t tile0)
{
    return (size_t)((job >> 8) & 0xFFu) * N + tile0 + (job & 0xFFu);
}

// The warp writes the first NB bytes of up to 32 bit rows (row r of the shared array -> NB bytes at
// base + job_row(jobs[r]) * NB).  Four lanes share a row: quad Q takes rows Q, Q+8, Q+16, Q+24 one after the other
// (a slice of 8 rows with no valid job is skipped, so a small batch costs a quarter), lane q of the quad the 16-byte
// destination windows q, q+4, q+8, ... of the row, so a store instruction writes 64 contiguous bytes per quad, every
// row is in flight in 1/4 of the instructions a row-at-a-time loop needs, and — the window loop being fully unrolled
// — a window costs two shared loads, a funnel shift, the expansion and the store: all offsets are immediates off
// per-row base registers.  The < 16 bytes before the first and after the last whole window of a row are written by
// lane r itself with at most four power-of-two stores each.
//
// What bounds this phase is HBM: in steady state every byte written evicts a dirty byte of the previous step from L2,
// and 123 MB per step at the ~6.4 TB/s a pure store kernel reaches here is 19 us.  (Tried: the same rows through the
// copy engine — `cp.async.bulk` from a staging buffer, 64 copies of ~700 B per tile — no faster, and its staging
// buffers cost a third of the resident tiles.)
// Slices [s_lo, s_hi) of the batch only (a slice = 8 rows): the per-step kernel shares a batch between its two warps.
template <int NB, bool LUT>
__device__ __forceinline__ void write_rows(uint8_t *base, size_t N, uint32_t tile0, const uint32_t *jobs,
                                           const uint32_t *s_bits, const uint2 *lut, int lane, int s_lo = 0, int s_hi = 4)
{
    constexpr int SLOTS = (NB + 46) / 16;   // windows between the 32-byte boundary below a row and its end, worst case
    constexpr int ITERS = (SLOTS + 3) / 4;
    const int q = lane & 3;
#pragma unroll 1
    for (int s = s_lo; s < s_hi; s++) {
        const int r = 8 * s + (lane >> 2);
        const uint32_t job = jobs[r];
        const bool on = job & JOB_VALID;
        if (!__any_sync(0xffffffffu, on)) continue;
        uint8_t *dst = base + job_row(job, N, tile0) * NB;
        // first bit (= byte) of window q; the quad's 64-byte runs start on 32-byte boundaries, so that every store
        // request is made of whole sectors (runs starting mid-sector moved 21 instead of 32 bytes per request cycle)
        const int c = 16 * q - (int)(reinterpret_cast<uintptr_t>(dst) & 31u);
        const uint32_t *row = s_bits + r * ROW_WORDS + (c >> 5);   // c < 0: word -1 is read, never used
        const int sh = c & 31;
        uint8_t *d = dst + c;
        // all the words this lane's windows need, loaded up front: one exposed shared-memory latency per slice
        // instead of one per window (the compiler otherwise reuses two registers and serialises load -> shift -> store)
        uint32_t wv[2 * ITERS];
#pragma unroll
        for (int i = 0; i < 2 * ITERS; i++) wv[i] = row[i];
#pragma unroll
        for (int i = 0; i < ITERS; i++) {
            bool ok = on;
            if (i == 0) ok = ok && c >= 0;
            if (64 * i + 48 + 16 > NB) ok = ok && c + 64 * i + 16 <= NB;
            if (ok) {
                if constexpr (LUT) __stcs(reinterpret_cast<uint4 *>(d + 64 * i), expand16_lut(lut, __funnelshift_r(wv[2 * i], wv[2 * i + 1], sh)));
                else __stcs(reinterpret_cast<uint4 *>(d + 64 * i), expand16(__funnelshift_r(wv[2 * i], wv[2 * i + 1], sh)));
            }
        }
    }
    const uint32_t job = jobs[lane];
    if ((job & JOB_VALID) && lane >= 8 * s_lo && lane < 8 * s_hi) {
        uint64_t keep;
        asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
        uint8_t *dst = base + job_row(job, N, tile0) * NB;
        const uint32_t *row = s_bits + lane * ROW_WORDS;
        const int head = (16 - (int)(reinterpret_cast<uintptr_t>(dst) & 15u)) & 15;
        const int tail = (NB - head) & 15;
        uint32_t v = row[0];
        int at = 0;
#pragma unroll
        for (int n = 1; n <= 8; n *= 2)   // 1, 2, 4, 8 bytes: each store lands n-aligned
            if (head & n) { store_bits(dst + at, v >> at, n, keep); at += n; }
        at = NB - tail;
        v = __funnelshift_r(row[at >> 5], row[(at >> 5) + 1], at & 31);
        int used = 0;
#pragma unroll
        for (int n = 8; n >= 1; n /= 2)
            if (tail & n) { store_bits(dst + at + used, v >> used, n, keep); used += n; }
    }
}

// legal-move masks of the same rows: 20 x i32 per row as 5 aligned 128-bit stores, the 160 chunks dealt to the lanes
__device__ __forceinline__ void write_masks(int32_t *mask, size_t N, uint32_t tile0, const uint32_t *jobs,
                                            const uint32_t *legal, int lane)
{
#pragma unroll
    for (int it = 0; it < MOVES / 4; it++) {
        const int idx = it * 32 + lane, r = idx / (MOVES / 4), c = idx % (MOVES / 4);
        const uint32_t job = jobs[r];
        if (job & JOB_VALID) {
            const uint32_t m = legal[r] >> (4 * c);
            __stcs(reinterpret_cast<int4 *>(mask + job_row(job, N, tile0) * MOVES + 4 * c),
                   make_int4(m & 1, (m >> 1) & 1, (m >> 2) & 1, (m >> 3) & 1));
        }
    }
}

struct Buffers {
    uint32_t *game;
    const int32_t *actions;
    uint8_t *obs, *state;
    int32_t *mask, *active, *done;
    float *reward;
    unsigned long long *trace;
    int32_t *actions_out;   // the action tensor again, writable (random rollouts draw the next move on the device)
    uint32_t N;
};

__device__ __forceinline__ void trace_stamp(const Buffers &B, uint32_t tile, int slot)
{
    if (B.trace && (threadIdx.x & 31) == 0) {
        unsigned long long t;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        B.trace[(size_t)tile * 8 + slot] = t;
    }
}

constexpr int CTA_THREADS = 64;   // two warps per tile
// per-step kernel: the writing warp writes the state rows of slices [0, HB_STEP_SPLIT) of the play batch (a slice = 8
// rows), the game warp — once it is through with its own work — the others (4 = the writing warp writes them all)
#ifndef HB_STEP_SPLIT
#define HB_STEP_SPLIT 2
#endif

// ONE CTA PER SM holds up to TILES_MAX tiles: warps [0, K) are the K tiles' game warps, warps [K, 2K) their writing
// warps (see the fused rollout below for what that order buys); the two warps of a tile hand over through mbarriers in
// shared memory (hardware named barriers are a pool of 16 per CTA)
constexpr int TILES_MAX = 14;

struct RowBatch {   // up to 32 composed rows handed from the game warp to the writing warp
    uint32_t job[32];     // world | player << 8 | JOB_VALID
    uint32_t len[32];     // obs bits | all bits << 16
    uint32_t legal[32];   // legal-move bits
};
// shared memory of one tile of the per-step kernel: decks + the final records (handed to the writing warp for the
// write-back), bit rows of the worlds that play on, bit rows of the reset worlds
struct alignas(16) StepTile {
    uint32_t game[GAME_WORDS * 32];     // [word][lane]: DECK_WORDS of deck, then the final record
    uint32_t bits[2][32 * ROW_WORDS];   // rows of the worlds that play on; rows of the reset worlds (one round of <= 32)
    RowBatch play, reset;
    uint32_t done_mask;
    uint64_t m_count, m_rows, m_decks;   // mbarriers, 32 arrivals each: reset count known, play rows composed, decks landed
};

// uniform random legal move for the NEXT step, straight into the action tensor (b200env_random_rollout): the same
// draw policy::sample_kernel makes from the mask tensor
struct NextAction {
    int32_t *actions;   // nullptr = off
    uint32_t seed, step;
};

// The variable-length rows (information tokens above their maximum: every section after the board moves) are rare and
// their code is long (no bit offset is a constant): ONE out-of-line copy, the record passed by value.
template <bool STD>
__device__ __noinline__ uint32_t compose_long_row(GameRules R, Rec g, int swap, uint32_t agent, uint32_t *row)
{
    if (STD) { R.colors = 5; R.ranks = 5; R.max_info = 8; R.max_life = 3; }
    BitWriter bw{row, 0u, 0, 0};
    const int obs_bits = swap ? generate_bits<false>(R, g, g.oth, g.cur, agent, bw) : generate_bits<false>(R, g, g.cur, g.oth, agent, bw);
    const int all_bits = bw.finish();
    return (uint32_t)obs_bits | ((uint32_t)all_bits << 16);
}

// composes one row (bit stream + legal moves) of world record `g` for `agent`, whose hand is g.cur (swap = 0) or g.oth
// (swap = 1: the other player's view of a fresh game), into shared memory; returns obs bits | all bits << 16
template <bool STD>
__device__ __forceinline__ uint32_t compose_row(GameRules R, const Rec &g, int swap, uint32_t agent, uint32_t *row, uint32_t *legal)
{
    if (STD) { R.colors = 5; R.ranks = 5; R.max_info = 8; R.max_life = 3; }
    const Hand mine = swap ? g.oth : g.cur, theirs = swap ? g.cur : g.oth;
    *legal = generate_mask(R, g, mine, theirs);
    // nominal section lengths unless the token counters exceed their maxima (sim.cpp:675-677)
    const int max_size = 2 * R.ranks * R.colors - HAND * NP;
    if ((int)(g.board & 0xFFu) <= max_size && (int)((g.board >> 8) & 0xFFu) <= R.max_info && (int)((g.board >> 16) & 0xFFu) <= R.max_life) {
        BitWriter bw{row, 0u, 0, 0};
        const int obs_bits = generate_bits<true>(R, g, mine, theirs, agent, bw);
        const int all_bits = bw.finish();
        return (uint32_t)obs_bits | ((uint32_t)all_bits << 16);
    }
    return compose_long_row<STD>(R, g, swap, agent, row);
}

// One batch of composed rows -> bytes in HBM, one tensor at a time (TENSOR 0 = observation rows, 1 = state rows)
// E = element type of the observation / state tensors: uint8_t (the reference's int8 export), uint16_t (bfloat16 bits), float
template <class E, int TENSOR, bool LUT>
__device__ __forceinline__ void write_tensor_rows(uint8_t *base, size_t N, uint32_t tile0, const RowBatch &rb, const uint32_t *bits,
                                                  const uint2 *lut, int lane, int s_lo = 0, int s_hi = 4)
{
    constexpr int NB = TENSOR ? STATE : OBS;
    if constexpr (sizeof(E) == 1) {
        write_rows<NB, LUT>(base, N, tile0, rb.job, bits, lut, lane, s_lo, s_hi);
    } else {   // one row at a time, the whole warp on it
#pragma unroll 1
        for (int r = 8 * s_lo; r < 8 * s_hi; r++) {
            const uint32_t job = rb.job[r];
            if (!(job & JOB_VALID)) continue;
            write_row_typed(reinterpret_cast<E *>(base) + job_row(job, N, tile0) * NB, bits + r * ROW_WORDS, 0, NB, lane);
        }
    }
    // information_tokens above its maximum made a row longer than 658 / 783 bytes: the excess lands on byte 0.. of the
    // NEXT player's rows, exactly where the reference's out-of-row writes go (dropped for the last player)
    const uint32_t job = rb.job[lane], lens = rb.len[lane];
    const int a = (int)((job >> 8) & 0xFFu);
    for (unsigned rem = __ballot_sync(0xffffffffu, (job & JOB_VALID) && (int)(lens & 0xFFFFu) > OBS && a + 1 < NP &&
                                                       lane >= 8 * s_lo && lane < 8 * s_hi); rem;) {
        const int r = __ffs(rem) - 1;
        rem &= rem - 1;
        const size_t row = job_row(rb.job[r], N, tile0) + N;
        const int end = TENSOR ? (int)(rb.len[r] >> 16) : (int)(rb.len[r] & 0xFFFFu);
        if constexpr (sizeof(E) == 1) write_row(base + row * NB, bits + r * ROW_WORDS, NB, end - NB, lane);
        else write_row_typed(reinterpret_cast<E *>(base) + row * NB, bits + r * ROW_WORDS, NB, end - NB, lane);
    }
}

// legal-move masks and (random rollouts, one launch per step) the next step's actions of the rows of one batch
__device__ __forceinline__ void write_batch_masks(const Buffers &B, const NextAction &next, uint32_t tile0, const RowBatch &rb, int lane)
{
    write_masks(B.mask, B.N, tile0, rb.job, rb.legal, lane);
    const uint32_t job = rb.job[lane];
    if (next.actions && (job & JOB_VALID)) {
        const size_t i = job_row(job, B.N, tile0);
        const uint32_t legal = rb.legal[lane], count = __popc(legal);
        const uint32_t r = policy::mix(next.seed, next.step, (uint32_t)i);
        next.actions[i] = count ? (int32_t)__fns(legal, 0, (int)(((uint64_t)r * count) >> 32) + 1) : 0;
    }
}

// the writing warp of the per-step kernel turns one batch into bytes in HBM: observation and state rows, masks, next actions
// (state rows: slices [0, state_slices) only — the game warp takes the others of the play batch, see step_kernel)
template <class E>
__device__ __forceinline__ void write_batch(const Buffers &B, const NextAction &next, uint32_t tile0, const RowBatch &rb,
                                            const uint32_t *bits, const uint2 *lut, int lane, int state_slices = 4)
{
    write_tensor_rows<E, 0, HB_LUT>(B.obs, B.N, tile0, rb, bits, lut, lane);
    write_tensor_rows<E, 1, HB_LUT>(B.state, B.N, tile0, rb, bits, lut, lane, 0, state_slices);
    write_batch_masks(B, next, tile0, rb, lane);
}

// ONE copy of each tensor's row writer for the fused rollout (out of line: the unrolled window loops are ~10 KB each
// and the kernel's two warps run different phases against one instruction cache).  Pointers and sizes travel as scalar
// arguments (registers): behind a reference the callee fetched them from the caller's stack frame, four
// long-scoreboard round trips per call.
template <class E, int TENSOR>
__device__ __noinline__ void write_tensor_rows_shared(uint8_t *base, uint32_t N, uint32_t tile0, const RowBatch *rb,
                                                      const uint32_t *bits, int lane)
{
    write_tensor_rows<E, TENSOR, false>(base, N, tile0, *rb, bits, nullptr, lane);
}

// STD = the standard game (5 colours, 5 ranks, 8 information and 3 life tokens): the rules become compile-time
// constants, so divisions by `ranks`, section lengths and (through generate_bits<true>) every bit offset fold away.
//
// A tile of 32 worlds = two warps.  The GAME warp (lane = world) holds the records in registers and touches shared
// memory for the deck and the composed rows only: action, score / done, reset count, bit rows + legal moves of the
// worlds that play on — handed to the WRITING warp, which turns them into bytes in HBM for the rest of the kernel.
// Then, while that warp streams, the game warp does the step's tail itself: episode numbers, resets, the records and
// players-to-move flags back to HBM, both players' rows of the reset worlds (composed and written), and the state
// rows of the play batch's last two slices.  (The stores of all the SM's warps share one path to the crossbar; what
// the split buys is that neither warp of a tile is left waiting for the other at the end.)
template <bool STD, class E>
__global__ void __launch_bounds__(CTA_THREADS * TILES_MAX, 1) step_kernel(const Rules Rp, const Buffers B, const EpisodeScan scan,
                                                                        const NextAction next, const uint32_t num_tiles,
                                                                        const int init_only)
{
    GameRules R = Rp;   // (the fresh deck, Rp.fresh, stays in the parameter bank)
    if (STD) { R.colors = 5; R.ranks = 5; R.max_info = 8; R.max_life = 3; }
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ uint32_t s_cta[3];
    __shared__ uint2 s_lut[256];
    expand_lut_fill(s_lut, threadIdx.x, blockDim.x);   // (before the block-wide barrier below)
    const uint2 *lut = s_lut;
    const int K = (int)(blockDim.x >> 6);   // tiles of this CTA
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool is_game = warp < K;
    const int k = is_game ? warp : warp - K;
    StepTile &sh = reinterpret_cast<StepTile *>(smem)[k];

    if (is_game && lane == 0) {
        mbar_init(&sh.m_count, 32); mbar_init(&sh.m_rows, 32); mbar_init(&sh.m_decks, 32);
        fence_mbar_init();
    }
    // (programmatic dependent launch: the previous step's last tile rewinds the ticket counter, so wait first)
    asm volatile("griddepcontrol.wait;" ::: "memory");
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    // a ticket per CTA: its tiles are ticket * K ..., so every lower-numbered tile is already running or done
    if (threadIdx.x == 0) {
        s_cta[0] = atomicAdd(scan.ticket, 1u);
        s_cta[1] = *(volatile uint32_t *)scan.epoch;
        s_cta[2] = *(volatile uint32_t *)scan.cur_episode;
    }
    __syncthreads();
    EpisodeScanTile st;
    st.tile = s_cta[0] * (uint32_t)K + (uint32_t)k;
    st.epoch = s_cta[1];
    st.base = s_cta[2];
    if (st.tile >= num_tiles) return;   // (the last CTA may hold fewer tiles)
    trace_stamp(B, st.tile, 0);
    const uint32_t tile0 = st.tile * 32u, n = tile0 + lane;
    const bool live = n < B.N;

    uint32_t *s_game = sh.game;
    uint32_t *s_rec = s_game + DECK_WORDS * 32;
    uint32_t *s_bits = sh.bits[0], *s_tail = sh.bits[1];

    if (!is_game) {
        // ---------------- the writing warp ----------------
        if (live && !init_only) {   // the decks (the game warp loads the records into its registers meanwhile)
            for (int w = 0; w < DECK_WORDS; w++)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s_game + w * 32 + lane)),
                             "l"(B.game + (size_t)w * B.N + n)
                             : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        mbar_arrive(&sh.m_decks);
        mbar_wait(&sh.m_count, 0);      // the atomic's round trip is this warp's to wait for, it has nothing else to do
        if (lane == 0) st.publish_count(scan, __popc(sh.done_mask), num_tiles);
        mbar_wait(&sh.m_rows, 0);
        write_batch<E>(B, next, tile0, sh.play, s_bits, lut, lane, HB_STEP_SPLIT);
        trace_stamp(B, st.tile, 4);
        trace_stamp(B, st.tile, 7);
        return;
    }

    // ---------------- the game warp ----------------
    int act0 = 0, act1 = 0;
    Rec g{};
    if (live && !init_only) {   // both players' actions: which one counts is only known once the record has landed
        act0 = B.actions[n];
        act1 = B.actions[(size_t)B.N + n];
        rec_load(g, B.game + (size_t)DECK_WORDS * B.N + n, B.N);
    }
    mbar_wait(&sh.m_decks, 0);
    trace_stamp(B, st.tile, 1);
    const Deck deck{reinterpret_cast<uint8_t *>(s_game + lane)};
    int done = 0;
    float reward = 0.f;
    if (live) {
        if (init_only) {
            done = 1;
        } else {
            action_system(R, g, deck, (uint32_t)((g.turn & 1u) ? act1 : act0));
            done = check_done(R, g, reward);
        }
    }
    const unsigned done_mask = __ballot_sync(0xffffffffu, done);
    if (lane == 0) sh.done_mask = done_mask;
    mbar_arrive(&sh.m_count);
    trace_stamp(B, st.tile, 2);
    if (live) {
        B.reward[n] = reward;
        B.reward[(size_t)B.N + n] = reward;
        B.done[n] = init_only ? 0 : done;
    }

    // the new current player's rows of the worlds that play on: nine tenths of the step's output, none of it needs
    // an episode number
    {
        uint32_t job = 0, lens = 0;
        if (live && !done) {
            const uint32_t agent = g.turn & 1u;
            lens = compose_row<STD>(R, g, 0, agent, s_bits + lane * ROW_WORDS, &sh.play.legal[lane]);
            job = (uint32_t)lane | (agent << 8) | JOB_VALID;
        }
        sh.play.job[lane] = job;
        sh.play.len[lane] = lens;
        mbar_arrive(&sh.m_rows);
        trace_stamp(B, st.tile, 3);
    }

    // episode numbers: every earlier tile published its count long ago, so this is one round trip without polling
    const uint32_t episode0 = st.exclusive_base_warp(scan, __popc(done_mask), num_tiles);
    trace_stamp(B, st.tile, 5);
    if (done) reset_game(R, Rp.fresh, g, deck, episode0 + __popc(done_mask & ((1u << lane) - 1u)));
    rec_store(g, s_rec + lane, 32);   // final: for the write-back, and for the lanes that compose the reset worlds' rows
    __syncwarp();
    // The game warp is through ~10 us into the step while the writing warp has ~15 us of play rows to stream, so the
    // step's tail is the game warp's: players to move, the records back to HBM, and both players' rows of the reset
    // worlds (composed and written by this warp, 32 rows at a time; job j = (j / 2-th reset world, player j % 2) on lane j).
    if (live) {
        const uint32_t cur = g.turn & 1u;                     // 0 after a reset
        B.active[(size_t)cur * B.N + n] = 1;                   // sim.cpp:510 / 802-808
        B.active[(size_t)(1 - cur) * B.N + n] = 0;
        for (int w = 0; w < GAME_WORDS; w++) B.game[(size_t)w * B.N + n] = s_game[w * 32 + lane];
    }
    const int njobs = NP * __popc(done_mask);
    for (int first = 0; first < njobs; first += 32) {
        const int j = first + lane;
        uint32_t job = 0, lens = 0;
        if (j < njobs) {
            const int r = (int)__fns(done_mask, 0, j / NP + 1), a = j % NP;
            Rec f;
            rec_load(f, s_rec + r, 32);
            lens = compose_row<STD>(R, f, a, (uint32_t)a, s_tail + lane * ROW_WORDS, &sh.reset.legal[lane]);   // player 0 moves first: its hand is f.cur
            job = (uint32_t)r | ((uint32_t)a << 8) | JOB_VALID;
        }
        sh.reset.job[lane] = job;
        sh.reset.len[lane] = lens;
        __syncwarp();
        write_batch<E>(B, next, tile0, sh.reset, s_tail, lut, lane);
        __syncwarp();
    }
    trace_stamp(B, st.tile, 6);
    // The game warp is done ~10 us into the step and the writing warp has ~10 more to go: it takes the state rows of
    // the last slices of the play batch (its own rows, composed above; nobody else writes them).
    if (HB_STEP_SPLIT < 4) write_tensor_rows<E, 1, HB_LUT>(B.state, B.N, tile0, sh.play, s_bits, lut, lane, HB_STEP_SPLIT, 4);
}

// ---------------------------------------------------------------------------------------------------------------
// Fused random rollout: T x {draw a random legal move -> step} in ONE launch (SURVEY 8(f) row 1 for Hanabi).
//
// Same two warps per tile, but the records never leave the game warp's registers (the decks stay in shared memory) and
// the game warp runs AHEAD of the writing warp: while the writing warp streams step s to HBM, the game warp has already
// fetched the episode numbers, reset the finished games, composed their rows and played step s+1; only composing the
// next rows waits for the row buffer.
//
// ONE row per world and step: the game warp plays, publishes its reset count, learns the episode numbers, resets the
// finished games and only then composes — the player to move in the worlds that play on, player 0 in the fresh games —
// so a step is one compose, one hand-over and one pass of the writing warp (flags, rewards, 32 rows).  Player
// 1's row of a fresh game is always rewritten by a later step before anything can observe it (after player 0's first
// move player 1 is the player to move, whose row every step rewrites; a game cannot end within its first moves), so it
// is composed for the resets of the LAST step only (iteration T of the game warp's loop, through the same code).
//
// Episode numbers couple the tiles once per step: number = launch base + resets of all tiles in earlier steps +
// resets of lower tiles in this step.  Per step a tile publishes its count (status word + group accumulator, as in
// EpisodeScan, tagged with the global step index, in a ring of 4 step slots) and, before it resets, reads the
// previous step's group sums (=> every tile has played step s-1: tiles never drift more than one step apart, which is
// what makes a ring of 4 slots safe) and this step's lower tiles / groups.  All tiles must be resident at once
// (the host checks occupancy and otherwise launches step by step).
// ---------------------------------------------------------------------------------------------------------------
// (RolloutScan, roll_slot and the ring layout live in common.cuh: the simple environments' fused rollouts share them)

// Row buffers of the fused rollout, a ring of ROLL_BUFS per tile: the game warp composes up to ROLL_BUFS steps ahead of
// the writing warp.  Each buffer has a "full" and an "empty" mbarrier in shared memory (32 arrivals: every lane of the
// producing / consuming warp), the classic pipeline.
#ifndef HB_ROLL_BUFS
#define HB_ROLL_BUFS 3
#endif
constexpr int ROLL_BUFS = HB_ROLL_BUFS;
// ONE CTA PER SM holds up to ROLL_TILES_MAX tiles: warps [0, K) are the K tiles' game warps, warps [K, 2K) their writing
// warps.  With one two-warp CTA per tile (14 CTAs per SM) the SM's schedulers, which serve older warps first, let the
// writing warps of the older CTAs — always busy, ~3 000 issues per step — starve the GAME warps of the youngest CTAs
// (4x slower: 6 us to play a move instead of 1.5), and since the ordered episode numbering makes every tile wait for
// the slowest tile's move every step, the whole GPU ran at the pace of the most starved game warp of some SM.  Here
// the game warps are the CTA's first warps: the latency-critical work gets the issue slots first, the writing warps
// (throughput work) what is left; and warp w issues on scheduler w % 4, so each scheduler holds both kinds.
constexpr int ROLL_TILES_MAX = 16;   // (1 024 threads: 64 registers each, a dozen words of the game warp spill)
// Not every SM stores equally fast (sm_rates_host.h: 32 of a B200's 148 SMs are ~18 % slower), and the episode
// numbering makes every tile wait for the slowest SM's tiles every step.  So tiles are dealt by SM speed: every SM
// gets a quota of tiles proportional to its measured store rate; a rollout CTA looks its quota up under its %smid and claims that many tiles with
// an atomic; after a grid-wide barrier whatever the quotas left over is claimed by the CTAs with warps to spare, so
// every tile finds a home whatever the placement of the CTAs (capacity = grid x 16 tiles >= tiles, host-checked).
struct RollBalance {
    const uint32_t *quota;   // [256] tiles for the CTA on SM %smid; nullptr = blockIdx.x * K + k (equal shares)
    uint32_t *ctl;           // [0] barrier arrivals, [1] barrier generation, [2 + parity] next tile to claim
};
struct RollStep {   // what the game warp hands over with the rows of one step
    uint32_t done_mask;
    uint32_t pending;     // worlds whose game ended in this step and is dealt afresh a step later: no row, player 0 to move
    float reward[32];
    RowBatch play;
};
struct alignas(16) RollTile {   // shared memory of one tile
    uint32_t deck[DECK_WORDS * 32];                // [word][lane]
    uint32_t bits[ROLL_BUFS][32 * ROW_WORDS];      // bit rows
    RollStep step[ROLL_BUFS];
    uint64_t full[ROLL_BUFS], empty[ROLL_BUFS];    // mbarriers: rows handed over / buffer free
};
static_assert(sizeof(RollTile) % 16 == 0, "tiles stay 16-byte aligned");

// debug timeline of the fused rollout: 16 %globaltimer stamps per (step, tile) for the first `steps` steps.  Game warp:
// 0 step begins, 1 played + count published, 4 earlier steps' resets known, 5 lower tiles' resets known + games reset,
// 2 row buffer free, 3 rows composed.  Writing warp: 8 rows handed over, 11 rows written.
struct RolloutTrace { unsigned long long *buf; int steps; };
__device__ __forceinline__ void roll_stamp(const RolloutTrace &tr, uint32_t num_tiles, uint32_t tile, int s, int slot)
{
    if (tr.buf && s < tr.steps && (threadIdx.x & 31) == 0) {
        // (a warp may run ahead of a barrier wait to its next memory instruction: a shared-memory load the timer value
        // depends on pins the stamp behind it)
        unsigned long long t;
        uint32_t pin;
        asm volatile("ld.volatile.shared.u32 %0, [%1];" : "=r"(pin) : "r"(0u) : "memory");   // any valid shared address
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t) : "r"(pin));
        tr.buf[((size_t)s * num_tiles + tile) * 16 + slot] = t;
    }
}

template <bool STD, class E>
__global__ void __launch_bounds__(CTA_THREADS * ROLL_TILES_MAX, 1) rollout_kernel(const Rules Rp, const Buffers B, const RolloutScan S,
                                                                                 const uint32_t seed, const uint32_t first_step,
                                                                                 const uint32_t num_tiles, const int T,
                                                                                 const RolloutTrace tr, const RollBalance bal)
{
    GameRules R = Rp;   // (the fresh deck, Rp.fresh, stays in the parameter bank)
    if (STD) { R.colors = 5; R.ranks = 5; R.max_info = 8; R.max_life = 3; }
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ uint32_t s_claim[4];   // first tile / tiles of the quota claim, of the left-over claim
    const int K = (int)(blockDim.x >> 6);   // tile slots of this CTA
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const bool is_game = warp < K;
    const int k = is_game ? warp : warp - K;
    uint32_t tile = blockIdx.x * (uint32_t)K + (uint32_t)k;
    if (bal.quota) {
        if (threadIdx.x == 0) {
            uint32_t smid;
            asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
            const uint32_t gen = *(volatile uint32_t *)(bal.ctl + 1);
            uint32_t *next = bal.ctl + 2 + (gen & 1u);
            const uint32_t want = min(bal.quota[smid & 255u], (uint32_t)K);
            const uint32_t f1 = atomicAdd(next, want);
            const uint32_t n1 = f1 < num_tiles ? min(want, num_tiles - f1) : 0u;
            // grid-wide barrier (all CTAs are resident: cooperative launch): every quota claim is in
            __threadfence();
            if (atomicAdd(bal.ctl, 1u) == gridDim.x - 1u) {
                *(volatile uint32_t *)(bal.ctl + 2 + ((gen + 1u) & 1u)) = 0u;   // the next launch's counter
                *(volatile uint32_t *)bal.ctl = 0u;
                __threadfence();
                atomicAdd(bal.ctl + 1, 1u);
            } else {
                while (*(volatile uint32_t *)(bal.ctl + 1) == gen) __nanosleep(200);
            }
            const uint32_t spare = (uint32_t)K - n1;
            const uint32_t f2 = spare ? atomicAdd(next, spare) : num_tiles;
            s_claim[0] = f1; s_claim[1] = n1;
            s_claim[2] = f2; s_claim[3] = f2 < num_tiles ? min(spare, num_tiles - f2) : 0u;
        }
        __syncthreads();
        const uint32_t kk = (uint32_t)k;
        tile = kk < s_claim[1] ? s_claim[0] + kk : (kk - s_claim[1] < s_claim[3] ? s_claim[2] + (kk - s_claim[1]) : num_tiles);
    }
    const uint32_t tile0 = tile * 32u, n = tile0 + lane;
    const bool live = tile < num_tiles && n < B.N;
    RollTile &sh = reinterpret_cast<RollTile *>(smem)[k];
    uint32_t *s_deck = sh.deck;

    if (is_game && lane == 0) {
        for (int i = 0; i < ROLL_BUFS; i++) { mbar_init(&sh.full[i], 32); mbar_init(&sh.empty[i], 32); }
        fence_mbar_init();
    }
    int act0 = 0, act1 = 0;
    Rec g{};
    if (live) {
        for (int w = is_game ? 0 : 1; w < DECK_WORDS; w += 2)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s_deck + w * 32 + lane)),
                         "l"(B.game + (size_t)w * B.N + n)
                         : "memory");
        if (is_game) {
            act0 = B.actions[n];
            act1 = B.actions[(size_t)B.N + n];
            rec_load(g, B.game + (size_t)DECK_WORDS * B.N + n, B.N);
        }
    }
    uint32_t base = *(volatile uint32_t *)S.cur_episode;
    const uint32_t step_base = *(volatile uint32_t *)S.step_base;
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncthreads();
    if (tile >= num_tiles) return;   // (the last CTA may hold fewer tiles)

    if (!is_game) {
        // ---------------- the writing warp: ONE pass per step ----------------
        int b = 0;
        uint32_t parity = 0;
        for (int s = 0; s <= T; s++) {
            const RollStep &st = sh.step[b];
            mbar_wait(&sh.full[b], parity);
            roll_stamp(tr, num_tiles, tile, s, 8);
            if (s < T && live) {
                const float reward = st.reward[lane];
                B.reward[n] = reward;
                B.reward[(size_t)B.N + n] = reward;
                B.done[n] = (st.done_mask >> lane) & 1u;
                // player to move after the step = the one whose row was composed; player 0 in a fresh game
                const uint32_t cur = ((st.pending >> lane) & 1u) ? 0u : (st.play.job[lane] >> 8) & 0xFFu;   // sim.cpp:510 / 802-808
                B.active[(size_t)cur * B.N + n] = 1;
                B.active[(size_t)(1 - cur) * B.N + n] = 0;
            }
            // s == T: player 1's view of the games that were reset in the last step (if any)
            if (s < T || st.done_mask) {
                write_tensor_rows_shared<E, 0>(B.obs, B.N, tile0, &st.play, sh.bits[b], lane);
                write_tensor_rows_shared<E, 1>(B.state, B.N, tile0, &st.play, sh.bits[b], lane);
                write_masks(B.mask, B.N, tile0, st.play.job, st.play.legal, lane);
            }
            if (s < T) mbar_arrive(&sh.empty[b]);
            roll_stamp(tr, num_tiles, tile, s, 11);
            if (++b == ROLL_BUFS) { b = 0; parity ^= 1u; }
        }
        if (live)   // the game warp is done with the decks: back to HBM, once per rollout
            for (int w = 0; w < DECK_WORDS; w++) B.game[(size_t)w * B.N + n] = s_deck[w * 32 + lane];
        return;
    }

    // ---------------- the game warp ----------------
    const Deck deck{reinterpret_cast<uint8_t *>(s_deck + lane)};
    const uint32_t group = tile / SCAN_GROUP;
    const uint32_t group_size = min(SCAN_GROUP, num_tiles - group * SCAN_GROUP);
    int act = 0;
    if (live) act = (g.turn & 1u) ? act1 : act0;
    unsigned done_mask = 0;
    unsigned pending = 0;    // worlds whose game ended in the previous step and whose fresh game has not been dealt yet
    int counted = 0;         // `base` holds the launch base + the resets of all tiles in steps [0, counted)
    int b = 0;
    uint32_t parity = 1;   // of the "empty" phase to wait for: a fresh barrier counts as having completed phase 1
    // resets of all tiles in the steps before `upto` into `base` (waits until every tile has played them)
    auto count_steps = [&](int upto) {
        for (; counted < upto; counted++) base += roll_step_total(S, num_tiles, step_base + (uint32_t)counted, lane);
    };
    // the games that ended in step r get their episode numbers (base + resets of the lower tiles in step r + rank in
    // the tile) and are dealt afresh
    auto deal_fresh = [&](int r, unsigned mask) {
        count_steps(r);
        const uint32_t lower = roll_lower_tiles(S, num_tiles, tile, step_base + (uint32_t)r, lane);
        if ((mask >> lane) & 1u) reset_game(R, Rp.fresh, g, deck, base + lower + __popc(mask & ((1u << lane) - 1u)));
        __syncwarp();
    };
    for (int s = 0; s <= T; s++) {
        float reward = 0.f;
        if (s < T) {
            roll_stamp(tr, num_tiles, tile, s, 0);
            // ---- the games that ended in the previous step: dealt now, a step late (DEFERRED, see below), and their
            // first move drawn from the fresh game's legal moves ----
            if (s >= 2) count_steps(s - 1);   // (every step: keeps the tiles within two steps of each other, the ring's bound)
            roll_stamp(tr, num_tiles, tile, s, 4);
            if (pending) {
                deal_fresh(s - 1, pending);
                if ((pending >> lane) & 1u) {
                    const uint32_t legal = generate_mask(R, g, g.cur, g.oth), cnt = __popc(legal);
                    const uint32_t r = policy::mix(seed, first_step + (uint32_t)s, (uint32_t)n);   // player 0's row index = n
                    act = cnt ? (int)__fns(legal, 0, (int)(((uint64_t)r * cnt) >> 32) + 1) : 0;
                }
                pending = 0;
            }
            roll_stamp(tr, num_tiles, tile, s, 5);
            // ---- play ----
            int done = 0;
            if (live) {
                action_system(R, g, deck, (uint32_t)act);
                done = check_done(R, g, reward);
            }
            done_mask = __ballot_sync(0xffffffffu, done);
            const uint32_t count = __popc(done_mask), gstep = step_base + (uint32_t)s, tag = (gstep + 1u) & 0xFFFFu;
            const RollSlot now = roll_slot(S, num_tiles, gstep);
            if (lane == 0) {   // publish this step's reset count
                *(volatile uint32_t *)(now.status + tile) = (tag << 16) | count;
                const uint32_t old = atomicAdd(now.acc + group, (1u << 16) | count);
                if ((old >> 16) + 1u == group_size) {   // this tile completes its group
                    // the accumulator of three steps ahead (last used a step ago, complete since) is zeroed BEFORE
                    // anybody can learn from the group sum that this step is complete
                    *(volatile uint32_t *)(roll_slot(S, num_tiles, gstep + 3u).acc + group) = 0u;
                    __threadfence();
                    *(volatile uint32_t *)(now.gsum + group) = (tag << 16) | ((old & 0xFFFFu) + count);
                }
            }
            roll_stamp(tr, num_tiles, tile, s, 1);
            // DEFERRED RESETS.  A game that ends in step s needs an episode number that depends on every lower tile's
            // step s: dealt at once, each step of every tile waits for the slowest tile of the GPU.  But a fresh game
            // cannot end within its first three moves (three lives, 40 cards in the deck), so the tile's reset counts
            // of the next steps do not depend on the deal, and player 0's first row of a fresh game is rewritten two
            // steps later before anything can observe it — like player 1's (above).  So the deal waits until the
            // start of the tile's NEXT step, when the lower tiles have long played: no row for the ended games in
            // this step, and tiles may drift up to two steps apart.  The last two steps of a launch deal at once and
            // write the fresh games' rows: what a launch leaves behind is exactly what one launch per step leaves.
            if (s + 2 >= T) {
                if (done_mask) deal_fresh(s, done_mask);
            } else {
                pending = done_mask;
            }
        }
        // ---- ONE row per world: the player to move in the worlds that play on (player 0 in the fresh games of the last
        // steps); in iteration T player 1's view of the games that were reset in the last step ----
        Rec v = g;
        int swap = 0;
        uint32_t world = (uint32_t)lane;
        bool valid = live && !((pending >> lane) & 1u);
        if (s == T) {
            const int cnt = __popc(done_mask);
            valid = lane < cnt;
            world = valid ? __fns(done_mask, 0, lane + 1) : 0u;
            v = rec_shfl(g, (int)world);
            swap = 1;
        }
        RollStep &st = sh.step[b];
        mbar_wait(&sh.empty[b], parity);   // the writing warp has read this buffer's last rows
        roll_stamp(tr, num_tiles, tile, s, 2);
        uint32_t job = 0, lens = 0;
        if (valid) {
            const uint32_t agent = swap ? 1u : (v.turn & 1u);   // 0 after a reset
            lens = compose_row<STD>(R, v, swap, agent, sh.bits[b] + lane * ROW_WORDS, &st.play.legal[lane]);
            job = world | (agent << 8) | JOB_VALID;
        }
        st.play.job[lane] = job;
        st.play.len[lane] = lens;
        st.reward[lane] = reward;
        if (lane == 0) { st.done_mask = done_mask; st.pending = pending; }
        mbar_arrive(&sh.full[b]);
        roll_stamp(tr, num_tiles, tile, s, 3);
        if (s < T && job) {   // next step's move: uniform over the legal ones (policy::sample_kernel's draw)
            const uint32_t legal = st.play.legal[lane], cnt = __popc(legal);
            const uint32_t r = policy::mix(seed, first_step + (uint32_t)s + 1u, (uint32_t)job_row(job, B.N, tile0));
            act = cnt ? (int)__fns(legal, 0, (int)(((uint64_t)r * cnt) >> 32) + 1) : 0;
        }
        if (++b == ROLL_BUFS) { b = 0; parity ^= 1u; }
    }
    if (live) {
        B.actions_out[(size_t)(g.turn & 1u) * B.N + n] = act;   // the move step T would play
        rec_store(g, B.game + (size_t)DECK_WORDS * B.N + n, B.N);   // the records: back to HBM, once per rollout
    }
    if (tile == 0) {   // counters for the next launch: every tile read them before it played its first step
        count_steps(T);
        if (lane == 0) {
            *(volatile uint32_t *)S.cur_episode = base;
            *(volatile uint32_t *)S.step_base = step_base + (uint32_t)T;
        }
    }
}

// rules + the deck of a freshly reset game
static Rules make_rules(const LaunchArgs &a)
{
    Rules R{};
    R.colors = a.colors; R.ranks = a.ranks; R.max_info = a.max_info; R.max_life = a.max_life;
    uint8_t f[DECK_WORDS * 4];
    fresh_deck(R, f);
    memcpy(R.fresh, f, sizeof(f));
    return R;
}

// one instantiation per (standard game, element type of the observation / state rows)
template <bool STD, class E>
static cudaError_t launch_step_se(cudaLaunchConfig_t &cfg, const Rules &R, const Buffers &B, const EpisodeScan &scan,
                                  const NextAction &next, uint32_t tiles, int init_only)
{
    if (cudaError_t e = ensure_dynamic_smem(step_kernel<STD, E>, (int)(TILES_MAX * sizeof(StepTile)))) return e;
    return cudaLaunchKernelEx(&cfg, step_kernel<STD, E>, R, B, scan, next, tiles, init_only);
}
template <bool STD, class E>
static cudaError_t launch_rollout_se(cudaLaunchConfig_t &cfg, const Rules &R, const Buffers &B, const RolloutScan &S,
                                     uint32_t seed, uint32_t first_step, uint32_t tiles, int num_steps, const RolloutTrace &tr,
                                     const RollBalance &bal)
{
    if (cudaError_t e = ensure_dynamic_smem(rollout_kernel<STD, E>, (int)(ROLL_TILES_MAX * sizeof(RollTile)))) return e;
    return cudaLaunchKernelEx(&cfg, rollout_kernel<STD, E>, R, B, S, seed, first_step, tiles, num_steps, tr, bal);
}
#define HANABI_DISPATCH(fn, std_game, elem, ...)                                                          \
    ((elem) == 1 ? ((std_game) ? fn<true, uint8_t>(__VA_ARGS__) : fn<false, uint8_t>(__VA_ARGS__))        \
     : (elem) == 2 ? ((std_game) ? fn<true, uint16_t>(__VA_ARGS__) : fn<false, uint16_t>(__VA_ARGS__))    \
     : (elem) == 4 ? ((std_game) ? fn<true, float>(__VA_ARGS__) : fn<false, float>(__VA_ARGS__))          \
                   : cudaErrorInvalidValue)

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    const bool std_game = a.colors == 5 && a.ranks == 5 && a.max_info == 8 && a.max_life == 3;
    const Rules R = make_rules(a);
    Buffers B{a.game, a.actions, a.obs, a.state, a.mask, a.active, a.done, a.reward, a.trace,
              const_cast<int32_t *>(a.actions), a.N};
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    NextAction next{a.sample_next ? const_cast<int32_t *>(a.actions) : nullptr, a.sample_seed, a.sample_step};
    const uint32_t tiles = (a.N + 31) / 32;
    // tiles per CTA: the fewest that let one CTA per SM hold them all, at most TILES_MAX (more tiles than that: several
    // waves of CTAs, which the ticket order of the episode scan allows).  Programmatic dependent launch: the launch
    // latency overlaps the previous step's tail.
    const uint32_t sms = (uint32_t)device_info().sms;
    uint32_t per_cta = (tiles + sms - 1) / sms;
    if (per_cta > (uint32_t)TILES_MAX) per_cta = TILES_MAX;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((tiles + per_cta - 1) / per_cta);
    cfg.blockDim = dim3(CTA_THREADS * per_cta);
    cfg.dynamicSmemBytes = per_cta * sizeof(StepTile);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return HANABI_DISPATCH(launch_step_se, std_game, a.obs_elem, cfg, R, B, scan, next, tiles, init_only);
}

size_t rollout_ring_words(uint32_t N) { return roll_ring_words(num_tiles(N)); }

// Fills a.roll_quota: tiles per SM in proportion to its measured store rate (sm_rates_host.h), summing to the tiles of
// the batch, at most ROLL_TILES_MAX each.  Leaves equal shares by CTA index when the batch is small or nothing was measured.
cudaError_t prepare_rollout_balance(LaunchArgs &a, void *scratch, size_t scratch_bytes)
{
    a.roll_balanced = 0;
    if (!a.roll_quota || !a.roll_ctl || getenv("B200ENV_HANABI_NO_BALANCE")) return cudaSuccess;
    const uint32_t tiles = num_tiles(a.N);
    const int sms = device_info().sms;
    if (tiles < 8u * (uint32_t)sms || tiles > (uint32_t)(ROLL_TILES_MAX * sms)) return cudaSuccess;
    cudaError_t e;
    const float *rates = sm_store_rates(scratch, scratch_bytes, &e);
    if (e != cudaSuccess) return e;
    const char *pe = getenv("B200ENV_HANABI_BALANCE_EXP");   // weights = rate ^ p (default 1: measured best of 0.7 / 1 / 1.5 / 2)
    uint32_t quota[256];
    if (!sm_quotas(rates, tiles, ROLL_TILES_MAX, pe ? atof(pe) : 1.0, quota)) return cudaSuccess;
    e = cudaMemcpy(a.roll_quota, quota, sizeof(quota), cudaMemcpyHostToDevice);
    a.roll_balanced = e == cudaSuccess;
    return e;
}

// Cooperative launch: it fails (cudaErrorCooperativeLaunchTooLarge) instead of deadlocking when the tiles cannot all
// be resident; the caller then falls back to one launch per step.
cudaError_t launch_rollout(const LaunchArgs &a, int num_steps, uint32_t seed, uint32_t first_step, cudaStream_t stream)
{
    if (num_steps <= 0) return cudaSuccess;
    const bool std_game = a.colors == 5 && a.ranks == 5 && a.max_info == 8 && a.max_life == 3;
    Rules R = make_rules(a);
    Buffers B{a.game, a.actions, a.obs, a.state, a.mask, a.active, a.done, a.reward, nullptr,
              const_cast<int32_t *>(a.actions), a.N};
    RolloutScan S{a.roll_ring, a.roll_step, a.cur_episode};
    const uint32_t tiles = num_tiles(a.N);
    // tiles per CTA: the fewest that let one CTA per SM hold them all (at most ROLL_TILES_MAX: beyond that the tiles
    // cannot all be resident and the caller launches step by step)
    const uint32_t sms = (uint32_t)device_info().sms;
    uint32_t per_cta = (tiles + sms - 1) / sms;
    if (per_cta > (uint32_t)ROLL_TILES_MAX) return cudaErrorCooperativeLaunchTooLarge;
    RollBalance bal{nullptr, a.roll_ctl};
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((tiles + per_cta - 1) / per_cta);
    if (a.roll_balanced && tiles >= 8 * sms && !getenv("B200ENV_HANABI_NO_BALANCE")) {   // tiles dealt by SM speed: one full-size CTA on every SM
        bal.quota = a.roll_quota;
        per_cta = ROLL_TILES_MAX;
        cfg.gridDim = dim3(sms);
    }
    cfg.blockDim = dim3(CTA_THREADS * per_cta);
    cfg.dynamicSmemBytes = per_cta * sizeof(RollTile);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeCooperative;
    attr[0].val.cooperative = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const RolloutTrace tr{a.trace, a.trace_steps};
    return HANABI_DISPATCH(launch_rollout_se, std_game, a.obs_elem, cfg, R, B, S, seed, first_step, tiles, num_steps, tr, bal);
}

}   // namespace hanabi
}   // namespace b200
