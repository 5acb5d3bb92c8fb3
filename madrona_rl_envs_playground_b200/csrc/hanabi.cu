// 2-player Hanabi step kernel (sm_100a): action apply, deck draw, obs / state / legal-move-mask encoding, score, done,
// in-step reset.  Behaviour follows src/hanabi_env/sim.cpp:45-891 of the reference including its quirks (cited inline);
// the structure does not: one warp owns a tile of 32 consecutive worlds, lane = world, the packed game state of the
// tile sits in shared memory as [word][lane] (conflict free), and the freshly encoded rows are staged in shared
// memory congruent (mod 16) to their destination and drained by the copy engine (cp.async.bulk).
//
// HBM layout:
//   game   u32 [GAME_WORDS][N]    packed per-world game state, word-major so a warp's 32 worlds coalesce
//   obs    u8  [2][N][658]   state u8 [2][N][783]   mask i32 [2][N][20]   active i32 [2][N]
//   reward f32 [2][N]        done  i32 [N]          actions i32 [2][N]
// Only the NEW current player's obs/state/mask rows are rewritten each step (sim.cpp:794-810); the other player's
// rows keep their previous content, exactly like the reference's exported columns.  On a reset both are rewritten.
#include "common.cuh"
#include "hanabi_host.h"

namespace b200 {
namespace hanabi {

constexpr int HAND = 5, NP = 2, MOVES = 20;
constexpr int OBS = 658, STATE = 783;
enum : int { kDiscard = 0, kPlay, kRevealColor, kRevealRank, kInvalid };   // sim.hpp:83

// byte offsets inside the packed game record
constexpr int OFF_DECK = 0;        // 50 cards
constexpr int OFF_DISCARD = 52;    // 25 counts
constexpr int OFF_FIREWORKS = 80;  // 5
constexpr int OFF_DECKSIZE = 85, OFF_INFO = 86, OFF_LIFE = 87, OFF_CUR = 88, OFF_TURNS = 89, OFF_SCORE = 90;
constexpr int OFF_LM_MOVE = 91, OFF_LM_PLAYER = 92, OFF_LM_TARGET = 93, OFF_LM_CARD = 94, OFF_LM_SCORED = 95,
              OFF_LM_INFOTOK = 96, OFF_LM_COLOR = 97, OFF_LM_RANK = 98, OFF_LM_REVEAL = 99;
constexpr int OFF_HAND = 100;      // per player 16 bytes: cards[5], known_color[5], known_rank[5], size
constexpr int HAND_CARDS = 0, HAND_KCOLOR = 5, HAND_KRANK = 10, HAND_SIZE = 15;
constexpr int OFF_PLAUS = 132;     // u32 [2][5] plausible-card masks (25 bits used)
constexpr int OFF_RNG = 172;       // u32
static_assert(GAME_WORDS * 4 >= OFF_RNG + 4, "game record too small");

struct Rules {
    int colors, ranks, max_info, max_life;
};

// one world's packed record inside the [word][lane] shared tile
struct Game {
    uint8_t *base;   // shared tile + lane*4
    __device__ __forceinline__ uint8_t &u8(int off) const { return base[(off >> 2) * 128 + (off & 3)]; }
    __device__ __forceinline__ int i8(int off) const { return (int)(int8_t)u8(off); }
    __device__ __forceinline__ uint32_t &u32(int off) const { return *reinterpret_cast<uint32_t *>(base + (off >> 2) * 128); }
    __device__ __forceinline__ int hand(int p, int field) const { return OFF_HAND + 16 * p + field; }
};

__device__ __forceinline__ int draw_deck(const Game &g)   // sim.cpp:45-52
{
    uint32_t rng = g.u32(OFF_RNG);
    const int size = g.u8(OFF_DECKSIZE);
    const int swaploc = (int)((float)size * rng_rand(rng));
    g.u32(OFF_RNG) = rng;
    const int ret = g.u8(OFF_DECK + swaploc);
    g.u8(OFF_DECK + swaploc) = g.u8(OFF_DECK + size - 1);
    g.u8(OFF_DECKSIZE) = (uint8_t)(size - 1);
    return ret;
}

// Row sink: bytes [0,658) go to the obs row and the state row, [658,783) to the state row only.  Offsets beyond
// the row (information_tokens above its maximum shifts the encoding, see generate_rows) spill into the next
// agent's row exactly where the reference's out-of-row write lands, or are dropped for the last agent.
struct Sink {
    uint8_t *obs, *state;             // staging rows of the agent being encoded
    uint8_t *obs_spill, *state_spill; // byte 0.. of the next agent's GLOBAL rows (nullptr for the last agent)
    __device__ __forceinline__ void both(int off, int v) const
    {
        if (off < OBS) obs[off] = (uint8_t)v;
        else if (obs_spill) obs_spill[off - OBS] = (uint8_t)v;
        st(off, v);
    }
    __device__ __forceinline__ void st(int off, int v) const
    {
        if (off < STATE) state[off] = (uint8_t)v;
        else if (state_spill) state_spill[off - STATE] = (uint8_t)v;
    }
};

// generateObsState, sim.cpp:367-379 (encodeHands 54-90, encodeBoard 92-135, encodeDiscards 137-156,
// encodeLastAction 158-289, encodeCardKnowledge 291-331, copyObsToState 333-341, encodeOwnHand 343-365)
__device__ void generate_rows(const Rules &R, const Game &g, int agent, const Sink &s)
{
    const int bits = R.colors * R.ranks;
    int off = 0;
    // hands: partner's cards one-hot, then "hand is short" flags for self-relative order
    {
        const int partner = (agent + 1) % NP;
        const int psize = g.u8(g.hand(partner, HAND_SIZE));
        for (int c = 0; c < HAND; c++) {
            const int card = c < psize ? (int)g.u8(g.hand(partner, HAND_CARDS) + c) : -1;
            for (int b = 0; b < bits; b++) s.both(off++, b == card);
        }
        for (int i = 0; i < NP; i++) s.both(off++, g.u8(g.hand((agent + i) % NP, HAND_SIZE)) < HAND);
    }
    // board
    {
        const int size = g.u8(OFF_DECKSIZE);
        const int max_size = (4 + (R.ranks - 2) * 2) * R.colors - HAND * NP;
        for (int i = 0; i < size; i++) s.both(off++, 1);
        for (int i = size; i < max_size; i++) s.both(off++, 0);
        for (int c = 0; c < R.colors; c++)
            for (int i = 0; i < R.ranks; i++) s.both(off++, i + 1 == (int)g.u8(OFF_FIREWORKS + c));
        const int info = g.u8(OFF_INFO), life = g.u8(OFF_LIFE);
        for (int i = 0; i < info; i++) s.both(off++, 1);               // may exceed max_info: no cap at sim.cpp:675-677
        for (int i = info; i < R.max_info; i++) s.both(off++, 0);
        for (int i = 0; i < life; i++) s.both(off++, 1);
        for (int i = life; i < R.max_life; i++) s.both(off++, 0);
    }
    // discards
    {
        int id = 0;
        for (int c = 0; c < R.colors; c++)
            for (int r = 0; r < R.ranks; r++) {
                const int cr = (r == 0 ? 3 : r == R.ranks - 1 ? 1 : 2);
                const int cnt = g.u8(OFF_DISCARD + id);
                for (int i = 0; i < cr; i++) s.both(off++, cnt > i);
                id++;
            }
    }
    // last action.  QUIRK: reveal moves never update lastmove.move (sim.cpp:695-788 set move.type only), so `move` is
    // the last discard / play / invalid; the reveal-only sections are therefore encoded from that stale value too.
    {
        const int move = g.u8(OFF_LM_MOVE);
        const int lm_player = g.i8(OFF_LM_PLAYER), lm_target = g.i8(OFF_LM_TARGET);
        const int lm_color = g.i8(OFF_LM_COLOR), lm_rank = g.i8(OFF_LM_RANK);
        const bool reveal = (move == kRevealColor || move == kRevealRank);
        const bool playdisc = (move == kPlay || move == kDiscard);
        const int rel = lm_player == -1 ? -1 : (agent - lm_player + NP) % NP;
        for (int i = 0; i < NP; i++) s.both(off++, i == rel);
        s.both(off + 0, move == kPlay);
        s.both(off + 1, move == kDiscard);
        s.both(off + 2, move == kRevealColor);
        s.both(off + 3, move == kRevealRank);
        off += 4;
        const int rel_target = (int)(int8_t)((agent - lm_target + NP) % NP);
        for (int i = 0; i < NP; i++) s.both(off++, reveal && i == rel_target);
        for (int i = 0; i < R.colors; i++) s.both(off++, move == kRevealColor && i == (int)(uint8_t)lm_color);
        for (int i = 0; i < R.ranks; i++) s.both(off++, move == kRevealRank && i == (int)(uint8_t)lm_rank);
        const int mask = g.u8(OFF_LM_REVEAL);
        for (int i = 0; i < HAND; i++) s.both(off++, reveal && ((mask >> i) & 1));
        const int lm_card = g.i8(OFF_LM_CARD);
        for (int i = 0; i < HAND; i++) s.both(off++, playdisc && i == lm_card);
        // uint32 arithmetic like the reference (int8 colour * uint32 ranks + int8 rank), sim.cpp:267
        const uint32_t played = (uint32_t)lm_color * (uint32_t)R.ranks + (uint32_t)lm_rank;
        for (uint32_t i = 0; i < (uint32_t)bits; i++) s.both(off++, playdisc && i == played);
        s.both(off++, move == kPlay && g.u8(OFF_LM_SCORED));
        s.both(off++, move == kPlay && g.u8(OFF_LM_INFOTOK));
    }
    // card knowledge.  QUIRK sim.cpp:311: the plausibility bit tested is the PLAYER offset i, for all `bits` slots.
    for (int i = 0; i < NP; i++) {
        const int pl = (agent + i) % NP;
        const int psize = g.u8(g.hand(pl, HAND_SIZE));
        for (int c = 0; c < HAND; c++) {
            if (c < psize) {
                const int plaus = (g.u32(OFF_PLAUS + 4 * (pl * HAND + c)) >> i) & 1;
                for (int v = 0; v < bits; v++) s.both(off++, plaus);
                const int kc = (int)(int8_t)g.u8(g.hand(pl, HAND_KCOLOR) + c);
                const int kr = (int)(int8_t)g.u8(g.hand(pl, HAND_KRANK) + c);
                for (int v = 0; v < R.colors; v++) s.both(off++, kc == v);
                for (int v = 0; v < R.ranks; v++) s.both(off++, kr == v);
            } else {
                for (int v = 0; v < bits + R.colors + R.ranks; v++) s.both(off++, 0);
            }
        }
    }
    // own hand (state only)
    {
        const int size = g.u8(g.hand(agent, HAND_SIZE));
        for (int c = 0; c < HAND; c++) {
            const int card = c < size ? (int)g.u8(g.hand(agent, HAND_CARDS) + c) : -1;
            for (int b = 0; b < bits; b++) s.st(off++, b == card);
        }
    }
}

// generateActionMask, sim.cpp:381-444; written straight to the agent's global mask row (80 B, 16-byte aligned)
__device__ void generate_mask(const Rules &R, const Game &g, int agent, int32_t *mask_row)
{
    int m[MOVES];
    const int size = g.u8(g.hand(agent, HAND_SIZE));
    const int info = g.u8(OFF_INFO);
    const int partner = (agent + 1) % NP;
    int off = 0;
    for (int i = 0; i < HAND; i++) m[off++] = (i < size && info < R.max_info);
    for (int i = 0; i < HAND; i++) m[off++] = (i < size);
    for (int c = 0; c < R.colors; c++) {
        int has = 0;   // QUIRK sim.cpp:416-418: all 5 slots are scanned, stale ones after an empty-deck shift included
        for (int n = 0; n < HAND; n++) has |= ((int)g.u8(g.hand(partner, HAND_CARDS) + n) / R.ranks == c);
        m[off++] = (info > 0 && has);
    }
    for (int r = 0; r < R.ranks; r++) {
        int has = 0;
        for (int n = 0; n < HAND; n++) has |= ((int)g.u8(g.hand(partner, HAND_CARDS) + n) % R.ranks == r);
        m[off++] = (info > 0 && has);
    }
    for (; off < MOVES; off++) m[off] = 0;
    int4 *dst = reinterpret_cast<int4 *>(mask_row);
#pragma unroll
    for (int i = 0; i < MOVES / 4; i++) dst[i] = make_int4(m[4 * i], m[4 * i + 1], m[4 * i + 2], m[4 * i + 3]);
}

__device__ __forceinline__ void remove_from_hand(const Rules &R, const Game &g, int p, int index)   // sim.cpp:567-594
{
    const int hc = g.hand(p, HAND_CARDS), hk = g.hand(p, HAND_KCOLOR), hr = g.hand(p, HAND_KRANK);
    const int size = g.u8(g.hand(p, HAND_SIZE));
    if (g.u8(OFF_DECKSIZE) == 0) {
        for (int i = index + 1; i < size; i++) {   // shift left; the vacated last slot keeps its stale card
            g.u8(hc + i - 1) = g.u8(hc + i);
            g.u8(hk + i - 1) = g.u8(hk + i);
            g.u8(hr + i - 1) = g.u8(hr + i);
            g.u32(OFF_PLAUS + 4 * (p * HAND + i - 1)) = g.u32(OFF_PLAUS + 4 * (p * HAND + i));
        }
        g.u8(g.hand(p, HAND_SIZE)) = (uint8_t)(size - 1);
    } else {
        g.u8(hc + index) = (uint8_t)draw_deck(g);
        g.u8(hk + index) = 0xFF;
        g.u8(hr + index) = 0xFF;
        g.u32(OFF_PLAUS + 4 * (p * HAND + index)) = (1u << (R.colors * R.ranks)) - 1u;
    }
}

__device__ void action_system(const Rules &R, const Game &g, uint32_t uid)   // sim.cpp:596-792
{
    if (g.u8(OFF_DECKSIZE) == 0) g.u8(OFF_TURNS) = (uint8_t)(g.i8(OFF_TURNS) - 1);
    const int agent = g.u8(OFF_CUR);
    const uint32_t ranks = R.ranks, colors = R.colors;

    g.u8(OFF_LM_PLAYER) = (uint8_t)agent;
    g.u8(OFF_LM_TARGET) = 0xFF; g.u8(OFF_LM_CARD) = 0xFF; g.u8(OFF_LM_SCORED) = 0; g.u8(OFF_LM_INFOTOK) = 0;
    g.u8(OFF_LM_COLOR) = 0xFF; g.u8(OFF_LM_RANK) = 0xFF; g.u8(OFF_LM_REVEAL) = 0;
    g.u8(OFF_CUR) = (uint8_t)((agent + 1) % NP);

    if (uid < (uint32_t)HAND) {   // discard
        g.u8(OFF_LM_MOVE) = kDiscard; g.u8(OFF_LM_CARD) = (uint8_t)uid;
        const uint32_t card = g.u8(g.hand(agent, HAND_CARDS) + uid);
        g.u8(OFF_LM_COLOR) = (uint8_t)(card / ranks); g.u8(OFF_LM_RANK) = (uint8_t)(card % ranks);
        g.u8(OFF_DISCARD + card)++;
        g.u8(OFF_INFO)++;
        remove_from_hand(R, g, agent, (int)uid);
        return;
    }
    uid -= HAND;
    if (uid < (uint32_t)HAND) {   // play
        g.u8(OFF_LM_MOVE) = kPlay; g.u8(OFF_LM_CARD) = (uint8_t)uid;
        const uint32_t card = g.u8(g.hand(agent, HAND_CARDS) + uid);
        const uint32_t color = card / ranks, rank = card % ranks;
        g.u8(OFF_LM_COLOR) = (uint8_t)color; g.u8(OFF_LM_RANK) = (uint8_t)rank;
        if (g.u8(OFF_FIREWORKS + color) == rank) {
            const uint32_t fw = g.u8(OFF_FIREWORKS + color) + 1u;
            g.u8(OFF_FIREWORKS + color) = (uint8_t)fw;
            if (fw == ranks) { g.u8(OFF_INFO)++; g.u8(OFF_LM_INFOTOK) = 1; }   // no cap at max, sim.cpp:675-677
            g.u8(OFF_LM_SCORED) = 1;
        } else {
            g.u8(OFF_DISCARD + card)++;
            g.u8(OFF_LIFE)--;
        }
        remove_from_hand(R, g, agent, (int)uid);
        return;
    }
    uid -= HAND;
    const int partner = (agent + 1) % NP;   // target_offset is always 1 with two players
    const int psize = g.u8(g.hand(partner, HAND_SIZE));
    const int hc = g.hand(partner, HAND_CARDS);
    g.u8(OFF_INFO)--;
    g.u8(OFF_LM_TARGET) = (uint8_t)partner;
    uint32_t reveal = 0, newmask = 0;
    if (uid < colors) {           // reveal colour (lastmove.move deliberately untouched, see generate_rows)
        const uint32_t color = uid % colors;
        g.u8(OFF_LM_COLOR) = (uint8_t)color;
        for (uint32_t i = 0; i < ranks; i++) newmask |= 1u << (color * ranks + i);
        for (int i = 0; i < psize; i++) {
            const int pw = OFF_PLAUS + 4 * (partner * HAND + i);
            if (g.u8(hc + i) / ranks == color) {
                reveal |= 1u << i;
                g.u8(g.hand(partner, HAND_KCOLOR) + i) = (uint8_t)color;
                g.u32(pw) &= newmask;
            } else {
                g.u32(pw) &= ~newmask;
            }
        }
    } else {                      // reveal rank
        uid -= colors;
        const uint32_t rank = uid % ranks;
        g.u8(OFF_LM_RANK) = (uint8_t)rank;
        for (uint32_t i = 0; i < ranks; i++) newmask |= 1u << (i * ranks + rank);
        for (int i = 0; i < psize; i++) {
            const int pw = OFF_PLAUS + 4 * (partner * HAND + i);
            if (g.u8(hc + i) % ranks == rank) {
                reveal |= 1u << i;
                g.u8(g.hand(partner, HAND_KRANK) + i) = (uint8_t)rank;
                g.u32(pw) &= newmask;
            } else {
                g.u32(pw) &= ~newmask;
            }
        }
    }
    g.u8(OFF_LM_REVEAL) = (uint8_t)reveal;
}

__device__ void reset_game(const Rules &R, const Game &g, uint32_t episode_idx)   // sim.cpp:446-526
{
    g.u32(OFF_RNG) = rng_make(episode_idx);
    int ci = 0;
    for (int c = 0; c < R.colors; c++)
        for (int r = 0; r < R.ranks; r++) {
            const int id = R.ranks * c + r;
            const int cr = (r == 0 ? 3 : r == R.ranks - 1 ? 1 : 2);
            for (int i = 0; i < cr; i++) g.u8(OFF_DECK + ci++) = (uint8_t)id;
            g.u8(OFF_DISCARD + id) = 0;
        }
    g.u8(OFF_DECKSIZE) = (uint8_t)ci;
    for (int i = 0; i < R.colors; i++) g.u8(OFF_FIREWORKS + i) = 0;
    g.u8(OFF_INFO) = (uint8_t)R.max_info; g.u8(OFF_LIFE) = (uint8_t)R.max_life; g.u8(OFF_CUR) = 0;
    g.u8(OFF_TURNS) = NP; g.u8(OFF_SCORE) = 0;
    g.u8(OFF_LM_MOVE) = kInvalid; g.u8(OFF_LM_PLAYER) = 0xFF; g.u8(OFF_LM_TARGET) = 0xFF; g.u8(OFF_LM_CARD) = 0xFF;
    g.u8(OFF_LM_SCORED) = 0; g.u8(OFF_LM_INFOTOK) = 0; g.u8(OFF_LM_COLOR) = 0xFF; g.u8(OFF_LM_RANK) = 0xFF;
    g.u8(OFF_LM_REVEAL) = 0;
    for (int p = 0; p < NP; p++) {   // deal: agent-major, 5 draws each (sim.cpp:508-519)
        for (int j = 0; j < HAND; j++) {
            g.u8(g.hand(p, HAND_CARDS) + j) = (uint8_t)draw_deck(g);
            g.u8(g.hand(p, HAND_KCOLOR) + j) = 0xFF;
            g.u8(g.hand(p, HAND_KRANK) + j) = 0xFF;
            g.u32(OFF_PLAUS + 4 * (p * HAND + j)) = (1u << (R.colors * R.ranks)) - 1u;
        }
        g.u8(g.hand(p, HAND_SIZE)) = HAND;
    }
}

struct Buffers {
    uint32_t *game;
    const int32_t *actions;
    uint8_t *obs, *state;
    int32_t *mask, *active, *done;
    float *reward;
    uint32_t N;
};

constexpr int OBS_PITCH = 688, STATE_PITCH = 800;   // >= row + 15, multiples of 16
constexpr size_t SMEM_BYTES = (size_t)GAME_WORDS * 128 + 32 * OBS_PITCH + 32 * STATE_PITCH;

__global__ void __launch_bounds__(32) step_kernel(const Rules R, const Buffers B, const EpisodeScan scan,
                                                  const uint32_t num_tiles, const int init_only)
{
    extern __shared__ __align__(16) uint8_t smem[];
    __shared__ uint32_t s_scan[4];
    const int lane = threadIdx.x;

    EpisodeScanTile st;
    st.begin(scan, s_scan);
    const uint32_t n = st.tile * 32u + lane;
    const bool live = n < B.N;

    uint32_t *s_game = reinterpret_cast<uint32_t *>(smem);
    uint8_t *s_obs = smem + GAME_WORDS * 128;
    uint8_t *s_state = s_obs + 32 * OBS_PITCH;
    const Game g{reinterpret_cast<uint8_t *>(s_game + lane)};

    if (live && !init_only) {
        for (int w = 0; w < GAME_WORDS; w++)
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(s_game + w * 32 + lane)),
                         "l"(B.game + (size_t)w * B.N + n)
                         : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    {
        uint4 *s4 = reinterpret_cast<uint4 *>(s_obs);
        for (int i = lane; i < (32 * OBS_PITCH + 32 * STATE_PITCH) / 16; i += 32) s4[i] = make_uint4(0, 0, 0, 0);
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    __syncwarp();

    int done = 0;
    float reward = 0.f;
    if (live) {
        if (init_only) {
            done = 1;
        } else {
            const int cur = g.u8(OFF_CUR);
            action_system(R, g, (uint32_t)B.actions[(size_t)cur * B.N + n]);
            // checkDone, sim.cpp:812-849 (score / new_rew are int8 in the reference)
            const int old_score = g.i8(OFF_SCORE);
            int score = 0;
            if (g.u8(OFF_LIFE) > 0)
                for (int i = 0; i < R.colors; i++) score = (int)(int8_t)(score + g.u8(OFF_FIREWORKS + i));
            g.u8(OFF_SCORE) = (uint8_t)score;
            reward = (float)(int)(int8_t)(score - old_score);
            done = (g.u8(OFF_LIFE) < 1) || ((uint32_t)score >= (uint32_t)(R.colors * R.ranks)) || (g.i8(OFF_TURNS) <= 0);
        }
    }
    const unsigned done_mask = __ballot_sync(0xffffffffu, done);
    const uint32_t episode0 = st.exclusive_base(scan, __popc(done_mask), num_tiles, s_scan + 3);
    if (done) reset_game(R, g, episode0 + __popc(done_mask & ((1u << lane) - 1u)));

    const int cur = live ? (int)g.u8(OFF_CUR) : 0;   // 0 after a reset
    if (live) {
        B.reward[n] = reward;
        B.reward[(size_t)B.N + n] = reward;
        B.done[n] = init_only ? 0 : done;
        B.active[(size_t)cur * B.N + n] = 1;               // sim.cpp:510 / 802-808
        B.active[(size_t)(1 - cur) * B.N + n] = 0;
    }

    // pass 0: the (new) current player's rows; pass 1: the other player's rows of worlds that were reset
    for (int pass = 0; pass < 2; pass++) {
        const unsigned rows = pass == 0 ? __ballot_sync(0xffffffffu, live) : done_mask;
        if (!rows) break;
        if (pass == 1) {
            if (lane == 0) bulk_wait_read<0>();
            __syncwarp();
        }
        const int agent = pass == 0 ? cur : 1 - cur;
        uint8_t *obs_dst = B.obs + ((size_t)agent * B.N + n) * OBS;
        uint8_t *state_dst = B.state + ((size_t)agent * B.N + n) * STATE;
        uint8_t *obs_row = s_obs + lane * OBS_PITCH + (reinterpret_cast<uintptr_t>(obs_dst) & 15u);
        uint8_t *state_row = s_state + lane * STATE_PITCH + (reinterpret_cast<uintptr_t>(state_dst) & 15u);
        if ((rows >> lane) & 1u) {
            Sink s{obs_row, state_row, nullptr, nullptr};
            if (agent + 1 < NP) {
                s.obs_spill = B.obs + ((size_t)(agent + 1) * B.N + n) * OBS;
                s.state_spill = B.state + ((size_t)(agent + 1) * B.N + n) * STATE;
            }
            generate_rows(R, g, agent, s);
            generate_mask(R, g, agent, B.mask + ((size_t)agent * B.N + n) * MOVES);
        }
        fence_proxy_async_smem();
        __syncwarp();
        for (unsigned rem = rows; rem;) {
            const int r = __ffs(rem) - 1;
            rem &= rem - 1;
            const int a = __shfl_sync(0xffffffffu, agent, r);
            const size_t row = (size_t)a * B.N + (st.tile * 32u + r);
            uint8_t *od = B.obs + row * OBS;
            uint8_t *sd = B.state + row * STATE;
            slab_store(od, s_obs + r * OBS_PITCH + (reinterpret_cast<uintptr_t>(od) & 15u), OBS, lane, 32);
            slab_store(sd, s_state + r * STATE_PITCH + (reinterpret_cast<uintptr_t>(sd) & 15u), STATE, lane, 32);
        }
    }

    if (live)
        for (int w = 0; w < GAME_WORDS; w++) B.game[(size_t)w * B.N + n] = s_game[w * 32 + lane];
    if (lane == 0) bulk_wait_read<0>();
}

cudaError_t launch_step(const LaunchArgs &a, int init_only, cudaStream_t stream)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(step_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    Rules R{a.colors, a.ranks, a.max_info, a.max_life};
    Buffers B{a.game, a.actions, a.obs, a.state, a.mask, a.active, a.done, a.reward, a.N};
    EpisodeScan scan{a.scan_status, a.scan_ticket, a.scan_epoch, a.cur_episode};
    const uint32_t tiles = (a.N + 31) / 32;
    step_kernel<<<tiles, 32, SMEM_BYTES, stream>>>(R, B, scan, tiles, init_only);
    return cudaGetLastError();
}

}   // namespace hanabi
}   // namespace b200
