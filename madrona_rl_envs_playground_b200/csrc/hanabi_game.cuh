// 2-player Hanabi game logic on a REGISTER-RESIDENT packed record (sm_100a kernels in hanabi.cu; the same code compiles
// for the host, where tests/test_hanabi_game_host.py checks it against the oracle without a GPU).
//
// Behaviour follows src/hanabi_env/sim.cpp:45-891 of the reference including its quirks (cited inline).  The layout
// does not: a world is 12 words that live in the registers of its lane plus the deck (50 bytes, the one thing that is
// indexed at random) in shared memory.  Everything a move touches is a bit field of a register, so a move is a few
// dozen ALU instructions instead of ~60 dependent shared-memory byte accesses, and composing a row reads registers.
// Only what can reach an exported tensor is stored:
//   * of each card's plausibility mask the reference's encoder reads bit 0 and bit 1 only (sim.cpp:311 tests the
//     PLAYER offset, not the card value), and the bits of a mask evolve independently (and / and-not), so two bits per
//     card are kept;
//   * reveal moves never update lastmove.move (sim.cpp:695-788 set move.type only), so the reveal-only sections of the
//     last-action encoding (target, colour, rank, reveal outcome) are always zero and target / reveal outcome are not
//     stored;
//   * the discard pile is kept in the unary format of the observation (3 / 2 / ... / 1 bits per card type), which
//     saturates exactly where the encoder's `count > i` comparisons do.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define HB_FN __host__ __device__ __forceinline__
#else
#define HB_FN inline
#endif

namespace b200 {
namespace hanabi {

constexpr int HAND = 5, NP = 2, MOVES = 20;
constexpr int OBS = 658, STATE = 783;
enum : int { kDiscard = 0, kPlay, kRevealColor, kRevealRank, kInvalid };   // sim.hpp:83

struct GameRules {
    int colors, ranks, max_info, max_life;
};

// one player's hand
struct Hand {
    uint32_t cards;   // card id of slot i in bits [5i, 5i+5) (stale ids stay in vacated slots, like the reference's arrays); size in bits [25, 28)
    uint32_t know;    // known colour + 1 of slot i in bits [3i, 3i+3), known rank + 1 in bits [16+3i, 16+3i+3); 0 = unknown
    uint32_t plaus;   // bit i = bit 0 of slot i's plausibility mask, bit 8+i = its bit 1
};
// one world
struct Rec {
    uint32_t board;     // deck size | information tokens << 8 | life tokens << 16 | score << 24
    uint32_t turn;      // player to move | turns_to_play << 8 | fireworks, 3 bits per colour, << 16
    uint32_t lm;        // last move: type [0,3) | player+1 [3,5) | card index+1 [7,10) | scored [10] | information token [11] | colour+1 [12,15) | rank+1 [15,18)
    uint32_t disc_lo, disc_hi;   // discard pile, observation format: colour-major, per colour 3 + 2 * (ranks - 2) + 1 unary bits
    uint32_t rng;
    Hand cur, oth;      // hand of the player to move, hand of the other player (they swap after every move)
};
constexpr int REC_WORDS = 12;
constexpr int DECK_WORDS = 13;   // 50 card ids, one byte each
static_assert(sizeof(Rec) == REC_WORDS * 4, "Rec is REC_WORDS words");

// The deck: bytes of DECK_WORDS words that are STRIDE bytes apart (128 in a [word][lane] shared tile, 4 on the host)
template <int STRIDE>
struct DeckT {
    uint8_t *base;
    HB_FN uint8_t &at(int i) const { return base[(i >> 2) * STRIDE + (i & 3)]; }
};

// ---- RNG: bit-for-bit the reference generator (src/cartpole_env/rng.hpp:7-36): TEA hash of the episode index, LCG ----
HB_FN uint32_t hb_rng_make(uint32_t episode_idx)
{
    uint32_t v0 = episode_idx, v1 = 0, s0 = 0;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int n = 0; n < 8; n++) {
        s0 += 0x9e3779b9u;
        v0 += ((v1 << 4) + 0xa341316cu) ^ (v1 + s0) ^ ((v1 >> 5) + 0xc8013ea4u);
        v1 += ((v0 << 4) + 0xad90777du) ^ (v0 + s0) ^ ((v0 >> 5) + 0x7e95761eu);
    }
    return v0;
}
HB_FN float hb_rng_rand(uint32_t &v)
{
    v = 1664525u * v + 1013904223u;
    return (float)(v & 0x00FFFFFFu) / 16777216.0f;   // exact: 24-bit integer scaled by 2^-24
}

template <class D>
HB_FN uint32_t draw_deck(Rec &g, const D &deck)   // sim.cpp:45-52
{
    const int size = (int)(g.board & 0xFFu);
    const int swaploc = (int)((float)size * hb_rng_rand(g.rng));
    if (size == 0) return 0u;   // (a deck smaller than the deal: undefined in the reference, refused at create)
    const uint32_t ret = deck.at(swaploc);
    deck.at(swaploc) = deck.at(size - 1);
    g.board -= 1u;
    return ret;
}

// card id = colour * ranks + rank; the standard 5-rank game divides by a constant
HB_FN void split_card(const GameRules &R, uint32_t card, uint32_t &color, uint32_t &rank)
{
    if (R.ranks == 5) { color = card / 5u; rank = card - 5u * color; }
    else { color = card / (uint32_t)R.ranks; rank = card - (uint32_t)R.ranks * color; }
}

// ---- bit stream of one agent's observation + state row -------------------------------------------------------------
// The reference's rows are one-hot / unary BYTE vectors; a lane composes its world's row as BITS into its own row of a
// shared [32][ROW_WORDS] array (ROW_WORDS odd => conflict free both for the writing lane and for the warp that later
// reads one row across lanes); the writing warp then expands bits to bytes 16 at a time with aligned 128-bit stores.
constexpr int ROW_WORDS = 27;   // 25 words hold 783 + 5 spill bits; one guard word for the 16-bit window reads
struct BitWriter {
    uint32_t *row;
    uint32_t acc;
    int nacc, widx;   // nacc in [0, 32)
    // append the low n (0..32) bits of v; FITS = the caller guarantees v < 2^n
    template <bool FITS = false>
    HB_FN void put(uint32_t v, int n)
    {
        if (n <= 0) return;
        if (!FITS && n < 32) v &= (1u << n) - 1u;
        acc |= v << nacc;
        const int tot = nacc + n;
        if (tot >= 32) {
            row[widx++] = acc;
#if defined(__CUDA_ARCH__)
            acc = __funnelshift_rc(v, 0u, 32 - nacc);   // the bits of v that did not fit (0 when nacc == 0)
#else
            acc = nacc ? v >> (32 - nacc) : 0u;
#endif
            nacc = tot - 32;
        } else {
            nacc = tot;
        }
    }
    HB_FN void ones_then_zeros(int ones, int total)   // unary code, `total` >= 0 bits in all
    {
        for (int done = 0; done < total; done += 32) {
            const int n = total - done < 32 ? total - done : 32;
            int o = ones - done;
            o = o < 0 ? 0 : o > n ? n : o;
            put<true>(o >= 32 ? 0xFFFFFFFFu : ((1u << o) - 1u), n);
        }
    }
    HB_FN int finish()   // returns the number of bits written
    {
        const int bits = widx * 32 + nacc;
        row[widx] = acc;
        for (int i = widx + 1; i < ROW_WORDS; i++) row[i] = 0;
        return bits;
    }
};

// generateObsState, sim.cpp:367-379 (encodeHands 54-90, encodeBoard 92-135, encodeDiscards 137-156,
// encodeLastAction 158-289, encodeCardKnowledge 291-331, copyObsToState 333-341, encodeOwnHand 343-365) for the agent
// `agent` whose hand is `mine` (the partner's is `theirs`).
// Returns the length in bits of the obs part (658, or more when information_tokens exceeds its maximum and the board
// section grows — sim.cpp:675-677 has no cap — which shifts everything after it); the state part follows it.
// FIXED = the caller guarantees deck size <= its maximum and tokens <= their maxima, so every section has its nominal
// length: with compile-time rules (standard game) all bit offsets then fold to constants and a put() is a shift + or.
// Every section is written with the same put() sequence whatever the data (values are selected, not branched on), so
// the writer's position never depends on control flow.
template <bool FIXED>
HB_FN int generate_bits(const GameRules &R, const Rec &g, const Hand &mine, const Hand &theirs, uint32_t agent, BitWriter &bw)
{
    const int bits = R.colors * R.ranks;
    const int msize = (int)(mine.cards >> 25), tsize = (int)(theirs.cards >> 25);
    // hands: partner's cards one-hot, then "hand is short" flags in self-relative order
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < HAND; c++) bw.put<true>(c < tsize ? 1u << ((theirs.cards >> (5 * c)) & 31u) : 0u, bits);
    bw.put<true>((msize < HAND ? 1u : 0u) | (tsize < HAND ? 2u : 0u), NP);
    // board
    {
        const int size = (int)(g.board & 0xFFu), info = (int)((g.board >> 8) & 0xFFu), life = (int)((g.board >> 16) & 0xFFu);
        const int max_size = 2 * R.ranks * R.colors - HAND * NP;
        bw.ones_then_zeros(size, FIXED ? max_size : (size > max_size ? size : max_size));
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int c = 0; c < 5; c++)
            if (c < R.colors) bw.put<true>((1u << ((g.turn >> (16 + 3 * c)) & 7u)) >> 1, R.ranks);   // one-hot of fireworks - 1
        bw.ones_then_zeros(info, FIXED ? R.max_info : (info > R.max_info ? info : R.max_info));   // no cap at sim.cpp:675-677
        bw.ones_then_zeros(life, FIXED ? R.max_life : (life > R.max_life ? life : R.max_life));
    }
    // discards: unary count per card type, 3 / 2 / ... / 1 copies — the record keeps them in this very format
    {
        const int nd = 2 * R.ranks * R.colors;
        bw.put<true>(g.disc_lo, nd < 32 ? nd : 32);
        bw.put<true>(g.disc_hi, nd - 32);
    }
    // last action.  QUIRK: reveal moves never update lastmove.move (sim.cpp:695-788 set move.type only), so `move` is
    // the last discard / play / invalid: the reveal-only sections (target, colour, rank, reveal outcome) are always 0,
    // while the colour / rank a reveal stores DO show up in the "card played" section of a stale discard / play.
    {
        const uint32_t move = g.lm & 7u, player1 = (g.lm >> 3) & 3u, card1 = (g.lm >> 7) & 7u;
        const uint32_t scored = (g.lm >> 10) & 1u, infotok = (g.lm >> 11) & 1u;
        const int lm_color = (int)((g.lm >> 12) & 7u) - 1, lm_rank = (int)((g.lm >> 15) & 7u) - 1;
        const bool playdisc = move <= (uint32_t)kPlay;
        bw.put<true>(player1 ? 1u << ((agent ^ (player1 - 1u)) & 1u) : 0u, NP);   // (agent - player + 2) % 2
        bw.put<true>((move == (uint32_t)kPlay ? 1u : 0u) | (move == (uint32_t)kDiscard ? 2u : 0u), 4);
        bw.put<true>(0u, NP + R.colors + R.ranks + HAND);
        bw.put<true>(playdisc ? (1u << card1) >> 1 : 0u, HAND);
        // uint32 arithmetic like the reference (int8 colour * uint32 ranks + int8 rank), sim.cpp:267
        const uint32_t played = (uint32_t)lm_color * (uint32_t)R.ranks + (uint32_t)lm_rank;
        bw.put<true>((playdisc && played < (uint32_t)bits) ? (1u << played) : 0u, bits);
        bw.put<true>(move == (uint32_t)kPlay ? (scored | (infotok << 1)) : 0u, 2);
    }
    // card knowledge.  QUIRK sim.cpp:311: the plausibility bit tested is the PLAYER offset i, for all `bits` slots.
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < NP; i++) {
        const Hand &h = i ? theirs : mine;
        const int hsize = i ? tsize : msize;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int c = 0; c < HAND; c++) {
            const bool present = c < hsize;
            const uint32_t plaus = (h.plaus >> (8 * i + c)) & 1u;
            bw.put((present && plaus) ? 0xFFFFFFFFu : 0u, bits);
            const uint32_t kc = (1u << ((h.know >> (3 * c)) & 7u)) >> 1;          // one-hot of known colour, none when unknown
            const uint32_t kr = (1u << ((h.know >> (16 + 3 * c)) & 7u)) >> 1;
            bw.put<true>(present ? (kc | (kr << R.colors)) : 0u, R.colors + R.ranks);
        }
    }
    const int obs_bits = bw.widx * 32 + bw.nacc;
    // own hand (state only)
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int c = 0; c < HAND; c++) bw.put<true>(c < msize ? 1u << ((mine.cards >> (5 * c)) & 31u) : 0u, bits);
    return obs_bits;
}

// generateActionMask, sim.cpp:381-444: legal moves as 20 bits (discard 0-4, play 5-9, reveal colour 10..,
// reveal rank 10+colors..) of the agent whose hand is `mine`
HB_FN uint32_t generate_mask(const GameRules &R, const Rec &g, const Hand &mine, const Hand &theirs)
{
    const int size = (int)(mine.cards >> 25), info = (int)((g.board >> 8) & 0xFFu);
    const uint32_t in_hand = (1u << size) - 1u;
    uint32_t legal = (info < R.max_info ? in_hand : 0u) | (in_hand << HAND);
    if (info > 0) {
        uint32_t colors_seen = 0, ranks_seen = 0;
        // QUIRK sim.cpp:416-418,431-433: all 5 slots are scanned, stale ones after an empty-deck shift included
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int n = 0; n < HAND; n++) {
            uint32_t c, r;
            split_card(R, (theirs.cards >> (5 * n)) & 31u, c, r);
            colors_seen |= 1u << c;
            ranks_seen |= 1u << r;
        }
        legal |= (colors_seen & ((1u << R.colors) - 1u)) << (2 * HAND);
        legal |= (ranks_seen & ((1u << R.ranks) - 1u)) << (2 * HAND + R.colors);
    }
    return legal;
}

// composes one row (bit stream + legal moves); returns obs bits | all bits << 16
template <bool STD>
HB_FN uint32_t compose_bits(GameRules R, const Rec &g, const Hand &mine, const Hand &theirs, uint32_t agent, uint32_t *row,
                            uint32_t *legal)
{
    if (STD) { R.colors = 5; R.ranks = 5; R.max_info = 8; R.max_life = 3; }
    BitWriter bw{row, 0u, 0, 0};
    // nominal section lengths unless the deck / token counters exceed their maxima (tokens can: sim.cpp:675-677)
    const int max_size = 2 * R.ranks * R.colors - HAND * NP;
    int obs_bits;
    if ((int)(g.board & 0xFFu) <= max_size && (int)((g.board >> 8) & 0xFFu) <= R.max_info && (int)((g.board >> 16) & 0xFFu) <= R.max_life)
        obs_bits = generate_bits<true>(R, g, mine, theirs, agent, bw);
    else
        obs_bits = generate_bits<false>(R, g, mine, theirs, agent, bw);
    const int all_bits = bw.finish();
    *legal = generate_mask(R, g, mine, theirs);
    return (uint32_t)obs_bits | ((uint32_t)all_bits << 16);
}

// ---- moves ------------------------------------------------------------------------------------------------------------
// slots [idx, size-1) of a word of `b`-bit slots take the value of the slot above; everything else (the stale last slot
// included) stays: the reference's shift loop, sim.cpp:570-578
HB_FN uint32_t shift_slots(uint32_t w, int b, int idx, int size, uint32_t replicate)
{
    const uint32_t low = (1u << (b * idx)) - 1u, upto = (1u << (b * (size > 0 ? size - 1 : 0))) - 1u;
    const uint32_t mid = (upto & ~low) * replicate;   // (0 when idx >= size - 1)
    return (w & ~mid) | ((w >> b) & mid);
}

template <class D>
HB_FN void remove_from_hand(const GameRules &R, Rec &g, Hand &h, const D &deck, int index)   // sim.cpp:567-594
{
    if ((g.board & 0xFFu) == 0u) {
        const int size = (int)(h.cards >> 25);
        h.cards = shift_slots(h.cards, 5, index, size, 1u) - (size > 0 ? 1u << 25 : 0u);
        h.know = shift_slots(h.know, 3, index, size, 0x10001u);
        h.plaus = shift_slots(h.plaus, 1, index, size, 0x101u);
    } else {
        const uint32_t card = draw_deck(g, deck);
        h.cards = (h.cards & ~(31u << (5 * index))) | (card << (5 * index));
        h.know &= ~(0x70007u << (3 * index));
        h.plaus |= 0x101u << index;
    }
    (void)R;
}

// one more copy of card (color, rank) on the discard pile (unary, saturating at the number of copies)
HB_FN void discard_add(const GameRules &R, Rec &g, uint32_t color, uint32_t rank)
{
    const int k = 2 * R.ranks * (int)color + (rank ? 2 * (int)rank + 1 : 0);
    const uint64_t field = rank == 0 ? 7u : rank == (uint32_t)R.ranks - 1u ? 1u : 3u;
    uint64_t d = (uint64_t)g.disc_lo | ((uint64_t)g.disc_hi << 32);
    const uint64_t room = ~(d >> k) & field;
    d |= (room & (0 - room)) << k;
    g.disc_lo = (uint32_t)d;
    g.disc_hi = (uint32_t)(d >> 32);
}

// actionSystem, sim.cpp:596-792, for the move `uid` of the player to move; afterwards g.cur is the NEXT player's hand
template <class D>
HB_FN void action_system(const GameRules &R, Rec &g, const D &deck, uint32_t uid)
{
    if ((g.board & 0xFFu) == 0u) g.turn -= 1u << 8;   // turns_to_play-- (it is >= 1 here: the game ends when it reaches 0)
    const uint32_t agent = g.turn & 1u;
    g.turn ^= 1u;
    uint32_t lm = (g.lm & 7u) | ((agent + 1u) << 3);   // player = the mover, everything else back to its default; the type stays
    if (uid < (uint32_t)(2 * HAND)) {   // discard / play
        const bool play = uid >= (uint32_t)HAND;
        const int idx = (int)(play ? uid - HAND : uid);
        uint32_t color, rank;
        split_card(R, (g.cur.cards >> (5 * idx)) & 31u, color, rank);
        lm = (play ? (uint32_t)kPlay : (uint32_t)kDiscard) | ((agent + 1u) << 3) | ((uint32_t)(idx + 1) << 7) | ((color + 1u) << 12) | ((rank + 1u) << 15);
        bool to_pile = true;
        if (!play) {
            // Legal discards happen below max_info only; plays that complete a stack can push the counter to max_info +
            // colors (no cap, sim.cpp:675-677) and the rows then grow by that many bytes.  ILLEGAL discards (the caller
            // ignored the mask; undefined in the reference, whose rows would run further and further out of bounds) stop
            // counting there, so that a row never outgrows its ROW_WORDS of shared memory.
            if ((int)((g.board >> 8) & 0xFFu) < R.max_info + R.colors) g.board += 1u << 8;
        } else {
            const uint32_t fw = (g.turn >> (16 + 3 * color)) & 7u;
            if (fw == rank) {
                g.turn += 1u << (16 + 3 * color);
                lm |= 1u << 10;
                if (fw + 1u == (uint32_t)R.ranks) { g.board += 1u << 8; lm |= 1u << 11; }   // no cap at max, sim.cpp:675-677
                to_pile = false;
            } else {
                g.board -= 1u << 16;   // a life token
            }
        }
        if (to_pile) discard_add(R, g, color, rank);
        remove_from_hand(R, g, g.cur, deck, idx);
    } else {   // reveal colour / rank to the partner (target_offset is always 1 with two players)
        uint32_t v = uid - 2 * HAND;
        const bool by_color = v < (uint32_t)R.colors;
        if (!by_color) v = (v - (uint32_t)R.colors) % (uint32_t)R.ranks;
        if ((g.board >> 8) & 0xFFu) g.board -= 1u << 8;   // (an ILLEGAL hint without a token would wrap the counter: ignored, as above)
        lm |= (v + 1u) << (by_color ? 12 : 15);
        const int psize = (int)(g.oth.cards >> 25);
        uint32_t match = 0, know = g.oth.know;
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int i = 0; i < HAND; i++) {
            uint32_t cc, cr;
            split_card(R, (g.oth.cards >> (5 * i)) & 31u, cc, cr);
            if (i < psize && (by_color ? cc : cr) == v) {
                match |= 1u << i;
                const int pos = 3 * i + (by_color ? 0 : 16);
                know = (know & ~(7u << pos)) | ((v + 1u) << pos);
            }
        }
        g.oth.know = know;
        // plausibility masks: matching cards keep the bits inside the hint's mask, the others the bits outside it.
        // Bit 0 is card (colour 0, rank 0), bit 1 is card (colour 0, rank 1) (ranks >= 2): inside a colour hint's mask
        // when the colour is 0, inside a rank hint's when the rank is 0 / 1 (sim.cpp:721-737, 770-786)
        const uint32_t present = (1u << psize) - 1u;
        const bool in0 = v == 0u, in1 = by_color ? v == 0u : v == 1u;
        const uint32_t keep0 = ~present | (in0 ? match : ~match), keep1 = ~present | (in1 ? match : ~match);
        g.oth.plaus &= (keep0 & 31u) | ((keep1 & 31u) << 8) | ~0x1F1Fu;
    }
    g.lm = lm;
    const Hand t = g.cur;
    g.cur = g.oth;
    g.oth = t;
}

// checkDone, sim.cpp:812-849 (score / new_rew are int8 in the reference): updates the score, returns done
HB_FN bool check_done(const GameRules &R, Rec &g, float &reward)
{
    const int old_score = (int)(g.board >> 24);
    int score = 0;
    if ((g.board >> 16) & 0xFFu) {
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
        for (int c = 0; c < 5; c++)
            if (c < R.colors) score += (int)((g.turn >> (16 + 3 * c)) & 7u);
    }
    g.board = (g.board & 0x00FFFFFFu) | ((uint32_t)score << 24);
    reward = (float)(score - old_score);
    return ((g.board >> 16) & 0xFFu) < 1u || score >= R.colors * R.ranks || (int)(int8_t)((g.turn >> 8) & 0xFFu) <= 0;
}

// the deck of a fresh game: (colour, rank) order with 3 / 2 / ... / 1 copies (sim.cpp:446-470); returns its size
inline int fresh_deck(const GameRules &R, uint8_t bytes[DECK_WORDS * 4])
{
    int ci = 0;
    for (int i = 0; i < DECK_WORDS * 4; i++) bytes[i] = 0;
    for (int c = 0; c < R.colors; c++)
        for (int r = 0; r < R.ranks; r++) {
            const int cr = (r == 0 ? 3 : r == R.ranks - 1 ? 1 : 2);
            for (int i = 0; i < cr; i++) bytes[ci++] = (uint8_t)(R.ranks * c + r);
        }
    return ci;
}

// resetWorld, sim.cpp:446-526; `fresh` = the words of fresh_deck().  Player 0 moves first: g.cur is player 0's hand.
template <class D>
HB_FN void reset_game(const GameRules &R, const uint32_t *fresh, Rec &g, const D &deck, uint32_t episode_idx)
{
    g.rng = hb_rng_make(episode_idx);
#if defined(__CUDA_ARCH__)
#pragma unroll
#endif
    for (int i = 0; i < DECK_WORDS; i++) *reinterpret_cast<uint32_t *>(&deck.at(4 * i)) = fresh[i];
    g.board = (uint32_t)(2 * R.ranks * R.colors) | ((uint32_t)R.max_info << 8) | ((uint32_t)R.max_life << 16);
    g.turn = (uint32_t)NP << 8;
    g.lm = (uint32_t)kInvalid;
    g.disc_lo = g.disc_hi = 0u;
    Hand h[NP];
    for (int p = 0; p < NP; p++) {   // deal: agent-major, 5 draws each (sim.cpp:508-519)
        uint32_t cards = (uint32_t)HAND << 25;
        for (int j = 0; j < HAND; j++) cards |= draw_deck(g, deck) << (5 * j);
        h[p].cards = cards;
        h[p].know = 0u;
        h[p].plaus = 0x1F1Fu;
    }
    g.cur = h[0];
    g.oth = h[1];
}

}   // namespace hanabi
}   // namespace b200
