/* TEST INFRASTRUCTURE — CPU oracle, 2-player Hanabi.  See oracle.h.
 * Restates src/hanabi_env/sim.cpp:45-891 including its quirks (each cited inline), with the reference's own
 * structs (sim.hpp:57-119) flattened into one World record.  Card id = colour*ranks + rank. */
#include "oracle.h"
#include "oracle_rng.h"

#include <stdlib.h>
#include <string.h>

#define MAX_CARDS 50
#define HAND 5
#define NP 2
enum { kDiscard, kPlay, kRevealColor, kRevealRank, kInvalid };     /* sim.hpp:83 */

typedef struct {                                                   /* sim.hpp:108-115 */
    uint8_t cards[HAND];
    uint64_t plausible[HAND];
    uint8_t size;
    int8_t known_color[HAND], known_rank[HAND];
} Hand;

typedef struct {
    /* Deck, sim.hpp:57-81 */
    uint8_t cards[MAX_CARDS], size, discard_counts[25], fireworks[5], info, life, cur_player;
    int8_t turns_to_play, score, new_rew;
    /* LastMove, sim.hpp:85-97 */
    int lm_move; int8_t lm_player, lm_target, lm_card_index; uint8_t lm_scored, lm_info_token;
    int8_t lm_color, lm_rank; uint8_t lm_reveal, lm_newly;
    Hand hand[NP];
    uint32_t rng;
    int32_t done;
    float reward;
    int32_t active[NP], mask[NP][ORACLE_HANABI_MOVES];
    uint8_t obs[NP][ORACLE_HANABI_OBS], state[NP][ORACLE_HANABI_STATE];
} World;

typedef struct {
    uint32_t N, cur_episode, colors, ranks, players, max_info, max_life, hand_size;
    int max_info_encoded;   /* test aid: largest information-token count ever encoded (> max_info = long rows) */
    World *w;
} Sim;

static uint8_t *oracle_hanabi_row_override = 0;   /* see generate_obs_state; the oracle is single-threaded */

static uint8_t draw_deck(World *w)                                 /* sim.cpp:45-52 */
{
    int32_t swaploc = (int32_t)(w->size * oracle_rng_rand(&w->rng));
    uint8_t ret = w->cards[swaploc];
    w->cards[swaploc] = w->cards[w->size - 1];
    w->size--;
    return ret;
}

static int encode_hands(const Sim *s, World *w, int agent, int off) /* sim.cpp:54-90 */
{
    uint8_t *o = oracle_hanabi_row_override ? oracle_hanabi_row_override : w->obs[agent];
    int bits = s->colors * s->ranks;
    for (uint32_t i = 1; i < s->players; i++) {
        const Hand *ph = &w->hand[(agent + i) % s->players];
        for (uint32_t c = 0; c < ph->size; c++)
            for (int b = 0; b < bits; b++) o[off++] = (b == ph->cards[c]);
        for (uint32_t c = ph->size; c < s->hand_size; c++)
            for (int b = 0; b < bits; b++) o[off++] = 0;
    }
    for (uint32_t i = 0; i < s->players; i++)
        o[off++] = (w->hand[(agent + i) % s->players].size < s->hand_size);
    return off;
}

static int encode_board(const Sim *s, World *w, int agent, int off) /* sim.cpp:92-135 */
{
    uint8_t *o = oracle_hanabi_row_override ? oracle_hanabi_row_override : w->obs[agent];
    for (int i = 0; i < w->size; i++) o[off++] = 1;
    int max_size = (4 + (s->ranks - 2) * 2) * s->colors - s->hand_size * s->players;
    for (int i = w->size; i < max_size; i++) o[off++] = 0;
    for (uint32_t c = 0; c < s->colors; c++)
        for (uint32_t i = 0; i < s->ranks; i++) o[off++] = (i + 1 == w->fireworks[c]);
    if (w->info > ((Sim *)s)->max_info_encoded) ((Sim *)s)->max_info_encoded = w->info;
    for (int i = 0; i < w->info; i++) o[off++] = 1;
    for (uint32_t i = w->info; i < s->max_info; i++) o[off++] = 0;
    for (int i = 0; i < w->life; i++) o[off++] = 1;
    for (uint32_t i = w->life; i < s->max_life; i++) o[off++] = 0;
    return off;
}

static int encode_discards(const Sim *s, World *w, int agent, int off) /* sim.cpp:137-156 */
{
    uint8_t *o = oracle_hanabi_row_override ? oracle_hanabi_row_override : w->obs[agent];
    int id = 0;
    for (uint32_t c = 0; c < s->colors; c++)
        for (uint32_t r = 0; r < s->ranks; r++) {
            int cr = (r == 0 ? 3 : r == s->ranks - 1 ? 1 : 2);
            for (int i = 0; i < cr; i++) o[off++] = (w->discard_counts[id] > i);
            id++;
        }
    return off;
}

static int encode_last_action(const Sim *s, World *w, int agent, int off) /* sim.cpp:158-289 */
{
    uint8_t *o = oracle_hanabi_row_override ? oracle_hanabi_row_override : w->obs[agent];
    int np = s->players;
    int reveal = (w->lm_move == kRevealColor || w->lm_move == kRevealRank);
    int playdisc = (w->lm_move == kPlay || w->lm_move == kDiscard);
    int rel = (w->lm_player == -1 ? -1 : (agent - w->lm_player + np) % np);
    for (int i = 0; i < np; i++) o[off++] = (i == rel);
    for (int i = 0; i < 4; i++) o[off + i] = 0;
    switch (w->lm_move) {
    case kPlay: o[off] = 1; break;
    case kDiscard: o[off + 1] = 1; break;
    case kRevealColor: o[off + 2] = 1; break;
    case kRevealRank: o[off + 3] = 1; break;
    default: break;
    }
    off += 4;
    int8_t rel_target = reveal ? (int8_t)((agent - w->lm_target + np) % np) : -1;
    for (int i = 0; i < np; i++) o[off + i] = reveal ? (i == rel_target) : 0;
    off += np;
    for (uint32_t i = 0; i < s->colors; i++)
        o[off + i] = (w->lm_move == kRevealColor) ? (i == (uint8_t)w->lm_color) : 0;
    off += s->colors;
    for (uint32_t i = 0; i < s->ranks; i++)
        o[off + i] = (w->lm_move == kRevealRank) ? (i == (uint8_t)w->lm_rank) : 0;
    off += s->ranks;
    for (uint32_t i = 0, m = 1; i < s->hand_size; i++, m <<= 1) o[off + i] = reveal ? ((w->lm_reveal & m) > 0) : 0;
    off += s->hand_size;
    for (uint32_t i = 0; i < s->hand_size; i++) o[off + i] = playdisc ? ((int)i == w->lm_card_index) : 0;
    off += s->hand_size;
    for (uint32_t i = 0; i < s->colors * s->ranks; i++)
        o[off + i] = playdisc ? (i == (uint32_t)(w->lm_color * s->ranks + w->lm_rank)) : 0;
    off += s->colors * s->ranks;
    o[off] = (w->lm_move == kPlay) ? w->lm_scored : 0;
    o[off + 1] = (w->lm_move == kPlay) ? w->lm_info_token : 0;
    return off + 2;
}

static int encode_card_knowledge(const Sim *s, World *w, int agent, int off) /* sim.cpp:291-331 */
{
    uint8_t *o = oracle_hanabi_row_override ? oracle_hanabi_row_override : w->obs[agent];
    uint32_t bits = s->colors * s->ranks;
    for (uint32_t i = 0; i < s->players; i++) {
        const Hand *ph = &w->hand[(agent + i) % s->players];
        for (int c = 0; c < ph->size; c++) {
            /* QUIRK sim.cpp:311: the tested bit is the PLAYER offset i, not the card value v */
            for (uint32_t v = 0; v < bits; v++) o[off++] = ((ph->plausible[c] & (1 << i)) != 0);
            for (uint32_t v = 0; v < s->colors; v++) o[off++] = (ph->known_color[c] == (int)v);
            for (uint32_t v = 0; v < s->ranks; v++) o[off++] = (ph->known_rank[c] == (int)v);
        }
        for (uint32_t c = ph->size; c < s->hand_size; c++)
            for (uint32_t v = 0; v < bits + s->colors + s->ranks; v++) o[off++] = 0;
    }
    return off;
}

static void generate_obs_state(const Sim *s, World *w, int agent)   /* sim.cpp:333-379 */
{
    /* The encoders write through w->obs[agent]; when information_tokens exceeds its maximum (a firework completed
     * at 8 tokens, sim.cpp:675-677 has no cap) encodeBoard emits one extra byte (sim.cpp:119-124) and everything after
     * it shifts by one, so the reference writes 659 / 784 bytes into 658 / 783-byte rows.  Rows of one world are
     * adjacent in its table, so the spilled byte lands on byte 0 of the NEXT agent's row (and past the table for the
     * last agent, where it is never exported).  Reproduced here through a scratch row instead of undefined behaviour. */
    uint8_t scratch_obs[ORACLE_HANABI_OBS + 8], scratch_state[ORACLE_HANABI_STATE + 8];
    int off;
    oracle_hanabi_row_override = scratch_obs;   /* the encoders below write into the scratch row */
    off = encode_hands(s, w, agent, 0);
    off = encode_board(s, w, agent, off);
    off = encode_discards(s, w, agent, off);
    off = encode_last_action(s, w, agent, off);
    off = encode_card_knowledge(s, w, agent, off);
    oracle_hanabi_row_override = NULL;
    memcpy(scratch_state, scratch_obs, off);                         /* copyObsToState :333-341 */
    int soff = off;
    const Hand *hd = &w->hand[agent];
    int bits = s->colors * s->ranks;
    for (int c = 0; c < hd->size; c++)                               /* encodeOwnHand :343-365 */
        for (int b = 0; b < bits; b++) scratch_state[soff++] = (b == hd->cards[c]);
    for (uint32_t c = hd->size; c < s->hand_size; c++)
        for (int b = 0; b < bits; b++) scratch_state[soff++] = 0;

    memcpy(w->obs[agent], scratch_obs, ORACLE_HANABI_OBS);
    memcpy(w->state[agent], scratch_state, ORACLE_HANABI_STATE);
    if ((uint32_t)agent + 1 < s->players) {
        if (off > ORACLE_HANABI_OBS) memcpy(w->obs[agent + 1], scratch_obs + ORACLE_HANABI_OBS, off - ORACLE_HANABI_OBS);
        if (soff > ORACLE_HANABI_STATE)
            memcpy(w->state[agent + 1], scratch_state + ORACLE_HANABI_STATE, soff - ORACLE_HANABI_STATE);
    }
}

static void generate_action_mask(const Sim *s, World *w, int agent) /* sim.cpp:381-444 */
{
    int32_t *m = w->mask[agent];
    const Hand *hd = &w->hand[agent];
    int off = 0;
    for (uint32_t i = 0; i < s->hand_size; i++) m[off++] = (i < hd->size && w->info < s->max_info);
    for (uint32_t i = 0; i < s->hand_size; i++) m[off++] = (i < hd->size);
    for (uint32_t p = 1; p < s->players; p++) {
        const Hand *ph = &w->hand[(agent + p) % s->players];
        for (uint32_t c = 0; c < s->colors; c++) {
            int has = 0;   /* QUIRK :416-418 scans all hand_size slots, stale ones included */
            for (uint32_t n = 0; n < s->hand_size; n++) has |= (ph->cards[n] / s->ranks == c);
            m[off++] = (w->info > 0 && has);
        }
    }
    for (uint32_t p = 1; p < s->players; p++) {
        const Hand *ph = &w->hand[(agent + p) % s->players];
        for (uint32_t r = 0; r < s->ranks; r++) {
            int has = 0;
            for (uint32_t n = 0; n < s->hand_size; n++) has |= (ph->cards[n] % s->ranks == r);
            m[off++] = (w->info > 0 && has);
        }
    }
    for (; off < 20; off++) m[off] = 0;
}

static void reset_world(Sim *s, World *w)                          /* sim.cpp:446-532 */
{
    w->rng = oracle_rng_make(s->cur_episode++);
    int ci = 0;
    for (uint32_t c = 0; c < s->colors; c++)
        for (uint32_t r = 0; r < s->ranks; r++) {
            int id = s->ranks * c + r;
            int cr = (r == 0 ? 3 : r == s->ranks - 1 ? 1 : 2);
            for (int i = 0; i < cr; i++) w->cards[ci++] = id;
            w->discard_counts[id] = 0;
        }
    w->size = ci;
    for (uint32_t i = 0; i < s->colors; i++) w->fireworks[i] = 0;
    w->info = s->max_info; w->life = s->max_life; w->cur_player = 0;
    w->turns_to_play = s->players; w->score = 0; w->new_rew = 0;
    w->lm_move = kInvalid; w->lm_player = w->lm_target = w->lm_card_index = -1;
    w->lm_scored = w->lm_info_token = 0; w->lm_color = w->lm_rank = -1; w->lm_reveal = w->lm_newly = 0;
    uint64_t valid = ((uint64_t)1 << (s->colors * s->ranks)) - 1;
    for (uint32_t i = 0; i < s->players; i++) {
        w->active[i] = (i == 0);
        for (uint32_t j = 0; j < s->hand_size; j++) {
            w->hand[i].cards[j] = draw_deck(w);
            w->hand[i].plausible[j] = valid;
            w->hand[i].known_color[j] = -1;
            w->hand[i].known_rank[j] = -1;
        }
        w->hand[i].size = s->hand_size;
    }
    for (uint32_t i = 0; i < s->players; i++) { generate_obs_state(s, w, i); generate_action_mask(s, w, i); }
}

static void remove_from_hand(const Sim *s, World *w, Hand *hd, int8_t index)  /* sim.cpp:567-594 */
{
    uint64_t valid = ((uint64_t)1 << (s->colors * s->ranks)) - 1;
    if (w->size == 0) {
        for (int8_t i = index + 1; i < hd->size; i++) {
            hd->cards[i - 1] = hd->cards[i];
            hd->plausible[i - 1] = hd->plausible[i];
            hd->known_color[i - 1] = hd->known_color[i];
            hd->known_rank[i - 1] = hd->known_rank[i];
        }
        hd->size--;
    } else {
        hd->cards[index] = draw_deck(w);
        hd->plausible[index] = valid;
        hd->known_color[index] = -1;
        hd->known_rank[index] = -1;
    }
}

static void action_system(const Sim *s, World *w, const int32_t *act /* [players] */)  /* sim.cpp:596-792 */
{
    if (w->size == 0) w->turns_to_play--;
    int agent = w->cur_player;
    Hand *hd = &w->hand[agent];
    uint32_t uid = (uint32_t)act[agent];
    uint32_t ranks = s->ranks, colors = s->colors, hs = s->hand_size, np = s->players;

    w->lm_player = w->cur_player; w->lm_target = -1; w->lm_card_index = -1; w->lm_scored = 0;
    w->lm_info_token = 0; w->lm_color = -1; w->lm_rank = -1; w->lm_reveal = 0; w->lm_newly = 0;
    w->cur_player = (w->cur_player + 1) % np;

    if (uid < hs) {                                                  /* discard :634-654 */
        w->lm_move = kDiscard; w->lm_card_index = uid;
        uint8_t card = hd->cards[uid];
        w->lm_color = card / ranks; w->lm_rank = card % ranks;
        w->discard_counts[card]++;
        w->info++;
        remove_from_hand(s, w, hd, uid);
        return;
    }
    uid -= hs;
    if (uid < hs) {                                                  /* play :658-691 */
        w->lm_move = kPlay; w->lm_card_index = uid;
        uint8_t card = hd->cards[uid];
        w->lm_color = card / ranks; w->lm_rank = card % ranks;
        if (w->fireworks[w->lm_color] == w->lm_rank) {
            w->fireworks[w->lm_color]++;
            if (w->fireworks[w->lm_color] == ranks) { w->info++; w->lm_info_token = 1; }   /* no cap at max, :675-677 */
            else w->lm_info_token = 0;
            w->lm_scored = 1;
        } else {
            w->discard_counts[card]++;
            w->life--;
            w->lm_scored = 0;
        }
        remove_from_hand(s, w, hd, uid);
        return;
    }
    uid -= hs;
    if (uid < (np - 1) * colors) {                                   /* reveal colour :695-740 */
        /* QUIRK :697-711: lastmove.move is NOT updated on reveal moves; it keeps the last discard/play/invalid */
        int target_offset = 1 + uid / colors;
        w->info--;
        int partner = (agent + target_offset) % np;
        Hand *ph = &w->hand[partner];
        w->lm_target = partner; w->lm_color = uid % colors;
        for (int i = 0; i < ph->size; i++)
            if (ph->cards[i] / ranks == (uint32_t)w->lm_color) w->lm_reveal |= (uint8_t)1 << i;
        uint64_t newmask = 0;
        for (uint32_t i = 0; i < ranks; i++) newmask |= (uint8_t)1 << (w->lm_color * ranks + i);  /* int shift, :723 */
        for (int i = 0; i < ph->size; i++) {
            if (ph->cards[i] / ranks == (uint32_t)w->lm_color) {
                if (ph->known_color[i] == -1) w->lm_newly |= (uint8_t)1 << i;
                ph->known_color[i] = w->lm_color;
                ph->plausible[i] &= newmask;
            } else {
                ph->plausible[i] &= ~newmask;
            }
        }
        return;
    }
    uid -= (np - 1) * colors;
    {                                                                /* reveal rank :744-788 */
        /* QUIRK :746-760: lastmove.move is NOT updated here either */
        int target_offset = 1 + uid / ranks;
        w->info--;
        int partner = (agent + target_offset) % np;
        Hand *ph = &w->hand[partner];
        w->lm_target = partner; w->lm_rank = uid % ranks;
        for (int i = 0; i < ph->size; i++)
            if (ph->cards[i] % ranks == (uint32_t)w->lm_rank) w->lm_reveal |= (uint8_t)1 << i;
        uint64_t newmask = 0;
        for (uint32_t i = 0; i < ranks; i++) newmask |= (uint8_t)1 << (i * ranks + w->lm_rank);
        for (int i = 0; i < ph->size; i++) {
            if (ph->cards[i] % ranks == (uint32_t)w->lm_rank) {
                if (ph->known_color[i] == -1) w->lm_newly |= (uint8_t)1 << i;   /* QUIRK :776 known_color on rank path */
                ph->known_rank[i] = w->lm_rank;
                ph->plausible[i] &= newmask;
            } else {
                ph->plausible[i] &= ~newmask;
            }
        }
    }
}

void *oracle_hanabi_create(uint32_t num_worlds, uint32_t colors, uint32_t ranks, uint32_t players,
                           uint32_t max_information_tokens, uint32_t max_life_tokens)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->N = num_worlds; s->colors = colors; s->ranks = ranks; s->players = players;
    s->max_info = max_information_tokens; s->max_life = max_life_tokens;
    s->hand_size = (players < 4 ? 5 : 4);                            /* sim.cpp:875 */
    s->w = (World *)calloc(num_worlds, sizeof(World));
    for (uint32_t n = 0; n < num_worlds; n++) reset_world(s, &s->w[n]);   /* Sim::Sim :866-891 */
    return s;
}

void oracle_hanabi_destroy(void *h) { Sim *s = (Sim *)h; free(s->w); free(s); }

void oracle_hanabi_step(void *h, const int32_t *actions)
{
    Sim *s = (Sim *)h;
    int32_t act[NP];
    for (uint32_t n = 0; n < s->N; n++) {
        World *w = &s->w[n];
        for (uint32_t p = 0; p < s->players; p++) act[p] = actions[(size_t)p * s->N + n];
        action_system(s, w, act);
        for (uint32_t i = 0; i < s->players; i++) {                   /* observationSystem :794-810 */
            if (i == w->cur_player) { w->active[i] = 1; generate_obs_state(s, w, i); generate_action_mask(s, w, i); }
            else w->active[i] = 0;
        }
        /* checkDone :812-849 */
        w->done = 0;
        int8_t old_score = w->score;
        w->score = 0;
        if (w->life > 0) for (uint32_t i = 0; i < s->colors; i++) w->score += w->fireworks[i];
        w->new_rew = w->score - old_score;
        w->reward = w->new_rew;
        if (w->life < 1) w->done = 1;
        if ((uint32_t)w->score >= s->colors * s->ranks) w->done = 1;
        if (w->turns_to_play <= 0) w->done = 1;
        if (w->done) reset_world(s, w);
    }
}

void oracle_hanabi_read(void *h, uint8_t *obs, uint8_t *state, int32_t *mask, int32_t *active, float *rewards,
                        int32_t *done)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) {
        const World *w = &s->w[n];
        for (uint32_t p = 0; p < s->players; p++) {
            size_t r = (size_t)p * s->N + n;
            if (obs) memcpy(obs + r * ORACLE_HANABI_OBS, w->obs[p], ORACLE_HANABI_OBS);
            if (state) memcpy(state + r * ORACLE_HANABI_STATE, w->state[p], ORACLE_HANABI_STATE);
            if (mask) memcpy(mask + r * ORACLE_HANABI_MOVES, w->mask[p], sizeof(int32_t) * ORACLE_HANABI_MOVES);
            if (active) active[r] = w->active[p];
            if (rewards) rewards[r] = w->reward;
        }
        if (done) done[n] = w->done;
    }
}

int oracle_hanabi_max_info_encoded(void *h) { return ((Sim *)h)->max_info_encoded; }
