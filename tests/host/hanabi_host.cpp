// TEST HARNESS (CPU): the product's Hanabi game logic (csrc/hanabi_game.cuh, the code the sm_100a kernels inline) compiled
// for the host and driven one world after the other in the reference's order, so that tests/test_hanabi_game_host.py can
// compare it with the oracle without a GPU.  What the kernels add on top — tiles, the ordered episode scan, the row
// writer — is reproduced here in its simplest form (a loop over worlds, a counter, a bit -> byte loop).
#include <cstdlib>
#include <cstring>
#include <vector>

#include "hanabi_game.cuh"

using namespace b200::hanabi;
typedef DeckT<4> HostDeck;

struct HostSim {
    uint32_t N;
    GameRules R;
    uint32_t fresh[DECK_WORDS];
    uint32_t cur_episode;
    std::vector<Rec> rec;
    std::vector<uint32_t> deck;   // [N][DECK_WORDS]
    std::vector<uint8_t> obs, state;   // [2][N][OBS], [2][N][STATE]
    std::vector<int32_t> mask, active, done;
    std::vector<float> reward;
};

static void write_row(HostSim *s, uint32_t n, uint32_t agent, const Rec &g, const Hand &mine, const Hand &theirs)
{
    uint32_t row[ROW_WORDS + 1] = {}, legal;
    const bool std_game = s->R.colors == 5 && s->R.ranks == 5 && s->R.max_info == 8 && s->R.max_life == 3;
    const uint32_t lens = std_game ? compose_bits<true>(s->R, g, mine, theirs, agent, row, &legal)
                                   : compose_bits<false>(s->R, g, mine, theirs, agent, row, &legal);
    const int ob = (int)(lens & 0xFFFFu), ab = (int)(lens >> 16);
    auto bit = [&](int b) { return (uint8_t)((row[b >> 5] >> (b & 31)) & 1u); };
    const size_t r = (size_t)agent * s->N + n;
    for (int b = 0; b < OBS; b++) s->obs[r * OBS + b] = bit(b);
    for (int b = 0; b < STATE; b++) s->state[r * STATE + b] = bit(b);
    if (agent + 1 < (uint32_t)NP) {   // longer rows spill onto the next player's (the reference's out-of-row writes)
        for (int b = OBS; b < ob; b++) s->obs[(r + s->N) * OBS + (b - OBS)] = bit(b);
        for (int b = STATE; b < ab; b++) s->state[(r + s->N) * STATE + (b - STATE)] = bit(b);
    }
    for (int m = 0; m < MOVES; m++) s->mask[r * MOVES + m] = (legal >> m) & 1u;
}

static void reset_world(HostSim *s, uint32_t n)
{
    Rec &g = s->rec[n];
    const HostDeck deck{reinterpret_cast<uint8_t *>(&s->deck[(size_t)n * DECK_WORDS])};
    reset_game(s->R, s->fresh, g, deck, s->cur_episode++);
    write_row(s, n, 0, g, g.cur, g.oth);
    write_row(s, n, 1, g, g.oth, g.cur);
    s->active[n] = 1;
    s->active[s->N + n] = 0;
}

extern "C" {

void *hh_create(uint32_t N, int colors, int ranks, int max_info, int max_life)
{
    HostSim *s = new HostSim;
    s->N = N;
    s->R = GameRules{colors, ranks, max_info, max_life};
    uint8_t bytes[DECK_WORDS * 4];
    fresh_deck(s->R, bytes);
    memcpy(s->fresh, bytes, sizeof(bytes));
    s->cur_episode = 0;
    s->rec.resize(N);
    s->deck.assign((size_t)N * DECK_WORDS, 0u);
    s->obs.assign((size_t)2 * N * OBS, 0);
    s->state.assign((size_t)2 * N * STATE, 0);
    s->mask.assign((size_t)2 * N * MOVES, 0);
    s->active.assign((size_t)2 * N, 0);
    s->done.assign(N, 0);
    s->reward.assign((size_t)2 * N, 0.f);
    for (uint32_t n = 0; n < N; n++) reset_world(s, n);
    return s;
}

void hh_destroy(void *h) { delete (HostSim *)h; }

void hh_step(void *h, const int32_t *actions)
{
    HostSim *s = (HostSim *)h;
    for (uint32_t n = 0; n < s->N; n++) {
        Rec &g = s->rec[n];
        const HostDeck deck{reinterpret_cast<uint8_t *>(&s->deck[(size_t)n * DECK_WORDS])};
        const uint32_t mover = g.turn & 1u;
        action_system(s->R, g, deck, (uint32_t)actions[(size_t)mover * s->N + n]);
        float reward;
        const bool done = check_done(s->R, g, reward);
        s->reward[n] = s->reward[s->N + n] = reward;
        s->done[n] = done;
        if (done) {
            reset_world(s, n);
        } else {
            const uint32_t agent = g.turn & 1u;
            write_row(s, n, agent, g, g.cur, g.oth);
            s->active[(size_t)agent * s->N + n] = 1;
            s->active[(size_t)(1 - agent) * s->N + n] = 0;
        }
    }
}

void hh_read(void *h, uint8_t *obs, uint8_t *state, int32_t *mask, int32_t *active, float *reward, int32_t *done)
{
    HostSim *s = (HostSim *)h;
    memcpy(obs, s->obs.data(), s->obs.size());
    memcpy(state, s->state.data(), s->state.size());
    memcpy(mask, s->mask.data(), s->mask.size() * 4);
    memcpy(active, s->active.data(), s->active.size() * 4);
    memcpy(reward, s->reward.data(), s->reward.size() * 4);
    memcpy(done, s->done.data(), s->done.size() * 4);
}

}   // extern "C"

#ifdef HH_FUZZ_MAIN
// Robustness run (tests/test_hanabi_game_host.py, built with -fsanitize=address,undefined): ARBITRARY action values —
// what a caller that ignores the legal-move mask may send; undefined behaviour in the reference — must not take the
// game logic out of its memory (deck bytes, the 27-word row, the record's bit fields) or out of its counters' ranges.
#include <cstdio>
int main()
{
    const int configs[4][4] = {{5, 5, 8, 3}, {5, 2, 8, 3}, {3, 3, 4, 2}, {2, 4, 2, 3}};
    uint32_t lcg = 12345u;
    for (const int *c : configs) {
        const uint32_t N = 64;
        HostSim *s = (HostSim *)hh_create(N, c[0], c[1], c[2], c[3]);
        std::vector<int32_t> act(2 * N);
        for (int t = 0; t < 3000; t++) {
            for (auto &a : act) {
                lcg = 1664525u * lcg + 1013904223u;
                a = (int32_t)((lcg >> 8) % 48u) - 8;   // -8 .. 39: mostly illegal
            }
            hh_step(s, act.data());
            for (uint32_t n = 0; n < N; n++) {
                const Rec &g = s->rec[n];
                const int size = g.board & 0xFF, info = (g.board >> 8) & 0xFF, life = (g.board >> 16) & 0xFF;
                if (size > 2 * c[0] * c[1] || info > c[2] + c[0] || life > c[3] || (g.cur.cards >> 25) > 5 || (g.oth.cards >> 25) > 5) {
                    printf("counter out of range: config %d %d step %d world %u: deck %d info %d life %d\n", c[0], c[1], t, n, size, info, life);
                    return 1;
                }
            }
        }
        hh_destroy(s);
    }
    printf("ok\n");
    return 0;
}
#endif
