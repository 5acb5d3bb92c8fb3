/* TEST INFRASTRUCTURE — CPU oracle, Balance Beam.  See oracle.h.
 * Restates src/balance_beam_env/sim.cpp:46-205 (2 agents, 5 spaces, TIME 3). */
#include "oracle.h"
#include "oracle_rng.h"

#include <stdlib.h>
#include <string.h>

#define BUFFER 2
#define NUM_SPACES 5
#define TIME 3
#define SCALE 0.2

typedef struct {
    int32_t loc[2];
    int32_t obs[2][2 * TIME + 1];   /* Observation{ x[6], time } is 7 contiguous int32 (sim.hpp:50-53) */
    int32_t time, done;
    float reward;
} World;
typedef struct { uint32_t N, cur_episode; World *w; } Sim;

/* sim.cpp:46-81 */
static void reset_world(Sim *s, World *w)
{
    uint32_t rng = oracle_rng_make(s->cur_episode++);
    w->time = TIME - 1;
    for (int i = 0; i < 2; i++) {
        w->loc[i] = (int32_t)(NUM_SPACES * oracle_rng_rand(&rng));
        for (int t = 0; t < 2 * TIME; t++) w->obs[i][t] = 0;
        w->obs[i][2 * TIME] = w->time;
    }
    for (int i = 0; i < 2; i++) {
        w->obs[i][0] = w->loc[i] + BUFFER;
        w->obs[i][TIME] = w->loc[1 - i] + BUFFER;
    }
}

void *oracle_balance_create(uint32_t num_worlds)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->N = num_worlds;
    s->w = (World *)calloc(num_worlds, sizeof(World));
    for (uint32_t n = 0; n < num_worlds; n++) reset_world(s, &s->w[n]);     /* Sim::Sim, sim.cpp:178-205 */
    return s;
}

void oracle_balance_destroy(void *h) { Sim *s = (Sim *)h; free(s->w); free(s); }

void oracle_balance_step(void *h, const int32_t *actions)
{
    Sim *s = (Sim *)h;
    static const int delta[4] = { -2, -1, 1, 2 };
    for (uint32_t n = 0; n < s->N; n++) {
        World *w = &s->w[n];
        for (int i = 0; i < 2; i++) {                                         /* actionSystem :83-99 */
            int32_t a = actions[(size_t)i * s->N + n];
            if (a >= 0 && a < 4) w->loc[i] += delta[a];
        }
        w->time -= 1;                                                         /* timeSystem :101-104 */
        for (int i = 0; i < 2; i++) {                                         /* observationSystem :106-118 */
            int32_t *x = w->obs[i];
            for (int k = TIME * 2; k > 0; k--) x[k] = x[k - 1];              /* also shifts into the time slot */
            x[TIME] = w->loc[1 - i] + BUFFER;
            x[0] = w->loc[i] + BUFFER;
            x[2 * TIME] = w->time;
        }
        /* checkDone :120-155 */
        w->done = 0;
        int d = w->loc[0] - w->loc[1];
        float reward = (w->loc[0] == w->loc[1] ? 1.0 : -abs(d) * SCALE);
        for (int i = 0; i < 2; i++) {
            if (w->loc[i] < 0 || w->loc[i] >= NUM_SPACES) {
                w->done = 1;
                reward = -NUM_SPACES * (w->time + 1) * SCALE;
            }
        }
        w->reward = reward;
        if (w->time == 0) w->done = 1;
        if (w->done) reset_world(s, w);
    }
}

void oracle_balance_read(void *h, int32_t *obs, float *rewards, int32_t *done)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) {
        const World *w = &s->w[n];
        for (int i = 0; i < 2; i++) {
            if (obs) memcpy(obs + ((size_t)i * s->N + n) * 7, w->obs[i], 7 * sizeof(int32_t));
            if (rewards) rewards[(size_t)i * s->N + n] = w->reward;
        }
        if (done) done[n] = w->done;
    }
}
