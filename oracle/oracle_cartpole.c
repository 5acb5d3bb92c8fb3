/* TEST INFRASTRUCTURE — CPU oracle, Cartpole.  See oracle.h.
 * Restates src/cartpole_env/sim.cpp:9-145.  The physical constants are double literals in the reference, so each
 * expression is evaluated in double and rounded to float on assignment; build with -ffp-contract=off. */
#include "oracle.h"
#include "oracle_rng.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

#define GRAVITY 9.8
#define MASSCART 1.0
#define MASSPOLE 0.1
#define TOTAL_MASS (MASSPOLE + MASSCART)
#define LENGTH 0.5
#define POLEMASS_LENGTH (MASSPOLE * LENGTH)
#define FORCE_MAG 10
#define TAU 0.02
#define X_THRESHOLD 2.4
#define MA_PI 3.141592653589793238463
#define THETA_THRESHOLD_RADIANS (12 * 2 * MA_PI / 360)

typedef struct { float x, x_dot, theta, theta_dot; } State;
typedef struct {
    uint32_t N, cur_episode;
    State *state;
    float *reward;
    int32_t *done;
} Sim;

/* sim.cpp:48-66 */
static void reset_world(Sim *s, uint32_t n)
{
    uint32_t rng = oracle_rng_make(s->cur_episode++);
    const float lo = -0.05f, hi = 0.05f;
    float diff = hi - lo;
    State st;
    st.x = lo + oracle_rng_rand(&rng) * diff;
    st.x_dot = lo + oracle_rng_rand(&rng) * diff;
    st.theta = lo + oracle_rng_rand(&rng) * diff;
    st.theta_dot = lo + oracle_rng_rand(&rng) * diff;
    s->state[n] = st;
}

void *oracle_cartpole_create(uint32_t num_worlds)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->N = num_worlds;
    s->state = (State *)calloc(num_worlds, sizeof(State));
    s->reward = (float *)calloc(num_worlds, sizeof(float));
    s->done = (int32_t *)calloc(num_worlds, sizeof(int32_t));
    for (uint32_t n = 0; n < num_worlds; n++) reset_world(s, n);   /* Sim::Sim, sim.cpp:123-144 */
    return s;
}

void oracle_cartpole_destroy(void *h)
{
    Sim *s = (Sim *)h;
    free(s->state); free(s->reward); free(s->done); free(s);
}

void oracle_cartpole_step(void *h, const int32_t *actions)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) {
        State *st = &s->state[n];
        /* actionSystem, sim.cpp:68-84 */
        float force = (actions[n] == 1 ? FORCE_MAG : -FORCE_MAG);
        float costheta = cosf(st->theta);
        float sintheta = sinf(st->theta);
        float temp = (force + POLEMASS_LENGTH * st->theta_dot * st->theta_dot * sintheta) / TOTAL_MASS;
        float thetaacc = (GRAVITY * sintheta - costheta * temp) /
                         (LENGTH * (4.0 / 3.0 - MASSPOLE * costheta * costheta / TOTAL_MASS));
        float xacc = temp - POLEMASS_LENGTH * thetaacc * costheta / TOTAL_MASS;
        st->x = st->x + TAU * st->x_dot;
        st->x_dot = st->x_dot + TAU * xacc;
        st->theta = st->theta + TAU * st->theta_dot;
        st->theta_dot = st->theta_dot + TAU * thetaacc;
        s->reward[n] = 1.f;
        /* checkDone, sim.cpp:86-96 */
        float x = st->x, theta = st->theta;
        s->done[n] = x < -X_THRESHOLD || x > X_THRESHOLD || theta < -THETA_THRESHOLD_RADIANS ||
                     theta > THETA_THRESHOLD_RADIANS;
        if (s->done[n]) reset_world(s, n);
    }
}

void oracle_cartpole_read(void *h, float *state, float *rewards, int32_t *done)
{
    Sim *s = (Sim *)h;
    if (state) memcpy(state, s->state, sizeof(State) * s->N);
    if (rewards) memcpy(rewards, s->reward, sizeof(float) * s->N);
    if (done) memcpy(done, s->done, sizeof(int32_t) * s->N);
}

void oracle_cartpole_write_state(void *h, const float *state)
{
    Sim *s = (Sim *)h;
    memcpy(s->state, state, sizeof(State) * s->N);
}
