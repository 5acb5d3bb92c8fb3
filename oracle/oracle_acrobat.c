/* TEST INFRASTRUCTURE — CPU oracle, Acrobat.  See oracle.h.
 * Restates src/acrobat_env/sim.cpp:12-240 (one world = one two-link pendulum, RK4 over one 0.2 s step).
 *
 * Two things a restatement must not "fix":
 *  * precision: everything is float except the three sub-expressions the reference writes with double operands —
 *    `cos(theta2)` inside d2 (the unqualified call resolves to ::cos(double), the built reference imports `cos`),
 *    `pi / 2.0` inside phi2, and `dt / 2.0`, `dt / 6.0` in rk4.  Build with -ffp-contract=off.
 *  * episode length: EpisodeManager::episodeLength (init.hpp:9) is ONE counter shared by all worlds of a Manager.
 *    Every world-step increments it (sim.cpp:175), any reset zeroes it (:57), and a world is truncated when it finds
 *    the counter above 500 (:210).  With the deterministic one-worker order (worlds ascending, whole task graph per
 *    world, external/madrona/include/madrona/mw_cpu.inl:60-66) it counts world-steps since the last reset anywhere. */
#include "oracle.h"
#include "oracle_rng.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>

static const float MA_PI = 3.14159265358979323846264338327950288f;   /* madrona::math::pi, math.hpp:26 */

typedef struct { float theta1, theta2, omega1, omega2; } State;
typedef struct { float v[5]; } Vec5;
typedef struct {
    uint32_t N, cur_episode, episode_length;
    State *state;
    float *reward;
    int32_t *done;
} Sim;

/* sim.cpp:46-69 */
static void reset_world(Sim *s, uint32_t n)
{
    uint32_t rng = oracle_rng_make(s->cur_episode++);
    s->episode_length = 0;
    const float lo = -0.1f, hi = 0.1f;
    float diff = hi - lo;
    State st;
    st.theta1 = lo + oracle_rng_rand(&rng) * diff;
    st.theta2 = lo + oracle_rng_rand(&rng) * diff;
    st.omega1 = lo + oracle_rng_rand(&rng) * diff;
    st.omega2 = lo + oracle_rng_rand(&rng) * diff;
    s->state[n] = st;
}

/* sim.cpp:71-97; powf(x, 2) is x * x (exact in double, rounded once) */
static Vec5 ds_dt(Vec5 y)
{
    const float m1 = 1.0f, m2 = 1.0f, l1 = 1.0f, lc1 = 0.5f, lc2 = 0.5f, I1 = 1.0f, I2 = 1.0f, g = 9.8f;
    float theta1 = y.v[0], theta2 = y.v[1], dtheta1 = y.v[2], dtheta2 = y.v[3], a = y.v[4];
    float d1 = (m1 * (lc1 * lc1) + m2 * ((l1 * l1) + (lc2 * lc2) + 2 * l1 * lc2 * cosf(theta2)) + I1 + I2);
    float d2 = m2 * ((lc2 * lc2) + l1 * lc2 * cos(theta2)) + I2;                      /* double cos, see header */
    float phi2 = m2 * lc2 * g * cosf(theta1 + theta2 - MA_PI / 2.0);                  /* double subtraction */
    float phi1 = (-m2 * l1 * lc2 * (dtheta2 * dtheta2) * sinf(theta2) - 2 * m2 * l1 * lc2 * dtheta2 * dtheta1 * sinf(theta2) +
                  (m1 * lc1 + m2 * l1) * g * cosf(theta1 - MA_PI / 2) + phi2);
    float ddtheta2 = (a + d2 / d1 * phi1 - m2 * l1 * lc2 * (dtheta1 * dtheta1) * sinf(theta2) - phi2) /
                     (m2 * (lc2 * lc2) + I2 - (d2 * d2) / d1);
    float ddtheta1 = -(d2 * ddtheta2 + phi1) / d1;
    Vec5 out = {{dtheta1, dtheta2, ddtheta1, ddtheta2, 0.f}};
    return out;
}

static Vec5 mul_const(Vec5 x, float c) { for (int i = 0; i < 5; i++) x.v[i] *= c; return x; }
static Vec5 add(Vec5 a, Vec5 b) { for (int i = 0; i < 5; i++) a.v[i] += b.v[i]; return a; }

static float wrap(float x, float m, float M)
{
    float diff = M - m;
    while (x > M) x -= diff;
    while (x < m) x += diff;
    return x;
}
static float bound(float x, float m, float M) { return fminf(fmaxf(x, m), M); }

void *oracle_acrobat_create(uint32_t num_worlds)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->N = num_worlds;
    s->state = (State *)calloc(num_worlds, sizeof(State));
    s->reward = (float *)calloc(num_worlds, sizeof(float));
    s->done = (int32_t *)calloc(num_worlds, sizeof(int32_t));
    for (uint32_t n = 0; n < num_worlds; n++) reset_world(s, n);   /* Sim::Sim, sim.cpp:216-236 */
    return s;
}

void oracle_acrobat_destroy(void *h)
{
    Sim *s = (Sim *)h;
    free(s->state); free(s->reward); free(s->done); free(s);
}

void oracle_acrobat_step(void *h, const int32_t *actions)
{
    Sim *s = (Sim *)h;
    const float MAX_VEL_1 = 4 * MA_PI, MAX_VEL_2 = 9 * MA_PI;
    for (uint32_t n = 0; n < s->N; n++) {
        State *st = &s->state[n];
        /* actionSystem, sim.cpp:172-196 */
        s->episode_length++;
        const float torques[3] = {-1.0f, 0.0f, 1.0f};
        Vec5 y0 = {{st->theta1, st->theta2, st->omega1, st->omega2, torques[actions[n]]}};
        /* rk4 over {0, dt}, sim.cpp:117-147 */
        const float t0 = 0.0f, t1 = 0.2f;
        float dt = t1 - t0;
        float dt2 = dt / 2.0;
        float dt6 = dt / 6.0;
        Vec5 k1 = ds_dt(y0);
        Vec5 k2 = ds_dt(add(y0, mul_const(k1, dt2)));
        Vec5 k3 = ds_dt(add(y0, mul_const(k2, dt2)));
        Vec5 k4 = ds_dt(add(y0, mul_const(k3, dt)));
        Vec5 ns = add(y0, mul_const(add(k1, add(mul_const(k2, 2), add(mul_const(k3, 2), k4))), dt6));
        st->theta1 = wrap(ns.v[0], -MA_PI, MA_PI);
        st->theta2 = wrap(ns.v[1], -MA_PI, MA_PI);
        st->omega1 = bound(ns.v[2], -MAX_VEL_1, MAX_VEL_1);
        st->omega2 = bound(ns.v[3], -MAX_VEL_2, MAX_VEL_2);
        s->reward[n] = -1.f;
        /* checkDone, sim.cpp:198-214 */
        int termination = -cosf(st->theta1) - cosf(st->theta2 + st->theta1) > 1.0;
        int truncation = s->episode_length > 500;
        s->done[n] = termination || truncation;
        if (s->done[n]) reset_world(s, n);
    }
}

void oracle_acrobat_read(void *h, float *state, float *rewards, int32_t *done)
{
    Sim *s = (Sim *)h;
    if (state) memcpy(state, s->state, sizeof(State) * s->N);
    if (rewards) memcpy(rewards, s->reward, sizeof(float) * s->N);
    if (done) memcpy(done, s->done, sizeof(int32_t) * s->N);
}

void oracle_acrobat_write_state(void *h, const float *state)
{
    Sim *s = (Sim *)h;
    memcpy(s->state, state, sizeof(State) * s->N);
}

uint32_t oracle_acrobat_episode_length(void *h) { return ((Sim *)h)->episode_length; }
