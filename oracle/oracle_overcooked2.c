/* TEST INFRASTRUCTURE — CPU oracle, Overcooked2 ("Simplecooked", the JS-compatible rules).  See oracle.h.
 *
 * Restates src/overcooked2_env/sim.cpp of the reference as flat arrays (no ECS): the 13 task-graph nodes run in
 * insertion order (sim.cpp:419-450), each over its entities in creation order, as the Madrona CPU backend executes
 * them.  Differences from src/overcooked_env that matter (all restated below):
 *   - terrain enum order AIR, POT, COUNTER, ONION_SOURCE, DISH_SOURCE, SERVING, TOMATO_SOURCE (sim.hpp:40);
 *   - 5P+10 observation channels: terrain one-hot at 5P+t-1 — so a TOMATO_SOURCE cell's bit lands on channel 5P+5,
 *     which observationSystem overwrites on every pass (sim.cpp:79,555-560); 5P+5 onions in a pot's soup, 5P+6 its
 *     cooking tick, 5P+7 soup elsewhere / held, 5P+8 dish, 5P+9 onion; tomatoes and urgency have no channel;
 *   - ALL interactions are resolved by one sequential pass over the players in id order (sim.cpp:198-286);
 *   - a pot starts cooking by itself on its third ingredient, never by an empty-handed interact (sim.cpp:270-272);
 *   - dish-pickup shaping reward: num_dishes_out == 0 and fewer held dishes (players 0 and 1) than pots that are
 *     non-empty and cooking or not full, counted BEFORE the pass (sim.cpp:174-196,244-248).
 * The observation is kept incrementally (past_player / current_player bookkeeping) like the reference.
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

enum { A_NORTH = 0, A_SOUTH = 1, A_EAST = 2, A_WEST = 3, A_STAY = 4, A_INTERACT = 5 };       /* sim.hpp:39 */
enum { T_AIR, T_POT, T_COUNTER, T_ONION_SOURCE, T_DISH_SOURCE, T_SERVING, T_TOMATO_SOURCE }; /* sim.hpp:40 */
enum { O_NONE, O_TOMATO, O_ONION, O_DISH, O_SOUP };                                          /* sim.hpp:42 */
#define MAX_INGREDIENTS 3

typedef struct { uint8_t name, onions, tomatoes; int8_t tick; } Obj;                       /* sim.hpp:59-74 */
static const Obj OBJ_NONE = { O_NONE, 0, 0, -1 };

typedef struct {                                                                              /* sim.hpp:133-144 */
    Obj object;
    int32_t past_player, past_orientation, current_player, future_player;
    int32_t interacting[4], num_interacting;
} Cell;

typedef struct {                                                                              /* sim.hpp:149-184 */
    uint8_t position, orientation, proposed_position, proposed_orientation;
    Obj held;
    int8_t interaction_index;
} Player;

typedef struct {
    int32_t timestep, reset_now, calculated_reward, should_update_pos;
    uint8_t num_dishes_out;                                                                   /* sim.hpp:107 */
    Player *players;  /* [P] */
    Cell *cells;      /* [S] */
    uint8_t *obs;     /* [S][P][C] : row (cell, observer) like LocationXObservation, sim.cpp:546-552 */
} World;

typedef struct {
    OracleOvercookedCfg cfg;
    int P, S, C, W, H, npots;
    int pots[256];
    /* WorldState narrows these to uint8_t (sim.hpp:94-96, sim.cpp:525-527) */
    uint8_t placement_rew, dish_rew, soup_rew;
    uint32_t N;
    World *worlds;
} Sim;

static int recipe_of(const Obj *o) { return (MAX_INGREDIENTS + 1) * o->onions + o->tomatoes; } /* sim.hpp:70-73 */
static int get_time(const Sim *s, const Obj *o) { return s->cfg.recipe_times[recipe_of(o)]; }   /* sim.cpp:53-56 */
static int is_cooking(const Sim *s, const Obj *o) { return o->tick >= 0 && o->tick < get_time(s, o); }  /* :58-61 */
static int is_ready(const Sim *s, const Obj *o) { return o->tick >= 0 && o->tick >= get_time(s, o); }   /* :63-66 */

static int move_in_direction(int point, int dir, int width)                                   /* sim.cpp:160-172 */
{
    if (dir == A_NORTH) return point - width;
    if (dir == A_SOUTH) return point + width;
    if (dir == A_EAST) return point + 1;
    if (dir == A_WEST) return point - 1;
    return point;
}

static int relative_player(int other, int observer)                                           /* sim.cpp:106-128 */
{
    if (other == observer) return 0;
    return other < observer ? other + 1 : other;
}

/* sim.cpp:68-142, one (cell, observer) row */
static void observe_row(const Sim *s, World *w, int loc, int observer)
{
    const int P = s->P, shift = 5 * P;
    uint8_t *x = w->obs + ((size_t)loc * P + observer) * s->C;
    const Cell *dat = &w->cells[loc];
    const Obj *obj = &dat->object;

    for (int c = 5; c <= 9; c++) x[shift + c] = 0;

    if (obj->name == O_SOUP) {
        if (s->cfg.terrain[loc] == T_POT) {
            x[shift + 5] = obj->onions;
            x[shift + 6] = (obj->tick < 0) ? 0 : (uint8_t)obj->tick;
        } else {
            x[shift + 7] = 1;
        }
    } else if (obj->name == O_DISH) {
        x[shift + 8] = 1;
    } else if (obj->name == O_ONION) {
        x[shift + 9] = 1;
    }

    if (dat->past_player != -1) {
        int r = relative_player(dat->past_player, observer);
        x[r] = 0;
        x[P + 4 * r + dat->past_orientation] = 0;
    }
    if (dat->current_player != -1) {
        int other = dat->current_player;
        int i = relative_player(other, observer);
        const Player *ps = &w->players[other];
        x[i] = 1;
        x[P + 4 * i + ps->orientation] = 1;
        if (ps->held.name == O_SOUP) x[shift + 7] = 1;
        else if (ps->held.name == O_DISH) x[shift + 8] = 1;
        else if (ps->held.name == O_ONION) x[shift + 9] = 1;
    }
}

/* LocationXPlayer rows are created cell-major, player-minor with id = cell + player*S (sim.cpp:546-552); the
 * observation node visits them in that creation order. */
static void observe_world(const Sim *s, World *w)
{
    for (int loc = 0; loc < s->S; loc++)
        for (int p = 0; p < s->P; p++)
            observe_row(s, w, loc, p);
}

/* sim.cpp:361-368, 371-377, 389-403 as called by resetWorld (sim.cpp:454-468) and by the task graph */
static void reset_world_system(World *w)
{
    w->should_update_pos = 1;
    if (w->reset_now) { w->timestep = 0; w->num_dishes_out = 0; }
}
static void reset_objects_system(World *w, Cell *c) { c->num_interacting = 0; if (w->reset_now) c->object = OBJ_NONE; }
static void reset_actors_system(const Sim *s, World *w, int i, int32_t *reward_out)
{
    Player *p = &w->players[i];
    if (reward_out) *reward_out = w->calculated_reward;
    if (w->reset_now) {
        p->position = (uint8_t)(s->cfg.start_y[i] * (uint8_t)s->W + s->cfg.start_x[i]);
        w->cells[p->position].current_player = i;
        p->orientation = A_NORTH;
        p->proposed_position = p->position;
        p->proposed_orientation = p->orientation;
        p->held = OBJ_NONE;
    }
}

void *oracle_overcooked2_create(const OracleOvercookedCfg *cfg, uint32_t num_worlds)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->cfg = *cfg;
    s->P = cfg->num_players; s->W = cfg->width; s->H = cfg->height;
    s->S = cfg->width * cfg->height; s->C = 5 * s->P + 10; s->N = num_worlds;
    s->placement_rew = (uint8_t)cfg->placement_in_pot_rew;
    s->dish_rew = (uint8_t)cfg->dish_pickup_rew;
    s->soup_rew = (uint8_t)cfg->soup_pickup_rew;
    for (int x = 0; x < s->S; x++) if (cfg->terrain[x] == T_POT) s->pots[s->npots++] = x;   /* sim.cpp:493-512 */
    s->worlds = (World *)calloc(num_worlds, sizeof(World));
    for (uint32_t n = 0; n < num_worlds; n++) {                                              /* Sim::Sim, sim.cpp:471-576 */
        World *w = &s->worlds[n];
        w->players = (Player *)calloc(s->P, sizeof(Player));
        w->cells = (Cell *)calloc(s->S, sizeof(Cell));
        w->obs = (uint8_t *)calloc((size_t)s->S * s->P * s->C, 1);
        for (int i = 0; i < s->P; i++) { w->players[i].held = OBJ_NONE; w->players[i].interaction_index = -1; }
        for (int c = 0; c < s->S; c++) {
            Cell *cell = &w->cells[c];
            cell->object = OBJ_NONE;
            cell->past_player = cell->past_orientation = cell->current_player = cell->future_player = -1;
            int t = cfg->terrain[c];
            if (t) for (int p = 0; p < s->P; p++) w->obs[((size_t)c * s->P + p) * s->C + t - 1 + 5 * s->P] = 1;
        }
        w->reset_now = 1;
        reset_world_system(w);
        for (int c = 0; c < s->S; c++) reset_objects_system(w, &w->cells[c]);
        for (int i = 0; i < s->P; i++) reset_actors_system(s, w, i, NULL);
        w->reset_now = 0;
        observe_world(s, w);
    }
    return s;
}

void oracle_overcooked2_destroy(void *h)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) { free(s->worlds[n].players); free(s->worlds[n].cells); free(s->worlds[n].obs); }
    free(s->worlds);
    free(s);
}

static void step_world(const Sim *s, World *w, const int32_t *act /* [P], this world's actions */, int32_t *rew /* [P] */)
{
    const int P = s->P, S = s->S, W = s->W;
    Player *pl = w->players;
    Cell *cells = w->cells;

    /* 1 resolve_interacts :198-286 — one sequential pass over the players */
    int non_empty_pots = 0;                                                   /* get_pot_states :174-185 */
    for (int k = 0; k < s->npots; k++) {
        const Obj *o = &cells[s->pots[k]].object;
        if (o->name != O_NONE && (o->tick >= 0 || (uint8_t)(o->onions + o->tomatoes) < MAX_INGREDIENTS)) non_empty_pots++;
    }
    int rew_acc = 0;
    for (int i = 0; i < P; i++) {
        Player *p = &pl[i];
        if (act[i] != A_INTERACT) continue;
        int i_pos = move_in_direction(p->position, p->orientation, W);
        int t = s->cfg.terrain[i_pos];
        Obj *soup = &cells[i_pos].object;
        if (t == T_COUNTER) {
            if (p->held.name != O_NONE && soup->name == O_NONE) {
                *soup = p->held; p->held = OBJ_NONE;
                if (soup->name == O_DISH) w->num_dishes_out++;
            } else if (p->held.name == O_NONE && soup->name != O_NONE) {
                if (soup->name == O_DISH) w->num_dishes_out--;
                p->held = *soup; *soup = OBJ_NONE;
            }
        } else if (t == T_ONION_SOURCE) {
            if (p->held.name == O_NONE) { p->held = OBJ_NONE; p->held.name = O_ONION; }
        } else if (t == T_TOMATO_SOURCE) {
            if (p->held.name == O_NONE) { p->held = OBJ_NONE; p->held.name = O_TOMATO; }
        } else if (t == T_DISH_SOURCE) {
            if (p->held.name == O_NONE) {
                int player_dishes = 0;                                        /* is_dish_pickup_useful :187-196 */
                for (int q = 0; q < 2 && q < P; q++) if (pl[q].held.name == O_DISH) player_dishes++;
                if (w->num_dishes_out == 0 && player_dishes < non_empty_pots) rew_acc += s->dish_rew;
                p->held = OBJ_NONE; p->held.name = O_DISH;
            }
        } else if (t == T_POT) {
            if (p->held.name == O_DISH && soup->name == O_SOUP && is_ready(s, soup)) {
                p->held = *soup; *soup = OBJ_NONE;
                rew_acc += s->soup_rew;
            } else if (p->held.name == O_ONION || p->held.name == O_TOMATO) {
                if (soup->name == O_NONE) { *soup = OBJ_NONE; soup->name = O_SOUP; }
                if (!(soup->tick >= 0 || (uint8_t)(soup->onions + soup->tomatoes) == MAX_INGREDIENTS)) {
                    Obj obj = p->held; p->held = OBJ_NONE;
                    if (obj.name == O_ONION) soup->onions++; else soup->tomatoes++;
                    rew_acc += s->placement_rew;
                }
                /* soup_to_be_cooked_at_location && full, sim.cpp:270-272 */
                if (soup->name == O_SOUP && !is_cooking(s, soup) && !is_ready(s, soup) &&
                    (uint8_t)(soup->onions + soup->tomatoes) == MAX_INGREDIENTS)
                    soup->tick = 0;
            }
        } else if (t == T_SERVING) {
            if (p->held.name == O_SOUP) {                                     /* deliver_soup :144-148 */
                Obj so = p->held; p->held = OBJ_NONE;
                rew_acc += s->cfg.recipe_values[recipe_of(&so)];
            }
        }
    }
    w->calculated_reward = rew_acc;

    for (int i = 0; i < P; i++) {                                             /* 2 _move_if_direction :290-306 */
        Player *p = &pl[i];
        if (act[i] == A_INTERACT) {
            p->proposed_position = p->position; p->proposed_orientation = p->orientation;
        } else {
            int new_pos = move_in_direction(p->position, act[i], W);
            int new_or = (act[i] == A_STAY) ? p->orientation : act[i];
            p->proposed_position = (uint8_t)(s->cfg.terrain[new_pos] != T_AIR ? p->position : new_pos);
            p->proposed_orientation = (uint8_t)new_or;
        }
        cells[p->proposed_position].future_player = i;
    }
    for (int i = 0; i < P; i++) {                                             /* 3 _check_collisions :310-323 */
        Player *p = &pl[i];
        int comp = cells[p->proposed_position].current_player;
        if (cells[p->proposed_position].future_player != i ||
            (comp != -1 && comp != i && cells[p->position].future_player == comp))
            w->should_update_pos = 0;
    }
    for (int i = 0; i < P; i++) {                                             /* 4 _unset_loc_info :327-333 */
        Player *p = &pl[i];
        cells[p->position].current_player = -1;
        cells[p->proposed_position].future_player = -1;
        cells[p->position].past_player = i;
        cells[p->position].past_orientation = p->orientation;
    }
    for (int i = 0; i < P; i++) {                                             /* 5 _handle_collisions :337-346 */
        Player *p = &pl[i];
        if (w->should_update_pos) p->position = p->proposed_position;
        p->orientation = p->proposed_orientation;
        cells[p->position].current_player = i;
    }
    for (int k = 0; k < s->npots; k++) {                                      /* 6 step_pot_effects :350-358 */
        Obj *o = &cells[s->pots[k]].object;
        if (o->name == O_SOUP && is_cooking(s, o)) o->tick++;
    }
    w->timestep += 1;                                                         /* 7 check_reset_system :406-410 */
    w->reset_now = (w->timestep >= s->cfg.horizon);
    reset_world_system(w);                                                    /* 8 :361-368 */
    for (int c = 0; c < S; c++) reset_objects_system(w, &cells[c]);           /* 9 :371-377 */
    if (w->reset_now)                                                         /* 10 _pre_reset_actors_system :380-385 */
        for (int i = 0; i < P; i++) cells[pl[i].position].current_player = -1;
    for (int i = 0; i < P; i++) reset_actors_system(s, w, i, &rew[i]);        /* 11 :389-403 */
    observe_world(s, w);                                                      /* 12 observationSystem :68-142 */
    for (int c = 0; c < S; c++) cells[c].past_player = cells[c].past_orientation = -1;   /* 13 :412-416 */
}

/* [S][P][C] rows -> canonical [P][N][W][H][C] (envs/overcooked2_env.py:91-103: scatter by LocationXID, reshape
 * (P,H,W,N,C), transpose(1,3)) */
static void export_obs(const Sim *s, const World *w, uint32_t n, uint8_t *obs)
{
    const int P = s->P, C = s->C, W = s->W, H = s->H;
    for (int p = 0; p < P; p++)
        for (int x = 0; x < W; x++)
            for (int y = 0; y < H; y++) {
                int cell = y * W + x;
                memcpy(obs + ((((size_t)p * s->N + n) * W + x) * H + y) * C,
                       w->obs + ((size_t)cell * P + p) * C, C);
            }
}

void oracle_overcooked2_observe(void *h, uint8_t *obs)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) export_obs(s, &s->worlds[n], n, obs);
}

void oracle_overcooked2_step(void *h, const int32_t *actions, uint8_t *obs, int32_t *rewards, int32_t *done)
{
    Sim *s = (Sim *)h;
    int32_t act[64], rew[64];
    for (uint32_t n = 0; n < s->N; n++) {
        for (int p = 0; p < s->P; p++) act[p] = actions[(size_t)p * s->N + n];
        step_world(s, &s->worlds[n], act, rew);
        if (rewards) for (int p = 0; p < s->P; p++) rewards[(size_t)p * s->N + n] = rew[p];
        if (done) done[n] = s->worlds[n].reset_now;
        if (obs) export_obs(s, &s->worlds[n], n, obs);
    }
}
