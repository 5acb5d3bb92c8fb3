/* TEST INFRASTRUCTURE — CPU oracle, Overcooked.  See oracle.h for scope and parity status.
 *
 * Restates src/overcooked_env/sim.cpp of the reference as flat arrays (no ECS): the 19 task-graph nodes run in
 * insertion order, each over its entities in creation order, exactly as the Madrona CPU backend executes them
 * (sim.cpp:498-537; external/madrona/src/core/taskgraph.cpp:129-135).  The observation is kept incrementally
 * (past_player / current_player bookkeeping) like the reference, so that tests comparing it with the CUDA
 * kernel's from-scratch encoding also prove the two formulations agree.
 */
#include "oracle.h"

#include <stdlib.h>
#include <string.h>

enum { A_NORTH = 0, A_SOUTH = 1, A_EAST = 2, A_WEST = 3, A_STAY = 4, A_INTERACT = 5 };       /* sim.hpp:40 */
enum { T_AIR, T_POT, T_COUNTER, T_ONION_SOURCE, T_TOMATO_SOURCE, T_DISH_SOURCE, T_SERVING }; /* sim.hpp:41 */
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
    Player *players;  /* [P] */
    Cell *cells;      /* [S] */
    uint8_t *obs;     /* [S][P][C] : row (cell, observer) like LocationXObservation, sim.cpp:624-636 */
} World;

typedef struct {
    OracleOvercookedCfg cfg;
    int P, S, C, W, H, npots;
    int pots[256];
    /* WorldState narrows these to uint8_t (sim.hpp:93-95, sim.cpp:603-605) */
    uint8_t placement_rew, dish_rew, soup_rew;
    uint32_t N;
    World *worlds;
} Sim;

static int recipe_of(const Obj *o) { return (MAX_INGREDIENTS + 1) * o->onions + o->tomatoes; } /* sim.hpp:70-73 */
static int get_time(const Sim *s, const Obj *o) { return s->cfg.recipe_times[recipe_of(o)]; }   /* sim.cpp:53-56 */
static int is_cooking(const Sim *s, const Obj *o) { return o->tick >= 0 && o->tick < get_time(s, o); }  /* :58-61 */
static int is_ready(const Sim *s, const Obj *o) { return o->tick >= 0 && o->tick >= get_time(s, o); }   /* :63-66 */

static int move_in_direction(int point, int dir, int width)                                   /* sim.cpp:185-197 */
{
    if (dir == A_NORTH) return point - width;
    if (dir == A_SOUTH) return point + width;
    if (dir == A_EAST) return point + 1;
    if (dir == A_WEST) return point - 1;
    return point;
}

static int relative_player(int other, int observer)                                           /* sim.cpp:122-145 */
{
    if (other == observer) return 0;
    return other < observer ? other + 1 : other;
}

/* sim.cpp:68-167, one (cell, observer) row */
static void observe_row(const Sim *s, World *w, int loc, int observer)
{
    const int P = s->P, shift = 5 * P;
    uint8_t *x = w->obs + ((size_t)loc * P + observer) * s->C;
    const Cell *dat = &w->cells[loc];
    const Obj *obj = &dat->object;

    x[shift + 15] = (s->cfg.horizon - w->timestep < 40) ? 1 : 0;
    for (int c = 6; c <= 14; c++) x[shift + c] = 0;

    if (obj->name == O_SOUP) {
        if (s->cfg.terrain[loc] == T_POT) {
            if (obj->tick < 0) {
                x[shift + 6] = obj->onions;
                x[shift + 7] = obj->tomatoes;
            } else {
                x[shift + 8] = obj->onions;
                x[shift + 9] = obj->tomatoes;
                x[shift + 10] = (uint8_t)(get_time(s, obj) - obj->tick);
                if (is_ready(s, obj)) x[shift + 11] = 1;
            }
        } else {
            x[shift + 8] = obj->onions;
            x[shift + 9] = obj->tomatoes;
            x[shift + 10] = 0;
            x[shift + 11] = 1;
        }
    } else if (obj->name == O_DISH) {
        x[shift + 12] = 1;
    } else if (obj->name == O_ONION) {
        x[shift + 13] = 1;
    } else if (obj->name == O_TOMATO) {
        x[shift + 14] = 1;
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
        if (ps->held.name != O_NONE) {
            const Obj *o2 = &ps->held;
            if (o2->name == O_SOUP) {
                x[shift + 8] = o2->onions;
                x[shift + 9] = o2->tomatoes;
                x[shift + 10] = 0;
                x[shift + 11] = 1;
            } else if (o2->name == O_DISH) {
                x[shift + 12] = 1;
            } else if (o2->name == O_ONION) {
                x[shift + 13] = 1;
            } else if (o2->name == O_TOMATO) {
                x[shift + 14] = 1;
            }
        }
    }
}

/* LocationXPlayer rows are created cell-major, player-minor with id = cell + player*S (sim.cpp:624-636); the
 * observation node visits them in that creation order. */
static void observe_world(const Sim *s, World *w)
{
    for (int loc = 0; loc < s->S; loc++)
        for (int p = 0; p < s->P; p++)
            observe_row(s, w, loc, p);
}

/* sim.cpp:441-447, 450-456, 468-482 as called by resetWorld (sim.cpp:539-553) and by the task graph */
static void reset_world_system(World *w) { w->should_update_pos = 1; if (w->reset_now) w->timestep = 0; }
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

void *oracle_overcooked_create(const OracleOvercookedCfg *cfg, uint32_t num_worlds)
{
    Sim *s = (Sim *)calloc(1, sizeof(Sim));
    s->cfg = *cfg;
    s->P = cfg->num_players; s->W = cfg->width; s->H = cfg->height;
    s->S = cfg->width * cfg->height; s->C = 5 * s->P + 16; s->N = num_worlds;
    s->placement_rew = (uint8_t)cfg->placement_in_pot_rew;
    s->dish_rew = (uint8_t)cfg->dish_pickup_rew;
    s->soup_rew = (uint8_t)cfg->soup_pickup_rew;
    for (int x = 0; x < s->S; x++) if (cfg->terrain[x] == T_POT) s->pots[s->npots++] = x;   /* sim.cpp:588-595 */
    s->worlds = (World *)calloc(num_worlds, sizeof(World));
    for (uint32_t n = 0; n < num_worlds; n++) {                                              /* Sim::Sim, sim.cpp:556-659 */
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

void oracle_overcooked_destroy(void *h)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) { free(s->worlds[n].players); free(s->worlds[n].cells); free(s->worlds[n].obs); }
    free(s->worlds);
    free(s);
}

/* sim.cpp:286-338 */
static void counter_pot_interaction(const Sim *s, World *w, Player *player, int iternum)
{
    if (player->interaction_index != iternum) return;
    int i_pos = move_in_direction(player->position, player->orientation, s->W);
    int terrain = s->cfg.terrain[i_pos];
    Obj *soup = &w->cells[i_pos].object;
    if (terrain == T_COUNTER) {
        if (player->held.name != O_NONE && soup->name == O_NONE) {
            *soup = player->held; player->held = OBJ_NONE;
        } else if (player->held.name == O_NONE && soup->name != O_NONE) {
            player->held = *soup; *soup = OBJ_NONE;
        }
    } else if (terrain == T_POT) {
        if (player->held.name == O_NONE) {
            /* soup_to_be_cooked_at_location, sim.cpp:175-178 */
            if (soup->name == O_SOUP && !is_cooking(s, soup) && !is_ready(s, soup) &&
                (uint8_t)(soup->onions + soup->tomatoes) > 0)
                soup->tick = 0;
        } else if (player->held.name == O_DISH && soup->name == O_SOUP && is_ready(s, soup)) {
            player->held = *soup; *soup = OBJ_NONE;
            w->calculated_reward += s->soup_rew;
        } else if (player->held.name == O_ONION || player->held.name == O_TOMATO) {
            if (soup->name == O_NONE) { *soup = OBJ_NONE; soup->name = O_SOUP; }
            if (!(soup->tick >= 0 || (uint8_t)(soup->onions + soup->tomatoes) == MAX_INGREDIENTS)) {
                Obj obj = player->held; player->held = OBJ_NONE;
                if (obj.name == O_ONION) soup->onions++; else soup->tomatoes++;
                w->calculated_reward += s->placement_rew;
            }
        }
    }
}

static void step_world(const Sim *s, World *w, const int32_t *act /* [P], this world's actions */, int32_t *rew /* [P] */)
{
    const int P = s->P, S = s->S, W = s->W;
    Player *pl = w->players;
    Cell *cells = w->cells;

    w->calculated_reward = 0;                                                 /* 1 pre_resolve_interacts :201 */

    for (int i = 0; i < P; i++) {                                             /* 2 resolve_interacts :208-255 */
        Player *p = &pl[i];
        p->interaction_index = -1;
        if (act[i] != A_INTERACT) continue;
        int i_pos = move_in_direction(p->position, p->orientation, W);
        Cell *dat = &cells[i_pos];
        int t = s->cfg.terrain[i_pos];
        if (t == T_COUNTER || t == T_POT) {
            dat->interacting[dat->num_interacting++] = i;
        } else if (t == T_ONION_SOURCE) {
            if (p->held.name == O_NONE) { p->held = OBJ_NONE; p->held.name = O_ONION; }
        } else if (t == T_TOMATO_SOURCE) {
            if (p->held.name == O_NONE) { p->held = OBJ_NONE; p->held.name = O_TOMATO; }
        } else if (t == T_DISH_SOURCE) {
            if (p->held.name == O_NONE) { p->held = OBJ_NONE; p->held.name = O_DISH; }
        } else if (t == T_SERVING) {
            if (p->held.name == O_SOUP) {                                     /* deliver_soup :169-173 */
                Obj soup = p->held; p->held = OBJ_NONE;
                w->calculated_reward += s->cfg.recipe_values[recipe_of(&soup)];
            }
        }
    }
    for (int i = 0; i < P; i++) {                                             /* 3 setup_interact_time :259-282 */
        Player *p = &pl[i];
        if (act[i] != A_INTERACT) continue;
        int i_pos = move_in_direction(p->position, p->orientation, W);
        int t = s->cfg.terrain[i_pos];
        if (t == T_COUNTER || t == T_POT) {
            p->interaction_index = 0;
            for (int k = 0; k < cells[i_pos].num_interacting; k++)
                if (cells[i_pos].interacting[k] < i) p->interaction_index++;
        }
    }
    for (int r = 0; r < 4; r++)                                               /* 4-7 do_counter_pot_int0..3 :340-358 */
        for (int i = 0; i < P; i++) counter_pot_interaction(s, w, &pl[i], r);

    for (int i = 0; i < P; i++) {                                             /* 8 _move_if_direction :363-379 */
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
    for (int i = 0; i < P; i++) {                                             /* 9 _check_collisions :383-396 */
        Player *p = &pl[i];
        int comp = cells[p->proposed_position].current_player;
        if (cells[p->proposed_position].future_player != i ||
            (comp != -1 && comp != i && cells[p->position].future_player == comp))
            w->should_update_pos = 0;
    }
    for (int i = 0; i < P; i++) {                                             /* 10 _unset_loc_info :400-413 */
        Player *p = &pl[i];
        cells[p->position].current_player = -1;
        cells[p->proposed_position].future_player = -1;
        cells[p->position].past_player = i;
        cells[p->position].past_orientation = p->orientation;
    }
    for (int i = 0; i < P; i++) {                                             /* 11 _handle_collisions :417-426 */
        Player *p = &pl[i];
        if (w->should_update_pos) p->position = p->proposed_position;
        p->orientation = p->proposed_orientation;
        cells[p->position].current_player = i;
    }
    for (int k = 0; k < s->npots; k++) {                                      /* 12 step_pot_effects :430-438 */
        Obj *o = &cells[s->pots[k]].object;
        if (o->name == O_SOUP && is_cooking(s, o)) o->tick++;
    }
    w->timestep += 1;                                                         /* 13 check_reset_system :485-489 */
    w->reset_now = (w->timestep >= s->cfg.horizon);
    reset_world_system(w);                                                    /* 14 :441-447 */
    for (int c = 0; c < S; c++) reset_objects_system(w, &cells[c]);           /* 15 :450-456 */
    if (w->reset_now)                                                         /* 16 _pre_reset_actors_system :459-464 */
        for (int i = 0; i < P; i++) cells[pl[i].position].current_player = -1;
    for (int i = 0; i < P; i++) reset_actors_system(s, w, i, &rew[i]);        /* 17 :468-482 */
    observe_world(s, w);                                                      /* 18 observationSystem :68-167 */
    for (int c = 0; c < S; c++) cells[c].past_player = cells[c].past_orientation = -1;   /* 19 :491-495 */
}

/* [S][P][C] rows -> canonical [P][N][W][H][C] (envs/overcooked_env.py:92-102: scatter by LocationXID, reshape
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

void oracle_overcooked_observe(void *h, uint8_t *obs)
{
    Sim *s = (Sim *)h;
    for (uint32_t n = 0; n < s->N; n++) export_obs(s, &s->worlds[n], n, obs);
}

void oracle_overcooked_step(void *h, const int32_t *actions, uint8_t *obs, int32_t *rewards, int32_t *done)
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
